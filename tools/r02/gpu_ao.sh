#!/bin/bash
# r02ao (1 GPU): last verification of the committed tree: full GPU suite, smoke, short N=1 bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02ao_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r02ao_pytest.log | cut -c1-200
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02ao_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r02ao_smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-configs > gpurun_out/r02ao_bench.json 2> gpurun_out/r02ao_bench.err; echo "bench rc=$?"
python - <<'P'
import json
j=json.loads(open("gpurun_out/r02ao_bench.json").read().strip().splitlines()[-1])
print("N=1", round(j["value"],3), "frac", round(j["roofline"]["frac"],4), "e2e", round(j["e2e"]["value"],3), {k:(round(v,3) if isinstance(v,float) else v) for k,v in j["stage_ms"].items()})
P
