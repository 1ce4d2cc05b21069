#!/bin/bash
# r02t (1 GPU): whole GPU suite on the tree with the van der Waals term, smoke, bench (configs incl. the pair-term legs)
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/r02t_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02t_pytest.log
timeout 600 python __graft_entry__.py smoke > gpurun_out/r02t_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r02t_smoke.log
timeout 1500 python bench.py --steps 10 --warmup 3 > gpurun_out/r02t_bench_n1.json 2> gpurun_out/r02t_bench_n1.err; echo "bench rc=$?"
python - <<'P'
import json
j=json.loads(open("gpurun_out/r02t_bench_n1.json").read().strip().splitlines()[-1])
print("N=1", round(j["value"],3), "frac", j["roofline"]["frac"], "e2e", round(j["e2e"]["value"],3))
for k,v in j.get("small_configs", j.get("configs", {})).items(): print(k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items()})
P
