#!/bin/bash
# r02m: after the throttle overflow fix -- small-cell probe, density-path tests, N = 1 bench
mkdir -p gpurun_out
timeout 300 python tools/r02/probe_small_density.py > gpurun_out/r02m_probe_small_density.log 2>&1; cat gpurun_out/r02m_probe_small_density.log
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_edge_gpu.py tests/test_density_pair_gpu.py tests/test_fullsize_gpu.py tests/test_step_gpu.py -m gpu -q > gpurun_out/r02m_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02m_pytest.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02m_bench.json 2> gpurun_out/r02m_bench.err; echo "bench rc=$?"
python - <<'P'
import json
try:
    j=json.loads(open("gpurun_out/r02m_bench.json").read().strip().splitlines()[-1])
    print(round(j["value"],3), {k:(round(v,3) if v else v) for k,v in j["stage_ms"].items()}, round(j["roofline"]["frac"],4), "e2e", round(j["e2e"]["value"],3))
    print({k:(round(v["staged_ms"],3),round(v["step_ms"],3),round(v["graph_ms"],3)) for k,v in j["configs"].items()})
except Exception as e: print("no json", e); print(open("gpurun_out/r02m_bench.err").read()[-2000:])
P
