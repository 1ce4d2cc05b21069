#!/bin/bash
# r02z (1 GPU): dense update clears only the blocks the new pattern drops -- step / eigen / parity / edge suites + N=1 bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_step_gpu.py tests/test_eigen_gpu.py tests/test_parity_gpu.py tests/test_edge_gpu.py tests/test_pettifor_gpu.py tests/test_dftb_gpu.py -m gpu -q -x > gpurun_out/r02z_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r02z_pytest.log | cut -c1-300
timeout 1500 python bench.py --steps 10 --warmup 3 > gpurun_out/r02z_bench_n1.json 2> gpurun_out/r02z_bench_n1.err; echo "bench rc=$?"
python - <<'P'
import json
j=json.loads(open("gpurun_out/r02z_bench_n1.json").read().strip().splitlines()[-1])
print("N=1", round(j["value"],3), "frac", round(j["roofline"]["frac"],4), "e2e", round(j["e2e"]["value"],3))
print({k:(round(v,3) if isinstance(v,float) else v) for k,v in j["stage_ms"].items()})
print(j.get("dense_update")); print(j.get("variant_dense_full_write"))
for k,v in j.get("configs", {}).items(): print(k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a.endswith("_ms")})
P
