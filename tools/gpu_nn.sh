#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_edge_gpu.py tests/test_parity_gpu.py tests/test_pettifor_gpu.py tests/test_fullsize_gpu.py tests/test_dftb_gpu.py -m gpu -x -q > gpurun_out/r01n_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r01n_pytest.log
timeout 900 python bench.py --no-cpu --steps 5 --warmup 3 > gpurun_out/r01n_bench_nocpu.json 2> gpurun_out/r01n_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r01n_bench.err
python - <<P
import json
j=json.loads(open("gpurun_out/r01n_bench_nocpu.json").read().strip().splitlines()[-1])
print(j["value"], j["stage_ms"], j["e2e"]["value"], j["roofline"]["frac"])
P
