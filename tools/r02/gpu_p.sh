#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_density_tm_gpu.py tests/test_eigen_gpu.py -m gpu -x -q > gpurun_out/r02p_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02p_pytest.log
for lay in state_minor fortran; do
XTB_BENCH_EIGEN_LAYOUT=$lay timeout 300 python bench.py --no-cpu --no-e2e --no-configs --steps 3 --warmup 2 > gpurun_out/r02p_$lay.json 2> gpurun_out/r02p_$lay.err
python - <<P
import json
try:
    j=json.loads(open("gpurun_out/r02p_$lay.json").read().strip().splitlines()[-1])
    print("$lay step", round(j["value"],2), "density", round(j["stage_ms"]["density"],2), "kernel", round(j["stage_ms"]["density_kernel"],2), "retile", round(j["stage_ms"]["retile_kernel"],2), "frac", round(j["roofline"]["frac"],3))
except Exception as e: print("$lay no json", e); print(open("gpurun_out/r02p_$lay.err").read()[-1500:])
P
done
