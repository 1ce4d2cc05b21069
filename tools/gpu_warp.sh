#!/bin/bash
# small-basis density kernel: parity tests of the branches that use it, then the per-config stage times (GPU only)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pettifor_gpu.py tests/test_dftb_gpu.py tests/test_edge_gpu.py tests/test_repulsive_gpu.py -m gpu -x -q > gpurun_out/r01k_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r01k_pytest.log
timeout 900 python tools/bench_configs.py --no-cpu > gpurun_out/r01k_configs.json 2> gpurun_out/r01k_configs.err; echo "rc=$?"
cut -c1-700 gpurun_out/r01k_configs.err | tail -8
