#!/bin/bash
NG=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_edge_gpu.py tests/test_parity_gpu.py tests/test_pettifor_gpu.py tests/test_coupling_gpu.py -m gpu -x -q > gpurun_out/r01u_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r01u_pytest.log
MODES=packed bash tools/gpu_mg.sh $NG
