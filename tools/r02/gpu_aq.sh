#!/bin/bash
# r02aq (1 GPU): 64-bit cuSOLVER interface (cusolverDnXsyevd) forced at small sizes -- eigensolve suite
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_eigen_gpu.py -m gpu -q -x > gpurun_out/r02aq_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r02aq_pytest.log | cut -c1-300
