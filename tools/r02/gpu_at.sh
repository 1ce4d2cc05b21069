#!/bin/bash
# r02at (1 GPU): last sanity of the committed library: coupling (default kernel), eigensolve and no-alloc suites
mkdir -p gpurun_out
timeout 35 python -m pytest tests/test_coupling_gpu.py tests/test_eigen_gpu.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r02at_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02at_pytest.log | cut -c1-300
