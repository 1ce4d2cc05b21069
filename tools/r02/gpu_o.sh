#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_density_tm_gpu.py -m gpu -x -q > gpurun_out/r02o_pytest.log 2>&1; echo "pytest rc=$?"; tail -30 gpurun_out/r02o_pytest.log
