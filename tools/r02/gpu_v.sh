#!/bin/bash
# r02v (2 GPUs): the C driver with 2 ranks on the tree with the image-cell pair terms (vdW leg in the partitioned step)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_abi_driver_gpu.py -m gpu -q -s > gpurun_out/r02v_pytest.log 2>&1; echo "pytest rc=$?"; grep -E "OK rank|OK vdw|passed|failed" gpurun_out/r02v_pytest.log | cut -c1-330
