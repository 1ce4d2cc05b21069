#!/bin/bash
# r02u (1 GPU): Coulomb term -- parity vs the oracle + the suites sharing its code (vdW, step, wall, driver)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_coulomb_gpu.py tests/test_vdw_gpu.py tests/test_abi_driver_gpu.py tests/test_step_gpu.py tests/test_wall_gpu.py -m gpu -q > gpurun_out/r02u_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/r02u_pytest.log | cut -c1-300
