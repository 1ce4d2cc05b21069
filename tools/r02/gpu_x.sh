#!/bin/bash
# r02x (1 GPU): the General short-range class -- parity vs the oracle + the suites sharing wall.cu / step.cu
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_short_gpu.py tests/test_wall_gpu.py tests/test_step_gpu.py tests/test_vdw_gpu.py tests/test_coulomb_gpu.py tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r02x_pytest.log 2>&1; echo "pytest rc=$?"; tail -30 gpurun_out/r02x_pytest.log | cut -c1-400
