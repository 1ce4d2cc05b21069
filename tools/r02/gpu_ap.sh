#!/bin/bash
# r02ap (1 GPU): sum kernels with a run-time number of atoms per CTA -- force / supercell-force suites
mkdir -p gpurun_out
timeout 170 python -m pytest tests/test_edge_gpu.py tests/test_parity_gpu.py tests/test_step_gpu.py -m gpu -q -x > gpurun_out/r02ap_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02ap_pytest.log | cut -c1-200
