#!/bin/bash
# r02as (1 GPU): tiled S*Ha kernel of the coupling matrix -- parity suite, then A/B at au4096
mkdir -p gpurun_out
timeout 40 python -m pytest tests/test_coupling_gpu.py -m gpu -q -x > gpurun_out/r02as_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02as_pytest.log | cut -c1-300
timeout 50 python tools/r02/coupling_ab.py > gpurun_out/r02as_coupling_ab.json 2> gpurun_out/r02as_coupling_ab.err; echo "ab rc=$?"; tail -1 gpurun_out/r02as_coupling_ab.json | cut -c1-700; tail -2 gpurun_out/r02as_coupling_ab.err | cut -c1-300
