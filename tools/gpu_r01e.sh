#!/bin/bash
# round-1 (e): GPU parity suite incl. repulsive terms
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r01e_pytest.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/r01e_pytest.log
