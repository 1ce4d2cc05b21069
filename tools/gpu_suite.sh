#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r01v_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r01v_pytest.log
timeout 200 python __graft_entry__.py smoke > gpurun_out/r01v_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r01v_smoke.log
