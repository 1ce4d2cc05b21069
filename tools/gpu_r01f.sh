#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r01f_pytest.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/r01f_pytest.log
timeout 600 python tools/bench_coupling.py au4096 > gpurun_out/r01f_coupling.json 2> gpurun_out/r01f_coupling.err; echo "coupling rc=$?"
tail -3 gpurun_out/r01f_coupling.err; cat gpurun_out/r01f_coupling.json
