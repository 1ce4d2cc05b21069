#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_fullsize_gpu.py -m gpu -x -q > gpurun_out/r01h_fullsize.log 2>&1; echo "fullsize rc=$?"
tail -30 gpurun_out/r01h_fullsize.log
