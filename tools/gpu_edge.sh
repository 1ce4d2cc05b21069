#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_edge_gpu.py -m gpu -x -q > gpurun_out/r01k_edge.log 2>&1; echo "rc=$?"; tail -30 gpurun_out/r01k_edge.log
