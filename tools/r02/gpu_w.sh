#!/bin/bash
# r02w (1 GPU): the allocation-free entry point -- its test alone in a fresh process, then the whole GPU suite (it is collected last)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r02w_pytest_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -15 gpurun_out/r02w_pytest_noalloc_alone.log | cut -c1-300
timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/r02w_pytest_gpu_full.log 2>&1; echo "full rc=$?"; tail -6 gpurun_out/r02w_pytest_gpu_full.log | cut -c1-300
