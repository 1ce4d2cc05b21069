#!/bin/bash
# r02e: ncu --set full of the pair density kernel (one launch) with source import
mkdir -p gpurun_out
XTB_DENSITY_IMPL=2 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'density_pair_kernel' -c 1 \
  -f -o gpurun_out/r02e_pair python bench.py --steps 1 --warmup 0 --no-e2e --no-cpu > gpurun_out/r02e_ncu.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/r02e_ncu.log
ls -la gpurun_out/r02e_pair.ncu-rep
