#!/bin/bash
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'density_dmma_kernel' -c 1 \
  -f -o gpurun_out/r01h_density python bench.py --steps 1 --warmup 0 --no-e2e --no-cpu > gpurun_out/r01h_ncu_density.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/*.ncu-rep
