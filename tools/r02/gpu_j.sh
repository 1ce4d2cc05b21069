#!/bin/bash
# r02j: ncu evidence of the round-2 tree.  (1) launch list of the bench command, (2) one --set full capture of the dominant
# kernel and of the stage kernels, (3) launch list of smoke() (what the driver's own capture sees)
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-configs"
timeout 600 $B > gpurun_out/r02j_bench_plain.json 2> gpurun_out/r02j_bench_plain.err; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02j_launches_bench_au4096.csv $B > gpurun_out/r02j_ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'density_dmma_kernel' --launch-skip 1 --launch-count 1 \
  -f -o gpurun_out/r02j_density $B > gpurun_out/r02j_ncu_density.log 2>&1; echo "ncu density rc=$?"
timeout 1200 ncu --set full --clock-control none -k regex:'ct_tile_kernel|dense_scatter_kernel|dense_zero_prev_kernel|nrl_forces_kernel|nrl_blocks_kernel|nn_build_kernel|mirror_blocks_kernel|nn_rev_kernel|nn_rowptr_kernel' \
  --launch-skip 9 --launch-count 9 -f -o gpurun_out/r02j_stages $B > gpurun_out/r02j_ncu_stages.log 2>&1; echo "ncu stages rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 100 --csv --log-file gpurun_out/r02j_launches_smoke.csv python __graft_entry__.py smoke > gpurun_out/r02j_ncu_smoke.log 2>&1; echo "smoke launch list rc=$?"
ls -la gpurun_out/*.ncu-rep
