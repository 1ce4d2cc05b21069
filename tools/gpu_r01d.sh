#!/bin/bash
# round-1 (d): GPU parity suite, bench, per-stage ncu --set full captures (assembly, neighbour, forces, pressure kernels)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r01d_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/r01d_pytest.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r01d_bench.json 2> gpurun_out/r01d_bench.err; echo "bench rc=$?"
cat gpurun_out/r01d_bench.json | cut -c1-600
timeout 600 ncu --set full --clock-control none --import-source on \
  -k regex:'nn_build_kernel|nrl_blocks_kernel|dense_panel_kernel|ct_tile_kernel|mirror_blocks_kernel|nrl_forces_kernel' -c 6 \
  -f -o gpurun_out/r01d_stages python bench.py --steps 1 --warmup 0 --no-e2e --no-cpu > gpurun_out/r01d_ncu_stages.log 2>&1; echo "ncu rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'nrl_pressure_kernel|pett_|dftb_' -c 8 \
  -f -o gpurun_out/r01d_press python -m pytest tests/test_parity_gpu.py -m gpu -q -k "al500" > gpurun_out/r01d_ncu_press.log 2>&1; echo "ncu2 rc=$?"
ls -la gpurun_out
