#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02k_bench.json 2> gpurun_out/r02k_bench.err; echo "bench rc=$?"
tail -c 1500 gpurun_out/r02k_bench.json; tail -5 gpurun_out/r02k_bench.err
bash tools/r02/gpu_j.sh
