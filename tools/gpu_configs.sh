#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python tools/bench_configs.py > gpurun_out/r01j_configs.json 2> gpurun_out/r01j_configs.err; echo "rc=$?"
tail -4 gpurun_out/r01j_configs.err | cut -c1-600
