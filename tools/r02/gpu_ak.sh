#!/bin/bash
# r02ak (1 GPU): per-stage table of configs 1-4 + the Au size ladder on the final tree (GPU part of tools/bench_configs.py)
mkdir -p gpurun_out
timeout 900 python tools/bench_configs.py --no-cpu > gpurun_out/r02ak_configs_gpu.json 2> gpurun_out/r02ak_configs_gpu.err; echo "rc=$?"; tail -3 gpurun_out/r02ak_configs_gpu.err
python - <<'P'
import json
j=json.loads(open("gpurun_out/r02ak_configs_gpu.json").read().strip().splitlines()[-1])
for k,v in j.items():
    g=v.get("gpu",{})
    print(k, {a:round(b,4) for a,b in g.get("stage_ms",{}).items()}, round(g.get("ms_per_step_graft",0),4), g.get("pair_blocks_per_s"))
P
