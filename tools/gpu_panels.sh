#!/bin/bash
# N-GPU run of the dense-layout variant (owned column panels instead of replicated dense H/S); NG = number of GPUs
NG=${1:-8}
mkdir -p gpurun_out
XTB_BENCH_DENSE_LAYOUT=panels timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29511 \
   bench.py --gpus $NG --steps 5 --warmup 3 > gpurun_out/r01k_bench_panels_n$NG.json 2> gpurun_out/r01k_bench_panels_n$NG.err; echo "rc=$?"
tail -3 gpurun_out/r01k_bench_panels_n$NG.err | cut -c1-300
python - <<P
import json
l=[x for x in open("gpurun_out/r01k_bench_panels_n$NG.json").read().splitlines() if x.startswith("{")][-1]
j=json.loads(l); print(j["n_gpus"], j["value"], j["stage_ms"], j["e2e"]["value"], j["config"]["dense_layout"])
P
