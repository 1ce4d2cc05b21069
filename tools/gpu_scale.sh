#!/bin/bash
# scaling run: bench.py at N = 4, 8 on one 8-GPU box (N = 1, 2 are measured on their own boxes)
mkdir -p gpurun_out
for n in ${NS:-4 8}; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) \
     bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err; echo "n$n rc=$?"
  tail -2 gpurun_out/scale_n$n.err
  python - <<P
import json
try:
    j=json.loads(open("gpurun_out/scale_n$n.json").read().strip().splitlines()[-1]); print($n, j["value"], j["stage_ms"], j["e2e"])
except Exception as e: print($n, "no line", e)
P
done
