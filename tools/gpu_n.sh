#!/bin/bash
# default N-GPU bench line (headline replicated layout + the column-panel variant inside the same line)
NG=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29512 \
   bench.py --gpus $NG --steps 5 --warmup 3 > gpurun_out/r01k_bench_n$NG.json 2> gpurun_out/r01k_bench_n$NG.err; echo "rc=$?"
tail -3 gpurun_out/r01k_bench_n$NG.err | cut -c1-300
python - <<P
import json
l=[x for x in open("gpurun_out/r01k_bench_n$NG.json").read().splitlines() if x.startswith("{")][-1]
j=json.loads(l); print(j["n_gpus"], j["value"], j["stage_ms"], j["e2e"]["value"], j["config"]["dense_layout"], j.get("variant_dense_column_panels"))
P
