#!/bin/bash
mkdir -p gpurun_out
N=${N:-2}
timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "partition" 2>&1 | tail -2
XTB_BENCH_PHASES=1 timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29502 \
     bench.py --gpus $N --steps 5 --warmup 3 --no-e2e > gpurun_out/phases_n$N.json 2> gpurun_out/phases_n$N.err; echo "phases rc=$?"
grep phases gpurun_out/phases_n$N.err
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29503 \
     bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err; echo "bench rc=$?"
python - <<P
import json
j=json.loads(open("gpurun_out/scale_n$N.json").read().strip().splitlines()[-1]); print($N, j["value"], j["stage_ms"], j["e2e"]["value"])
P
