#!/bin/bash
# N-GPU bench: packed (default) and direct (XTB_MG_GATHER=direct) block exchange, each with the multi- vs single-GPU result check
NG=${1:-2}
mkdir -p gpurun_out
for mode in ${MODES:-packed direct}; do
XTB_MG_GATHER=$mode timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29513 \
   bench.py --gpus $NG --steps 5 --warmup 3 > gpurun_out/r01p_bench_n${NG}_$mode.json 2> gpurun_out/r01p_bench_n${NG}_$mode.err; echo "rc=$?"
tail -3 gpurun_out/r01p_bench_n${NG}_$mode.err | cut -c1-300
python - <<P
import json
l=[x for x in open("gpurun_out/r01p_bench_n${NG}_$mode.json").read().splitlines() if x.startswith("{")][-1]
j=json.loads(l); print("$mode", j["n_gpus"], j["value"], j["stage_ms"], j["e2e"]["value"], j.get("variant_dense_column_panels",{}).get("ms_per_step"), j.get("multi_gpu_vs_single_gpu_max_abs_diff"))
P
done
