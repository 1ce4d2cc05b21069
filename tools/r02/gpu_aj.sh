#!/bin/bash
# r02aj (8 GPUs): final-tree scaling points N = 8 and N = 4 (N = 1: r02ai, N = 2: r02ae)
mkdir -p gpurun_out
for n in 8 4; do
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2953$n bench.py --gpus $n --steps 5 --warmup 3 --no-cpu > gpurun_out/r02aj_bench_n$n.json 2> gpurun_out/r02aj_bench_n$n.err; echo "bench n=$n rc=$?"; tail -2 gpurun_out/r02aj_bench_n$n.err
python - $n <<'P'
import json,sys
n=sys.argv[1]
j=json.loads(open("gpurun_out/r02aj_bench_n%s.json"%n).read().strip().splitlines()[-1])
print("N=",n, round(j["value"],3), {k:(round(v,3) if isinstance(v,float) else v) for k,v in j["stage_ms"].items()})
print({k:v for k,v in j.items() if "multi" in k}); print(j.get("e2e",{}).get("value"))
P
done
