#!/bin/bash
# r02ae (2 GPUs): the pair-once block / force kernels in a partitioned run -- C driver with 2 ranks, N=2 bench (multi-GPU vs single-GPU diffs)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_abi_driver_gpu.py -m gpu -q -x > gpurun_out/r02ae_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02ae_pytest.log | cut -c1-300
n=2
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n bench.py --gpus $n --steps 5 --warmup 3 --no-cpu > gpurun_out/r02ae_bench_n$n.json 2> gpurun_out/r02ae_bench_n$n.err; echo "bench n=$n rc=$?"; tail -3 gpurun_out/r02ae_bench_n$n.err
python - <<'P'
import json
j=json.loads(open("gpurun_out/r02ae_bench_n2.json").read().strip().splitlines()[-1])
print("N=2", round(j["value"],3), {k:(round(v,3) if isinstance(v,float) else v) for k,v in j["stage_ms"].items()})
print({k:v for k,v in j.items() if "diff" in k or "multi" in k})
print(j.get("e2e"))
P
