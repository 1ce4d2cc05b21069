#!/bin/bash
# r02i (2 GPUs): the C driver as two "MPI ranks", then the N = 2 bench through torchrun (exchanges inside the library)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_abi_driver_gpu.py -m gpu -q -s > gpurun_out/r02i_pytest.log 2>&1; echo "pytest rc=$?"; grep -n "OK \|passed\|failed\|FAIL" gpurun_out/r02i_pytest.log | tail -20
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02i_bench_n2.json 2> gpurun_out/r02i_bench_n2.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/r02i_bench_n2.json; tail -8 gpurun_out/r02i_bench_n2.err
XTB_RHO_EXCHANGE=allgather XTB_HS_OVERLAP=0 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3 --no-e2e > gpurun_out/r02i_bench_n2_allgather.json 2> gpurun_out/r02i_bench_n2_allgather.err; echo "bench(allgather) rc=$?"
tail -c 1200 gpurun_out/r02i_bench_n2_allgather.json
