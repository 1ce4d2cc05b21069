#!/bin/bash
# r02q: final-tree verification on one GPU: full GPU suite, smoke, N = 1 bench, reference arm, then the ncu evidence (gpu_j.sh)
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q --durations=12 > gpurun_out/r02q_pytest.log 2>&1; echo "pytest rc=$?"; tail -18 gpurun_out/r02q_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02q_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r02q_smoke.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02q_bench.json 2> gpurun_out/r02q_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r02q_bench.err
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02q_bench_reference_arm.json 2> gpurun_out/r02q_bench_reference_arm.err; echo "reference arm rc=$?"
sed -i 's/r02j_/r02q_/g' tools/r02/gpu_j.sh
bash tools/r02/gpu_j.sh
