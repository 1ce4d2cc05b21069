#!/bin/bash
# r02b: pair density kernel -- correctness against the one-atom kernel, then timing at au4096
mkdir -p gpurun_out
timeout 420 python -m pytest tests/test_density_pair_gpu.py -m gpu -x -q > gpurun_out/r02b_pair_pytest.log 2>&1; echo "pair pytest rc=$?"; tail -15 gpurun_out/r02b_pair_pytest.log
for impl in 2 1; do
  XTB_DENSITY_IMPL=$impl timeout 300 python bench.py --no-cpu --no-e2e --steps 3 --warmup 2 > gpurun_out/r02b_bench_impl$impl.json 2> gpurun_out/r02b_bench_impl$impl.err; echo "bench impl=$impl rc=$?"; tail -2 gpurun_out/r02b_bench_impl$impl.err
  python - <<P
import json
try:
    j=json.loads(open("gpurun_out/r02b_bench_impl$impl.json").read().strip().splitlines()[-1])
    print("impl $impl", j["value"], j["stage_ms"]["density"], j["stage_ms"]["density_kernel"], j["roofline"]["frac"])
except Exception as e: print("no json", e)
P
done
