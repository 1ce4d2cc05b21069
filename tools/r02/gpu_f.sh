#!/bin/bash
# r02f: density kernel experiments (stagger of the second consumer warp per sub-partition; always-load ay)
mkdir -p gpurun_out
run() { # tag, env...
  tag=$1; shift
  env "$@" timeout 300 python bench.py --no-cpu --no-e2e --steps 2 --warmup 1 > gpurun_out/r02f_$tag.json 2> gpurun_out/r02f_$tag.err; rc=$?
  python - <<P
import json
try:
    j=json.loads(open("gpurun_out/r02f_$tag.json").read().strip().splitlines()[-1])
    print("$tag rc=$rc", round(j["value"],2), round(j["stage_ms"]["density"],2), round(j["stage_ms"]["density_kernel"],2), round(j["roofline"]["frac"],3))
except Exception as e: print("$tag rc=$rc no json", e)
P
}
for v in st600 st1200 st2400; do
  run old_$v XTB_LIB_VARIANT=$v XTB_DENSITY_IMPL=1
  run pair_$v XTB_LIB_VARIANT=$v XTB_DENSITY_IMPL=2
done
run pair_loady XTB_LIB_VARIANT=loady XTB_DENSITY_IMPL=2
run old_base XTB_DENSITY_IMPL=1
run pair_base XTB_DENSITY_IMPL=2
