#!/bin/bash
# r02n: density consumer loop, instruction-order variants (library variants built with tools/r02/build_variants.sh)
mkdir -p gpurun_out
run() { tag=$1; shift
  env "$@" timeout 300 python bench.py --no-cpu --no-e2e --no-configs --steps 3 --warmup 2 > gpurun_out/r02n_$tag.json 2> gpurun_out/r02n_$tag.err; rc=$?
  python - <<P
import json
try:
    j=json.loads(open("gpurun_out/r02n_$tag.json").read().strip().splitlines()[-1])
    print("$tag rc=$rc step", round(j["value"],2), "density", round(j["stage_ms"]["density"],2), "kernel", round(j["stage_ms"]["density_kernel"],2), "frac", round(j["roofline"]["frac"],3))
except Exception as e: print("$tag rc=$rc no json", e)
P
}
run base A=1
for v in o1 o2 o3 o4 u2; do run $v XTB_LIB_VARIANT=$v; done
run base2 A=1
