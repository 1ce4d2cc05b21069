#!/bin/bash
# r02af (1 GPU): density-kernel helper-warp placement variants, same box: product | B warp last | 2 producers | both
mkdir -p gpurun_out
B="python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-configs"
for v in "" bl np2 blnp2 ""; do
XTB_LIB_VARIANT=$v timeout 600 $B > gpurun_out/r02af_bench_${v:-product}.json 2> gpurun_out/r02af_bench_${v:-product}.err; echo "variant '${v:-product}' rc=$?"
python - ${v:-product} <<'P'
import json,sys
j=json.loads(open("gpurun_out/r02af_bench_%s.json"%sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], "step", round(j["value"],3), "density_kernel", round(j["stage_ms"]["density_kernel"],3), "frac", round(j["roofline"]["frac"],4))
P
done
