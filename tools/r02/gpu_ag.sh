#!/bin/bash
# r02ag (1 GPU): timing experiment -- half of the A-fragment loads of the density kernel removed (wrong results) vs product
mkdir -p gpurun_out
B="python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-configs"
for v in "" hl ""; do
XTB_LIB_VARIANT=$v timeout 600 $B > gpurun_out/r02ag_bench_${v:-product}.json 2> gpurun_out/r02ag_bench_${v:-product}.err; echo "variant '${v:-product}' rc=$?"
python - ${v:-product} <<'P'
import json,sys
j=json.loads(open("gpurun_out/r02ag_bench_%s.json"%sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], "step", round(j["value"],3), "density_kernel", round(j["stage_ms"]["density_kernel"],3), "frac", round(j["roofline"]["frac"],4))
P
done
