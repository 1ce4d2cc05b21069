#!/bin/bash
# r02am (1 GPU): lock-step lead sweep of the density kernel (XTB_DENSITY_LEAD, chunks a CTA may run ahead of the slowest; 64 = default)
mkdir -p gpurun_out
B="python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-configs"
for v in 64 8 16 32 128 256 1024 0; do
XTB_DENSITY_LEAD=$v timeout 600 $B > gpurun_out/r02am_bench_lead$v.json 2> gpurun_out/r02am_bench_lead$v.err; echo "lead $v rc=$?"
python - $v <<'P'
import json,sys
j=json.loads(open("gpurun_out/r02am_bench_lead%s.json"%sys.argv[1]).read().strip().splitlines()[-1])
print("lead", sys.argv[1], "step", round(j["value"],3), "density_kernel", round(j["stage_ms"]["density_kernel"],3), "frac", round(j["roofline"]["frac"],4))
P
done
