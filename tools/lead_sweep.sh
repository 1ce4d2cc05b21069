#!/bin/bash
# density kernel time vs lock-step lead (XTB_DENSITY_LEAD, 0 = off)
for l in ${LEADS:-0 32 64 128 256}; do
  XTB_DENSITY_LEAD=$l timeout 200 python bench.py --no-cpu --no-e2e --steps 3 --warmup 2 > gpurun_out/lead$l.log 2>&1
  echo "lead=$l rc=$? $(python -c "import json;j=json.loads(open('gpurun_out/lead$l.log').read().strip().splitlines()[-1]);print(j['stage_ms']['density_kernel'], j['stage_ms']['density'])")"
done
