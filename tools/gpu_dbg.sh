#!/bin/bash
mkdir -p gpurun_out
run() { env "$@" timeout 200 python bench.py --no-cpu --no-e2e --steps 2 --warmup 1 > gpurun_out/x.log 2>&1
  echo "$* rc=$? $(python -c "import json;j=json.loads(open('gpurun_out/x.log').read().strip().splitlines()[-1]);print(j['stage_ms']['density_kernel'], j['stage_ms']['density'])")"; }
run XTB_DENSITY_NCW=8
run XTB_DENSITY_NCW=12
run XTB_DENSITY_DBG=4
run XTB_DENSITY_DBG=16
