#!/bin/bash
mkdir -p gpurun_out
for ncw in 8 12; do
XTB_DENSITY_NCW=$ncw python bench.py --no-cpu --no-e2e --steps 5 --warmup 3 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('NCW=$ncw N=1', j['value'], j['stage_ms']['density_kernel'])"
XTB_DENSITY_NCW=$ncw python tools/bench_configs.py --no-cpu 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('NCW=$ncw', {k: round(v['gpu']['stage_ms']['density'],3) for k,v in j.items()})"
done
