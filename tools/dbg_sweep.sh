#!/bin/bash
# timing experiments on the density kernel: XTB_DENSITY_DBG bit0 = no math, bit1 = quarter of the tile copies
for w in ${NCWS:-8 12 16}; do
XTB_DENSITY_NCW=$w timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "density or forces" 2>&1 | tail -1
for d in ${DBGS:-0 1 2 3}; do
  XTB_DENSITY_NCW=$w XTB_DENSITY_DBG=$d timeout 200 python bench.py --no-cpu --no-e2e --steps 2 --warmup 1 > gpurun_out/dbg$d.log 2>&1
  echo "ncw=$w dbg=$d rc=$? $(python -c "import json;j=json.loads(open('gpurun_out/dbg$d.log').read().strip().splitlines()[-1]);print(j['stage_ms']['density_kernel'], j['stage_ms']['density'])")"
done; done
