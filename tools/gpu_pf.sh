#!/bin/bash
mkdir -p gpurun_out
for pf in ${PFS:-0 2 4 8 16}; do
XTB_DENSITY_PF=$pf timeout 300 python bench.py --no-cpu --no-e2e --steps 3 --warmup 2 > gpurun_out/pf$pf.json 2>/dev/null
python -c "import json;j=json.loads(open('gpurun_out/pf$pf.json').read().strip().splitlines()[-1]);print('pf=$pf', j['value'], j['stage_ms']['density_kernel'])"
done
XTB_DENSITY_PF=4 timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k density 2>&1 | tail -1
