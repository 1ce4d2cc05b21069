#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_fullsize_gpu.py tests/test_edge_gpu.py -m gpu -x -q 2>&1 | tail -2
for ncw in 8 12; do
XTB_DENSITY_NCW=$ncw timeout 300 python bench.py --no-cpu --no-e2e --steps 3 --warmup 2 > gpurun_out/bc$ncw.json 2>/dev/null
python -c "import json;j=json.loads(open('gpurun_out/bc$ncw.json').read().strip().splitlines()[-1]);print('ncw=$ncw', j['value'], j['stage_ms']['density_kernel'], j['roofline']['frac'])"
done
