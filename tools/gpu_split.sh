#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
timeout 300 python bench.py --no-cpu --no-e2e --steps 3 --warmup 2 > gpurun_out/sp.json 2>/dev/null
python -c "import json;j=json.loads(open('gpurun_out/sp.json').read().strip().splitlines()[-1]);print('default', j['value'], j['stage_ms']['density_kernel'])"
