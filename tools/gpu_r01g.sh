#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q 2>&1 | tail -2
for m in 0 1; do
XTB_FORCES_MODE=$m timeout 300 python bench.py --no-cpu --no-e2e --steps 3 --warmup 2 > gpurun_out/fm$m.json 2>/dev/null
python - <<P
import json
j=json.loads(open("gpurun_out/fm$m.json").read().strip().splitlines()[-1]); print("mode $m", j["value"], j["stage_ms"]["forces"])
P
done
