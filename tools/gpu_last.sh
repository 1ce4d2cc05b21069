#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r01t_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r01t_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/r01t_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r01t_smoke.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/r01t_bench.json 2> gpurun_out/r01t_bench.err; echo "bench rc=$?"
python -c "
import json; j=json.loads(open('gpurun_out/r01t_bench.json').read().strip().splitlines()[-1]); print(j['value'], j['e2e']['value'], j['roofline']['frac'], j['gpu_launches'], j['clocks'], j['cpu_baseline']['value'])"
