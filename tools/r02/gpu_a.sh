#!/bin/bash
# r02a: GPU suite with the new full-size parity legs + baseline bench of the round-1 tree
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader; nproc; free -g | head -2
timeout 1500 python -m pytest tests/test_fullsize_gpu.py -m gpu -x -q --durations=8 > gpurun_out/r02a_fullsize.log 2>&1; echo "fullsize rc=$?"; tail -15 gpurun_out/r02a_fullsize.log
timeout 900 python -m pytest tests -m gpu -x -q --deselect tests/test_fullsize_gpu.py > gpurun_out/r02a_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02a_pytest.log
timeout 600 python bench.py --no-cpu > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r02a_bench.err
python - <<P
import json
j=json.loads(open("gpurun_out/r02a_bench.json").read().strip().splitlines()[-1])
print(j["value"], j["stage_ms"], j["e2e"]["value"], j["roofline"]["frac"])
P
