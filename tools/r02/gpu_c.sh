#!/bin/bash
# r02c: pair density kernel with device-side matching; timing experiments (B weighting off, pairing off)
mkdir -p gpurun_out
timeout 420 python -m pytest tests/test_density_pair_gpu.py -m gpu -x -q > gpurun_out/r02c_pair_pytest.log 2>&1; echo "pair pytest rc=$?"; tail -5 gpurun_out/r02c_pair_pytest.log
run() { # tag, env...
  tag=$1; shift
  env "$@" timeout 300 python bench.py --no-cpu --no-e2e --steps 2 --warmup 1 > gpurun_out/r02c_$tag.json 2> gpurun_out/r02c_$tag.err; echo "$tag rc=$?"; tail -1 gpurun_out/r02c_$tag.err
  python - <<P
import json
try:
    j=json.loads(open("gpurun_out/r02c_$tag.json").read().strip().splitlines()[-1])
    print("$tag", round(j["value"],2), round(j["stage_ms"]["density"],2), round(j["stage_ms"]["density_kernel"],2), round(j["roofline"]["frac"],3))
except Exception as e: print("no json", e)
P
}
run pair XTB_DENSITY_IMPL=2
run pair_noB XTB_DENSITY_IMPL=2 XTB_DENSITY_DBG=16
run pair_singles XTB_DENSITY_IMPL=2 XTB_DENSITY_PAIRING=0
run pair_nolead XTB_DENSITY_IMPL=2 XTB_DENSITY_LEAD=0
python - <<'P'
import importlib, sys, os
sys.path.insert(0, ".")
os.environ["XTB_DENSITY_IMPL"]="2"
import numpy as np
wk = importlib.import_module("xtant-3_b200.workloads"); api = importlib.import_module("xtant-3_b200.api")
wl = wk.au4096(); c = api.TBContext(0); c.set_workload(wl); c.get_near_neighbours_device()
C_, E, fe = wk.synthetic_eigen(64*9, seed=1)
print("units", c.counter(11))
P
