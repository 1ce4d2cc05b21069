#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_edge_gpu.py tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/r01m_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r01m_pytest.log
# owned range of 512 atoms on one GPU (what a rank of the 8-GPU run executes): density kernel with and without the tail split
python - <<'P'
import importlib, os, sys, json
sys.path.insert(0, ".")
import torch
wk = importlib.import_module("xtant-3_b200.workloads"); api = importlib.import_module("xtant-3_b200.api")
sys.argv = ["bench.py"]
import bench
wl = wk.au4096(); N = wl["na"] * 9
c = api.TBContext(0); c.set_workload(wl)
Cm, dE, dfe, E, fe = bench.make_eigen_device(torch, wk, wl, N)
c.set_eigen_device(Cm.data_ptr(), dE.data_ptr(), dfe.data_ptr())
c.get_near_neighbours_device()
out = {}
for rng in ((1, 512), (1025, 1536)):
    c.set_partition(*rng)
    for tail in ("-1", None):
        if tail is None: os.environ.pop("XTB_DENSITY_TAIL", None)
        else: os.environ["XTB_DENSITY_TAIL"] = tail
        for _ in range(3): c.density()
        out[f"atoms {rng} tail={'off' if tail else 'auto'}"] = [c.stage_ms(7), c.stage_ms(3)]
print(json.dumps(out))
open("gpurun_out/r01m_tail.json", "w").write(json.dumps(out))
P
python bench.py --no-cpu --no-e2e --steps 5 --warmup 3 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N=1', j['value'], j['stage_ms'])"
