#!/bin/bash
# r02s (1 GPU): van der Waals term -- parity vs the oracle, the step / wall / driver suites that share its code paths, timing
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_vdw_gpu.py tests/test_abi_driver_gpu.py tests/test_step_gpu.py tests/test_wall_gpu.py tests/test_repulsive_gpu.py -m gpu -q > gpurun_out/r02s_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02s_pytest.log
python - <<'P' > gpurun_out/r02s_vdw_timing.log 2>&1
import importlib, sys, time
sys.path.insert(0, ".")
wk = importlib.import_module("xtant-3_b200.workloads"); api = importlib.import_module("xtant-3_b200.api")
for name in ("ac216", "ac1000"):
  for env in ("", "1"):
    import os
    os.environ.pop("XTB_VDW_NO_PRUNE", None)
    if env: os.environ["XTB_VDW_NO_PRUNE"] = env
    wl = getattr(wk, name)()
    c = api.TBContext(0); c.set_workload(wl)
    for rep in range(3):
        c.get_vdW_s(); e = c.stage_ms(17)
        c.vdW_forces(); f = c.stage_ms(17)
        c.vdW_Pressure(); p = c.stage_ms(17)
    print(name, "no_prune" if env else "pruned", "cells", c.get_mirror_cell_num(), "ms: energy %.4f forces %.4f pressure %.4f" % (e, f, p))
    c.close()
P
cat gpurun_out/r02s_vdw_timing.log
