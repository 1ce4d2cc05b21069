#!/bin/bash
# r02al (1 GPU): pressure sum kernel reads the recorded source entry -- touched suites + per-config stage table
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_edge_gpu.py tests/test_parity_gpu.py tests/test_step_gpu.py tests/test_fullsize_gpu.py tests/test_abi_driver_gpu.py tests/test_eigen_gpu.py -m gpu -q -x > gpurun_out/r02al_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02al_pytest.log | cut -c1-300
timeout 900 python tools/bench_configs.py --no-cpu > gpurun_out/r02al_configs_gpu.json 2> gpurun_out/r02al_configs_gpu.err; echo "rc=$?"; tail -3 gpurun_out/r02al_configs_gpu.err
python - <<'P'
import json
j=json.loads(open("gpurun_out/r02al_configs_gpu.json").read().strip().splitlines()[-1])
for k,v in j.items():
    g=v.get("gpu",{})
    print(k, {a:round(b,4) for a,b in g.get("stage_ms",{}).items()}, round(g.get("ms_per_step_graft",0),4), g.get("pair_blocks_per_s"))
P
