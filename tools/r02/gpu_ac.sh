#!/bin/bash
# r02ac (1 GPU): CTA-per-atom block kernel that also writes the dense matrices, warp-per-atom embedded density -- touched suites, bench, ncu
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_edge_gpu.py tests/test_parity_gpu.py tests/test_step_gpu.py tests/test_fullsize_gpu.py tests/test_abi_driver_gpu.py tests/test_eigen_gpu.py tests/test_coupling_gpu.py tests/test_density_tm_gpu.py -m gpu -q -x > gpurun_out/r02ac_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r02ac_pytest.log | cut -c1-300
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r02ac_bench_n1.json 2> gpurun_out/r02ac_bench_n1.err; echo "bench rc=$?"; tail -3 gpurun_out/r02ac_bench_n1.err
python - <<'P'
import json
j=json.loads(open("gpurun_out/r02ac_bench_n1.json").read().strip().splitlines()[-1])
print("N=1", round(j["value"],3), "frac", round(j["roofline"]["frac"],4), "e2e", round(j["e2e"]["value"],3), "pair_blocks/s", j["pair_blocks_per_s"])
print({k:(round(v,3) if isinstance(v,float) else v) for k,v in j["stage_ms"].items()})
for k in ("roofline_hbm_stage","roofline_hbm_assembly"):
    r=j.get(k) or {}
    print(k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in r.items() if a in ("achieved","frac","ms","fp64_floor_ms")})
for k,v in j.get("configs", {}).items(): print(k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a.endswith("_ms")})
P
B="python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-configs"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'dense_remember_kernel|dense_zero_prev_kernel|nrl_force_pairs_kernel|nrl_force_sum_kernel|nrl_blocks_kernel|nrl_onsite_w_kernel|nrl_dens_kernel' \
  --launch-skip 7 --launch-count 7 -f -o gpurun_out/r02ac_stages $B > gpurun_out/r02ac_ncu_stages.log 2>&1; echo "ncu stages rc=$?"
ls -la gpurun_out/*.ncu-rep
