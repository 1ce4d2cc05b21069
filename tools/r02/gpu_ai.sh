#!/bin/bash
# r02ai: final-tree verification on one GPU: full GPU suite, smoke, N = 1 bench, reference arm, then the ncu evidence:
# (1) launch list of the bench command, (2) one --set full capture of the dominant kernel and of the stage kernels, (3) launch list of smoke()
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q --durations=8 > gpurun_out/r02ai_pytest.log 2>&1; echo "pytest rc=$?"; tail -14 gpurun_out/r02ai_pytest.log | cut -c1-200
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02ai_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r02ai_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r02ai_bench.json 2> gpurun_out/r02ai_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r02ai_bench.err
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02ai_bench_reference_arm.json 2> gpurun_out/r02ai_bench_reference_arm.err; echo "reference arm rc=$?"
B="python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-configs"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02ai_launches_bench_au4096.csv $B > gpurun_out/r02ai_ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'density_dmma_kernel' --launch-skip 1 --launch-count 1 \
  -f -o gpurun_out/r02ai_density $B > gpurun_out/r02ai_ncu_density.log 2>&1; echo "ncu density rc=$?"
timeout 1200 ncu --set full --clock-control none -k regex:'dense_remember_kernel|dense_zero_prev_kernel|nrl_force_pairs_kernel|nrl_force_sum_kernel|nrl_blocks_kernel|nrl_dens_kernel|nrl_onsite_w_kernel|nn_build_kernel|mirror_blocks_kernel|nn_rev_kernel|nn_rowptr_kernel' \
  --launch-skip 11 --launch-count 11 -f -o gpurun_out/r02ai_stages $B > gpurun_out/r02ai_ncu_stages.log 2>&1; echo "ncu stages rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 100 --csv --log-file gpurun_out/r02ai_launches_smoke.csv python __graft_entry__.py smoke > gpurun_out/r02ai_ncu_smoke.log 2>&1; echo "smoke launch list rc=$?"
python - <<'P'
import json
j=json.loads(open("gpurun_out/r02ai_bench.json").read().strip().splitlines()[-1])
print("N=1", round(j["value"],3), "frac", round(j["roofline"]["frac"],4), "e2e", round(j["e2e"]["value"],3), "pair_blocks/s", j["pair_blocks_per_s"], "launches", j["gpu_launches"])
print({k:(round(v,3) if isinstance(v,float) else v) for k,v in j["stage_ms"].items()})
print("cpu", j["cpu_baseline"]["value"], j["cpu_baseline"]["cores"])
for k,v in j.get("configs", {}).items(): print(k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a.endswith("_ms")})
r=json.loads(open("gpurun_out/r02ai_bench_reference_arm.json").read().strip().splitlines()[-1]); print("reference arm", r.get("value"), r.get("cpu_baseline",{}).get("cores"))
P
ls -la gpurun_out/*.ncu-rep
