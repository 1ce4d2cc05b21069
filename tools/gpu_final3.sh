#!/bin/bash
# round-end style verification: GPU suite, smoke, bench (own + reference arm), ncu launch list of the bench command
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r01r_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r01r_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/r01r_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r01r_smoke.log
timeout 900 python bench.py > gpurun_out/r01r_bench.json 2> gpurun_out/r01r_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r01r_bench.err
timeout 900 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/r01r_bench_reference.json 2> gpurun_out/r01r_ref.err; echo "ref rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01r_launches.csv \
   python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e > gpurun_out/r01r_ncu_launch.log 2>&1; echo "ncu rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'nrl_forces_kernel|spmm_blocks_kernel|antisym_kernel|rep_forces_kernel' -c 4 \
  -f -o gpurun_out/r01r_misc python tools/bench_coupling.py au_small > gpurun_out/r01r_ncu_misc.log 2>&1; echo "ncu2 rc=$?"
python - <<P
import json
j=json.loads(open("gpurun_out/r01r_bench.json").read().strip().splitlines()[-1])
print(j["value"], j["stage_ms"], j["e2e"]["value"], j["roofline"]["frac"], j.get("variant_density_fcut_1e-30"), j["cpu_baseline"]["value"])
P
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'density_dmma_kernel' -c 1 \
  -f -o gpurun_out/r01r_density python bench.py --steps 1 --warmup 0 --no-e2e --no-cpu > gpurun_out/r01r_ncu_density.log 2>&1; echo "ncu3 rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'nn_build_kernel|nrl_blocks_kernel|dense_panel_kernel|ct_tile_kernel|mirror_blocks_kernel|nrl_forces_kernel' -c 6 \
  -f -o gpurun_out/r01r_stages python bench.py --steps 1 --warmup 0 --no-e2e --no-cpu > gpurun_out/r01r_ncu_stages.log 2>&1; echo "ncu4 rc=$?"
cat > /tmp/pet.py <<'P'
import importlib, sys
sys.path.insert(0, ".")
wk = importlib.import_module("xtant-3_b200.workloads"); api = importlib.import_module("xtant-3_b200.api")
wl = wk.ac1000(); c = api.TBContext(0); c.set_workload(wl)
for _ in range(2):
    c.get_near_neighbours_device(); c.assemble_HS(to_host=False); Ei = c.eigensolve()[0]
    fe, mu = wk.occupations(Ei, wl["n_electrons"], wl["Te_eV"]); c.set_occupations(fe); c.density(); c.forces_device()
print("ok", c.stage_ms(3), c.stage_ms(7))
P
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'density_warp_kernel' -c 2 \
  -f -o gpurun_out/r01r_warp python /tmp/pet.py > gpurun_out/r01r_ncu_warp.log 2>&1; echo "ncu5 rc=$?"; tail -1 gpurun_out/r01r_ncu_warp.log
