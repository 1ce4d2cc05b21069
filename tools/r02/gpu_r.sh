#!/bin/bash
# r02r (8 GPUs): final tree -- the C driver with 2 and 8 ranks, N = 2 / 4 / 8 bench lines
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_abi_driver_gpu.py -m gpu -q > gpurun_out/r02r_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r02r_pytest.log
python - <<'P'
import importlib, subprocess, sys, tempfile, pathlib
sys.path.insert(0, "."); sys.path.insert(0, "tests")
wk = importlib.import_module("xtant-3_b200.workloads")
import test_abi_driver_gpu as t
d = pathlib.Path(tempfile.mkdtemp())
wl = wk.au_fcc((3, 3, 16), sigma=0.08, seed=35)          # 576 atoms, eight z-slabs of two cells
r = subprocess.run([t.DRIVER] + list(t._write_problem(d, wl)) + ["8"], capture_output=True, text=True, timeout=900)
open("gpurun_out/r02r_cdriver_8ranks.log", "w").write(r.stdout + "\n--- stderr ---\n" + r.stderr[-3000:])
print("c driver 8 ranks rc =", r.returncode, r.stdout.count("OK rank"), "rank lines;", "OK all" in r.stdout)
P
for n in 2 4 8; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/r02r_bench_n$n.json 2> gpurun_out/r02r_bench_n$n.err; echo "bench n=$n rc=$?"
python - <<P
import json
try:
    j=json.loads(open("gpurun_out/r02r_bench_n$n.json").read().strip().splitlines()[-1])
    print("N=$n", round(j["value"],3), {k:(round(v,3) if v else v) for k,v in j["stage_ms"].items()}, "e2e", round(j["e2e"]["value"],3), "variant", round(j["variant_dense_column_panels"]["ms_per_step"],3))
    print(j["exchange"]); print(j["multi_gpu_vs_single_gpu_max_abs_diff"])
except Exception as e: print("no json", e); print(open("gpurun_out/r02r_bench_n$n.err").read()[-2000:])
P
done
