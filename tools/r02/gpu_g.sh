#!/bin/bash
# r02g: full GPU suite on the tree with xtb_step / in-library exchanges / wall / Fermi / Verlet, smoke, N = 1 bench
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q --durations=20 > gpurun_out/r02g_pytest.log 2>&1; echo "pytest rc=$?"
tail -30 gpurun_out/r02g_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02g_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r02g_smoke.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02g_bench.json 2> gpurun_out/r02g_bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/r02g_bench.json; tail -5 gpurun_out/r02g_bench.err
