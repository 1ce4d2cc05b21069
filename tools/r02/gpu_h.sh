#!/bin/bash
# r02h: step / graph / coupling tests, N = 1 bench with the configs key
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_step_gpu.py tests/test_coupling_gpu.py tests/test_edge_gpu.py -m gpu -q --durations=8 > gpurun_out/r02h_pytest.log 2>&1; echo "pytest rc=$?"
tail -25 gpurun_out/r02h_pytest.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02h_bench.json 2> gpurun_out/r02h_bench.err; echo "bench rc=$?"
tail -c 2500 gpurun_out/r02h_bench.json; tail -5 gpurun_out/r02h_bench.err
timeout 600 python tools/bench_coupling.py > gpurun_out/r02h_coupling.json 2> gpurun_out/r02h_coupling.err; echo "coupling rc=$?"; tail -c 1500 gpurun_out/r02h_coupling.json; tail -3 gpurun_out/r02h_coupling.err
