#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_fullsize_gpu.py tests/test_edge_gpu.py tests/test_eigen_gpu.py -m gpu -x -q > gpurun_out/r01q_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r01q_pytest.log
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_write.sum,dram__bytes_read.sum --clock-control none -k regex:'nrl_blocks_kernel|dense_panel_kernel' -c 6 --csv --log-file gpurun_out/r01q_blocks.csv \
   python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e > gpurun_out/r01q_ncu.log 2>&1; echo "ncu rc=$?"
grep -v "^==" gpurun_out/r01q_blocks.csv | cut -d, -f5,13- | head -20
python bench.py --no-cpu --no-e2e --steps 5 --warmup 3 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N=1', j['value'], j['stage_ms'], j['pair_blocks_per_s'])"
