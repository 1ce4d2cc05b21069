#!/bin/bash
# r02ad (1 GPU): CTA-per-atom mirror kernel; A/B of the dense update written by the block kernel (default) vs the separate scatter kernel
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_edge_gpu.py tests/test_parity_gpu.py tests/test_step_gpu.py tests/test_density_tm_gpu.py tests/test_density_pair_gpu.py tests/test_pettifor_gpu.py tests/test_dftb_gpu.py tests/test_fullsize_gpu.py -m gpu -q -x > gpurun_out/r02ad_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r02ad_pytest.log | cut -c1-300
B="python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-configs"
for f in 1 0; do
XTB_DENSE_FUSED=$f timeout 900 $B > gpurun_out/r02ad_bench_fused$f.json 2> gpurun_out/r02ad_bench_fused$f.err; echo "bench fused=$f rc=$?"; tail -2 gpurun_out/r02ad_bench_fused$f.err
python - $f <<'P'
import json,sys
j=json.loads(open("gpurun_out/r02ad_bench_fused%s.json"%sys.argv[1]).read().strip().splitlines()[-1])
print("fused",sys.argv[1],"N=1", round(j["value"],3), "frac", round(j["roofline"]["frac"],4), "pair_blocks/s", j["pair_blocks_per_s"])
print({k:(round(v,3) if isinstance(v,float) else v) for k,v in j["stage_ms"].items()})
P
done
