#!/bin/bash
# r02an (1 GPU): density kernel with the first on-site row tile K-split over the four sub-partitions (variant library "ks"): bench, then parity suites
mkdir -p gpurun_out
B="python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-configs"
for v in "" ks; do for lead in 64 1024; do
XTB_LIB_VARIANT=$v XTB_DENSITY_LEAD=$lead timeout 600 $B > gpurun_out/r02an_bench_${v:-product}_$lead.json 2> gpurun_out/r02an_bench_${v:-product}_$lead.err; echo "variant '${v:-product}' lead $lead rc=$?"
python - ${v:-product}_$lead <<'P'
import json,sys
j=json.loads(open("gpurun_out/r02an_bench_%s.json"%sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], "step", round(j["value"],3), "density_kernel", round(j["stage_ms"]["density_kernel"],3), "frac", round(j["roofline"]["frac"],4))
P
done; done
XTB_LIB_VARIANT=ks timeout 900 python -m pytest tests/test_density_tm_gpu.py tests/test_edge_gpu.py tests/test_parity_gpu.py tests/test_fullsize_gpu.py -m gpu -q > gpurun_out/r02an_pytest_ks.log 2>&1; echo "pytest ks rc=$?"; tail -12 gpurun_out/r02an_pytest_ks.log | cut -c1-250
