#!/bin/bash
# experiment: states per chunk of the density kernel (XTB_KC), rebuilt on the box
mkdir -p gpurun_out
for kc in 32 8; do
  make -C xtant-3_b200/csrc clean > /dev/null 2>&1
  make -C xtant-3_b200/csrc -j16 PTXAS=-DXTB_KC=$kc > gpurun_out/kc_build_$kc.log 2>&1; echo "build kc=$kc rc=$?"
  timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "density" 2>&1 | tail -1
  timeout 300 python bench.py --no-cpu --no-e2e --steps 3 --warmup 2 > gpurun_out/kc$kc.json 2>gpurun_out/kc$kc.err
  python -c "import json;j=json.loads(open('gpurun_out/kc$kc.json').read().strip().splitlines()[-1]);print('kc=$kc', j['value'], j['stage_ms'])"
done
