#!/bin/bash
# A/B on one box: density kernel of the current library vs a library built with the previous density.cu
mkdir -p gpurun_out
run() { python bench.py --no-cpu --no-e2e --steps 5 --warmup 3 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', j['value'], j['stage_ms']['density_kernel'], j['stage_ms']['density'])"; }
run new1
cp xtant-3_b200/libxtb_b200.so /tmp/new.so; cp xtant-3_b200/libxtb_b200_prevdens.so xtant-3_b200/libxtb_b200.so
run prev1
cp /tmp/new.so xtant-3_b200/libxtb_b200.so
run new2
cp xtant-3_b200/libxtb_b200_prevdens.so xtant-3_b200/libxtb_b200.so
run prev2
cp /tmp/new.so xtant-3_b200/libxtb_b200.so
