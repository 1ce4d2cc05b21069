#!/bin/bash
# r02aa (1 GPU): pair-once NRL block and force kernels -- full GPU suite + N=1 bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/r02aa_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02aa_pytest.log | cut -c1-300
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r02aa_bench_n1.json 2> gpurun_out/r02aa_bench_n1.err; echo "bench rc=$?"; tail -3 gpurun_out/r02aa_bench_n1.err
python - <<'P'
import json
j=json.loads(open("gpurun_out/r02aa_bench_n1.json").read().strip().splitlines()[-1])
print("N=1", round(j["value"],3), "frac", round(j["roofline"]["frac"],4), "e2e", round(j["e2e"]["value"],3), "pair_blocks/s", j["pair_blocks_per_s"])
print({k:(round(v,3) if isinstance(v,float) else v) for k,v in j["stage_ms"].items()})
for k in ("roofline_hbm_stage","roofline_hbm_assembly"):
    r=j.get(k) or {}
    print(k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in r.items() if a in ("achieved","frac","ms","fp64_floor_ms")})
for k,v in j.get("configs", {}).items(): print(k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a.endswith("_ms")})
P
