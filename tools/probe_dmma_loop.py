#!/usr/bin/env python
"""DMMA consumer-loop probes (xtb_debug_dmma_loop, csrc/microbench.cu): TFLOP/s vs warps per SM sub-partition,
row tiles per warp and operand source.  Writes gpurun_out/probe_dmma_loop.json."""
import ctypes as C, importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
api = importlib.import_module("xtant-3_b200.api")
ctx = api.TBContext(0)
import ctypes as _C, os as _os
_probe = _C.CDLL(_os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "xtant-3_b200", "libxtb_b200_probe.so"), mode=_C.RTLD_GLOBAL)
f = _probe.xtb_debug_dmma_loop; f.restype = C.c_double; f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
out = {}
for var in (0, 1, 2, 3):
    for w in (2, 3):
        for mt in (4, 6):
            out[f"var{var}_w{w}_mt{mt}"] = round(f(ctx.h, var, w, mt, 2), 2)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "probe_dmma_loop.json"), "w"), indent=1)
print(json.dumps(out, indent=1))
