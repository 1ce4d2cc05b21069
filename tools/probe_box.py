#!/usr/bin/env python
"""One-off probe of the GPU box: toolchain, host memory, and the FP64 roofline denominators
(cuBLAS DGEMM via torch.matmul, DMMA.8x8x4 and DFMA issue-rate kernels).  Writes gpurun_out/probe.json."""
import ctypes as C, importlib, json, os, shutil, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
out = {"nproc": os.cpu_count(), "gfortran": shutil.which("gfortran"), "ifx": shutil.which("ifx"),
       "nvfortran": shutil.which("nvfortran")}
try:
    out["mem_total_gb"] = int(open("/proc/meminfo").read().split("MemTotal:")[1].split()[0]) / 1e6
except Exception as e:
    out["mem_total_gb"] = str(e)
out["gpu"] = torch.cuda.get_device_name(0)
out["gpu_mem_gb"] = torch.cuda.get_device_properties(0).total_memory / 1e9
n = 8192
a = torch.randn(n, n, dtype=torch.float64, device="cuda"); b = torch.randn(n, n, dtype=torch.float64, device="cuda")
best = 0
for i in range(6):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); c = a @ b; e1.record(); torch.cuda.synchronize()
    if i: best = max(best, 2 * n ** 3 / (e0.elapsed_time(e1) * 1e-3) * 1e-12)
out["dgemm_8192_tflops"] = best
del a, b, c
api = importlib.import_module("xtant-3_b200.api")
ctx = api.TBContext(0)
import ctypes as _C, os as _os
_probe = _C.CDLL(_os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "xtant-3_b200", "libxtb_b200_probe.so"), mode=_C.RTLD_GLOBAL)
f = _probe.xtb_debug_fp64_peak; f.restype = C.c_double; f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
for w in (4, 8, 16, 32):
    out[f"dmma884_tflops_w{w}"] = f(ctx.h, 0, w, 3)
    out[f"dfma_tflops_w{w}"] = f(ctx.h, 1, w, 3)
    out[f"mixed_dmma_plus_dfma_tflops_w{w}"] = f(ctx.h, 2, w, 3)
    out[f"split_dmma_dfma_tflops_w{w}"] = f(ctx.h, 3, w, 3)
# copy bandwidth (same method as MEASURED_PEAKS.json)
x = torch.empty(1 << 30, dtype=torch.bfloat16, device="cuda"); y = torch.empty_like(x)
best = 0
for i in range(6):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); y.copy_(x); e1.record(); torch.cuda.synchronize()
    if i: best = max(best, 2 * x.numel() * 2 / (e0.elapsed_time(e1) * 1e-3) * 1e-9)
out["copy_gbs"] = best
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "probe.json"), "w"), indent=1)
print(json.dumps(out, indent=1))
