#!/usr/bin/env python
"""Time xtb_coupling (Electron_ion_coupling_Mij) on the north-star cell au4096 (N = 36864).
Synthetic resident eigenvectors as in bench.py (timing does not depend on their values); Ha0 = a device copy.
One JSON line: total / S*Ha part / DGEMM part [ms], useful flop rates, and the dense-S alternative (2 DGEMMs)."""
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    wk = importlib.import_module("xtant-3_b200.workloads")
    api = importlib.import_module("xtant-3_b200.api")
    name = sys.argv[1] if len(sys.argv) > 1 else "au4096"
    wl = wk.CONFIGS[name]() if name in wk.CONFIGS else wk.au_fcc((4, 4, 4))
    na, N = wl["na"], wl["na"] * 9
    ctx = api.TBContext(0)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    ctx.set_workload(wl)
    ctx.get_near_neighbours_device()
    ctx.assemble_blocks()                       # S blocks only; no dense matrices needed
    Cm, dE, dfe, E, fe = bench.make_eigen_device(torch, wk, wl, N)
    ctx.set_eigen_device(Cm.data_ptr(), dE.data_ptr(), dfe.data_ptr())
    ctx.save_eigen_prev()
    res = []
    for _ in range(3):
        ctx.Electron_ion_coupling_Mij(0, to_host=False)
        res.append((ctx.stage_ms(11), ctx.stage_ms(12), ctx.stage_ms(13)))
    tot, spmm, gemm = res[-1]
    nblk = ctx.counter(4)
    out = {"workload": wl["name"], "N": N, "n_blocks": nblk, "runs_ms(total, norms+S*Ha, dgemm)": res,
           "total_ms": tot, "spmm_ms": spmm, "dgemm_ms": gemm, "antisym_ms": tot - spmm - gemm,
           "spmm_useful_tflops": 2 * 81 * nblk * N / (spmm * 1e-3) * 1e-12,
           "dgemm_tflops": 2 * N ** 3 / (gemm * 1e-3) * 1e-12,
           "antisym_GBps": 2 * 8 * N * N / ((tot - spmm - gemm) * 1e-3) * 1e-9,
           "dense_S_alternative_ms": tot + gemm - spmm,
           "note": "reference: 2 N^3 multiply-adds in scalar OpenMP loops (Nonadiabatic.f90:889-936)"}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
