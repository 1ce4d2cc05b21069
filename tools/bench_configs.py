#!/usr/bin/env python
"""ms per MD step and pair-blocks/s versus atom count: the BASELINE.json configs 1-4 plus an Au fcc size ladder,
GPU stages (device-timed, through the C ABI, eigenvectors from the on-device eigensolve) next to the CPU restatement
of the reference algorithm (oracle, all host threads; `verbatim` = the reference's dense per-atom force loop where it
is affordable).  One JSON object on stdout; feeds BASELINE.md section 4 and profiles/.

    python tools/bench_configs.py [--no-cpu]
"""
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def gpu_case(api, wk, wl, reps=3):
    c = api.TBContext(0)
    c.set_workload(wl)
    ortho = wl["tb"] == "pettifor"
    out = {}
    for _ in range(reps):
        c.get_near_neighbours_device()
        c.assemble_HS(to_host=False)
        Ei = c.eigensolve()[0]
        fe, mu = wk.occupations(Ei, wl["n_electrons"], wl["Te_eV"])
        c.set_occupations(fe)
        c.density()
        c.forces_device()
        c.pressure_device()
        st = {k: c.stage_ms(s) for k, s in (("neighbours", 1), ("assembly", 2), ("density", 3), ("forces", 4), ("pressure", 5),
                                            ("eigensolve", 8))}
        if ortho and wl.get("rep_params"):
            c.dErdr_s(); c.dErdr_Pressure_s()
            st["repulsive_forces"] = c.stage_ms(9); st["repulsive_pressure"] = c.stage_ms(10)
        out = st
    nblk = c.counter(4)
    graft = out["neighbours"] + out["assembly"] + out["density"] + out["forces"]
    res = {"n_blocks": nblk, "n_pairs": c.counter(5), "stage_ms": out, "ms_per_step_graft": graft,
           "ms_per_step_with_pressure": graft + out["pressure"], "ms_per_step_incl_eigensolve": graft + out["eigensolve"],
           "pair_blocks_per_s": nblk / (out["assembly"] * 1e-3)}
    c.close()
    return res


def cpu_case(orc, wk, wl, verbatim):
    t = {}
    nrl = wl["tb"] == "nrl"
    t0 = time.perf_counter(); nb = orc.neighbours(wl); t["neighbours"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    if nrl:
        lmn, xrr, x1 = orc.geometry(wl, nb); rad = orc.nrl_radial(wl, nb)
        H, S = orc.nrl_assemble_HS(wl, nb, lmn, rad)
    else:
        H = orc.pet_assemble_H(wl, nb)
    t["assembly"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    if nrl:
        Ha, Ei, _ = orc.eigensolve_nonorthogonal(H, S)
    else:
        Ha, Ei = orc.eigensolve_orthogonal(H)
    t["eigensolve"] = time.perf_counter() - t0
    fe, mu = wk.occupations(Ei, wl["n_electrons"], wl["Te_eV"])
    no = 9 if nrl else 4
    t0 = time.perf_counter()
    Aij = orc.density_dense(Ha, fe); AijE = orc.density_dense(Ha, fe, Ei)          # what the reference does: two dgemm
    t["density_dense_dgemm"] = time.perf_counter() - t0
    t0 = time.perf_counter(); rho, rhoE = orc.density_blocks(nb, no, Ha, fe, Ei); t["density_blocks"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    if nrl:
        orc.nrl_forces_sparse(wl, nb, lmn, rad, rho, rhoE)
    else:
        orc.pet_forces_sparse(wl, nb, rho)
    t["forces_sparse"] = time.perf_counter() - t0
    if verbatim:
        t0 = time.perf_counter()
        if nrl:
            orc.nrl_forces_verbatim(wl, nb, lmn, rad, Aij, AijE)
        else:
            orc.pet_forces_verbatim(wl, nb, Aij)
        t["forces_verbatim"] = time.perf_counter() - t0
    ms = {k: 1e3 * v for k, v in t.items()}
    f = ms.get("forces_verbatim", ms["forces_sparse"])
    return {"cores": orc.num_threads(), "stage_ms": ms, "forces_form": "verbatim" if verbatim else "sparse-equivalent",
            "ms_per_step_graft": ms["neighbours"] + ms["assembly"] + ms["density_dense_dgemm"] + f,
            "pair_blocks_per_s": nb.n_blocks() / t["assembly"]}


def main():
    wk = importlib.import_module("xtant-3_b200.workloads")
    api = importlib.import_module("xtant-3_b200.api")
    import oracle as orc
    no_cpu = "--no-cpu" in sys.argv
    cases = [("diamond64", wk.diamond64, True), ("si216", wk.si216, True), ("al500", wk.al500, False), ("ac1000", wk.ac1000, True),
             ("au256", lambda: wk.au_fcc((4, 4, 4), sigma=0.05, seed=256), False),
             ("au864", lambda: wk.au_fcc((6, 6, 6), sigma=0.05, seed=864), False),
             ("au2048", lambda: wk.au_fcc((8, 8, 8), sigma=0.05, seed=2048), False)]
    out = {}
    gpu_case(api, wk, wk.au_fcc((2, 2, 2)), reps=2)            # warm-up: module load, cuSOLVER / cuBLAS handles
    for name, mk, verbatim in cases:
        wl = mk()
        N = wl["na"] * (9 if wl["tb"] == "nrl" else 4)
        r = {"na": wl["na"], "N": N, "tb": wl["tb"], "gpu": gpu_case(api, wk, wl, reps=3 if N <= 5000 else 2)}
        if not no_cpu and wl["na"] <= 1000:
            r["cpu"] = cpu_case(orc, wk, wl, verbatim)
        out[name] = r
        print(name, json.dumps(r), file=sys.stderr)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
