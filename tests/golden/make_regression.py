#!/usr/bin/env python
"""Regression values of the ORACLE itself (tests/golden/oracle_regression.json).

These are NOT reference outputs (the Fortran reference cannot be built here; parity is unpinned,
see oracle/xtant_oracle.c).  They freeze the oracle's results at the commit where its two
algorithmic forms, its invariants and its finite-difference checks all passed, so that later edits
cannot drift silently.   Run:  python tests/golden/make_regression.py
"""
import importlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle as orc  # noqa: E402

wk = importlib.import_module("xtant-3_b200.workloads")
CASES = {
    "au_fcc_2x2x2": dict(workload="au_fcc", args=[[2, 2, 2]], kwargs=dict(sigma=0.05, seed=1)),
    "si_dia_2x2x2": dict(workload="si_dia", args=[[2, 2, 2]], kwargs=dict(sigma=0.05, seed=2)),
    "au108": dict(workload="au108"),
}
out = {}
for name, c in CASES.items():
    wl = getattr(wk, c["workload"])(*[tuple(a) for a in c.get("args", [])], **c.get("kwargs", {}))
    o = orc.nrl_step(wl)
    Ei = o["Ei"]
    out[name] = dict(c, n_pairs=int(o["nb"].size.sum()), H_abs_sum=float(np.abs(o["H"]).sum()),
                     S_abs_sum=float(np.abs(o["S"]).sum()), Ei_pick=[float(Ei[0]), float(Ei[len(Ei) // 2]), float(Ei[-1])],
                     F_first4=[float(x) for x in o["F"][:4].ravel()], P=[float(x) for x in o["P"].ravel(order="F")])
    print(name, out[name]["n_pairs"], out[name]["Ei_pick"])
json.dump(out, open(os.path.join(HERE, "oracle_regression.json"), "w"), indent=1)

# DFTB branch: tests/golden/dftb_regression.json (same status: the oracle's own outputs, frozen)
dout = {}
for name, fn in (("sial", wk.sial_zb), ("ch", wk.ch_zb)):
    o = orc.dftb_step(fn(), verbatim=False)
    dout[name] = dict(nn_size=[int(x) for x in o["nb"].size], H_head=[float(x) for x in o["H"][:9, :18].ravel()],
                      S_head=[float(x) for x in o["S"][:9, :18].ravel()], Ei_head=[float(x) for x in o["Ei"][:16]],
                      F_head=[float(x) for x in o["F"][:8].ravel()], P=[float(x) for x in o["P"].ravel()])
    print(name, sum(dout[name]["nn_size"]), dout[name]["Ei_head"][:2])
json.dump(dout, open(os.path.join(HERE, "dftb_regression.json"), "w"), indent=0)
