#!/usr/bin/env python
"""On-device eigensolve at the north-star size (au4096, N = 36 864) through cusolverDnXsyevd, then the rest of the step on its
eigenvectors: one JSON object with the stage times and invariants (no CPU oracle exists at this size)."""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
wk = importlib.import_module("xtant-3_b200.workloads"); api = importlib.import_module("xtant-3_b200.api")
wl = wk.CONFIGS["au4096"]()
c = api.TBContext(0); c.set_workload(wl)
t0 = time.time()
c.get_near_neighbours_device(); c.assemble_HS(to_host=False)
Ei, wS = c.eigensolve()
t_eig = c.stage_ms(8)
fe, mu = wk.occupations(Ei, wl["n_electrons"], wl["Te_eV"])
c.set_occupations(fe); c.density(); F = c.forces()
out = {"workload": "au4096", "N": int(len(Ei)), "eigensolve_ms": t_eig, "wall_s": time.time() - t0,
       "stage_ms": {k: c.stage_ms(s) for k, s in (("neighbours", 1), ("assembly", 2), ("density", 3), ("forces", 4))},
       "Ei_min": float(Ei.min()), "Ei_max": float(Ei.max()), "Ei_ascending": bool(np.all(np.diff(Ei) >= 0)),
       "eigen_S_min": float(wS.min()), "eigen_S_max": float(wS.max()), "mu_eV": float(mu), "sum_fe": float(fe.sum()),
       "n_electrons": float(wl["n_electrons"]), "E_band_eV_per_atom": float((fe * Ei).sum() / wl["na"]),
       "sum_forces_eV": [float(x) for x in F.sum(axis=0)], "max_abs_force_eV": float(np.abs(F).max())}
print(json.dumps(out))
