"""Slater-Koster file (.skf) reader for the DFTB branch -- host side of xtb_set_params_dftb.

Restates what the reference does between the file and type(TB_H_DFTB) (Objects.f90:202-222):
    read_skf_file               Source_files/Dealing_with_DFTB.f90:48-149   (grid, on-site line, table, unit conversion)
    distribute_parameters_DFTB  :335-362   (file columns Hdd0 Hdd1 Hdd2 Hpd0 Hpd1 Hpp0 Hpp1 Hsd0 Hsp0 Hss0 | S... ->
                                             channels ss, sp, sd, pp_sigma, pp_pi, pd_sigma, pd_pi, dd_sigma, dd_pi, dd_delta)
    idnetify_basis_size         :302-331   (basis index 0 = s, 1 = sp3, 2 = sp3d5; empty levels shifted to 80 / 100 eV)
    read_DFTB_TB_Params         Source_files/Read_input_data.f90:4162-4265  (rcut, d come from the material's parameter file)
The repulsive part (polynomial / spline) is outside the hot path and is skipped.
"""
import os
import re

import numpy as np

G_AU2A = 0.529177249          # Universal_constants.f90:103
G_AU2EV = 27.211396131788     # Universal_constants.f90:105

# 0-based file column of channel c for the H table; the S table is the same + 10   (Dealing_with_DFTB.f90:341-360)
_COL = (9, 8, 7, 5, 6, 3, 4, 0, 1, 2)

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")


def _record(line, need=None):
    """one list-directed Fortran record: blanks or commas separate, `n*v` repeats; stops after `need` values
    (the reference reads only the leading items of the on-site line, which may continue with non-numeric flags)"""
    out = []
    for tok in re.split(r"[,\s]+", line.strip()):
        if need is not None and len(out) >= need:
            break
        if not tok:
            continue
        if "*" in tok:
            n, v = tok.split("*")
            out.extend([float(v.lower().replace("d", "e")) if v else 0.0] * int(n))
        else:
            out.append(float(tok.lower().replace("d", "e")))
    return out


def parse_skf(path, same_element):
    """-> dict(Ed, Ep, Es [eV], Rr [A] (n_r), Vr [eV] (n_r,10), Sr (n_r,10)); f-orbital files ('@') are refused."""
    with open(path) as f:
        lines = f.read().splitlines()
    if lines[0].lstrip().startswith("@"):
        raise ValueError("DFTB parameterisation with f orbitals is not supported (Dealing_with_DFTB.f90:66-69)")
    t = _record(lines[0], 2)
    dr, ngrid = t[0], int(t[1]) - 1                 # :74-84 ; Ngrid = Ngrid - 1
    dr = dr * G_AU2A
    pos = 1
    if same_element:
        t = _record(lines[pos], 7); pos += 1
        Ed, Ep, Es = t[0] * G_AU2EV, t[1] * G_AU2EV, t[2] * G_AU2EV
    else:
        Ed = Ep = Es = 0.0
    pos += 1                                         # polynomial repulsive line (skipped)
    Rr = np.empty(ngrid)
    Vr = np.empty((ngrid, 10)); Sr = np.empty((ngrid, 10))
    acc = dr
    for i in range(ngrid):
        if i > 0:
            acc = acc + dr                           # Rr(i) = Rr(i-1) + dr, accumulated as the reference does (:128)
        Rr[i] = acc
        rec = _record(lines[pos]); pos += 1
        while len(rec) < 20:                         # a list-directed read continues on the next line
            rec += _record(lines[pos]); pos += 1
        for c in range(10):
            Vr[i, c] = rec[_COL[c]]
            Sr[i, c] = rec[_COL[c] + 10]
    Vr = Vr * G_AU2EV
    return dict(Ed=Ed, Ep=Ep, Es=Es, Rr=Rr, Vr=Vr, Sr=Sr)


def finish_set(pairs, rcut, d):
    """pairs: list over (KOA1,KOA2) column-major of parse_skf dicts.  Applies idnetify_basis_size and the cut-off
    parameters of the material file; returns (basis_ind, pairs)."""
    nd, npp = 1, 1
    if max(abs(p["Ed"]) for p in pairs) < 1e-8:
        nd = 0
        if max(abs(p["Ep"]) for p in pairs) < 1e-8:
            npp = 0
    for p in pairs:
        if abs(p["Ed"]) < 1e-8:
            p["Ed"] = 80.0
        if abs(p["Ep"]) < 1e-8:
            p["Ep"] = 100.0
        p["rcut"], p["d"] = float(rcut), float(d)
    return npp + nd, pairs


def save_set(name, elements, pairs):
    arrs = {}
    for idx, p in enumerate(pairs):
        for k in ("Ed", "Ep", "Es"):
            arrs[f"{idx}_{k}"] = np.float64(p[k])
        for k in ("Rr", "Vr", "Sr"):
            arrs[f"{idx}_{k}"] = p[k]
    arrs["elements"] = np.array(elements)
    np.savez_compressed(os.path.join(_DATA, f"dftb_{name}.npz"), **arrs)


def load_set(name, rcut, d):
    """committed table set (tests/golden/make_fixtures.py) -> (elements, basis_ind, pairs)"""
    z = np.load(os.path.join(_DATA, f"dftb_{name}.npz"))
    elements = [str(e) for e in z["elements"]]
    nk = len(elements)
    pairs = []
    for idx in range(nk * nk):
        pairs.append(dict(Ed=float(z[f"{idx}_Ed"]), Ep=float(z[f"{idx}_Ep"]), Es=float(z[f"{idx}_Es"]),
                          Rr=np.ascontiguousarray(z[f"{idx}_Rr"]), Vr=np.asfortranarray(z[f"{idx}_Vr"]),
                          Sr=np.asfortranarray(z[f"{idx}_Sr"])))
    basis, pairs = finish_set(pairs, rcut, d)
    return elements, basis, pairs
