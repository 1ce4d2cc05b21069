"""Deterministic workloads of the hot path: the five BASELINE.json configs plus
small cells for tests.  Generator seeds are part of the contract.

A workload is a dict:
    name, tb ("nrl" | "pettifor"), n_orb, na, koa (int32 [na], 1-based kinds), S (float64 [na,3],
    relative coordinates), supce (float64 [3,3], supce[a,b] == Scell%supce(a+1,b+1)),
    params (list of parameter dicts, (KOA1,KOA2) column-major), n_kinds, n_electrons, Te_eV
"""
import json
import os

import numpy as np

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")
K_B_EV = 8.617333262e-5


def _load(name):
    with open(os.path.join(_DATA, name)) as f:
        return json.load(f)


def load_params(kind, element):
    return _load(f"params_{kind}_{element}.json")


_FCC = np.array([[0, 0, 0], [0.5, 0.5, 0], [0.5, 0, 0.5], [0, 0.5, 0.5]], dtype=np.float64)
_DIA = np.vstack([_FCC, _FCC + 0.25])


def lattice(basis, a, reps, sigma, seed):
    """Orthorhombic supercell of `reps` conventional cells, Gaussian displacements sigma [A]."""
    nx, ny, nz = reps
    cells = np.array([[i, j, k] for i in range(nx) for j in range(ny) for k in range(nz)], dtype=np.float64)
    frac = (cells[:, None, :] + basis[None, :, :]).reshape(-1, 3)
    L = np.array([nx * a, ny * a, nz * a])
    cart = frac * a
    if sigma > 0:
        rng = np.random.default_rng(seed)
        cart = cart + rng.normal(0.0, sigma, size=cart.shape)
    S = np.mod(cart / L, 1.0)
    supce = np.diag(L)
    return np.ascontiguousarray(S), supce


def _wl(name, tb, element, S, supce, n_el_per_atom, Te_eV, koa=None):
    na = S.shape[0]
    return dict(name=name, tb=tb, n_orb=9 if tb == "nrl" else 4, na=na,
                koa=np.ones(na, dtype=np.int32) if koa is None else np.asarray(koa, dtype=np.int32),
                S=np.ascontiguousarray(S, dtype=np.float64), supce=np.ascontiguousarray(supce, dtype=np.float64),
                params=[load_params(tb, element)], n_kinds=1, element=element,
                rep_params=[load_params("pettifor_rep", element)] if tb == "pettifor" else None,
                n_electrons=float(n_el_per_atom * na), Te_eV=float(Te_eV))


def _from_cell(name, tb, element, cellfile, n_el_per_atom, Te_eV):
    c = _load(cellfile)
    return _wl(name, tb, element, np.array(c["S"]), np.array(c["supce"]), n_el_per_atom, Te_eV, koa=c["koa"])


def diamond64():   # config 1
    return _from_cell("diamond64", "pettifor", "C", "cell_diamond64.json", 4, 300.0 * K_B_EV)


def si216():       # config 2
    S, h = lattice(_DIA, 5.44, (3, 3, 3), 0.05, 216)
    return _wl("si216", "nrl", "Si", S, h, 4, 2.0)


def al500():       # config 3 (triclinic cell from the reference's SAVE files)
    return _from_cell("al500", "nrl", "Al", "cell_al500.json", 3, 1000.0 * K_B_EV)


def load_short(setname):
    """'General' short-range parameter set xtant-3_b200/data/params_short_<setname>.json (readers.read_Short_Rep_TB of the reference file)"""
    import importlib
    p = dict(_load(f"params_short_{setname}.json"))
    p["exp"] = tuple(p["exp"]) if p.get("exp") else None
    p["inv_exp"] = tuple(p["inv_exp"]) if p.get("inv_exp") else None
    p["pow"] = [tuple(t) for t in p.get("pow", [])]
    if p.get("tab"):
        rd = importlib.import_module("xtant-3_b200.readers")
        t = p["tab"]
        a, b, c, d = rd.make_cubic_splines(t["R"], t["E"])
        p["tab"] = {"R": np.array(t["R"]), "E": np.array(t["E"]), "spline": bool(t["spline"]), "a": a, "b": b, "c": c, "d": d}
    return p


def with_vdw(wl, setname):
    """attach Scell%TB_Waals from xtant-3_b200/data/params_vdw_<setname>.json (one kind of atom)"""
    p = dict(_load(f"params_vdw_{setname}.json"))
    wl["vdw_kind"] = p.pop("kind")
    wl["vdw_params"] = [p]
    return wl


def ac1000():      # config 4
    """The reference's amorphous-carbon folder (INPUT_DATA/Amorphous_Carbon) is the Diamond Pettifor set plus an ACTIVE
    Girifalco van der Waals file (its class line carries no ':exclude'), so this configuration runs with the vdW term."""
    S, h = lattice(_DIA, 3.5588, (5, 5, 5), 0.25, 1000)
    return with_vdw(_wl("ac1000", "pettifor", "C", S, h, 4, 1.0), "C_amorphous")


def ac216():       # the reference's own 216-atom amorphous-carbon cell (INPUT_DATA/Amorphous_Carbon/SAVE_*.dat), with its vdW term
    return with_vdw(_from_cell("ac216", "pettifor", "C", "cell_ac216.json", 4, 1.0), "C_amorphous")


def au4096():      # config 5 (north star): 8 x 8 x 16 conventional cells
    """Atoms are numbered with the LONG axis (z, 16 cells) as the outermost loop, so that contiguous atom ranges -- the
    multi-GPU row partition -- are compact z-slabs (8 ranks: 2 cells = 8.2 A thick against a 8.7 A cut-off; numbered
    along x they would be 4.1 A thick and 7/8 of the cell would be halo).  Built as the 16 x 8 x 8 lattice with the
    axes rotated (x, y, z) -> (z, x, y); fcc is invariant under that rotation."""
    S, h = lattice(_FCC, 4.0782, (16, 8, 8), 0.05, 4096)
    S = np.ascontiguousarray(S[:, [1, 2, 0]])
    h = np.diag([h[1, 1], h[2, 2], h[0, 0]])
    return _wl("au4096", "nrl", "Au", S, h, 11, 0.1)


def au108():       # small NRL metal cell from the reference's SAVE files (tests)
    return _from_cell("au108", "nrl", "Au", "cell_au108.json", 11, 0.1)


def au_fcc(reps, sigma=0.05, seed=1):
    S, h = lattice(_FCC, 4.0782, reps, sigma, seed)
    w = _wl(f"au_fcc_{reps[0]}x{reps[1]}x{reps[2]}", "nrl", "Au", S, h, 11, 0.1)
    return w


def si_dia(reps, sigma=0.05, seed=2):
    S, h = lattice(_DIA, 5.44, reps, sigma, seed)
    return _wl(f"si_dia_{reps[0]}x{reps[1]}x{reps[2]}", "nrl", "Si", S, h, 4, 2.0)


def c_dia(reps, sigma=0.05, seed=3):
    S, h = lattice(_DIA, 3.5588, reps, sigma, seed)
    return _wl(f"c_dia_{reps[0]}x{reps[1]}x{reps[2]}", "pettifor", "C", S, h, 4, 1.0)


# ---- DFTB branch: tables from xtant-3_b200/data/dftb_*.npz (skf.py) ----
_ZB_A, _ZB_B = _FCC, _FCC + 0.25      # zincblende sublattices


def _wl_dftb(name, setname, rcut, d, S, supce, koa, n_el_per_kind, Te_eV, kinds=None, force_s_basis=False):
    import importlib
    skf = importlib.import_module("xtant-3_b200.skf")
    elements, basis, pairs = skf.load_set(setname, rcut, d)
    if kinds is not None:            # sub-set of the elements of the table set
        nk = len(elements)
        pairs = [pairs[a + nk * b] for b in kinds for a in kinds]
        elements = [elements[k] for k in kinds]
        if force_s_basis:            # test-only: drop the p level so that idnetify_basis_size picks the s basis
            for p in pairs:
                p["Ep"] = 0.0
                p["Ed"] = 0.0
        basis, pairs = skf.finish_set(pairs, rcut, d)
    koa = np.asarray(koa, dtype=np.int32)
    na = S.shape[0]
    return dict(name=name, tb="dftb", basis=basis, n_orb=(1, 4, 9)[basis], na=na, koa=koa,
                S=np.ascontiguousarray(S, dtype=np.float64), supce=np.ascontiguousarray(supce, dtype=np.float64),
                params=pairs, n_kinds=len(elements), element="-".join(elements),
                n_electrons=float(sum(n_el_per_kind[k - 1] for k in koa)), Te_eV=float(Te_eV))


def _zincblende(a, reps, sigma, seed):
    S, h = lattice(np.vstack([_ZB_A, _ZB_B]), a, reps, sigma, seed)
    koa = np.tile(np.array([1, 1, 1, 1, 2, 2, 2, 2], dtype=np.int32), reps[0] * reps[1] * reps[2])
    return S, h, koa


def sial_zb(reps=(2, 2, 2), sigma=0.05, seed=64):
    """Si/Al zincblende cell, matsci-0-3 sp3d5 tables, cut-off of INPUT_DATA/SiAl/TB_Hamiltonian_parameters.txt (5.0, 0.1 A)"""
    S, h, koa = _zincblende(5.6, reps, sigma, seed)
    return _wl_dftb(f"sial_zb_{reps[0]}x{reps[1]}x{reps[2]}", "matsci_SiAl", 5.0, 0.1, S, h, koa, (4, 3), 0.5)


def ch_zb(reps=(2, 2, 2), sigma=0.05, seed=65):
    """C/H zincblende-ordered cell, mio-1-1 tables (sp3 basis: H carries the shifted p level), cut-off 3.5 + 10*0.1 A"""
    S, h, koa = _zincblende(3.7, reps, sigma, seed)
    return _wl_dftb(f"ch_zb_{reps[0]}x{reps[1]}x{reps[2]}", "mio_CH", 3.5, 0.1, S, h, koa, (4, 1), 0.5)


def h_sc(reps=(3, 3, 3), a=2.4, sigma=0.05, seed=66):
    """hydrogen simple-cubic cell, mio-1-1 H-H table restricted to the s basis (test of the basis_ind = 0 branch)"""
    S, h = lattice(np.zeros((1, 3)), a, reps, sigma, seed)
    return _wl_dftb(f"h_sc_{reps[0]}x{reps[1]}x{reps[2]}", "mio_CH", 3.0, 0.1, S, h, np.ones(S.shape[0], dtype=np.int32),
                    (1,), 0.5, kinds=[1], force_s_basis=True)


CONFIGS = {"diamond64": diamond64, "si216": si216, "al500": al500, "ac1000": ac1000, "au4096": au4096}


# ---- electron occupations (input of the density stage; produced by Electron_tools in the reference) ----
def fermi(E, mu, Te_eV):
    """Fermi_distribution, Electron_tools.f90:1896-1909 (spin factor 2)."""
    with np.errstate(over="ignore"):
        return 2.0 / (1.0 + np.exp((E - mu) / Te_eV))


def occupations(E, n_electrons, Te_eV):
    """mu from N_e by bisection (as Electron_Fixed_Te, Electron_tools.f90:2660), then f_n."""
    lo, hi = float(E.min()) - 50.0 * Te_eV - 1.0, float(E.max()) + 50.0 * Te_eV + 1.0
    for _ in range(200):
        mu = 0.5 * (lo + hi)
        if fermi(E, mu, Te_eV).sum() < n_electrons:
            lo = mu
        else:
            hi = mu
    mu = 0.5 * (lo + hi)
    return fermi(E, mu, Te_eV), mu


def synthetic_eigen(N, filling=0.61, Te_eV=0.1, seed=7, dtype=np.float64):
    """Kernel-throughput stand-in for the eigensolve at sizes no CPU can diagonalise
    (SURVEY.md 8d item 5): C = N(0,1)/sqrt(N), E sorted uniform(-10,40), f = Fermi at `filling`."""
    rng = np.random.default_rng(seed)
    E = np.sort(rng.uniform(-10.0, 40.0, size=N))
    fe, mu = occupations(E, 2.0 * filling * N, Te_eV)
    C = rng.standard_normal((N, N)).astype(dtype) / np.sqrt(N)
    return np.asfortranarray(C), E, fe
