"""ORACLE (test infrastructure) -- Python driver of oracle/liboracle.so.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  It restates, on the CPU, the reference's per-MD-step tight-binding path
(see the header of oracle/xtant_oracle.c for the file:line map and the "parity unpinned" note).

The eigensolve between assembly and density follows the reference exactly:
  Loewdin_Orthogonalization   Source_files/TB_NRL.f90:985-1060   (dsyev(S), X = U s^-1/2, H' = X^T H X)
  sym_diagonalize(H')         Source_files/Algebra_tools.f90:882-966 (dsyevd)
  Ha = X V                    Source_files/TB_NRL.f90:421
using scipy's LAPACK (the reference links MKL; both are "a LAPACK", unpinned).
"""
import ctypes as C
import importlib
import os
import subprocess
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
_pkg = importlib.import_module("xtant-3_b200")
_abi = importlib.import_module("xtant-3_b200._abi")

G_RY, G_AU2A, G_A2AU = 13.6056981, 0.529177249, 1.8897259885789     # Universal_constants.f90:67,103,104
G_E = 1.602176634e-19                                                 # :51

_LIB = None
_pd = C.POINTER(C.c_double)
_pi = C.POINTER(C.c_int)


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    src = [os.path.join(_HERE, f) for f in ("xtant_oracle.c", "sk_tables.h", "pettifor_oracle.c", "dftb_oracle.c", "wall_oracle.c", "vdw_oracle.c", "short_oracle.c")]
    src = [s for s in src if os.path.exists(s)]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(so):
            build()
        _LIB = C.CDLL(so)
        _LIB.orc_neighbours.restype = C.c_int
        _LIB.orc_num_threads.restype = C.c_int
        _LIB.orc_pet_erep.restype = C.c_double
        _LIB.orc_wall_energy.restype = C.c_double
        _LIB.orc_fermi.restype = C.c_double
        _LIB.orc_vdw_energy.restype = C.c_double
        _LIB.orc_short_energy.restype = C.c_double
        _LIB.orc_short_V.restype = C.c_double
        _LIB.orc_short_dV.restype = C.c_double
        _LIB.orc_vdw_V.restype = C.c_double
        _LIB.orc_vdw_dV.restype = C.c_double
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def num_threads():
    return int(lib().orc_num_threads())


def set_num_threads(n):
    """pin the OpenMP thread count of the oracle (a launcher may have exported OMP_NUM_THREADS=1)"""
    lib().orc_set_num_threads(C.c_int(int(n)))
    return num_threads()


def nrl_params_array(params):
    arr = (_abi.NrlPair * len(params))()
    for i, d in enumerate(params):
        arr[i] = _abi.nrl_pair_from_dict(d)
    return arr


def cutoff_radius(wl):
    """get_near_neighbours cut-off, Atomic_tools.f90:574-586 (NRL: max(Rc)*g_au2A; Pettifor: max(rm))."""
    if wl["tb"] == "nrl":
        return max(p["Rc"] for p in wl["params"]) * G_AU2A
    if wl["tb"] == "pettifor":
        rm = max(p["rm"] for p in wl["params"])
        if wl.get("rep_params"):                       # repulsive part longer than the attractive: Atomic_tools.f90:609-613
            rm = max(rm, max(p["dm"] for p in wl["rep_params"]))
        return rm
    if wl["tb"] == "dftb":                             # :588-591
        return max(p["rcut"] for p in wl["params"]) + 10.0 * max(p["d"] for p in wl["params"])
    raise ValueError(wl["tb"])


class Neigh:
    """compact ELL neighbour list (0-based)"""

    def __init__(self, na, cap):
        self.na, self.cap = na, cap
        self.size = np.zeros(na, dtype=np.int32)
        self.list = np.zeros((na, cap), dtype=np.int32)
        self.dist = np.zeros((na, cap, 4), dtype=np.float64)
        self.dist_s = np.zeros((na, cap, 3), dtype=np.float64)

    @property
    def max_nn(self):
        return int(self.size.max()) if self.na else 0

    def n_blocks(self):
        return int(self.size.sum()) + self.na


def neighbours(wl, rm=None, cap=None):
    na = wl["na"]
    S = np.ascontiguousarray(wl["S"], dtype=np.float64)
    h = np.asfortranarray(wl["supce"], dtype=np.float64)
    rm = cutoff_radius(wl) if rm is None else rm
    cap = cap or min(max(na - 1, 1), 256)
    while True:
        nb = Neigh(na, cap)
        mx = lib().orc_neighbours(C.c_int(na), _p(S), _p(h), C.c_double(rm), C.c_int(cap),
                                  _p(nb.size), _p(nb.list), _p(nb.dist), _p(nb.dist_s))
        if mx <= cap:
            return nb
        cap = mx


def geometry(wl, nb):
    h = np.asfortranarray(wl["supce"], dtype=np.float64)
    lmn = np.zeros((nb.na, nb.cap, 3)); xrr = np.zeros_like(lmn); x1 = np.zeros_like(lmn)
    lib().orc_geometry(C.c_int(nb.na), C.c_int(nb.cap), _p(nb.size), _p(nb.dist), _p(h), _p(lmn), _p(xrr), _p(x1))
    return lmn, xrr, x1


def _koa0(wl):
    return np.ascontiguousarray(wl["koa"].astype(np.int32) - 1)


def nrl_radial(wl, nb):
    rad = np.zeros((nb.na, nb.cap, 40))
    TB = nrl_params_array(wl["params"])
    lib().orc_nrl_radial(C.c_int(nb.na), C.c_int(nb.cap), _p(nb.size), _p(nb.list), _p(nb.dist), _p(_koa0(wl)),
                         C.c_int(wl["n_kinds"]), TB, _p(rad))
    return rad


def nrl_embedded_density(wl, nb):
    rho = np.zeros(nb.na)
    TB = nrl_params_array(wl["params"])
    lib().orc_nrl_embedded_density(C.c_int(nb.na), C.c_int(nb.cap), _p(nb.size), _p(nb.list), _p(nb.dist),
                                   _p(_koa0(wl)), C.c_int(wl["n_kinds"]), TB, _p(rho))
    return rho


def nrl_assemble_HS(wl, nb, lmn, rad):
    N = nb.na * 9
    H = np.zeros((N, N), order="F"); S = np.zeros((N, N), order="F")
    TB = nrl_params_array(wl["params"])
    lib().orc_nrl_assemble_HS(C.c_int(nb.na), C.c_int(nb.cap), _p(nb.size), _p(nb.list), _p(nb.dist), _p(lmn), _p(rad),
                              _p(_koa0(wl)), C.c_int(wl["n_kinds"]), TB, _p(H), _p(S))
    return H, S


def eigensolve_nonorthogonal(H, S):
    """Loewdin_Orthogonalization + sym_diagonalize + back-transform (TB_NRL.f90:985-1060,409,421)."""
    from scipy.linalg import lapack
    w, U, info = lapack.dsyev(S, compute_v=1, lower=0)
    assert info == 0
    eps = 1e-8
    if (w < 0).any() or (np.abs(w) < eps).any():
        w = np.abs(w)
    s = np.where(np.abs(w) > eps, 1.0 / np.sqrt(np.where(np.abs(w) > eps, w, 1.0)), 0.0)
    X = U * s[None, :]
    Hp = X.T @ (H @ X)
    Hp = 0.5 * (Hp + Hp.T)
    Ei, V, info = lapack.dsyevd(Hp, compute_v=1, lower=0)
    assert info == 0
    Ha = X @ V
    return np.asfortranarray(Ha), Ei, w


# ---- short-range wall (Exponential_wall.f90, class TB_Exp_wall_simple) ----
def wall_params_array(params):
    arr = (_abi.WallPair * len(params))()
    for i, d in enumerate(params):
        arr[i] = _abi.WallPair(float(d["C"]), float(d["r0"]), float(d["d0"]), float(d["dd"]))
    return arr


# ---- the 'General' short-range class (Exponential_wall.f90, class TB_Short_Rep) ----
def _short_args(wl):
    arr, keep = _abi.short_pair_array(wl["short_params"])
    Z = np.ascontiguousarray(wl["Z"], dtype=np.float64)
    return arr, Z, keep


def short_pair_value(p, r, Z1=1.0, Z2=1.0, deriv=False):
    """Shortrange_pot / d_Short_range_pot of one pair (Exponential_wall.f90:130-203)"""
    arr, keep = _abi.short_pair_array([p])
    f = lib().orc_short_dV if deriv else lib().orc_short_V
    return float(f(arr, C.c_double(float(r)), C.c_double(float(Z1)), C.c_double(float(Z2))))


def make_cubic_splines(x, y):
    """make_cubic_splines, Algebra_tools.f90:115-190 (the oracle's own restatement, independent of readers.py)"""
    x = np.ascontiguousarray(x, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
    out = [np.zeros(len(x)) for _ in range(4)]
    lib().orc_make_cubic_splines(C.c_int(len(x)), _p(x), _p(y), *[_p(o) for o in out])
    return tuple(out)


def short_energy(wl, nb):
    """get_short_range_rep_s, Exponential_wall.f90:48-127 -> (E, per-atom halves added to Epot)"""
    arr, Z, keep = _short_args(wl)
    Epot = np.zeros(nb.na)
    E = lib().orc_short_energy(C.c_int(nb.na), C.c_int(nb.cap), C.c_int(wl["n_kinds"]), _p(_koa0(wl)), _p(Z), _p(nb.size), _p(nb.list),
                               _p(nb.dist), arr, _p(Epot))
    return float(E), Epot


def short_forces(wl, nb):
    """d_Short_range_pot_s, Exponential_wall.f90:519-659 (verbatim loop nest) -> the term added to forces%rep (na,3)"""
    arr, Z, keep = _short_args(wl)
    F = np.zeros((nb.na, 3))
    h = np.asfortranarray(wl["supce"], dtype=np.float64).ravel(order="F").copy()
    lib().orc_short_forces(C.c_int(nb.na), C.c_int(nb.cap), C.c_int(wl["n_kinds"]), _p(_koa0(wl)), _p(Z), _p(h), _p(nb.size), _p(nb.list),
                           _p(nb.dist), arr, _p(F))
    return F


def short_pressure(wl, nb):
    """d_Short_range_Pressure_s, Exponential_wall.f90:664-719 -> the term added to SCforce%rep(3,3)"""
    arr, Z, keep = _short_args(wl)
    P = np.zeros(9)
    lib().orc_short_pressure(C.c_int(nb.na), C.c_int(nb.cap), C.c_int(wl["n_kinds"]), _p(_koa0(wl)), _p(Z), _p(nb.size), _p(nb.list),
                             _p(nb.dist), _p(nb.dist_s), arr, _p(P))
    return P.reshape(3, 3, order="F")


# ---- van der Waals term (Van_der_Waals.f90, class TB_vdW) ----
_VDW_KIND = {"Girifalco": _abi.VDW_GIRIFALCO, "LJ": _abi.VDW_LJ_CUT, "ILJ": _abi.VDW_ILJ_CUT}


def vdw_params_array(params):
    arr = (_abi.VdwPair * len(params))()
    for i, d in enumerate(params):
        arr[i] = _abi.vdw_pair_from_dict(d)
    return arr


def vdw_pair_value(kind, p, r, deriv=False):
    """vdW_energy / dvdW of one pair at distance r (Van_der_Waals.f90:578-627)"""
    W = _abi.vdw_pair_from_dict(p)
    f = lib().orc_vdw_dV if deriv else lib().orc_vdw_V
    return float(f(C.c_int(_VDW_KIND[kind]), C.byref(W), C.c_double(float(r))))


def _vdw_geom(wl):
    S = np.ascontiguousarray(wl["S"], dtype=np.float64)
    h = np.asfortranarray(wl["supce"], dtype=np.float64).ravel(order="F").copy()
    return S, h


def vdw_cells(wl, rm_tb):
    """get_mirror_cell_num, Van_der_Waals.f90:132-157 -> ((Nx, Ny, Nz), R_cut); rm_tb = cut-off of the TB part (cutoff_radius)"""
    S, h = _vdw_geom(wl)
    N = np.zeros(3, dtype=np.int32); R = np.zeros(1)
    lib().orc_vdw_cells(C.c_int(wl["n_kinds"]), vdw_params_array(wl["vdw_params"]), C.c_double(float(rm_tb)), _p(S), _p(h), _p(N), _p(R))
    return tuple(int(x) for x in N), float(R[0])


def vdw_energy(wl, N):
    """get_vdW_s, Van_der_Waals.f90:158-250 -> (E_vdW, per-atom halves added to Epot)"""
    S, h = _vdw_geom(wl)
    Epot = np.zeros(len(S)); Nc = np.asarray(N, dtype=np.int32)
    E = lib().orc_vdw_energy(C.c_int(len(S)), C.c_int(_VDW_KIND[wl["vdw_kind"]]), C.c_int(wl["n_kinds"]), _p(_koa0(wl)), _p(S), _p(h),
                             vdw_params_array(wl["vdw_params"]), _p(Nc), _p(Epot))
    return float(E), Epot


def vdw_forces(wl, N, literal=False):
    """vdW_forces -> Construct_B + d_Forces_s (TB.f90:3599-3781) -> the term added to forces%rep (na,3).
    literal: the O(Na^3) loop nest as written; else the same sums without the pairs whose (dik - djk) is zero"""
    S, h = _vdw_geom(wl)
    F = np.zeros((len(S), 3)); Nc = np.asarray(N, dtype=np.int32)
    rc = lib().orc_vdw_forces(C.c_int(len(S)), C.c_int(_VDW_KIND[wl["vdw_kind"]]), C.c_int(wl["n_kinds"]), _p(_koa0(wl)), _p(S), _p(h),
                              vdw_params_array(wl["vdw_params"]), _p(Nc), C.c_int(1 if literal else 0), _p(F))
    if rc:
        raise MemoryError("orc_vdw_forces")
    return F


def vdw_pressure(wl, N):
    """vdW_forces -> d_Forces_Pressure (TB.f90:3785-3848) -> the term added to SCforce%rep(3,3)"""
    S, h = _vdw_geom(wl)
    P = np.zeros(9); Nc = np.asarray(N, dtype=np.int32)
    rc = lib().orc_vdw_pressure(C.c_int(len(S)), C.c_int(_VDW_KIND[wl["vdw_kind"]]), C.c_int(wl["n_kinds"]), _p(_koa0(wl)), _p(S), _p(h),
                                vdw_params_array(wl["vdw_params"]), _p(Nc), _p(P))
    if rc:
        raise MemoryError("orc_vdw_pressure")
    return P.reshape(3, 3, order="F")


# ---- Coulomb term of a charged supercell (Coulomb.f90, class TB_Coulomb_cut): same loops, kind 4 of vdw_oracle.c ----
_COULOMB = 4


def _coulomb_table(wl):
    """the record the C side evaluates: eps = k*Q*Q (left to right, Coulomb_potential), d0_cut = dm, dd_cut = dd"""
    Q = float(wl.get("Q", 0.0))
    return vdw_params_array([{"eps": float(p.get("k", _abi.COULOMB_K)) * Q * Q, "d0_cut": p["dm"], "dd_cut": p["dd"]} for p in wl["coulomb_params"]])


def coulomb_cells(wl):
    """get_mirror_cell_num_C, Coulomb.f90:435-452 -> (Nx, Ny, Nz)"""
    S, h = _vdw_geom(wl)
    N = np.zeros(3, dtype=np.int32)
    lib().orc_coulomb_cells(C.c_double(float(wl["coulomb_params"][0]["dm"])), _p(S), _p(h), _p(N))
    return tuple(int(x) for x in N)


def coulomb_energy(wl, N):
    """get_Coulomb_s, Coulomb.f90:454-550 -> (E_coul (not halved, as written), amounts added to Epot)"""
    S, h = _vdw_geom(wl)
    Epot = np.zeros(len(S)); Nc = np.asarray(N, dtype=np.int32)
    E = lib().orc_vdw_energy(C.c_int(len(S)), C.c_int(_COULOMB), C.c_int(wl["n_kinds"]), _p(_koa0(wl)), _p(S), _p(h), _coulomb_table(wl), _p(Nc), _p(Epot))
    return float(E), Epot


def coulomb_forces(wl, N, literal=False):
    """Coulomb_forces -> Construct_B_C + d_Forces_s (TB.f90:3560-3596,3638-3781); zero when Q <= 1e-12 (Coulomb.f90:762)"""
    S, h = _vdw_geom(wl)
    F = np.zeros((len(S), 3)); Nc = np.asarray(N, dtype=np.int32)
    if float(wl.get("Q", 0.0)) <= 1.0e-12:
        return F
    rc = lib().orc_vdw_forces(C.c_int(len(S)), C.c_int(_COULOMB), C.c_int(wl["n_kinds"]), _p(_koa0(wl)), _p(S), _p(h), _coulomb_table(wl), _p(Nc),
                              C.c_int(1 if literal else 0), _p(F))
    if rc:
        raise MemoryError("orc_vdw_forces")
    return F


def coulomb_pressure(wl, N):
    """Coulomb_forces -> d_Forces_Pressure (TB.f90:3785-3848); zero when Q <= 1e-12"""
    S, h = _vdw_geom(wl)
    P = np.zeros(9); Nc = np.asarray(N, dtype=np.int32)
    if float(wl.get("Q", 0.0)) > 1.0e-12:
        lib().orc_vdw_pressure(C.c_int(len(S)), C.c_int(_COULOMB), C.c_int(wl["n_kinds"]), _p(_koa0(wl)), _p(S), _p(h), _coulomb_table(wl), _p(Nc), _p(P))
    return P.reshape(3, 3, order="F")


def wall_energy(wl, nb):
    """get_Exp_wall_s, Exponential_wall.f90:990-1067 -> (E_wall, per-atom halves added to Epot)"""
    Epot = np.zeros(nb.na)
    E = lib().orc_wall_energy(C.c_int(nb.na), C.c_int(nb.cap), C.c_int(wl["n_kinds"]), _p(_koa0(wl)), _p(nb.size), _p(nb.list),
                              _p(nb.dist), wall_params_array(wl["wall_params"]), _p(Epot))
    return float(E), Epot


def wall_forces(wl, nb):
    """d_Exp_wall_pot_s, Exponential_wall.f90:1200-1329 (verbatim loop nest) -> the term added to forces%rep (na,3)"""
    F = np.zeros((nb.na, 3))
    h = np.asfortranarray(wl["supce"], dtype=np.float64)
    lib().orc_wall_forces(C.c_int(nb.na), C.c_int(nb.cap), C.c_int(wl["n_kinds"]), _p(_koa0(wl)), _p(h), _p(nb.size), _p(nb.list),
                          _p(nb.dist), wall_params_array(wl["wall_params"]), _p(F))
    return F


def wall_pressure(wl, nb):
    """d_Exp_wall_Pressure_s, Exponential_wall.f90:1334-1385 -> the term added to SCforce%rep(3,3)"""
    P = np.zeros(9)
    lib().orc_wall_pressure(C.c_int(nb.na), C.c_int(nb.cap), C.c_int(wl["n_kinds"]), _p(_koa0(wl)), _p(nb.size), _p(nb.list),
                            _p(nb.dist), _p(nb.dist_s), wall_params_array(wl["wall_params"]), _p(P))
    return P.reshape(3, 3, order="F")


# ---- Fermi occupations and the velocity-Verlet halves ----
def fermi(Ei, n_electrons, Te_eV):
    """Electron_Fixed_Te + set_Fermi, Electron_tools.f90:2660-2706,1913-1969 -> (fe, mu)"""
    Ei = np.ascontiguousarray(Ei, dtype=np.float64)
    fe = np.zeros_like(Ei)
    mu = lib().orc_fermi(C.c_int(Ei.shape[0]), _p(Ei), C.c_double(n_electrons), C.c_double(Te_eV), _p(fe))
    return fe, float(mu)


def md_half_step(ind, koa0, mass_kg, supce, dt, Ftot, S, SV, Vsupce=None):
    """save_last_timestep + make_time_step_atoms_SC(ind), Atomic_tools.f90:1177-1192,1897-1974 -> (S, S0, SV, SV0)"""
    na = S.shape[0]
    S = np.array(S, dtype=np.float64); SV = np.array(SV, dtype=np.float64)
    S0 = np.zeros_like(S); SV0 = np.zeros_like(S)
    h = np.asfortranarray(supce, dtype=np.float64)
    V = None if Vsupce is None else np.asfortranarray(Vsupce, dtype=np.float64)
    F = np.ascontiguousarray(Ftot, dtype=np.float64)
    m = np.ascontiguousarray(mass_kg, dtype=np.float64)
    k0 = np.ascontiguousarray(koa0, dtype=np.int32)
    lib().orc_md_half_step(C.c_int(na), C.c_int(ind), _p(k0), _p(m), _p(h), None if V is None else _p(V), C.c_double(dt),
                           _p(F), _p(S), _p(S0), _p(SV), _p(SV0))
    return S, S0, SV, SV0


def kinetic_energy_eV(koa0, mass_kg, supce, SV):
    """sum_k M_k/2 |V_k|^2 with V(x) = sum_k SV(k) supce(k,x) [A/fs] (velocities_rel_to_abs), in eV"""
    V = np.asarray(SV) @ np.asarray(supce)
    m = np.asarray(mass_kg)[np.asarray(koa0)]
    return float(0.5 * np.sum(m[:, None] * V * V) * 1.0e10 / G_E)


def eigensolve_orthogonal(H):
    """sym_diagonalize(Ha) directly, S = I (TB_Pettifor.f90:1455)."""
    from scipy.linalg import lapack
    Ei, V, info = lapack.dsyevd(np.asfortranarray(H), compute_v=1, lower=0)
    assert info == 0
    return np.asfortranarray(V), Ei


def density_dense(Ha, fe, Ei=None):
    """Construct_Aij / Construct_Aij_x_En (TB.f90:4328-4469): Atemp = w*Ha ; Aij = Atemp Ha^T (dgemm)."""
    w = fe if Ei is None else fe * Ei
    return np.asfortranarray((Ha * w[None, :]) @ Ha.T)


def density_blocks(nb, n_orb, Ha, fe, Ei, atom_first=0, atom_last=None):
    na = nb.na
    atom_last = na if atom_last is None else atom_last
    Ha = np.asfortranarray(Ha)
    K = Ha.shape[1]
    rho = np.zeros((na, nb.cap + 1, n_orb * n_orb)); rhoE = np.zeros_like(rho)
    lib().orc_density_blocks(C.c_int(na), C.c_int(n_orb), C.c_int(nb.cap), _p(nb.size), _p(nb.list), C.c_int(K), _p(Ha),
                             _p(np.ascontiguousarray(fe)), _p(np.ascontiguousarray(Ei)),
                             C.c_int(atom_first), C.c_int(atom_last), _p(rho), _p(rhoE))
    return rho, rhoE


def gather_blocks(nb, n_orb, A):
    blk = np.zeros((nb.na, nb.cap + 1, n_orb * n_orb))
    lib().orc_gather_blocks(C.c_int(nb.na), C.c_int(n_orb), C.c_int(nb.cap), _p(nb.size), _p(nb.list),
                            _p(np.asfortranarray(A)), _p(blk))
    return blk


def nrl_assemble_panels(wl, nb, lmn, rad, j_first, j_last):
    N = nb.na * 9
    w = (j_last - j_first) * 9
    Hp = np.zeros((N, w), order="F"); Sp = np.zeros((N, w), order="F")
    TB = nrl_params_array(wl["params"])
    lib().orc_nrl_assemble_panels(C.c_int(nb.na), C.c_int(nb.cap), _p(nb.size), _p(nb.list), _p(nb.dist), _p(lmn), _p(rad),
                                  _p(_koa0(wl)), C.c_int(wl["n_kinds"]), TB, C.c_int(j_first), C.c_int(j_last), _p(Hp), _p(Sp))
    return Hp, Sp


def density_blocks_rows(nb, n_orb, Ct_rows, rowmap, fe, Ei, a_first, a_last):
    """Ct_rows: (n_gathered_atoms*n_orb, K) C-contiguous; rowmap[atom] -> gathered index or -1."""
    K = Ct_rows.shape[1]
    rho = np.zeros((nb.na, nb.cap + 1, n_orb * n_orb)); rhoE = np.zeros_like(rho)
    lib().orc_density_blocks_rows(C.c_int(nb.na), C.c_int(n_orb), C.c_int(nb.cap), _p(nb.size), _p(nb.list), C.c_int(K),
                                  _p(np.ascontiguousarray(Ct_rows)), _p(np.ascontiguousarray(rowmap, dtype=np.int32)),
                                  _p(np.ascontiguousarray(fe)), _p(np.ascontiguousarray(Ei)),
                                  C.c_int(a_first), C.c_int(a_last), _p(rho), _p(rhoE))
    return rho, rhoE


def _common(wl, nb, lmn, rad):
    h = np.asfortranarray(wl["supce"], dtype=np.float64)
    TB = nrl_params_array(wl["params"])
    return h, TB, _koa0(wl)


def nrl_forces_verbatim(wl, nb, lmn, rad, Aij, AijE):
    h, TB, koa = _common(wl, nb, lmn, rad)
    F = np.zeros((nb.na, 3))
    lib().orc_nrl_forces_verbatim(C.c_int(nb.na), C.c_int(nb.cap), _p(nb.size), _p(nb.list), _p(nb.dist), _p(lmn), _p(rad),
                                  _p(koa), C.c_int(wl["n_kinds"]), TB, _p(h), _p(np.asfortranarray(Aij)),
                                  _p(np.asfortranarray(AijE)), _p(F))
    return F


def nrl_forces_sparse(wl, nb, lmn, rad, rho_blk, rhoE_blk, atom_first=0, atom_last=None):
    h, TB, koa = _common(wl, nb, lmn, rad)
    atom_last = nb.na if atom_last is None else atom_last
    F = np.zeros((nb.na, 3))
    lib().orc_nrl_forces_sparse(C.c_int(nb.na), C.c_int(nb.cap), _p(nb.size), _p(nb.list), _p(nb.dist), _p(lmn), _p(rad),
                                _p(koa), C.c_int(wl["n_kinds"]), TB, _p(h), _p(rho_blk), _p(rhoE_blk),
                                C.c_int(atom_first), C.c_int(atom_last), _p(F))
    return F


def nrl_pressure_verbatim(wl, nb, lmn, rad, Aij, AijE):
    h, TB, koa = _common(wl, nb, lmn, rad)
    att = np.zeros(9)
    lib().orc_nrl_pressure_verbatim(C.c_int(nb.na), C.c_int(nb.cap), _p(nb.size), _p(nb.list), _p(nb.dist), _p(nb.dist_s),
                                    _p(lmn), _p(rad), _p(koa), C.c_int(wl["n_kinds"]), TB, _p(h),
                                    _p(np.asfortranarray(Aij)), _p(np.asfortranarray(AijE)), _p(att))
    return att.reshape(3, 3, order="F")


def nrl_pressure_sparse(wl, nb, lmn, rad, rho_blk, rhoE_blk, atom_first=0, atom_last=None):
    h, TB, koa = _common(wl, nb, lmn, rad)
    atom_last = nb.na if atom_last is None else atom_last
    att = np.zeros(9)
    lib().orc_nrl_pressure_sparse(C.c_int(nb.na), C.c_int(nb.cap), _p(nb.size), _p(nb.list), _p(nb.dist), _p(nb.dist_s),
                                  _p(lmn), _p(rad), _p(koa), C.c_int(wl["n_kinds"]), TB, _p(h), _p(rho_blk), _p(rhoE_blk),
                                  C.c_int(atom_first), C.c_int(atom_last), _p(att))
    return att.reshape(3, 3, order="F")


def ks_block(V, l, m, n):
    ts = np.zeros(81)
    lib().orc_ks_block(_p(np.ascontiguousarray(V, dtype=np.float64)), C.c_double(l), C.c_double(m), C.c_double(n), _p(ts))
    return ts.reshape(9, 9, order="F")


def dks_block(V, dV, l, m, n, dl, dm, dn, test_variant=False):
    ts = np.zeros(81)
    lib().orc_dks_block(_p(np.ascontiguousarray(V, dtype=np.float64)), _p(np.ascontiguousarray(dV, dtype=np.float64)),
                        C.c_double(l), C.c_double(m), C.c_double(n), C.c_double(dl), C.c_double(dm), C.c_double(dn),
                        C.c_int(1 if test_variant else 0), _p(ts))
    return ts.reshape(9, 9, order="F")


# ---- unit conversions used by the tolerance statements of the tests ---------------------------
def forces_cartesian(F_s, supce):
    """dE/ds_k [eV] -> dE/dR_k [eV/A].  R = s.h  (x_c = sum_k s_k supce(k,c), Atomic_tools.f90:4101)
    so dE/ds_k(alpha) = sum_c dE/dR_c supce(alpha,c)  =>  dE/dR = solve(supce, dE/ds)."""
    return np.linalg.solve(np.asarray(supce), np.asarray(F_s).T).T


def pressure_GPa(att, supce):
    """Potential part of Get_pressure: Potential_super_cell_forces + super_cell_forces + stress
    (Atomic_tools.f90:2532-2634, TB.f90:5679-5692) with rep = 0 and no kinetic term; W_PR cancels.
       att' = att*g_e/1e10 ; total(j,i) = att'(i,j) ; total = -total ; sigma(i,j) = d det(h)/d h(i,j) ;
       stress(i,j) = total(i,j) * inv(sigma)(i,j) * 1e40 [Pa]  (ELEMENT-wise, as the Fortran `*`) ;
       P = trace/3.  Returns (P [GPa], stress [GPa])."""
    h = np.asarray(supce, dtype=np.float64)
    total = -(np.asarray(att) * G_E / 1e10).T
    sigma = np.linalg.det(h) * np.linalg.inv(h).T          # cofactor matrix (d_detH_d_h_a_b)
    stress = total * np.linalg.inv(sigma) * 1e40
    return np.trace(stress) / 3.0 * 1e-9, stress * 1e-9


# ---- whole path, for parity tests ----------------------------------------------------------------
def nrl_step(wl, with_pressure=True, verbatim=False):
    nb = neighbours(wl)
    lmn, xrr, x1 = geometry(wl, nb)
    rad = nrl_radial(wl, nb)
    H, S = nrl_assemble_HS(wl, nb, lmn, rad)
    Ha, Ei, wS = eigensolve_nonorthogonal(H, S)
    wk = importlib.import_module("xtant-3_b200.workloads")
    fe, mu = wk.occupations(Ei, wl["n_electrons"], wl["Te_eV"])
    out = dict(nb=nb, lmn=lmn, xrr=xrr, x1=x1, rad=rad, H=H, S=S, Ha=Ha, Ei=Ei, fe=fe, mu=mu, eigen_S=wS)
    if verbatim:
        Aij = density_dense(Ha, fe); AijE = density_dense(Ha, fe, Ei)
        out.update(Aij=Aij, AijE=AijE)
        out["rho"], out["rhoE"] = gather_blocks(nb, 9, Aij), gather_blocks(nb, 9, AijE)
        out["F"] = nrl_forces_verbatim(wl, nb, lmn, rad, Aij, AijE)
        if with_pressure:
            out["P"] = nrl_pressure_verbatim(wl, nb, lmn, rad, Aij, AijE)
    else:
        out["rho"], out["rhoE"] = density_blocks(nb, 9, Ha, fe, Ei)
        out["F"] = nrl_forces_sparse(wl, nb, lmn, rad, out["rho"], out["rhoE"])
        if with_pressure:
            out["P"] = nrl_pressure_sparse(wl, nb, lmn, rad, out["rho"], out["rhoE"])
    return out


# ---- Pettifor (orthogonal sp3) branch: oracle/pettifor_oracle.c ---------------------------------------
def pettifor_params_array(params):
    arr = (_abi.PettiforPair * len(params))()
    for i, d in enumerate(params):
        arr[i] = _abi.pettifor_pair_from_dict(d)
    return arr


def pet_assemble_H(wl, nb):
    """Hamil_tot, TB_Pettifor.f90:1658-1810: dense H (N,N) [eV], N = 4*na (before diagonalisation)."""
    N = 4 * nb.na
    H = np.zeros((N, N), order="F")
    S = np.ascontiguousarray(wl["S"], dtype=np.float64)
    h = np.asfortranarray(wl["supce"], dtype=np.float64)
    lib().orc_pet_assemble_H(C.c_int(nb.na), C.c_int(nb.cap), C.c_int(wl["n_kinds"]), _p(_koa0(wl)), _p(S), _p(h),
                             _p(nb.size), _p(nb.list), pettifor_params_array(wl["params"]), _p(H))
    return H


def pet_forces_verbatim(wl, nb, Aij):
    """dHij_s with dense per-atom derivative matrices (TB_Pettifor.f90:141-223,463-521,354-399)."""
    F = np.zeros((nb.na, 3))
    h = np.asfortranarray(wl["supce"], dtype=np.float64)
    A = np.asfortranarray(Aij)
    lib().orc_pet_forces_verbatim(C.c_int(nb.na), C.c_int(nb.cap), C.c_int(wl["n_kinds"]), _p(_koa0(wl)), _p(h), _p(nb.size),
                                  _p(nb.list), _p(nb.dist), pettifor_params_array(wl["params"]), _p(A), _p(F))
    return F


def pet_forces_sparse(wl, nb, rho_blk, atom_first=0, atom_last=None):
    F = np.zeros((nb.na, 3))
    h = np.asfortranarray(wl["supce"], dtype=np.float64)
    rho_blk = np.ascontiguousarray(rho_blk)
    atom_last = nb.na if atom_last is None else atom_last
    lib().orc_pet_forces_sparse(C.c_int(nb.na), C.c_int(nb.cap), C.c_int(wl["n_kinds"]), _p(_koa0(wl)), _p(h), _p(nb.size),
                                _p(nb.list), _p(nb.dist), pettifor_params_array(wl["params"]), _p(rho_blk),
                                C.c_int(atom_first), C.c_int(atom_last), _p(F))
    return F


def pet_pressure(wl, nb, rho_blk):
    """Attract_TB_Forces_Press, TB_Pettifor.f90:1134-1198 -> SCforce%att(3,3)."""
    att = np.zeros(9)
    rho_blk = np.ascontiguousarray(rho_blk)
    lib().orc_pet_pressure(C.c_int(nb.na), C.c_int(nb.cap), C.c_int(wl["n_kinds"]), _p(_koa0(wl)), _p(nb.size), _p(nb.list),
                           _p(nb.dist), _p(nb.dist_s), pettifor_params_array(wl["params"]), _p(rho_blk), _p(att))
    return att.reshape(3, 3, order="F")


def pettifor_rep_params_array(params):
    arr = (_abi.PettiforRepPair * len(params))()
    for i, d in enumerate(params):
        arr[i] = _abi.pettifor_rep_pair_from_dict(d)
    return arr


def pet_erep(wl, nb):
    """get_Erep_s, TB_Pettifor.f90:2074-2105 -> (E_rep, Epot(na))"""
    Epot = np.zeros(nb.na)
    E = lib().orc_pet_erep(C.c_int(nb.na), C.c_int(nb.cap), C.c_int(wl["n_kinds"]), _p(_koa0(wl)), _p(nb.size), _p(nb.list),
                           _p(nb.dist), pettifor_rep_params_array(wl["rep_params"]), _p(Epot))
    return float(E), Epot


def pet_rep_forces(wl, nb, verbatim=False, atom_first=0, atom_last=None):
    """dErdr_s, TB_Pettifor.f90:2162-2306 -> forces%rep (na,3); verbatim = the reference's O(na^2 nn) loop nest"""
    F = np.zeros((nb.na, 3))
    h = np.asfortranarray(wl["supce"], dtype=np.float64)
    args = (C.c_int(nb.na), C.c_int(nb.cap), C.c_int(wl["n_kinds"]), _p(_koa0(wl)), _p(h), _p(nb.size), _p(nb.list),
            _p(nb.dist), pettifor_rep_params_array(wl["rep_params"]))
    if verbatim:
        lib().orc_pet_rep_forces_verbatim(*args, _p(F))
    else:
        lib().orc_pet_rep_forces_sparse(*args, C.c_int(atom_first), C.c_int(nb.na if atom_last is None else atom_last), _p(F))
    return F


def pet_rep_pressure(wl, nb):
    """dErdr_Pressure_s, TB_Pettifor.f90:2634-2690 -> SCforce%rep(3,3)"""
    rep = np.zeros(9)
    lib().orc_pet_rep_pressure(C.c_int(nb.na), C.c_int(nb.cap), C.c_int(wl["n_kinds"]), _p(_koa0(wl)), _p(nb.size), _p(nb.list),
                               _p(nb.dist), _p(nb.dist_s), pettifor_rep_params_array(wl["rep_params"]), _p(rep))
    return rep.reshape(3, 3, order="F")


def eigensolve_orthogonal(H):
    """sym_diagonalize -> dsyevd (Algebra_tools.f90:882-966), as construct_TB_H_Pettifor calls it (TB_Pettifor.f90:1455)."""
    from scipy.linalg import lapack
    w, v, info = lapack.dsyevd(np.asfortranarray(H), compute_v=1, lower=0)
    if info != 0:
        raise RuntimeError(f"dsyevd info={info}")
    return np.asfortranarray(v), w


def pettifor_step(wl, with_pressure=True, verbatim=False):
    nb = neighbours(wl)
    H = pet_assemble_H(wl, nb)
    Ha, Ei = eigensolve_orthogonal(H)
    wk = importlib.import_module("xtant-3_b200.workloads")
    fe, mu = wk.occupations(Ei, wl["n_electrons"], wl["Te_eV"])
    out = dict(nb=nb, H=H, Ha=Ha, Ei=Ei, fe=fe, mu=mu)
    if verbatim:
        Aij = density_dense(Ha, fe)
        out["Aij"] = Aij
        out["rho"] = gather_blocks(nb, 4, Aij)
        out["F"] = pet_forces_verbatim(wl, nb, Aij)
    else:
        out["rho"], out["rhoE"] = density_blocks(nb, 4, Ha, fe, Ei)
        out["F"] = pet_forces_sparse(wl, nb, out["rho"])
    if with_pressure:
        out["P"] = pet_pressure(wl, nb, out["rho"])
    return out


# ---- DFTB branch: oracle/dftb_oracle.c ----------------------------------------------------------------
def dftb_cutoff(wl):
    """get_near_neighbours for TB_H_DFTB, Atomic_tools.f90:588-591: rm = max(rcut) + 10*max(d)."""
    return max(p["rcut"] for p in wl["params"]) + 10.0 * max(p["d"] for p in wl["params"])


def dftb_radial(wl, nb):
    """Construct_Vij_DFTB, TB_DFTB.f90:56-228 -> compact (na, cap, 4, 10): M_Vij, M_dVij, M_SVij, M_dSVij of entry (j, a2)."""
    rad = np.zeros((nb.na, nb.cap, 40))
    TB, keep = _abi.dftb_pair_array(wl["params"])
    lib().orc_dftb_radial(C.c_int(nb.na), C.c_int(nb.cap), _p(nb.size), _p(nb.list), _p(nb.dist), _p(_koa0(wl)),
                          C.c_int(wl["n_kinds"]), TB, C.c_int(wl["basis"]), _p(rad))
    return rad


def dftb_assemble_HS(wl, nb, lmn, rad):
    N = nb.na * wl["n_orb"]
    H = np.zeros((N, N), order="F"); S = np.zeros((N, N), order="F")
    TB, keep = _abi.dftb_pair_array(wl["params"])
    lib().orc_dftb_assemble_HS(C.c_int(nb.na), C.c_int(nb.cap), _p(nb.size), _p(nb.list), _p(nb.dist), _p(lmn), _p(rad),
                               _p(_koa0(wl)), C.c_int(wl["n_kinds"]), TB, C.c_int(wl["basis"]), _p(H), _p(S))
    return H, S


def dftb_forces_verbatim(wl, nb, lmn, rad, Aij, AijE):
    h = np.asfortranarray(wl["supce"], dtype=np.float64)
    TB, keep = _abi.dftb_pair_array(wl["params"])
    F = np.zeros((nb.na, 3))
    lib().orc_dftb_forces_verbatim(C.c_int(nb.na), C.c_int(nb.cap), _p(nb.size), _p(nb.list), _p(nb.dist), _p(lmn), _p(rad),
                                   _p(_koa0(wl)), C.c_int(wl["n_kinds"]), TB, C.c_int(wl["basis"]), _p(h),
                                   _p(np.asfortranarray(Aij)), _p(np.asfortranarray(AijE)), _p(F))
    return F


def dftb_forces_sparse(wl, nb, lmn, rad, rho_blk, rhoE_blk, atom_first=0, atom_last=None):
    h = np.asfortranarray(wl["supce"], dtype=np.float64)
    TB, keep = _abi.dftb_pair_array(wl["params"])
    atom_last = nb.na if atom_last is None else atom_last
    F = np.zeros((nb.na, 3))
    lib().orc_dftb_forces_sparse(C.c_int(nb.na), C.c_int(nb.cap), _p(nb.size), _p(nb.list), _p(nb.dist), _p(lmn), _p(rad),
                                 _p(_koa0(wl)), C.c_int(wl["n_kinds"]), TB, C.c_int(wl["basis"]), _p(h),
                                 _p(np.ascontiguousarray(rho_blk)), _p(np.ascontiguousarray(rhoE_blk)),
                                 C.c_int(atom_first), C.c_int(atom_last), _p(F))
    return F


def dftb_pressure(wl, nb, lmn, rad, rho_blk, rhoE_blk, atom_first=0, atom_last=None):
    TB, keep = _abi.dftb_pair_array(wl["params"])
    atom_last = nb.na if atom_last is None else atom_last
    att = np.zeros(9)
    lib().orc_dftb_pressure(C.c_int(nb.na), C.c_int(nb.cap), _p(nb.size), _p(nb.list), _p(nb.dist), _p(nb.dist_s), _p(lmn),
                            _p(rad), _p(_koa0(wl)), C.c_int(wl["n_kinds"]), TB, C.c_int(wl["basis"]),
                            _p(np.ascontiguousarray(rho_blk)), _p(np.ascontiguousarray(rhoE_blk)),
                            C.c_int(atom_first), C.c_int(atom_last), _p(att))
    return att.reshape(3, 3, order="F")


def dftb_step(wl, with_pressure=True, verbatim=False):
    nb = neighbours(wl, rm=dftb_cutoff(wl))
    lmn, xrr, x1 = geometry(wl, nb)
    rad = dftb_radial(wl, nb)
    H, S = dftb_assemble_HS(wl, nb, lmn, rad)
    Ha, Ei, wS = eigensolve_nonorthogonal(H, S)
    wk = importlib.import_module("xtant-3_b200.workloads")
    fe, mu = wk.occupations(Ei, wl["n_electrons"], wl["Te_eV"])
    no = wl["n_orb"]
    out = dict(nb=nb, lmn=lmn, rad=rad, H=H, S=S, Ha=Ha, Ei=Ei, fe=fe, mu=mu, eigen_S=wS)
    out["rho"], out["rhoE"] = density_blocks(nb, no, Ha, fe, Ei)
    if verbatim:
        Aij = density_dense(Ha, fe); AijE = density_dense(Ha, fe, Ei)
        out.update(Aij=Aij, AijE=AijE)
        out["F"] = dftb_forces_verbatim(wl, nb, lmn, rad, Aij, AijE)
    else:
        out["F"] = dftb_forces_sparse(wl, nb, lmn, rad, out["rho"], out["rhoE"])
    if with_pressure:
        out["P"] = dftb_pressure(wl, nb, lmn, rad, out["rho"], out["rhoE"])
    return out


# ---- nonadiabatic coupling matrix -------------------------------------------------------------
def coupling_Mij_loops(Ha, Ha0, Sij=None, kind_M=0):
    """Electron_ion_coupling_Mij, Nonadiabatic.f90:782-945, statement by statement (pure Python loops: small N only).
    kind_M = -1: the `case(-1)` arm (:816-874); anything else: Tully's expression, the `case default` arm (:876-938)."""
    N = Ha.shape[0]
    eps = 1.0e-13                                                         # :812
    Mij = np.zeros((N, N))
    Norm0 = np.array([np.sqrt(np.sum(Ha0[:, i] * Ha0[:, i])) for i in range(N)])      # :823 / :883
    Norm1 = np.array([np.sqrt(np.sum(Ha[:, i] * Ha[:, i])) for i in range(N)])
    tij = np.zeros((N, N))
    if Sij is not None:                                                   # :829-842 / :889-902
        for i in range(N):
            for k in range(N):
                tij[i, k] = 0.0 if Norm1[i] < eps else np.sum(Sij[k, :] * Ha[:, i] / Norm1[i])
    for i in range(N):
        psi0 = np.zeros(N) if Norm0[i] < eps else Ha0[:, i] / Norm0[i]
        psi_back = np.zeros(N) if Norm1[i] < eps else Ha[:, i] / Norm1[i]
        for j in range(i, N):
            if i == j:
                continue
            Norm_val1 = 0.0 if Norm1[j] < eps else 1.0 / Norm1[j]
            Norm_val0 = 0.0 if Norm0[j] < eps else 1.0 / Norm0[j]
            if kind_M == -1:
                Mij[i, j] = np.sum(psi0 * tij[j, :]) if Sij is not None else np.sum(psi0 * Ha[:, j] * Norm_val1)      # :862-866
            elif Sij is not None:
                Mij[i, j] = np.sum(psi0 * tij[j, :] - (Ha0[:, j] * Norm_val0) * tij[i, :])                            # :927
            else:
                Mij[i, j] = np.sum(psi0 * (Ha[:, j] * Norm_val1) - (Ha0[:, j] * Norm_val0) * psi_back)                # :930
            Mij[j, i] = Mij[i, j]
    if kind_M != -1:
        Mij = Mij * 0.5                                                   # :938
    return Mij


def coupling_Mij(Ha, Ha0, Sij=None, kind_M=0):
    """Same result through matrix products (what the loops amount to): G = psi0^T S psi ;
    Tully: M(i<j) = (G(i,j) - G(j,i))/2 ; case(-1): M(i<j) = G(i,j) ; M(j,i) = M(i,j) ; M(i,i) = 0."""
    eps = 1.0e-13
    n0 = np.sqrt((Ha0 * Ha0).sum(axis=0)); n1 = np.sqrt((Ha * Ha).sum(axis=0))
    i0 = np.where(n0 < eps, 0.0, 1.0 / np.where(n0 < eps, 1.0, n0))
    i1 = np.where(n1 < eps, 0.0, 1.0 / np.where(n1 < eps, 1.0, n1))
    P0 = Ha0 * i0[None, :]; P1 = Ha * i1[None, :]
    G = P0.T @ (P1 if Sij is None else Sij @ P1)
    U = np.triu(G if kind_M == -1 else 0.5 * (G - G.T), 1)
    return U + U.T
