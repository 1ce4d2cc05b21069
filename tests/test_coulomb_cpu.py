"""CPU tests of the oracle's Coulomb term of a charged supercell (oracle/vdw_oracle.c, kind 4: get_mirror_cell_num_C, get_Coulomb_s,
Construct_B_C, d_Forces_s, d_Forces_Pressure; Coulomb.f90:435-880, TB.f90:3560-3848) and of its file reader.

No reference vectors exist for this term.  Pins: the pair potential against its closed form; forces and supercell force against
finite differences of the energy -- with the reference's own normalisation: get_Coulomb_s does NOT halve the double sum
(Coulomb.f90:549, unlike get_vdW_s) while d_Forces_s does (TB.f90:3777), so F = 1/2 dE/ds as written; the transposed contraction
of Construct_B_C (:864-866) against the one of Construct_B on a sheared cell; zero forces for Q <= 1e-12 (:762)."""
import importlib
import os

import numpy as np
import pytest


def charged(wl, Q=0.05, dm=6.0, dd=0.1):
    wl = dict(wl)
    wl["coulomb_params"] = [{"dm": dm, "dd": dd}] * (wl["n_kinds"] ** 2)
    wl["Q"] = Q
    return wl


def sheared(wk):
    """64-atom carbon cell with a non-symmetric supce: the transposed contraction of Construct_B_C differs from Construct_B's"""
    wl = dict(wk.c_dia((2, 2, 2), sigma=0.1, seed=31))
    h = np.array(wl["supce"], dtype=np.float64)
    h[0, 1] = 0.9; h[2, 0] = -0.6; h[1, 2] = 0.4
    wl["supce"] = h
    return wl


def test_reader_class_line_semantics(tmp_path):
    rd = importlib.import_module("xtant-3_b200.readers")
    abi = importlib.import_module("xtant-3_b200._abi")
    p = tmp_path / "C_C_TB_Coulomb.txt"
    rd.write_Coulomb_TB(str(p), {"dm": 6.0, "dd": 0.1})
    assert rd.read_Coulomb_TB(str(p)) == {"k": rd.COULOMB_K, "dm": 6.0, "dd": 0.1}
    assert rd.COULOMB_K == abi.COULOMB_K and abs(rd.COULOMB_K - 14.3996454) < 1e-6        # e^2 / (4 pi eps0) [eV A]
    rd.write_Coulomb_TB(str(p), {"dm": 6.0, "dd": 0.1}, enabled=False)
    assert rd.read_Coulomb_TB(str(p)) is None
    ref = "/root/reference/INPUT_DATA"
    if os.path.isdir(ref):
        assert rd.read_Coulomb_TB(f"{ref}/Diamond/C_C_TB_Coulomb.txt") is None             # 'Coulomb_cut:exclude'
        assert rd.read_Coulomb_TB(f"{ref}/C60_3/C_C_TB_Coulomb.txt") == {"k": rd.COULOMB_K, "dm": 35.0, "dd": 0.1}


def test_pair_potential_closed_form(orc):
    abi = importlib.import_module("xtant-3_b200._abi")
    Q, dm, dd = 0.3, 6.0, 0.1
    W = {"eps": abi.COULOMB_K * Q * Q, "d0_cut": dm, "dd_cut": dd}
    lib = orc.lib()
    import ctypes as C
    rec = abi.vdw_pair_from_dict(W)
    for r in (1.4, 3.0, 5.9, 6.0, 6.5, 6.99):
        f = 1.0 / (1.0 + np.exp((r - dm) / dd))
        V = float(lib.orc_vdw_V(C.c_int(4), C.byref(rec), C.c_double(r)))
        dV = float(lib.orc_vdw_dV(C.c_int(4), C.byref(rec), C.c_double(r)))
        assert abs(V - W["eps"] / r * f) <= 1e-15 * abs(V)
        assert abs(dV - (-W["eps"] / r ** 2 * f - W["eps"] / r * f * (1 - f) / dd)) <= 1e-12 * abs(dV)
    assert float(lib.orc_vdw_V(C.c_int(4), C.byref(rec), C.c_double(7.0 + 1e-9))) == 0.0


def test_image_cells_use_dm_alone(orc, wk):
    wl = charged(wk.ac216())                          # 12.53 A cube: 2 dm = 12 -> 0 -> 1
    assert orc.coulomb_cells(wl) == (1, 1, 1)
    assert orc.coulomb_cells(charged(wk.ac216(), dm=20.0)) == (4, 4, 4)       # 3 * 12.53 < 40 <= 4 * 12.53
    assert orc.coulomb_cells(charged(wk.diamond64(), dm=6.0)) == (2, 2, 2)


@pytest.mark.parametrize("case", ["ac216", "sheared", "two_kinds"])
def test_forces_and_pressure_vs_finite_differences(orc, wk, case):
    wl = {"ac216": lambda: charged(wk.ac216()), "sheared": lambda: charged(sheared(wk), Q=0.2),
          "two_kinds": lambda: charged(wk.ch_zb(), Q=0.1, dm=5.0)}[case]()
    N = orc.coulomb_cells(wl)
    E, Epot = orc.coulomb_energy(wl, N)
    assert E > 0 and abs(Epot.sum() - E) <= 1e-12 * E
    F = orc.coulomb_forces(wl, N); P = orc.coulomb_pressure(wl, N)
    assert np.array_equal(F, orc.coulomb_forces(wl, N, literal=True)) if wl["na"] <= 64 else True
    h = np.array(wl["supce"], dtype=np.float64)
    sym = np.allclose(h, h.T)
    d = 1e-6
    G = np.zeros((3, 3))                              # dE/ds of three atoms, all components
    atoms = (0, 7, wl["na"] - 1)
    for a, k in enumerate(atoms):
        for al in range(3):
            Es = []
            for sgn in (+1, -1):
                w = dict(wl); S = wl["S"].copy(); S[k, al] += sgn * d; w["S"] = S
                Es.append(orc.coulomb_energy(w, N)[0])
            G[a, al] = (Es[0] - Es[1]) / (2 * d)
    if sym:
        assert np.abs(0.5 * G - F[list(atoms)]).max() <= 2e-6 * max(1.0, np.abs(G).max())
    else:
        # as written the force carries h^T . r instead of h . r (Construct_B_C :864-866): F = 1/2 (h^T h^-1) dE/ds
        M = h.T @ np.linalg.inv(h)
        assert np.abs(0.5 * G @ M.T - F[list(atoms)]).max() <= 2e-6 * max(1.0, np.abs(G).max())
        assert np.abs(0.5 * G - F[list(atoms)]).max() > 1e-3 * np.abs(G).max()
    for l, k in ((0, 0), (1, 1), (2, 2), (0, 1), (2, 0)):
        Es = []
        for sgn in (+1, -1):
            w = dict(wl); hh = h.copy(); hh[l, k] += sgn * d; w["supce"] = hh
            Es.append(orc.coulomb_energy(w, N)[0])
        fd = (Es[0] - Es[1]) / (2 * d)
        assert abs(0.5 * fd - P[l, k]) <= 2e-6 * max(1.0, abs(fd)), (l, k, fd, P[l, k])


def test_no_charge_no_force(orc, wk):
    wl = charged(wk.ac216(), Q=1.0e-12)
    N = orc.coulomb_cells(wl)
    assert not orc.coulomb_forces(wl, N).any() and not orc.coulomb_pressure(wl, N).any()
    assert 0.0 < orc.coulomb_energy(wl, N)[0] < 1e-18                       # the energy routine has no such switch
