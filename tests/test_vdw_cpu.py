"""CPU tests of the oracle's van der Waals term (oracle/vdw_oracle.c: get_mirror_cell_num, get_vdW_s, Construct_B, d_Forces_s,
d_Forces_Pressure; Van_der_Waals.f90:132-421,578-1134, TB.f90:3599-3848) and of its file reader (Read_input_data.f90:1709-1840).

The reference holds no test or golden vector for this term.  Pins available without them:
  * the Girifalco files the reference ships join their polynomial arms to the Lennard-Jones arm continuously in value and slope,
    and end at zero with zero slope -- true only if the 16 numbers are read in the right order into the right formulas;
  * forces and supercell force are the exact derivatives of the energy (finite differences), for all three classes and for a
    two-kind table, symmetric or not;
  * the O(Na^2) form of d_Forces_s the GPU tests compare against is bit-identical to the loop nest as written (O(Na^3))."""
import importlib
import os

import numpy as np
import pytest

LJ_PD = {"eps": 0.17137125, "r0": 3.5025827299294447, "n": 6.0, "d0_short": 2.0, "dd_short": 0.03, "d0_cut": 5.0, "dd_cut": 0.1}


def two_kind_lj(wk, asymmetric=False):
    """zincblende C/H cell of the DFTB tests carrying a Lennard-Jones table (koa1, koa2) column-major"""
    wl = dict(wk.ch_zb())
    a = dict(LJ_PD)
    b = dict(LJ_PD, eps=0.05, r0=3.1)
    c = dict(LJ_PD, eps=0.11, r0=3.3, n=5.0)
    c2 = dict(c, eps=0.13) if asymmetric else c
    wl["vdw_kind"] = "LJ"
    wl["vdw_params"] = [a, c, c2, b]
    return wl


def ilj_cell(wk):
    wl = dict(wk.c_dia((2, 2, 2), sigma=0.15, seed=21))
    wl["vdw_kind"] = "ILJ"
    wl["vdw_params"] = [{"eps": 0.02, "r0": 3.6, "n": 11.0, "m": 6.0, "d0_short": 2.2, "dd_short": 0.05, "d0_cut": 5.5, "dd_cut": 0.08}]
    return wl


def test_reader_class_line_semantics(tmp_path, wk):
    rd = importlib.import_module("xtant-3_b200.readers")
    g = wk.ac216()["vdw_params"][0]
    p = tmp_path / "C_C_TB_vdW.txt"
    rd.write_vdW_TB(str(p), "Girifalco", g)
    kind, q = rd.read_vdW_TB(str(p))
    assert kind == "Girifalco" and q == g and q["d0_cut"] == q["d_cut"] and q["dd_cut"] == 0.0
    rd.write_vdW_TB(str(p), "Girifalco", g, enabled=False)       # 'Girifalco:exclude' -> case default -> no vdW
    assert rd.read_vdW_TB(str(p)) is None
    p.write_text("Girifalco: exclude\n1.0\n")                     # list-directed read stops at the blank: 'Girifalco:' -> default
    assert rd.read_vdW_TB(str(p)) is None
    rd.write_vdW_TB(str(p), "LJ", LJ_PD)
    assert rd.read_vdW_TB(str(p)) == ("LJ", LJ_PD)
    p.write_text("LJ\nSE\n0.02\t4.0d0\t\t! epsylon, r0, n\n4.0d0 0.03d0 ! short\n6.0d0 0.1d0 ! long\n")
    kind, q = rd.read_vdW_TB(str(p))                              # third token is not a number: n = 6 (Read_input_data.f90:2600-2604)
    assert kind == "LJ" and q["n"] == 6.0 and q["eps"] == 0.02 / 8.0 and q["r0"] == 2.0 ** 0.5 * 4.0
    p.write_text("LJ\nAB\n100.0 4.0\n1.0 0.1\n6.0 0.1\n")
    q = rd.read_vdW_TB(str(p))[1]
    assert q["eps"] == 4.0 ** 2 / 400.0 and q["r0"] == 25.0 ** (1.0 / 6.0)
    ilj = {"eps": 0.02, "r0": 3.6, "n": 11.0, "m": 6.0, "d0_short": 2.2, "dd_short": 0.05, "d0_cut": 5.5, "dd_cut": 0.08}
    rd.write_vdW_TB(str(p), "ILJ", ilj)
    assert rd.read_vdW_TB(str(p)) == ("ILJ", ilj)
    p.write_text("ILJ\n0.02 3.6 ! two numbers only\n2.2 0.05\n5.5 0.08\n")
    q = rd.read_vdW_TB(str(p))[1]
    assert q["n"] == 12.0 and q["m"] == 6.0
    p.write_text("Dumitrica\n")
    with pytest.raises(NotImplementedError):
        rd.read_vdW_TB(str(p))


def test_reference_fixture_files_when_present(wk):
    ref = "/root/reference/INPUT_DATA"
    if not os.path.isdir(ref):
        pytest.skip("reference tree not present")
    rd = importlib.import_module("xtant-3_b200.readers")
    for f in ("Diamond/C_C_TB_vdW.txt", "Graphite/C_C_TB_vdW.txt", "Graphene/C_C_TB_vdW.txt", "C_layer/C_C_TB_vdW.txt"):
        assert rd.read_vdW_TB(os.path.join(ref, f)) is None, f     # shipped switched off
    kind, p = rd.read_vdW_TB(os.path.join(ref, "Amorphous_Carbon/C_C_TB_vdW.txt"))
    assert kind == "Girifalco" and p == wk.ac216()["vdw_params"][0]
    kind, p = rd.read_vdW_TB(os.path.join(ref, "C60_Pa3_2/C_C_TB_vdW.txt"))       # 'Girifalco<TAB>! From [Lu et al. ...]' is active
    assert kind == "Girifalco" and p["C12"] == 29000.387
    assert rd.read_vdW_TB(os.path.join(ref, "Te2Pd/Pd_Te_TB_vdW.txt"))[0] == "LJ"


@pytest.mark.parametrize("folder", ["Amorphous_Carbon", "C60", "Overdence_aC"])
def test_shipped_girifalco_sets_are_continuous(orc, wk, folder):
    """joins at dm and dsm, ends at d_cut and ds_cut: value and slope (the files carry ~10 digits)"""
    if folder == "Amorphous_Carbon":
        p = wk.ac216()["vdw_params"][0]
    else:
        path = f"/root/reference/INPUT_DATA/{folder}/C_C_TB_vdW.txt"
        if not os.path.exists(path):
            pytest.skip("reference tree not present")
        p = importlib.import_module("xtant-3_b200.readers").read_vdW_TB(path)[1]
    V = lambda r, d=False: orc.vdw_pair_value("Girifalco", p, r, d)
    e = 1e-9
    lj = lambda r: p["C12"] / r ** 12 - p["C6"] / r ** 6
    dlj = lambda r: -12 * p["C12"] / r ** 13 + 6 * p["C6"] / r ** 7
    mid = 0.5 * (p["dsm"] + p["dm"])
    assert abs(V(mid) - lj(mid)) <= 1e-14 * abs(lj(mid)) and abs(V(mid, True) - dlj(mid)) <= 1e-13 * abs(dlj(mid))
    scale, dscale = abs(lj(p["dm"])), abs(dlj(p["dm"]))
    assert abs(V(p["dm"] - e) - V(p["dm"] + e)) < 1e-6 * scale and abs(V(p["dm"] - e, True) - V(p["dm"] + e, True)) < 1e-5 * dscale
    assert abs(V(p["d_cut"] - e)) < 1e-6 * scale and abs(V(p["d_cut"] - e, True)) < 1e-5 * dscale
    assert V(p["d_cut"] + e) == 0.0 and V(p["ds_cut"] - e) == 0.0
    scale, dscale = abs(lj(p["dsm"])), abs(dlj(p["dsm"]))
    assert abs(V(p["dsm"] - e) - V(p["dsm"] + e)) < 1e-3 * scale and abs(V(p["dsm"] - e, True) - V(p["dsm"] + e, True)) < 1e-4 * dscale
    assert abs(V(p["ds_cut"] + e)) < 1e-3 * scale and abs(V(p["ds_cut"] + e, True)) < 1e-4 * dscale


def test_image_cells(orc, wk):
    wl = wk.ac216()                                    # 12.53 A cube, R_cut = max(2.6, 6.0): 12.53 >= 12 -> 0 -> at least 1
    assert orc.vdw_cells(wl, orc.cutoff_radius(wl)) == ((1, 1, 1), 6.0)
    wl = wk.with_vdw(wk.diamond64(), "C_amorphous")    # 7.13 A cube: 1 * 7.13 < 12, 2 * 7.13 >= 12
    assert orc.vdw_cells(wl, orc.cutoff_radius(wl))[0] == (2, 2, 2)
    wl = dict(wl); wl["supce"] = np.diag([7.13, 14.0, 3.9])
    assert orc.vdw_cells(wl, 2.6)[0] == (2, 1, 4)      # z: 4 * 3.9 = 15.6 >= 12 > 3 * 3.9
    wl = two_kind_lj(wk)                               # maxval(d0_cut) + 10 maxval(dd_cut) = 6.0, TB cut-off larger
    assert orc.vdw_cells(wl, 7.5)[1] == 7.5 and orc.vdw_cells(wl, 3.0)[1] == 6.0


def test_ac216_energy_is_attractive_and_forces_balance(orc, wk):
    wl = wk.ac216()
    N = (1, 1, 1)
    E, Epot = orc.vdw_energy(wl, N)
    assert -20.0 < E < -1.0 and abs(Epot.sum() - E) <= 1e-12 * abs(E)
    F = orc.vdw_forces(wl, N)
    assert np.abs(F).max() > 1e-2 and np.abs(F.sum(axis=0)).max() < 1e-12 * np.abs(F).max() * wl["na"]


@pytest.mark.parametrize("case", ["girifalco", "lj2", "lj2_asym", "ilj"])
def test_collapsed_force_loops_equal_the_loop_nest_as_written(orc, wk, case):
    wl = {"girifalco": lambda: wk.with_vdw(wk.diamond64(), "C_amorphous"), "lj2": lambda: two_kind_lj(wk),
          "lj2_asym": lambda: two_kind_lj(wk, True), "ilj": lambda: ilj_cell(wk)}[case]()
    N = orc.vdw_cells(wl, 2.6)[0]
    assert np.array_equal(orc.vdw_forces(wl, N), orc.vdw_forces(wl, N, literal=True))


@pytest.mark.parametrize("case", ["girifalco", "lj2", "lj2_asym", "ilj"])
def test_forces_and_pressure_are_derivatives_of_the_energy(orc, wk, case):
    wl = {"girifalco": wk.ac216, "lj2": lambda: two_kind_lj(wk), "lj2_asym": lambda: two_kind_lj(wk, True), "ilj": lambda: ilj_cell(wk)}[case]()
    N = orc.vdw_cells(wl, 2.6)[0]
    F = orc.vdw_forces(wl, N)
    P = orc.vdw_pressure(wl, N)
    assert np.abs(F).max() > 1e-4
    d = 1e-6
    for k, al in ((0, 0), (7, 1), (wl["na"] - 1, 2)):
        E = []
        for sgn in (+1, -1):
            w = dict(wl); S = wl["S"].copy(); S[k, al] += sgn * d; w["S"] = S
            E.append(orc.vdw_energy(w, N)[0])
        fd = (E[0] - E[1]) / (2 * d)
        assert abs(fd - F[k, al]) <= 2e-6 * max(1.0, abs(fd)), (k, al, fd, F[k, al])
    for l, k in ((0, 0), (1, 1), (2, 2), (0, 1), (2, 0)):
        E = []
        for sgn in (+1, -1):
            w = dict(wl); h = np.array(wl["supce"], dtype=np.float64); h[l, k] += sgn * d; w["supce"] = h
            E.append(orc.vdw_energy(w, N)[0])
        fd = (E[0] - E[1]) / (2 * d)
        assert abs(fd - P[l, k]) <= 2e-6 * max(1.0, abs(fd)), (l, k, fd, P[l, k])


def test_term_vanishes_when_every_pair_is_outside_its_window(orc, wk):
    wl = wk.with_vdw(wk.diamond64(), "C_amorphous")
    p = dict(wl["vdw_params"][0], d_cut=1.0, dm=0.9, dsm=0.8, ds_cut=0.7); p["d0_cut"] = p["d_cut"]
    wl["vdw_params"] = [p]
    assert orc.vdw_energy(wl, (1, 1, 1))[0] == 0.0 and not orc.vdw_forces(wl, (1, 1, 1)).any()
