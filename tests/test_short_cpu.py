"""CPU tests of the oracle's 'General' short-range class (oracle/short_oracle.c: get_short_range_rep_s, Shortrange_pot,
d_Short_range_pot, d_Short_range_pot_s, d_Short_range_Pressure_s and the ZBL / NLH / tabulated / spline helpers;
Exponential_wall.f90:48-967, ZBL_potential.f90, Algebra_tools.f90:115-273) and of its readers (Read_input_data.f90:2110-2312).

No reference test or vector exists for this class.  Pins: the shipped parameter files parse into the committed sets; each of the
six functions against an independent closed form; the spline coefficients of the host reader equal the oracle's own restatement
bit for bit and reproduce the table at its nodes; forces and supercell force are the derivatives of the energy (finite
differences) for every form, one and two kinds of atoms, symmetric and asymmetric pair tables."""
import importlib
import os

import numpy as np
import pytest

REF = "/root/reference/INPUT_DATA"


def pbi2_like(wk, asymmetric=False):
    """two-kind cell (geometry of the C/H zincblende DFTB cell, stretched) carrying the shipped PbI2 sets: Pow polynomial + ZBL"""
    wl = dict(wk.ch_zb())
    wl["supce"] = np.array(wl["supce"]) * 1.55            # first neighbours ~2.5 A, second ~4.1 A: inside / outside the cut-offs
    pp, pi, ii = wk.load_short("Pb_Pb"), wk.load_short("Pb_I"), wk.load_short("I_I")
    ip = dict(pi, pow=[(t[0] * 1.07, t[1], t[2]) for t in pi["pow"]]) if asymmetric else pi
    wl["short_params"] = [pp, ip, pi, ii]                 # (koa1, koa2) column-major: (1,1), (2,1), (1,2), (2,2)
    wl["Z"] = [82.0, 53.0]
    return wl


def one_kind(wk, p, Z=8.0, scale=1.0, reps=(2, 2, 2)):
    wl = dict(wk.c_dia(reps, sigma=0.12, seed=17))
    wl["supce"] = np.array(wl["supce"]) * scale
    wl["short_params"] = [p]
    wl["Z"] = [Z]
    return wl


def synthetic_all_forms(wk):
    """every function at once on one kind: Exp + Inv_exp + Pow + ZBL + NLH + a splined table"""
    R = np.linspace(0.6, 3.2, 40)
    E = 30.0 * np.exp(-1.3 * R) * (1.0 + 0.2 * np.sin(3.0 * R))
    rd = importlib.import_module("xtant-3_b200.readers")
    a, b, c, d = rd.make_cubic_splines(R, E)
    p = {"d0": 2.4, "dd": 0.07, "exp": (12.0, 1.8, 0.35), "inv_exp": (3.0, 0.2), "pow": [(-2.0, 1.5, 0.0), (0.7, 1.5, -2.5)],
         "zbl": True, "nlh": True, "nlh_a": (0.04216, 0.95784, 0.0), "nlh_b": (7.5, 2.1, 0.0),
         "tab": {"R": R, "E": E, "spline": True, "a": a, "b": b, "c": c, "d": d}}
    return one_kind(wk, p, Z=6.0)


def test_reader_keywords_and_class_line(tmp_path):
    rd = importlib.import_module("xtant-3_b200.readers")
    f = tmp_path / "A_B_TB_short.txt"
    f.write_text("General\n\nCut-off\n2.7d0\t0.1d0\t\t! d0 [A]; dd [A]\n\nPow\n2\n-465.23d0 1.0d0 0.0 ! E r0 m\n658.0933d0\t1.0d0\t1.0\n"
                 "\n!Pow\n1\n-2.0d0 1.5d0 0.0\nZBL\nInv_exp\n3.0 0.2\nExp\n10.0e0 2.0e0 0.1e0\nsome comment line\nCutoff_short\n0.5 0.01\n")
    p = rd.read_Short_Rep_TB(str(f))
    assert (p["d0"], p["dd"]) == (2.7, 0.1) and p["pow"] == [(-465.23, 1.0, 0.0), (658.0933, 1.0, 1.0)]     # '!Pow' and its lines are skipped
    assert p["zbl"] and not p["nlh"] and p["inv_exp"] == (3.0, 0.2) and p["exp"] == (10.0, 2.0, 0.1) and p["cut_inv"] == (0.5, 0.01)
    assert rd.read_Exponential_wall_TB(str(f)) is None                      # not the simple class
    f.write_text("General:exclude\nCut-off\n1.0 0.1\n")
    assert rd.read_Short_Rep_TB(str(f)) is None                             # case default: no short-range forces
    f.write_text("general\n")                                               # defaults: cut off at zero
    p = rd.read_Short_Rep_TB(str(f))
    assert (p["d0"], p["dd"]) == (0.0, 0.01) and p["exp"] is None and p["pow"] == [] and p["tab"] is None
    f.write_text("General\nCut-off\n3.4 0.1\nTab\n4\t spline\t! n, method\n1.0 8.0\n1.5 4.0\n2.0 2.0\n3.0 1.0\n")
    p = rd.read_Short_Rep_TB(str(f))
    assert p["tab"]["spline"] and list(p["tab"]["R"]) == [1.0, 1.5, 2.0, 3.0] and list(p["tab"]["a"]) == [8.0, 4.0, 2.0, 1.0]
    f.write_text("General\nTab\n3\n1.0 8.0\n1.5 4.0\n2.0 2.0\nNLH\n")
    with pytest.raises(ValueError):
        rd.read_Short_Rep_TB(str(f))                                        # NLH needs Z and the coefficient file
    nl = tmp_path / "nlh_coeffs.dat"
    nl.write_text("# Z1 Z2 a1 b1 a2 b2 a3 b3\n 1  1 -8.9 9.5 9.9 8.8 0.0 0.0 1% 3%\n 1  6 0.04 7.5 0.95 2.1 0.01 0.3 1% 1%\n")
    p = rd.read_Short_Rep_TB(str(f), Z=(1, 6), nlh_file=str(nl))
    assert not p["tab"]["spline"] and p["nlh_a"] == (0.04, 0.95, 0.01) and p["nlh_b"] == (7.5, 2.1, 0.3)
    p = rd.read_Short_Rep_TB(str(f), Z=(6, 1), nlh_file=str(nl))            # lower triangle as written: b(:) = a(:) (ZBL_potential.f90:133)
    assert p["nlh_a"] == (0.04, 0.95, 0.01) and p["nlh_b"] == (0.04, 0.95, 0.01)


def test_reference_fixture_files_when_present(wk):
    if not os.path.isdir(REF):
        pytest.skip("reference tree not present")
    rd = importlib.import_module("xtant-3_b200.readers")
    for name, path in (("I_I", "PbI2/I_I_TB_short.txt"), ("Pb_I", "PbI2/Pb_I_TB_short.txt"), ("Pb_Pb", "PbI2/Pb_Pb_TB_short.txt"),
                       ("Fe_Fe", "Fe/Fe_Fe_TB_short.txt"), ("O_O_tab", "PbO2/O_O_TB_short.txt")):
        p = rd.read_Short_Rep_TB(os.path.join(REF, path)); q = wk.load_short(name)
        assert (p["d0"], p["dd"], p["exp"], p["pow"], p["zbl"]) == (q["d0"], q["dd"], q["exp"], q["pow"], q["zbl"]), name
        if p["tab"] is not None:
            assert len(p["tab"]["R"]) == 301 and not p["tab"]["spline"]
            for k in ("R", "E", "a", "b", "c", "d"):
                assert np.array_equal(p["tab"][k], q["tab"][k]), (name, k)
    assert rd.read_Short_Rep_TB(os.path.join(REF, "BN/B_N_TB_short.txt")) is None          # 'General:exclude'
    a, b = rd.read_NLH_coefficients(os.path.join(REF, "!Short_range_potentials/nlh_coeffs.dat"), 1, 6)
    assert a == (0.04216, 0.95784, 0.0) or len(a) == 3


def test_spline_coefficients_host_reader_equals_oracle_and_reproduces_the_nodes(orc, wk):
    rd = importlib.import_module("xtant-3_b200.readers")
    t = wk.load_short("O_O_tab")["tab"]
    for R, E in ((t["R"], t["E"]), (np.array([1.0, 1.5, 2.0, 3.0, 3.1]), np.array([8.0, 4.0, 2.0, 1.0, 0.9]))):
        host = rd.make_cubic_splines(R, E); ref = orc.make_cubic_splines(R, E)
        for h, o in zip(host, ref):
            assert np.array_equal(h, o)
        a, b, c, d = host
        assert np.array_equal(a, E)                                # value at the left node of each interval
        n = len(R)
        h_ = R[1:n - 1] - R[:n - 2]                                # intervals 1 .. n-2 carry coefficients; they hit the next node's slope
        slope_end = b[:n - 2] + 2 * c[:n - 2] * h_ + 3 * d[:n - 2] * h_ ** 2
        fp = (E[2:] - E[1:n - 1]) / (R[2:] - R[1:n - 1])           # get_first_derivative at i+1
        assert np.abs(slope_end - fp).max() <= 1e-9 * np.abs(fp).max()


def test_pair_functions_against_closed_forms(orc, wk):
    r = np.array([0.7, 1.3, 1.9, 2.4, 2.69, 3.2])
    fcut = lambda r, d0, dd: 1.0 / (1.0 + np.exp((r - d0) / dd))
    fe = wk.load_short("Fe_Fe")                                    # Exp alone: 10 exp(-(r - 2)/0.1) f_cut(2.2, 0.05)
    for x in r:
        want = 10.0 * np.exp(-(x - 2.0) / 0.1) * fcut(x, 2.2, 0.05) if x < 2.2 + 0.5 else 0.0
        assert abs(orc.short_pair_value(fe, x) - want) <= 1e-14 * abs(want)
    ii = wk.load_short("I_I")                                      # Pow polynomial + ZBL, Z = 53
    k = 1.0 / (4.0 * np.pi * 8.854187817620e-12) * 1.602176634e-19 / 1.0e-10
    for x in r:
        a = 0.8854 * 0.5291772085936 / (2 * 53.0 ** 0.23)
        phi = sum(c * np.exp(e * x / a) for c, e in ((0.1818, -3.2), (0.5099, -0.9423), (0.2802, -0.4028), (0.02817, -0.2016)))
        poly = sum(P * x ** m for P, _, m in ii["pow"])
        want = (poly + k * 53.0 * 53.0 / x * phi) * fcut(x, 2.7, 0.1) if x < 3.7 else 0.0
        assert abs(orc.short_pair_value(ii, x, 53.0, 53.0) - want) <= 1e-12 * max(abs(want), 1.0), x
    tab = wk.load_short("O_O_tab")                                 # linear interpolation of the table under the cut-off
    R, E = tab["tab"]["R"], tab["tab"]["E"]
    for x in (0.65, 1.2345, 2.0001, 3.3):
        want = np.interp(x, R, E) * fcut(x, 3.4, 0.1)
        assert abs(orc.short_pair_value(tab, x) - want) <= 1e-12 * abs(want), x
    assert orc.short_pair_value(tab, R[-1] + 1e-6) == 0.0          # beyond the grid
    assert orc.short_pair_value(tab, 0.5) > 1e5                    # below the grid: extrapolated from the first interval
    inv = {"d0": 2.0, "dd": 0.05, "inv_exp": (3.0, 0.2)}           # the simple wall's function inside the General class
    w = {"C": 3.0, "r0": 0.2, "d0": 2.0, "dd": 0.05}
    assert abs(orc.short_pair_value(inv, 1.1) - 3.0 * np.exp(1.0 / 0.9) * fcut(1.1, 2.0, 0.05)) < 1e-13
    # derivative of every form = finite difference of its value
    allp = synthetic_all_forms(wk)["short_params"][0]
    for p, Z in ((fe, 26.0), (ii, 53.0), (tab, 8.0), (inv, 1.0), (allp, 6.0)):
        for x in (0.93, 1.57, 2.13):
            d = 1e-6
            fd = (orc.short_pair_value(p, x + d, Z, Z) - orc.short_pair_value(p, x - d, Z, Z)) / (2 * d)
            an = orc.short_pair_value(p, x, Z, Z, deriv=True)
            assert abs(fd - an) <= 3e-6 * max(1.0, abs(an)), (x, fd, an)


@pytest.mark.parametrize("case", ["pbi2", "pbi2_asym", "fe_exp", "tab_linear", "tab_spline", "all_forms"])
def test_forces_and_pressure_are_derivatives_of_the_energy(orc, wk, case):
    tabp = wk.load_short("O_O_tab")
    wl = {"pbi2": lambda: pbi2_like(wk), "pbi2_asym": lambda: pbi2_like(wk, True),
          "fe_exp": lambda: one_kind(wk, wk.load_short("Fe_Fe"), 26.0, scale=1.25),
          "tab_linear": lambda: one_kind(wk, tabp), "tab_spline": lambda: one_kind(wk, dict(tabp, tab=dict(tabp["tab"], spline=True))),
          "all_forms": lambda: synthetic_all_forms(wk)}[case]()
    rm = 4.6 if wl["tb"] == "dftb" else max(orc.cutoff_radius(wl), 4.0)
    nb = orc.neighbours(wl, rm=rm)
    E, Epot = orc.short_energy(wl, nb)
    assert abs(E) > 1e-3 and abs(Epot.sum() - E) <= 1e-12 * abs(E)
    F = orc.short_forces(wl, nb); P = orc.short_pressure(wl, nb)
    assert np.abs(F.sum(axis=0)).max() <= 1e-10 * np.abs(F).max() * wl["na"] or case == "pbi2_asym"
    d = 1e-6
    scale = max(1.0, np.abs(F).max())
    for k, al in ((0, 0), (5, 1), (wl["na"] - 1, 2)):
        Es = []
        for sgn in (+1, -1):
            w = dict(wl); S = wl["S"].copy(); S[k, al] += sgn * d; w["S"] = S
            Es.append(orc.short_energy(w, orc.neighbours(w, rm=rm))[0])
        fd = (Es[0] - Es[1]) / (2 * d)
        assert abs(fd - F[k, al]) <= 3e-6 * scale, (k, al, fd, F[k, al])
    for l, k in ((0, 0), (1, 1), (2, 2), (0, 1), (2, 0)):
        Es = []
        for sgn in (+1, -1):
            w = dict(wl); h = np.array(wl["supce"], dtype=np.float64); h[l, k] += sgn * d; w["supce"] = h
            Es.append(orc.short_energy(w, orc.neighbours(w, rm=rm))[0])
        fd = (Es[0] - Es[1]) / (2 * d)
        assert abs(fd - P[l, k]) <= 3e-6 * max(1.0, np.abs(P).max()), (l, k, fd, P[l, k])


def test_term_vanishes_beyond_its_cutoff(orc, wk):
    wl = one_kind(wk, dict(wk.load_short("Fe_Fe"), d0=0.9, dd=0.01), 26.0)      # nearest neighbours at ~1.5 A > 0.9 + 0.1
    nb = orc.neighbours(wl, rm=4.0)
    assert orc.short_energy(wl, nb)[0] == 0.0 and not orc.short_forces(wl, nb).any()
