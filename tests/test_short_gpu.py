"""GPU parity tests of the 'General' short-range class (wall.cu, functor ShortRep) through the C ABI vs the CPU oracle
(oracle/short_oracle.c).  Cases: the shipped PbI2 sets on a two-kind cell (Pow polynomial + ZBL; symmetric and asymmetric pair
tables), the shipped Fe set (Exp), the shipped PbO2 O-O table (linear interpolation; the same table splined), and a synthetic set
with all six functions at once.

Tolerances: E 1e-12 relative; forces max(1e-9 eV/A, 1e-11 of the largest force) -- the ZBL core and the degree-4 polynomials of the
shipped sets reach 1e3 .. 1e4 eV/A and cancel to a few per cent of their terms, and the device `pow` / `exp` differ from glibc's
in the last bit; supercell force 1e-11 of its largest element."""
import numpy as np
import pytest

from test_short_cpu import pbi2_like, one_kind, synthetic_all_forms

pytestmark = pytest.mark.gpu

CASES = ["pbi2", "pbi2_asym", "fe_exp", "tab_linear", "tab_spline", "all_forms"]


def _workload(name, wk):
    tabp = wk.load_short("O_O_tab")
    return {"pbi2": lambda: pbi2_like(wk), "pbi2_asym": lambda: pbi2_like(wk, True),
            "fe_exp": lambda: one_kind(wk, wk.load_short("Fe_Fe"), 26.0, scale=1.25),
            "tab_linear": lambda: one_kind(wk, tabp), "tab_spline": lambda: one_kind(wk, dict(tabp, tab=dict(tabp["tab"], spline=True))),
            "all_forms": lambda: synthetic_all_forms(wk)}[name]()


@pytest.fixture(scope="module", params=CASES)
def case(request, wk, orc, api):
    wl = _workload(request.param, wk)
    rm = 4.6 if wl["tb"] == "dftb" else max(orc.cutoff_radius(wl), 4.0)
    nb = orc.neighbours(wl, rm=rm)
    c = api.TBContext(0)
    c.set_workload(wl)
    c.get_near_neighbours_device(rm)
    yield wl, nb, c
    c.close()


def test_short_energy(case, orc):
    wl, nb, c = case
    E, Epot = orc.short_energy(wl, nb)
    Eg, Epg = c.get_short_range_rep_s()
    assert abs(E) > 1e-3
    assert abs(Eg - E) <= 1e-12 * abs(E), (Eg, E)
    assert np.abs(Epg - Epot).max() <= 1e-12 * np.abs(Epot).max()


def test_short_forces(case, orc):
    wl, nb, c = case
    F = orc.short_forces(wl, nb)
    Fg = c.d_Short_range_pot_s()
    assert np.abs(F).max() > 1e-3
    Fc, Fgc = orc.forces_cartesian(F, wl["supce"]), orc.forces_cartesian(Fg, wl["supce"])
    assert np.abs(Fgc - Fc).max() <= max(1e-9, 1e-11 * np.abs(Fc).max()), (np.abs(Fgc - Fc).max(), np.abs(Fc).max())
    assert np.array_equal(Fg, c.d_Short_range_pot_s())        # deterministic


def test_short_pressure(case, orc):
    wl, nb, c = case
    P = orc.short_pressure(wl, nb)
    Pg = c.d_Short_range_Pressure_s()
    assert np.abs(Pg - P).max() <= 1e-11 * np.abs(P).max()


def test_one_slot_two_classes_partition_and_step(case, api, orc):
    """TB_Expwall holds one class: setting the simple wall replaces the General class and back; partition; xtb_step reports the
    class that is set through its wall outputs (eager, steady, graph)"""
    abi = api._abi
    wl, nb, c = case
    full = c.d_Short_range_pot_s(); E = c.get_short_range_rep_s()[0]
    na = wl["na"]
    c.set_partition(na // 3 + 1, 2 * na // 3)
    assert np.array_equal(c.d_Short_range_pot_s()[na // 3:2 * na // 3], full[na // 3:2 * na // 3])
    c.set_partition(1, na)
    simple = [{"C": 50.0, "r0": 0.0, "d0": 2.6, "dd": 0.05}] * (wl["n_kinds"] ** 2)
    c.set_params_wall(simple)
    with pytest.raises(api.XtbError) as e:
        c.d_Short_range_pot_s()
    assert e.value.code == 6
    w2 = dict(wl, wall_params=simple)
    Ew = orc.wall_energy(w2, nb)[0]
    assert abs(c.get_Exp_wall_s()[0] - Ew) <= 1e-12 * abs(Ew)
    c.set_params_short(wl["short_params"], wl["Z"])
    with pytest.raises(api.XtbError):
        c.get_Exp_wall_s()
    assert np.array_equal(c.d_Short_range_pot_s(), full)
    if wl["tb"] == "pettifor":                                # the fused step needs eigenvectors only for the TB part: pair terms alone
        want = ("E_wall", "F_wall", "SCforce_wall")
        flags = abi.STEP_NEIGHBOURS | abi.STEP_PAIR_TERMS | abi.STEP_PRESSURE
        outs = []
        try:
            outs = [c.step(flags, S=wl["S"], want=want, graph=g) for g in (False, False, True, True)]
        except api.XtbError as ex:                            # pressure of the TB part needs the density: pair terms without it
            flags = abi.STEP_NEIGHBOURS | abi.STEP_PAIR_TERMS
            want = ("E_wall", "F_wall")
            outs = [c.step(flags, S=wl["S"], want=want, graph=g) for g in (False, False, True, True)]
        # the step's list uses the TB cut-off; compare with the staged calls on that same list
        Es = c.get_short_range_rep_s()[0]; Fs = c.d_Short_range_pot_s()
        for o in outs:
            assert o["E_wall"] == Es and np.array_equal(o["F_wall"], Fs)
    c.set_params_short(None, None)
    with pytest.raises(api.XtbError):
        c.get_short_range_rep_s()
    c.set_params_short(wl["short_params"], wl["Z"])
