"""GPU parity tests of the Coulomb term of a charged supercell (vdw.cu, class 4) through the C ABI vs the CPU oracle.

Cases: the reference's 216-atom amorphous-carbon cell (27 image cells), a sheared 64-atom cell (125 cells; the transposed
contraction of Construct_B_C matters there), a two-kind cell, config 3's triclinic 500-atom cell, and a long-range table
(dm = 20 A on the 12.5 A cell: 729 image cells).  Tolerances: E 1e-12 relative; forces 1e-9 eV/A (Cartesian) and 1e-11 of the
largest; supercell force 1e-11 of its largest element."""
import os

import numpy as np
import pytest

from test_coulomb_cpu import charged, sheared

pytestmark = pytest.mark.gpu

CASES = ["ac216", "sheared", "two_kinds", "al500_triclinic", "ac216_dm20"]


def _workload(name, wk):
    return {"ac216": lambda: charged(wk.ac216()), "sheared": lambda: charged(sheared(wk), Q=0.2),
            "two_kinds": lambda: charged(wk.ch_zb(), Q=0.1, dm=5.0), "al500_triclinic": lambda: charged(wk.al500(), Q=0.02),
            "ac216_dm20": lambda: charged(wk.ac216(), Q=0.03, dm=20.0)}[name]()


@pytest.fixture(scope="module", params=CASES)
def case(request, wk, orc, api):
    wl = _workload(request.param, wk)
    c = api.TBContext(0)
    c.set_workload(wl)
    yield wl, c.get_mirror_cell_num_C(), c
    c.close()


def test_image_cells(case, orc):
    wl, N, c = case
    assert N == orc.coulomb_cells(wl)


def test_coulomb_energy_forces_pressure(case, orc):
    wl, N, c = case
    E, Epot = orc.coulomb_energy(wl, N)
    Eg, Epg = c.get_Coulomb_s()
    assert E > 1e-3 and abs(Eg - E) <= 1e-12 * E, (Eg, E)
    assert np.abs(Epg - Epot).max() <= 1e-12 * np.abs(Epot).max()
    F = orc.coulomb_forces(wl, N)
    Fg = c.Coulomb_forces()
    assert np.abs(F).max() > 1e-4
    assert np.abs(orc.forces_cartesian(Fg, wl["supce"]) - orc.forces_cartesian(F, wl["supce"])).max() <= 1e-9
    assert np.abs(Fg - F).max() <= 1e-11 * np.abs(F).max()
    assert np.array_equal(Fg, c.Coulomb_forces())
    P = orc.coulomb_pressure(wl, N)
    Pg = c.Coulomb_Pressure()
    assert np.abs(Pg - P).max() <= 1e-11 * np.abs(P).max()


def test_skipped_image_cells_add_exact_zeros(case):
    wl, N, c = case
    E, Ep = c.get_Coulomb_s(); F = c.Coulomb_forces(); P = c.Coulomb_Pressure()
    os.environ["XTB_VDW_NO_PRUNE"] = "1"
    try:
        E2, Ep2 = c.get_Coulomb_s(); F2 = c.Coulomb_forces(); P2 = c.Coulomb_Pressure()
    finally:
        del os.environ["XTB_VDW_NO_PRUNE"]
    assert E2 == E and np.array_equal(Ep2, Ep) and np.array_equal(F2, F) and np.array_equal(P2, P)


def test_charge_switch_partition_and_removal(case, api, orc):
    wl, N, c = case
    full = c.Coulomb_forces(); E = c.get_Coulomb_s()[0]
    c.set_charge(1.0e-12)                                     # Construct_B_C: no charge, no force; the energy has no switch
    assert not c.Coulomb_forces().any() and not c.Coulomb_Pressure().any()
    assert 0.0 < c.get_Coulomb_s()[0] < 1e-15
    c.set_charge(2.0 * wl["Q"])                               # E and F scale with Q^2
    assert abs(c.get_Coulomb_s()[0] - 4.0 * E) <= 1e-14 * 4.0 * E
    assert np.abs(c.Coulomb_forces() - 4.0 * full).max() <= 1e-14 * 4.0 * np.abs(full).max()
    c.set_charge(wl["Q"])
    na = wl["na"]
    c.set_partition(na // 3 + 1, 2 * na // 3)
    assert np.array_equal(c.Coulomb_forces()[na // 3:2 * na // 3], full[na // 3:2 * na // 3])
    assert c.get_Coulomb_s()[0] == E
    c.set_partition(1, na)
    c.set_params_coulomb(None)
    with pytest.raises(api.XtbError) as e:
        c.Coulomb_forces()
    assert e.value.code == 6
    c.set_params_coulomb(wl["coulomb_params"])
    assert np.array_equal(c.Coulomb_forces(), full)


def test_step_carries_every_pair_term(wk, orc, api):
    """xtb_step | PAIR_TERMS on a charged amorphous-carbon cell: repulsive part + van der Waals + Coulomb in one call; eager, steady,
    graph; a new charge goes through the eager path once (the table is rebuilt) and then replays again"""
    abi = api._abi
    wl = charged(wk.ac216(), Q=0.04)
    o = orc.pettifor_step(wl, with_pressure=True)
    c = api.TBContext(0)
    c.set_workload(wl)
    Nv = c.get_mirror_cell_num()[0]; Nc = c.get_mirror_cell_num_C()
    c.get_near_neighbours_device(); c.assemble_HS(to_host=False)
    c.set_eigen(o["Ha"], o["Ei"], o["fe"])
    flags = abi.STEP_ALL | abi.STEP_PRESSURE | abi.STEP_PAIR_TERMS
    want = ("F_att", "E_rep", "F_rep", "E_vdw", "F_vdw", "E_coul", "F_coul", "SCforce_coul", "SCforce_vdw")
    outs = [c.step(flags, S=wl["S"], want=want, graph=g) for g in (False, False, True, True)]
    r = outs[0]
    E, _ = orc.coulomb_energy(wl, Nc); F = orc.coulomb_forces(wl, Nc); P = orc.coulomb_pressure(wl, Nc)
    assert abs(r["E_coul"] - E) <= 1e-12 * E and np.abs(r["F_coul"] - F).max() <= 1e-11 * np.abs(F).max()
    assert np.abs(r["SCforce_coul"] - P).max() <= 1e-11 * np.abs(P).max()
    assert abs(r["E_vdw"] - orc.vdw_energy(wl, Nv)[0]) <= 1e-12 * abs(r["E_vdw"])
    for other in outs[1:]:
        for k in want:
            assert np.array_equal(np.asarray(other[k]), np.asarray(r[k])), k
    g0 = c.counter(16)
    assert g0 >= 1
    c.set_charge(0.08)
    r2 = c.step(flags, S=wl["S"], want=want, graph=True)      # eager (table rebuilt)
    r3 = c.step(flags, S=wl["S"], want=want, graph=True)      # re-captured
    assert abs(r2["E_coul"] - 4.0 * E) <= 1e-13 * 4.0 * E and np.array_equal(r2["F_coul"], r3["F_coul"]) and r2["E_coul"] == r3["E_coul"]
    assert np.array_equal(r2["F_vdw"], r["F_vdw"])
    m = np.array([12.011 * 1.66053906660e-27])
    c.md_set_velocities(None)
    c.make_time_step_atoms(2, 0.5, m, force_terms=abi.FORCE_ATT | abi.FORCE_REP | abi.FORCE_VDW | abi.FORCE_COUL)
    S1, SV1, _ = c.md_get_state(None)
    koa0 = np.zeros(wl["na"], dtype=np.int32)
    So, _, SVo, _ = orc.md_half_step(2, koa0, m, wl["supce"], 0.5, r3["F_att"] + r3["F_rep"] + r3["F_vdw"] + r3["F_coul"], wl["S"], np.zeros_like(wl["S"]))
    assert np.abs(S1 - So).max() <= 1e-15 and np.abs(SV1 - SVo).max() <= 1e-17
    c.close()
