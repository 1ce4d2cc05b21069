"""GPU parity tests of the van der Waals term (vdw.cu) through the C ABI vs the CPU oracle (oracle/vdw_oracle.c).

Cases: the reference's own 216-atom amorphous-carbon cell with its shipped Girifalco file (27 image cells), the 64-atom diamond
cell with the same file (125 image cells), a two-kind Lennard-Jones table (symmetric and not), an improved-LJ cell, and config 4
(1000 atoms).  The kernel sums the (j, cell) pairs of an atom in a different order than the reference's loop nest.
Tolerances: E_vdW 1e-12 relative; forces 1e-9 eV/A (Cartesian); supercell force 1e-8 GPa and 1e-11 relative to its largest element."""
import numpy as np
import pytest

import os

from test_vdw_cpu import two_kind_lj, ilj_cell, LJ_PD

pytestmark = pytest.mark.gpu

CASES = ["ac216", "diamond64_125cells", "lj2", "lj2_asym", "ilj", "al500_triclinic_lj", "ac1000"]


def _al500_lj(wk):
    """config 3's triclinic 500-atom cell with a Lennard-Jones table: the image-cell pruning on a sheared cell"""
    wl = wk.al500()
    wl["vdw_kind"] = "LJ"
    wl["vdw_params"] = [dict(LJ_PD)]
    return wl


def _workload(name, wk):
    return {"ac216": wk.ac216, "diamond64_125cells": lambda: wk.with_vdw(wk.diamond64(), "C_amorphous"), "lj2": lambda: two_kind_lj(wk),
            "lj2_asym": lambda: two_kind_lj(wk, True), "ilj": lambda: ilj_cell(wk), "al500_triclinic_lj": lambda: _al500_lj(wk),
            "ac1000": wk.ac1000}[name]()


@pytest.fixture(scope="module", params=CASES)
def case(request, wk, orc, api):
    wl = _workload(request.param, wk)
    c = api.TBContext(0)
    c.set_workload(wl)
    N, R = c.get_mirror_cell_num()
    yield wl, N, c
    c.close()


def test_image_cells(case, orc):
    wl, N, c = case
    No, Ro = orc.vdw_cells(wl, orc.cutoff_radius(wl))
    assert (N, c.get_mirror_cell_num()[1]) == (No, Ro)


def test_vdw_energy(case, orc):
    wl, N, c = case
    E, Epot = orc.vdw_energy(wl, N)
    Eg, Epg = c.get_vdW_s()
    assert abs(E) > 1e-3
    assert abs(Eg - E) <= 1e-12 * abs(E), (Eg, E)
    assert np.abs(Epg - Epot).max() <= 1e-12 * np.abs(Epot).max()


def test_vdw_forces(case, orc):
    wl, N, c = case
    F = orc.vdw_forces(wl, N)
    Fg = c.vdW_forces()
    assert np.abs(F).max() > 1e-4
    d = np.abs(orc.forces_cartesian(Fg, wl["supce"]) - orc.forces_cartesian(F, wl["supce"])).max()
    assert d <= 1e-9, d
    assert np.abs(Fg - F).max() <= 1e-11 * np.abs(F).max()
    assert np.array_equal(Fg, c.vdW_forces())                 # deterministic


def test_vdw_pressure(case, orc):
    wl, N, c = case
    P = orc.vdw_pressure(wl, N)
    Pg = c.vdW_Pressure()
    d = np.abs(np.array(orc.pressure_GPa(Pg, wl["supce"])[0]) - np.array(orc.pressure_GPa(P, wl["supce"])[0])).max()
    assert d <= 1e-8, d
    assert np.abs(Pg - P).max() <= 1e-11 * np.abs(P).max()


def test_skipped_image_cells_add_exact_zeros(case):
    """cells whose slab lies beyond the range of every pair potential are not visited; visiting them changes no bit"""
    wl, N, c = case
    E, Ep = c.get_vdW_s(); F = c.vdW_forces(); P = c.vdW_Pressure()
    os.environ["XTB_VDW_NO_PRUNE"] = "1"
    try:
        E2, Ep2 = c.get_vdW_s(); F2 = c.vdW_forces(); P2 = c.vdW_Pressure()
    finally:
        del os.environ["XTB_VDW_NO_PRUNE"]
    assert E2 == E and np.array_equal(Ep2, Ep) and np.array_equal(F2, F) and np.array_equal(P2, P)


def test_vdw_partition_and_removal(case, api):
    wl, N, c = case
    full = c.vdW_forces()
    Pfull = c.vdW_Pressure()
    E = c.get_vdW_s()[0]
    na = wl["na"]
    c.set_partition(na // 3 + 1, 2 * na // 3)
    part = c.vdW_forces()
    assert np.array_equal(part[na // 3:2 * na // 3], full[na // 3:2 * na // 3])
    assert c.get_vdW_s()[0] == E                              # the energy is replicated: every atom, every rank
    P1 = c.vdW_Pressure()
    c.set_partition(1, na // 3); P0 = c.vdW_Pressure()
    c.set_partition(2 * na // 3 + 1, na); P2 = c.vdW_Pressure()
    assert np.abs(P0 + P1 + P2 - Pfull).max() <= 1e-12 * np.abs(Pfull).max()       # partial sums of the owned atoms
    c.set_partition(1, na)
    c.set_params_vdw(None, None)                              # "proceeding without van der Waals forces"
    with pytest.raises(api.XtbError) as e:
        c.vdW_forces()
    assert e.value.code == 6
    c.set_params_vdw(wl["vdw_kind"], wl["vdw_params"])
    assert np.array_equal(c.vdW_forces(), full)


def test_step_carries_the_vdw_term(wk, orc, api):
    """xtb_step with XTB_STEP_PAIR_TERMS: repulsive part + vdW of config-4-like carbon in one call, eager, steady and graph replay"""
    abi = api._abi
    wl = wk.ac216()
    o = orc.pettifor_step(wl, with_pressure=True)
    c = api.TBContext(0)
    c.set_workload(wl)
    N = c.get_mirror_cell_num()[0]
    E = orc.vdw_energy(wl, N)[0]; F = orc.vdw_forces(wl, N); P = orc.vdw_pressure(wl, N)
    c.get_near_neighbours_device()
    c.assemble_HS(to_host=False)
    c.set_eigen(o["Ha"], o["Ei"], o["fe"])
    flags = abi.STEP_ALL | abi.STEP_PRESSURE | abi.STEP_PAIR_TERMS
    want = ("F_att", "E_rep", "F_rep", "E_vdw", "F_vdw", "SCforce_vdw", "SCforce_rep")
    outs = []
    for graph in (False, False, True, True):
        outs.append(c.step(flags, S=wl["S"], want=want, graph=graph))
    r = outs[0]
    assert abs(r["E_vdw"] - E) <= 1e-12 * abs(E)
    assert np.abs(r["F_vdw"] - F).max() <= 1e-11 * np.abs(F).max()
    assert np.abs(r["SCforce_vdw"] - P).max() <= 1e-11 * np.abs(P).max()
    assert np.array_equal(r["F_vdw"], c.vdW_forces())
    for other in outs[1:]:
        for k in want:
            assert np.array_equal(np.asarray(other[k]), np.asarray(r[k])), k
    assert c.counter(16) >= 1                                 # a graph was captured with the vdW launches in it
    # the integrator takes the term: F_total = att + rep + vdW
    m = np.array([12.011 * 1.66053906660e-27])
    c.md_set_velocities(None)
    c.make_time_step_atoms(2, 0.5, m, force_terms=abi.FORCE_ATT | abi.FORCE_REP | abi.FORCE_VDW)
    S1, SV1, _ = c.md_get_state(None)
    koa0 = np.zeros(wl["na"], dtype=np.int32)
    So, _, SVo, _ = orc.md_half_step(2, koa0, m, wl["supce"], 0.5, r["F_att"] + r["F_rep"] + r["F_vdw"], wl["S"], np.zeros_like(wl["S"]))
    assert np.abs(S1 - So).max() <= 1e-15 and np.abs(SV1 - SVo).max() <= 1e-17
    c.close()
