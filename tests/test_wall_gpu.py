"""GPU parity tests of the short-range wall (wall.cu), the Fermi occupations and the velocity-Verlet halves (step.cu)
through the C ABI vs the CPU oracle (oracle/wall_oracle.c).

Tolerances: E_wall 1e-12 relative; forces 1e-9 eV/A (Cartesian); supercell force 1e-8 GPa; occupations 1e-12 (mu 1e-11 eV);
integrator 1e-15 (same operations, FMA contraction aside)."""
import numpy as np
import pytest

from test_wall_cpu import WALL_AL, AMU, squeezed_al

pytestmark = pytest.mark.gpu


def _al500_squeezed(wk):
    """config 3's triclinic 500-atom cell, compressed by 0.8 so that first neighbours (2.3 A) are inside the wall's range"""
    wl = wk.al500()
    wl["supce"] = np.ascontiguousarray(wl["supce"] * 0.8)
    wl["wall_params"] = [dict(WALL_AL)]
    return wl


@pytest.fixture(scope="module", params=["al108_squeezed", "al500_squeezed"])
def case(request, wk, orc, api):
    wl = squeezed_al(wk) if request.param == "al108_squeezed" else _al500_squeezed(wk)
    nb = orc.neighbours(wl)
    c = api.TBContext(0)
    c.set_workload(wl)
    c.get_near_neighbours_device()
    yield wl, nb, c
    c.close()


def test_wall_energy(case, orc):
    wl, nb, c = case
    E, Epot = orc.wall_energy(wl, nb)
    Eg, Epg = c.get_Exp_wall_s()
    assert E > 1e-3
    assert abs(Eg - E) <= 1e-12 * abs(E)
    assert np.abs(Epg - Epot).max() <= 1e-12 * np.abs(Epot).max()


def test_wall_forces(case, orc):
    wl, nb, c = case
    F = orc.wall_forces(wl, nb)
    Fg = c.d_Exp_wall_pot_s()
    assert np.abs(F).max() > 1e-3
    d = np.abs(orc.forces_cartesian(Fg, wl["supce"]) - orc.forces_cartesian(F, wl["supce"])).max()
    assert d <= 1e-9, d
    assert np.array_equal(Fg, c.d_Exp_wall_pot_s())           # deterministic


def test_wall_pressure(case, orc):
    wl, nb, c = case
    P = orc.wall_pressure(wl, nb)
    Pg = c.d_Exp_wall_Pressure_s()
    d = np.abs(np.array(orc.pressure_GPa(Pg, wl["supce"])[0]) - np.array(orc.pressure_GPa(P, wl["supce"])[0])).max()
    assert d <= 1e-8, d
    assert np.abs(Pg - P).max() <= 1e-11 * np.abs(P).max()


def test_wall_partition_removal_and_other_branches(case, api, wk, orc):
    wl, nb, c = case
    full = c.d_Exp_wall_pot_s()
    na = wl["na"]
    c.set_partition(na // 3 + 1, 2 * na // 3)
    part = c.d_Exp_wall_pot_s()
    assert np.array_equal(part[na // 3:2 * na // 3], full[na // 3:2 * na // 3])
    c.set_partition(1, na)
    c.set_params_wall(None)                                   # "proceeding without short-range forces"
    with pytest.raises(api.XtbError) as e:
        c.d_Exp_wall_pot_s()
    assert e.value.code == 6
    c.set_params_wall(wl["wall_params"])
    assert np.array_equal(c.d_Exp_wall_pot_s(), full)
    # the wall rides on the list of any TB branch (dispatch TB.f90:3533-3556): Pettifor-type carbon, compressed
    w2 = wk.c_dia((2, 2, 2), sigma=0.1, seed=4)
    w2["supce"] = w2["supce"] * 0.7
    w2["wall_params"] = [{"C": 100.0, "r0": 0.0, "d0": 1.165, "dd": 0.01}]      # INPUT_DATA/Diamond/C_C_TB_wall.txt:2-5
    c2 = api.TBContext(0)
    c2.set_workload(w2)
    c2.get_near_neighbours_device()
    nb2 = orc.neighbours(w2)
    E2 = orc.wall_energy(w2, nb2)[0]
    assert E2 > 0 and abs(c2.get_Exp_wall_s()[0] - E2) <= 1e-12 * E2
    d = np.abs(orc.forces_cartesian(c2.d_Exp_wall_pot_s(), w2["supce"]) - orc.forces_cartesian(orc.wall_forces(w2, nb2), w2["supce"])).max()
    assert d <= 1e-9, d
    c2.close()


def test_fermi_occupations_on_device(wk, orc, api):
    wl = wk.au_fcc((2, 2, 2), sigma=0.1, seed=21)
    o = orc.nrl_step(wl, with_pressure=False)
    c = api.TBContext(0)
    c.set_workload(wl)
    c.get_near_neighbours_device()
    c.assemble_HS(to_host=False)
    c.set_eigen(o["Ha"], o["Ei"], o["fe"])
    for Te in (0.05, 0.1, 2.0, 0.0):
        fe, mu = orc.fermi(o["Ei"], wl["n_electrons"], Te)
        fg, mug, eb = c.fermi_occupations(wl["n_electrons"], Te)
        assert abs(mug - mu) <= 1e-10, (Te, mug, mu)
        assert np.abs(fg - fe).max() <= 1e-10, Te
        assert abs(fg.sum() - wl["n_electrons"]) <= 3e-12 * wl["n_electrons"] or Te == 0.0
        assert abs(eb - float((fe * o["Ei"]).sum())) <= 1e-9 * abs(eb)
    # the occupations set on the device arm the density stage: forces equal those with the same fe uploaded from the host
    fg, mug, eb = c.fermi_occupations(wl["n_electrons"], wl["Te_eV"])
    c.density(); F1 = c.forces()
    c.set_occupations(fg)
    c.density(); F2 = c.forces()
    assert np.array_equal(F1, F2)
    c.close()


def test_verlet_halves_on_device(wk, orc, api):
    wl = wk.diamond64()
    na = wl["na"]
    rng = np.random.default_rng(9)
    SV = rng.normal(0.0, 2e-3, size=(na, 3))
    m = np.array([12.011 * AMU])
    o = orc.pettifor_step(wl, with_pressure=False)
    c = api.TBContext(0)
    c.set_workload(wl)
    c.get_near_neighbours_device()
    c.assemble_HS(to_host=False)
    c.set_eigen(o["Ha"], o["Ei"], o["fe"])
    c.density()
    Fa = c.forces(); Fr = c.dErdr_s()
    c.md_set_velocities(SV)
    abi = c.lib
    c.make_time_step_atoms(2, 0.5, m, force_terms=1 | 2)
    S1, SV1, ek = c.md_get_state(m)
    koa0 = np.zeros(na, dtype=np.int32)
    So, S0o, SVo, SV0o = orc.md_half_step(2, koa0, m, wl["supce"], 0.5, Fa + Fr, wl["S"], SV)
    assert np.abs(S1 - So).max() <= 1e-15 and np.abs(SV1 - SVo).max() <= 1e-17
    assert abs(ek - orc.kinetic_energy_eV(koa0, m, wl["supce"], SVo)) <= 1e-12 * ek
    # the list is stale after the atoms moved
    with pytest.raises(api.XtbError):
        c.density()
    c.get_near_neighbours_device()
    g = c.get_near_neighbours()
    w2 = dict(wl); w2["S"] = So
    nb2 = orc.neighbours(w2)
    assert np.array_equal(g["size"], nb2.size)
    c.close()
