"""GPU parity tests of the repulsive pair terms (pettifor_rep.cu) through the C ABI vs the CPU oracle.

Tolerances: E_rep 1e-12 relative; forces 1e-9 eV/A (Cartesian); supercell force 1e-8 GPa -- the north-star bars.
Run-to-run determinism (fixed reduction order, no atomics) is asserted bit-wise.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _wl(wk, name):
    return {"diamond64": wk.diamond64,
            "c216": lambda: wk.c_dia((3, 3, 3), sigma=0.2, seed=8),
            "ac1000": wk.ac1000}[name]()


@pytest.fixture(scope="module", params=["diamond64", "c216", "ac1000"])
def case(request, wk, orc, api):
    wl = _wl(wk, request.param)
    nb = orc.neighbours(wl)
    c = api.TBContext(0)
    c.set_workload(wl)
    c.get_near_neighbours_device()
    yield wl, nb, c
    c.close()


def test_energy(case, orc):
    wl, nb, c = case
    E, Epot = orc.pet_erep(wl, nb)
    Eg, Epg = c.get_Erep_s()
    assert abs(Eg - E) <= 1e-12 * abs(E)
    assert np.abs(Epg - Epot).max() <= 1e-12 * np.abs(Epot).max()


def test_forces(case, orc):
    wl, nb, c = case
    F = orc.pet_rep_forces(wl, nb)
    Fg = c.dErdr_s()
    d = np.abs(orc.forces_cartesian(Fg, wl["supce"]) - orc.forces_cartesian(F, wl["supce"])).max()
    assert d <= 1e-9, d
    assert np.array_equal(Fg, c.dErdr_s())                    # deterministic


def test_pressure(case, orc):
    wl, nb, c = case
    P = orc.pet_rep_pressure(wl, nb)
    Pg = c.dErdr_Pressure_s()
    d = np.abs(np.array(orc.pressure_GPa(Pg, wl["supce"])[0]) - np.array(orc.pressure_GPa(P, wl["supce"])[0])).max()
    assert d <= 1e-8, d
    assert np.abs(Pg - P).max() <= 1e-11 * np.abs(P).max()


def test_partition_and_errors(case, api, wk):
    wl, nb, c = case
    full = c.dErdr_s()
    na = wl["na"]
    c.set_partition(na // 3 + 1, 2 * na // 3)
    part = c.dErdr_s()
    assert np.array_equal(part[na // 3:2 * na // 3], full[na // 3:2 * na // 3])
    assert not part[:na // 3].any()
    c.set_partition(1, na)
    # NRL has no repulsive part (TB_NRL.f90:3236-3275): the entry points refuse
    c2 = api.TBContext(0)
    c2.set_workload(wk.au_fcc((2, 2, 2)))
    c2.get_near_neighbours_device()
    with pytest.raises(api.XtbError) as e:
        c2.dErdr_s()
    assert e.value.code == 6
    c2.close()


def test_full_force_vanishes_at_equilibrium(wk, orc, api):
    """att + rep on the perfect diamond lattice at the reference's lattice constant: zero by symmetry, and the total
    supercell force is small (the cell is at the E(a) minimum within 0.1 %, tests/test_repulsive_cpu.py)."""
    S, h = wk.lattice(wk._DIA, 3.5588, (2, 2, 2), 0.0, 0)
    wl = wk._wl("x", "pettifor", "C", S, h, 4, 300.0 * wk.K_B_EV)
    o = orc.pettifor_step(wl)
    c = api.TBContext(0)
    c.set_workload(wl)
    c.get_near_neighbours_device()
    c.assemble_HS(to_host=False)
    c.set_eigen(o["Ha"], o["Ei"], o["fe"])
    c.density()
    F = c.forces() + c.dErdr_s()
    assert np.abs(orc.forces_cartesian(F, wl["supce"])).max() < 1e-8
    Ptot = c.pressure() + c.dErdr_Pressure_s()
    Patt = c.pressure()
    assert np.abs(np.diag(Ptot)).max() < 0.02 * np.abs(np.diag(Patt)).max()
    c.close()
