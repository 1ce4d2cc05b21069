"""CPU tests of the oracle's short-range wall (oracle/wall_oracle.c: get_Exp_wall_s, d_Exp_wall_pot_s, d_Exp_wall_Pressure_s,
Exponential_wall.f90:990-1385), its file reader, the Fermi occupations (Electron_Fixed_Te + set_Fermi, Electron_tools.f90:
2660-2706,1913-1969) and the velocity-Verlet halves (make_time_step_atoms_SC, Atomic_tools.f90:1897-1974).

Pins available without reference vectors: forces and supercell force are the exact derivatives of the wall energy (finite
differences), translation invariance, electron-number conservation of the occupations, and the time-reversibility /
constant-force closed form of the integrator."""
import importlib
import os

import numpy as np
import pytest

WALL_AL = {"C": 200.0, "r0": 0.0, "d0": 2.05, "dd": 0.05}      # INPUT_DATA/Aluminum/Al_Al_TB_wall.txt:2-5
AMU = 1.66053906660e-27


def squeezed_al(wk, a=3.3, reps=(3, 3, 3), sigma=0.08, seed=11):
    """fcc Al-like cell compressed until first neighbours (a / sqrt 2 = 2.33 A) sit inside the wall's range d0 + 10 dd = 2.55 A"""
    from importlib import import_module
    S, h = wk.lattice(wk._FCC, a, reps, sigma, seed)
    wl = wk._wl("al_squeezed", "nrl", "Al", S, h, 3, 0.1)
    wl["wall_params"] = [dict(WALL_AL)]
    return wl


@pytest.fixture(scope="module")
def sq(wk, orc):
    wl = squeezed_al(wk)
    return wl, orc.neighbours(wl)


def test_reader_class_line_semantics(tmp_path):
    rd = importlib.import_module("xtant-3_b200.readers")
    p = tmp_path / "Al_Al_TB_wall.txt"
    rd.write_Exponential_wall_TB(str(p), WALL_AL)
    assert rd.read_Exponential_wall_TB(str(p)) == WALL_AL
    rd.write_Exponential_wall_TB(str(p), WALL_AL, enabled=False)       # 'Simple_wall:exclude' -> case default -> no wall
    assert rd.read_Exponential_wall_TB(str(p)) is None
    p.write_text("General\n")                                            # the other class of the slot: read_Short_Rep_TB's
    assert rd.read_Exponential_wall_TB(str(p)) is None and rd.read_Short_Rep_TB(str(p))["d0"] == 0.0


def test_reference_fixture_files_when_present():
    ref = "/root/reference/INPUT_DATA"
    if not os.path.isdir(ref):
        pytest.skip("reference tree not present")
    rd = importlib.import_module("xtant-3_b200.readers")
    # shipped with the ':exclde' / ':exclude' suffix: the reference's select case falls to default and runs WITHOUT a wall
    for f in ("Aluminum/Al_Al_TB_wall.txt", "Silicon/Si_Si_TB_wall.txt", "Diamond/C_C_TB_wall.txt"):
        assert rd.read_Exponential_wall_TB(os.path.join(ref, f)) is None, f
    assert rd.read_Exponential_wall_TB(os.path.join(ref, "Nickel/Ni_Ni_TB_wall.txt")) == {"C": 150.0, "r0": 0.0, "d0": 1.63, "dd": 0.05}


def test_wall_is_active_and_translation_invariant(sq, orc):
    wl, nb = sq
    E, Epot = orc.wall_energy(wl, nb)
    assert E > 1e-3 and abs(Epot.sum() - E) <= 1e-12 * E
    F = orc.wall_forces(wl, nb)
    assert np.abs(F).max() > 1e-3
    assert np.abs(F.sum(axis=0)).max() < 1e-9 * np.abs(F).max() * wl["na"]


def test_wall_forces_are_the_gradient_of_the_energy(sq, orc):
    wl, nb = sq
    F = orc.wall_forces(wl, nb)
    d = 1e-6
    for k, al in ((0, 0), (7, 1), (50, 2), (101, 0)):
        E = []
        for sgn in (+1, -1):
            w = dict(wl); S = wl["S"].copy(); S[k, al] += sgn * d; w["S"] = S
            E.append(orc.wall_energy(w, orc.neighbours(w))[0])
        fd = (E[0] - E[1]) / (2 * d)
        assert abs(fd - F[k, al]) <= 2e-6 * max(1.0, abs(fd)), (k, al, fd, F[k, al])


def test_wall_pressure_is_the_cell_derivative_of_the_energy(sq, orc):
    wl, nb = sq
    P = orc.wall_pressure(wl, nb)
    d = 1e-6
    for l, k in ((0, 0), (1, 1), (2, 2), (0, 1), (2, 0)):
        E = []
        for sgn in (+1, -1):
            w = dict(wl); h = wl["supce"].copy(); h[l, k] += sgn * d; w["supce"] = h
            E.append(orc.wall_energy(w, orc.neighbours(w, rm=orc.cutoff_radius(wl)))[0])
        fd = (E[0] - E[1]) / (2 * d)
        assert abs(fd - P[l, k]) <= 2e-6 * max(1.0, abs(fd)), (l, k, fd, P[l, k])


def test_wall_vanishes_beyond_its_range(wk, orc):
    wl = squeezed_al(wk, a=4.05, sigma=0.02)          # first neighbours at 2.86 A > 2.55 A
    nb = orc.neighbours(wl)
    assert orc.wall_energy(wl, nb)[0] == 0.0
    assert not orc.wall_forces(wl, nb).any()


def test_fermi_conserves_the_electron_number(wk, orc):
    rng = np.random.default_rng(3)
    E = np.sort(rng.uniform(-12.0, 30.0, size=800))
    for Te in (0.05, 0.5, 2.0):
        fe, mu = orc.fermi(E, 2.0 * 0.45 * 800, Te)
        assert abs(fe.sum() - 720.0) <= 720.0 * 2e-12
        f2, mu2 = wk.occupations(E, 720.0, Te)
        assert abs(mu - mu2) < 1e-9 and np.abs(fe - f2).max() < 1e-9
        assert np.all(np.diff(fe) <= 1e-15) and fe.max() <= 2.0
    fe0, mu0 = orc.fermi(E, 720.0, 0.0)               # step function
    assert set(np.unique(fe0)) <= {0.0, 2.0} and fe0.sum() == 720.0


def test_verlet_halves_closed_form_and_wrap(orc):
    h = np.diag([10.0, 11.0, 12.0])
    na = 5
    rng = np.random.default_rng(5)
    S = rng.uniform(0.05, 0.95, size=(na, 3)); SV = rng.normal(0, 1e-3, size=(na, 3))
    S[0, 0] = 0.9995; SV[0, 0] = 2e-3                 # crosses the boundary
    F = rng.normal(0, 1.0, size=(na, 3))              # dE/ds [eV]
    m = np.array([26.98 * AMU]); koa0 = np.zeros(na, dtype=np.int32)
    dt = 0.5
    S1, S0, SV1, SV0 = orc.md_half_step(2, koa0, m, h, dt, F, S, SV)
    g_inv = np.linalg.inv(h.T @ h)
    acc = -(1.0 / m[0]) * orc.G_E * 1e-10 * (F @ g_inv.T)
    Sx = S + dt * SV + acc * dt * dt / 2
    assert np.array_equal(S0, S) and np.array_equal(SV0, SV)
    assert np.abs(SV1 - (SV + acc * dt / 2)).max() < 1e-18
    assert np.abs(S1 - (Sx - np.floor(Sx))).max() < 1e-15 and 0.0 <= S1.min() and S1.max() < 1.0
    assert S1[0, 0] < 0.01                            # wrapped
    S2, _, SV2, _ = orc.md_half_step(1, koa0, m, h, dt, F, S1, SV1)
    assert np.array_equal(S2, S1) and np.abs(SV2 - (SV + acc * dt)).max() < 1e-18
    # reversibility: flip the velocities, step again with the same force, flip back
    Sb, _, SVb, _ = orc.md_half_step(2, koa0, m, h, dt, F, S1, -SV2)
    Sb2, _, SVb2, _ = orc.md_half_step(1, koa0, m, h, dt, F, Sb, SVb)
    d = Sb2 - S; d -= np.round(d)
    assert np.abs(d).max() < 1e-14 and np.abs(-SVb2 - SV).max() < 1e-16
    ek = orc.kinetic_energy_eV(koa0, m, h, SV)
    assert ek > 0
