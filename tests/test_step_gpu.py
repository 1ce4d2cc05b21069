"""GPU tests of xtb_step (step.cu): the fused call gives the same bits as the stage-by-stage calls, in its sizing form
(first step), its steady-state form (no list read-back) and its recovery form (a list that outgrew the armed sizes);
a device-resident NVE trajectory conserves the total energy (a pin of the whole force chain that does not involve the
oracle: wrong forces make the total drift)."""
import importlib

import numpy as np
import pytest

from test_wall_cpu import AMU

pytestmark = pytest.mark.gpu


def _staged(api, wl, o, S=None, supce=None):
    c = api.TBContext(0)
    w = dict(wl)
    if S is not None:
        w["S"] = S
    if supce is not None:
        w["supce"] = supce
    c.set_workload(w)
    c.get_near_neighbours_device()
    HS = c.assemble_HS(to_host=True)
    c.set_eigen(o["Ha"], o["Ei"], o["fe"])
    c.density()
    F = c.forces(); P = c.pressure()
    n = c.counter(3)
    c.close()
    _staged.HS = HS             # dense H_non, Sij of this configuration (fresh context: written from scratch)
    return F, P, n


@pytest.mark.parametrize("name", ["au32", "diamond64", "si64"])
def test_step_equals_staged_calls(name, wk, orc, api):
    abi = importlib.import_module("xtant-3_b200._abi")
    wl = {"au32": lambda: wk.au_fcc((2, 2, 2), sigma=0.1, seed=21), "diamond64": wk.diamond64,
          "si64": lambda: wk.si_dia((2, 2, 2), sigma=0.1, seed=5)}[name]()
    o = orc.nrl_step(wl, with_pressure=False) if wl["tb"] == "nrl" else orc.pettifor_step(wl, with_pressure=False)
    F0, P0, n0 = _staged(api, wl, o)
    c = api.TBContext(0)
    c.set_workload(wl)
    c.set_eigen(o["Ha"], o["Ei"], o["fe"])
    flags = abi.STEP_ALL | abi.STEP_PRESSURE
    out = c.step(flags, want=("F_att", "SCforce_att"))          # sizing form
    assert np.array_equal(out["F_att"], F0) and np.array_equal(out["SCforce_att"], P0)
    assert c.counter(3) == n0
    l0 = c.counter(6)
    out = c.step(flags, want=("F_att", "SCforce_att"))          # steady-state form: same bits, no list read-back
    assert np.array_equal(out["F_att"], F0) and np.array_equal(out["SCforce_att"], P0)
    launches_steady = c.counter(6) - l0
    # moved atoms through io->S: equals a fresh staged run on those positions
    rng = np.random.default_rng(2)
    S2 = np.mod(wl["S"] + rng.normal(0, 2e-3, size=wl["S"].shape), 1.0)
    F2, P2, n2 = _staged(api, wl, o, S=S2)
    HS2 = _staged.HS
    out = c.step(flags, S=S2, want=("F_att", "SCforce_att"))
    assert np.array_equal(out["F_att"], F2) and np.array_equal(out["SCforce_att"], P2)
    assert c.counter(3) == n2

    def dense_equal(ref):
        """the RESIDENT dense matrices (only the blocks of the old and of the new pattern were touched by the step) hold
        exactly what a fresh context writes from scratch"""
        import ctypes as C
        N = wl["na"] * c.n_orb
        H = np.full((N, N), np.nan, order="F"); S = np.full((N, N), np.nan, order="F") if ref[1] is not None else None
        for which, arr in ((abi.BUF_H, H), (abi.BUF_S, S)):
            if arr is None:
                continue
            ptr, nbytes = c.device_buffer(which)
            assert nbytes == 8 * N * N
            torch = importlib.import_module("torch")
            interop = importlib.import_module("xtant-3_b200.interop")
            arr[...] = interop.buffer(c, which).cpu().numpy().reshape(N, N, order="F")
        return np.array_equal(H, ref[0]) and (S is None or np.array_equal(S, ref[1]))

    assert dense_equal(HS2)
    # a compressed cell: more neighbours than the armed bound -> the step is re-run through the sizing path, transparently
    h3 = wl["supce"] * 0.9
    F3, P3, n3 = _staged(api, wl, o, S=S2, supce=h3)
    if name == "si64":
        assert n3 > n2 + 2                                      # (au32: every atom already lists all the others)
    HS3 = _staged.HS
    out = c.step(flags, supce=h3, want=("F_att", "SCforce_att"))
    assert np.array_equal(out["F_att"], F3) and np.array_equal(out["SCforce_att"], P3)
    assert c.counter(3) == n3
    assert dense_equal(HS3)                                     # also after the list changed its stride (capacity grew)
    out = c.step(flags, want=("F_att",))                        # and steady again
    assert np.array_equal(out["F_att"], F3)
    assert launches_steady > 0
    # CUDA-graph replay of the steady step: one launch, same bits; new positions go through the pinned staging block
    l0 = c.counter(6)
    for rep in range(3):
        out = c.step(flags, want=("F_att", "SCforce_att"), graph=True)
        assert np.array_equal(out["F_att"], F3) and np.array_equal(out["SCforce_att"], P3)
    assert c.counter(16) == 1 and c.counter(6) - l0 <= launches_steady + 3        # captured once, replayed as single launches
    S4 = np.mod(S2 + rng.normal(0, 1e-3, size=S2.shape), 1.0)
    F4, P4, n4 = _staged(api, wl, o, S=S4, supce=h3)
    out = c.step(flags, S=S4, want=("F_att", "SCforce_att"), graph=True)      # the input set changed (S given): one more capture
    assert np.array_equal(out["F_att"], F4) and np.array_equal(out["SCforce_att"], P4)
    out = c.step(flags, S=S2, want=("F_att", "SCforce_att"), graph=True)
    assert np.array_equal(out["F_att"], F3) and np.array_equal(out["SCforce_att"], P3)
    assert c.counter(16) == 2
    # back to the wider cell: blocks DROP out of the pattern; the resident matrices must lose them (only those are cleared,
    # the blocks that stay are overwritten by the scatter) -- in both forms of the clearing kernel
    import os
    # -- and with the dense values written by the block kernel itself (default) or by the separate scatter kernel
    for zero_all, fused in (("0", "1"), ("1", "1"), ("0", "0"), ("1", "0"), ("0", "1")):
        os.environ["XTB_DENSE_ZERO_ALL"] = zero_all
        os.environ["XTB_DENSE_FUSED"] = fused
        try:
            out = c.step(flags, S=S2, supce=h3, want=("F_att",))
            assert np.array_equal(out["F_att"], F3) and dense_equal(HS3)
            out = c.step(flags, S=S2, supce=wl["supce"], want=("F_att", "SCforce_att"))
            assert np.array_equal(out["F_att"], F2) and np.array_equal(out["SCforce_att"], P2)
            assert c.counter(3) == n2 and dense_equal(HS2)
        finally:
            del os.environ["XTB_DENSE_ZERO_ALL"]
            del os.environ["XTB_DENSE_FUSED"]
    c.close()


def test_step_with_device_eigensolve_and_fermi(wk, orc, api):
    abi = importlib.import_module("xtant-3_b200._abi")
    wl = wk.si_dia((2, 2, 2), sigma=0.1, seed=5)
    o = orc.nrl_step(wl, with_pressure=False)
    c = api.TBContext(0)
    c.set_workload(wl)
    out = c.step(abi.STEP_ALL | abi.STEP_EIGEN, n_electrons=wl["n_electrons"], Te_eV=wl["Te_eV"],
                 want=("Ei", "fe_out", "mu", "E_band", "F_att"))
    assert np.abs(out["Ei"] - o["Ei"]).max() <= 1e-9
    fe, mu = orc.fermi(o["Ei"], wl["n_electrons"], wl["Te_eV"])
    assert abs(out["mu"] - mu) <= 1e-8 and np.abs(out["fe_out"] - fe).max() <= 1e-8
    d = np.abs(orc.forces_cartesian(out["F_att"], wl["supce"]) - orc.forces_cartesian(o["F"], wl["supce"])).max()
    assert d <= 1e-8, d
    c.close()


def _entropy_term(fe, Te):
    x = np.clip(fe / 2.0, 1e-300, 1.0)
    y = np.clip(1.0 - fe / 2.0, 1e-300, 1.0)
    return 2.0 * Te * float(np.sum(x * np.log(x) + y * np.log(y)))       # = -Te * S_e


def test_nve_trajectory_conserves_energy(wk, api):
    """diamond 64, Pettifor-type sp3 + repulsive part (exact derivatives: no d orbitals, so the reference's dxz/dx2-y2 quirk
    plays no part), eigensolve and Fermi occupations on the device, velocity Verlet on the device.  The conserved quantity at
    fixed electron temperature is E_kin + sum f E - Te S_e + E_rep.  dt = 0.5 fs vs 0.25 fs: the fluctuation scales as dt^2."""
    abi = importlib.import_module("xtant-3_b200._abi")
    wl = wk.diamond64()
    na = wl["na"]
    m = np.array([12.011 * AMU])
    Te = 0.1
    rng = np.random.default_rng(12)
    v = rng.normal(0.0, 0.006, size=(na, 3)); v -= v.mean(axis=0)         # [A/fs], ~ 500 K
    SV = np.linalg.solve(wl["supce"].T, v.T).T                           # V = SV . supce
    flags = abi.STEP_ALL | abi.STEP_EIGEN | abi.STEP_PAIR_TERMS

    def run(dt, nsteps):
        c = api.TBContext(0)
        c.set_workload(wl)
        c.md_set_velocities(SV)
        tot = []
        out = c.step(flags, n_electrons=wl["n_electrons"], Te_eV=Te, want=("E_band", "E_rep", "fe_out"))
        for it in range(nsteps + 1):
            ek = c.md_get_state(m)[2]
            tot.append(ek + out["E_band"] + out["E_rep"] + _entropy_term(out["fe_out"], Te))
            if it == nsteps:
                break
            c.make_time_step_atoms(2, dt, m, force_terms=abi.FORCE_ATT | abi.FORCE_REP)
            out = c.step(flags, n_electrons=wl["n_electrons"], Te_eV=Te, want=("E_band", "E_rep", "fe_out"))
            c.make_time_step_atoms(1, dt, m, force_terms=abi.FORCE_ATT | abi.FORCE_REP)
        ek0 = tot and c.md_get_state(m)[2]
        c.close()
        return np.array(tot), ek0

    e1, ek = run(0.5, 120)
    e2, _ = run(0.25, 240)
    assert ek > 0.5                                                       # the atoms do move (kinetic energy of order eV)
    f1 = np.abs(e1 - e1[0]).max() / na; f2 = np.abs(e2 - e2[0]).max() / na
    assert f1 < 5e-4, f1                                                  # eV / atom over 60 fs (velocity Verlet: (w dt)^2 / 8 of E_kin per atom)
    assert f2 < 0.4 * f1 + 1e-9, (f1, f2)                                 # second-order integrator with exact forces: ~ 1/4
