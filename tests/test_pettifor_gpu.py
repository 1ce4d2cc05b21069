"""GPU parity tests, Pettifor (orthogonal sp3) branch: CUDA path through the C ABI vs the CPU oracle.

Tolerances: neighbour list bit-exact; H entries 1e-12 relative (exp/pow of CUDA's libm vs glibc: <= 2 ulp);
rho blocks 1e-11 * max|rho|; forces 1e-9 eV/A; supercell force as pressure 1e-8 GPa.  The eigenvectors come from
the oracle (dsyevd as the reference, TB_Pettifor.f90:1455) and are fed to both sides.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _wl(wk, name):
    return {"diamond64": wk.diamond64,
            "c64": lambda: wk.c_dia((2, 2, 2), sigma=0.12, seed=3),
            "c216": lambda: wk.c_dia((3, 3, 3), sigma=0.2, seed=8)}[name]()


@pytest.fixture(scope="module", params=["diamond64", "c64", "c216"])
def case(request, wk, orc, api):
    wl = _wl(wk, request.param)
    o = orc.pettifor_step(wl)
    c = api.TBContext(0)
    c.set_workload(wl)
    yield wl, o, c
    c.close()


def _assert_rel(got, ref, rel, what):
    scale = np.abs(ref).max()
    diff = np.abs(got - ref)
    big = np.abs(ref) >= 1e-3 * scale
    if big.any():
        r = (diff[big] / np.abs(ref[big])).max()
        assert r <= rel, f"{what}: elementwise relative error {r:.3e} > {rel}"
    assert diff.max() <= rel * scale, f"{what}: abs error {diff.max():.3e}"


def test_neighbours_bit_exact(case):
    wl, o, c = case
    nb = o["nb"]
    g = c.get_near_neighbours()
    np.testing.assert_array_equal(g["size"], nb.size)
    m = nb.max_nn
    mask = np.arange(m)[None, :] < nb.size[:, None]
    np.testing.assert_array_equal(g["list"][mask], nb.list[:, :m][mask])
    assert np.array_equal(g["dist"][mask], nb.dist[:, :m][mask])
    assert np.array_equal(g["dist_s"][mask], nb.dist_s[:, :m][mask])


def test_H_parity(case):
    wl, o, c = case
    c.get_near_neighbours()
    H, S = c.assemble_HS()
    assert S is None                                            # orthogonal basis: no overlap matrix
    _assert_rel(H, o["H"], 1e-12, "H (Pettifor)")
    assert np.array_equal(H, H.T)
    assert ((H == 0) == (o["H"] == 0)).all()


def test_density_forces_pressure(case, orc):
    wl, o, c = case
    nb = o["nb"]
    c.get_near_neighbours()
    c.assemble_HS(to_host=False)
    c.set_eigen(o["Ha"], o["Ei"], o["fe"])
    A, AE = c.density(dense=True)
    gA = orc.gather_blocks(nb, 4, A)
    mask = (np.arange(nb.cap + 1)[None, :] <= nb.size[:, None])
    d = np.abs(gA[mask] - o["rho"][mask]).max()
    assert d <= 1e-11 * np.abs(o["rho"][mask]).max(), d
    F = c.forces()
    dF = np.abs(orc.forces_cartesian(F, wl["supce"]) - orc.forces_cartesian(o["F"], wl["supce"])).max()
    assert dF <= 1e-9, f"forces differ by {dF:.3e} eV/A"
    P = c.pressure()
    p_gpu, s_gpu = orc.pressure_GPa(P, wl["supce"])
    p_ref, s_ref = orc.pressure_GPa(o["P"], wl["supce"])
    assert abs(p_gpu - p_ref) <= 1e-8 and np.abs(s_gpu - s_ref).max() <= 1e-8, (p_gpu, p_ref)
    assert np.array_equal(c.forces(), F) and np.array_equal(c.pressure(), P)      # deterministic


def test_density_warp_kernel_vs_per_element_kernel(case, orc, api):
    """n_orb = 4 runs the warp-per-block kernel (density.cu, density_warp_kernel); XTB_DENSITY_IMPL=0 selects the
    one-thread-per-element cross-check kernel.  Both against the oracle (rho and rho_E), and against each other."""
    import os
    wl, o, _ = case
    nb = o["nb"]
    mask = (np.arange(nb.cap + 1)[None, :] <= nb.size[:, None])
    got = {}
    for impl in (1, 0):
        os.environ["XTB_DENSITY_IMPL"] = str(impl)
        try:
            c = api.TBContext(0)
        finally:
            os.environ.pop("XTB_DENSITY_IMPL", None)
        c.set_workload(wl)
        c.get_near_neighbours()
        c.set_eigen(o["Ha"], o["Ei"], o["fe"])
        A, AE = c.density(dense=True)
        got[impl] = (A, AE)
        for M, key in ((A, "rho"), (AE, "rhoE")):
            g = orc.gather_blocks(nb, 4, M)
            d = np.abs(g[mask] - o[key][mask]).max()
            assert d <= 1e-11 * np.abs(o[key][mask]).max(), (impl, key, d)
        off = np.kron(~np.eye(wl["na"], dtype=bool), np.ones((4, 4), dtype=bool))
        assert np.array_equal(A[off], A.T[off]) and np.array_equal(AE[off], AE.T[off])
        A2, AE2 = c.density(dense=True)
        assert np.array_equal(A, A2) and np.array_equal(AE, AE2)              # deterministic
        c.close()
    for a, b in zip(got[1], got[0]):
        assert np.abs(a - b).max() <= 1e-12 * np.abs(b).max()


def test_ac1000_config4(wk, orc, api):
    """config 4 (1000-atom disordered carbon): list bit-exact, H parity on the full cell, forces against the
    sparse oracle for all atoms with oracle eigenvectors."""
    wl = wk.ac1000()
    o = orc.pettifor_step(wl, with_pressure=True)
    c = api.TBContext(0)
    c.set_workload(wl)
    g = c.get_near_neighbours()
    nb = o["nb"]
    np.testing.assert_array_equal(g["size"], nb.size)
    H, _ = c.assemble_HS()
    _assert_rel(H, o["H"], 1e-12, "H ac1000")
    c.set_eigen(o["Ha"], o["Ei"], o["fe"])
    c.density()
    F = c.forces()
    dF = np.abs(orc.forces_cartesian(F, wl["supce"]) - orc.forces_cartesian(o["F"], wl["supce"])).max()
    assert dF <= 1e-9, dF
    P = c.pressure()
    assert abs(orc.pressure_GPa(P, wl["supce"])[0] - orc.pressure_GPa(o["P"], wl["supce"])[0]) <= 1e-8
    c.close()
