"""GPU parity tests: the CUDA path, called through the C ABI, against the CPU oracle on the same inputs.

Tolerances (BASELINE.json north_star):
  neighbour lists / pair indexing     bit-exact
  H, S entries                        1e-12 relative  (elementwise where |ref| >= 1e-3 * max|ref|,
                                                       1e-12 * max|ref| absolute below that)
  rho, rho_E blocks                   1e-11 * max|ref| (the reference pins no tolerance; dgemm order differs)
  forces                              1e-9 eV/A  (dE/ds converted with the cell matrix, oracle.forces_cartesian)
  pressure                            1e-8 GPa   (potential part of Get_pressure, oracle.pressure_GPa)
Eigenvectors are computed once (oracle: Loewdin + dsyevd, as the reference) and fed to BOTH sides, so
the eigenvector gauge never enters.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ctx(api, wl):
    c = api.TBContext(0)
    c.set_workload(wl)
    return c


CASES = ["au32", "si64", "au108", "au144"]


def _wl(wk, name):
    return {"au32": lambda: wk.au_fcc((2, 2, 2), sigma=0.12, seed=11),
            "si64": lambda: wk.si_dia((2, 2, 2), sigma=0.05, seed=2),
            "au108": wk.au108,
            "au144": lambda: wk.au_fcc((3, 3, 4), sigma=0.08, seed=9)}[name]()


@pytest.fixture(scope="module", params=CASES)
def case(request, wk, orc, api):
    wl = _wl(wk, request.param)
    o = orc.nrl_step(wl, with_pressure=True)
    c = _ctx(api, wl)
    yield wl, o, c
    c.close()


def test_neighbour_list_bit_exact(case):
    wl, o, c = case
    nb = o["nb"]
    g = c.get_near_neighbours()
    assert g["max_nn"] == nb.max_nn
    np.testing.assert_array_equal(g["size"], nb.size)
    m = nb.max_nn
    mask = np.arange(m)[None, :] < nb.size[:, None]
    np.testing.assert_array_equal(g["list"][mask], nb.list[:, :m][mask])
    assert np.array_equal(g["dist"][mask], nb.dist[:, :m][mask])          # bitwise (x,y,z,r)
    assert np.array_equal(g["dist_s"][mask], nb.dist_s[:, :m][mask])


def test_neighbour_list_dense_reference_shapes(case):
    wl, o, c = case
    nb = o["nb"]
    d = c.get_near_neighbours(dense=True)
    na = wl["na"]
    np.testing.assert_array_equal(d["size"], nb.size)
    for i in (0, na // 2, na - 1):
        m = nb.size[i]
        np.testing.assert_array_equal(d["list"][i, :m], nb.list[i, :m] + 1)     # 1-based
        assert (d["list"][i, m:] == 0).all()
        assert np.array_equal(d["dist"][i, :m, :], nb.dist[i, :m, :])
        assert (d["dist"][i, m:, :] == 1e8).all()


def _assert_rel(got, ref, rel, what):
    scale = np.abs(ref).max()
    diff = np.abs(got - ref)
    big = np.abs(ref) >= 1e-3 * scale
    if big.any():
        r = (diff[big] / np.abs(ref[big])).max()
        assert r <= rel, f"{what}: elementwise relative error {r:.3e} > {rel}"
    assert diff.max() <= rel * scale, f"{what}: abs error {diff.max():.3e} > {rel}*{scale:.3e}"


def test_radial_tables(case):
    wl, o, c = case
    nb = o["nb"]
    c.get_near_neighbours()
    t = c.radial_tables()
    na = wl["na"]
    for i in (0, na - 1):
        for k in range(nb.size[i]):
            j = nb.list[i, k]
            for key, sl in (("M_Vij", slice(0, 10)), ("M_dVij", slice(10, 20)), ("M_SVij", slice(20, 30)), ("M_dSVij", slice(30, 40))):
                ref = o["rad"][i, k, sl]
                np.testing.assert_allclose(t[key][i, j, :], ref, rtol=1e-12, atol=1e-12 * max(1e-300, np.abs(ref).max()))
            np.testing.assert_allclose(t["M_lmn"][:, i, j], o["lmn"][i, k], rtol=1e-15, atol=0)
            np.testing.assert_allclose(t["M_xrr"][:, i, j], o["xrr"][i, k], rtol=1e-15, atol=0)
            np.testing.assert_allclose(t["M_x1"][:, i, j], o["x1"][i, k], rtol=1e-14, atol=1e-14)


def test_HS_parity(case):
    wl, o, c = case
    c.get_near_neighbours()
    H, S = c.assemble_HS()
    _assert_rel(H, o["H"], 1e-12, "H_non")
    _assert_rel(S, o["S"], 1e-12, "Sij")
    assert np.array_equal(H, H.T) and np.array_equal(S, S.T)
    assert ((S == 0) == (o["S"] == 0)).all()            # the |S| <= 1e-12 -> 0 rule and the zero pattern


@pytest.mark.parametrize("impl", [1, 0])
def test_density_blocks_parity(case, orc, api, impl):
    wl, o, _ = case
    os.environ["XTB_DENSITY_IMPL"] = str(impl)
    try:
        c = _ctx(api, wl)
    finally:
        os.environ.pop("XTB_DENSITY_IMPL", None)
    c.get_near_neighbours()
    c.set_eigen(o["Ha"], o["Ei"], o["fe"])
    A, AE = c.density(dense=True)
    nb = o["nb"]
    gA, gAE = orc.gather_blocks(nb, 9, A), orc.gather_blocks(nb, 9, AE)
    mask = (np.arange(nb.cap + 1)[None, :] <= nb.size[:, None])
    for got, ref, nm in ((gA, o["rho"], "rho"), (gAE, o["rhoE"], "rho_E")):
        d = np.abs(got[mask] - ref[mask]).max()
        assert d <= 1e-11 * np.abs(ref[mask]).max(), f"{nm} impl={impl}: {d:.3e}"
    # mirrored blocks are exact transposes; on-site blocks are symmetric to rounding only
    off = np.kron(~np.eye(wl["na"], dtype=bool), np.ones((9, 9), dtype=bool))
    assert np.array_equal(A[off], A.T[off]) and np.allclose(A, A.T, rtol=0, atol=1e-13)
    # outside the neighbour pattern the dense drop-in output is zero by contract
    N = A.shape[0]
    pat = np.zeros((wl["na"], wl["na"]), dtype=bool)
    for i in range(wl["na"]):
        pat[i, i] = True
        pat[i, nb.list[i, :nb.size[i]]] = True
    full = np.kron(pat, np.ones((9, 9), dtype=bool))
    assert (A[~full] == 0).all() and full.shape == (N, N)
    c.close()


def test_forces_and_pressure_parity(case, orc):
    wl, o, c = case
    c.get_near_neighbours()
    c.assemble_HS(to_host=False)                        # on-site forces need the embedded densities of S2
    c.set_eigen(o["Ha"], o["Ei"], o["fe"])
    c.density()
    F = c.forces()
    dF = np.abs(orc.forces_cartesian(F, wl["supce"]) - orc.forces_cartesian(o["F"], wl["supce"])).max()
    assert dF <= 1e-9, f"forces differ by {dF:.3e} eV/A"
    P = c.pressure()
    p_gpu, s_gpu = orc.pressure_GPa(P, wl["supce"])
    p_ref, s_ref = orc.pressure_GPa(o["P"], wl["supce"])
    assert abs(p_gpu - p_ref) <= 1e-8 and np.abs(s_gpu - s_ref).max() <= 1e-8, (p_gpu, p_ref)
    # determinism: same bits run to run (fixed reduction order, no atomics)
    assert np.array_equal(c.forces(), F) and np.array_equal(c.pressure(), P)


def test_state_cutoff_is_exact_for_zero_occupations(case, api):
    wl, o, _ = case
    fe = o["fe"].copy()
    fe[-40:] = 0.0                                      # exact zeros are skipped by default (f_cut = 0)
    c = _ctx(api, wl)
    c.get_near_neighbours(); c.assemble_HS(to_host=False)
    c.set_eigen(o["Ha"], o["Ei"], fe)
    c.density()
    assert c.counter(7) == len(fe) - 40
    F1 = c.forces()
    fe2 = fe.copy(); fe2[-40:] = 1e-300                 # now kept: contributes ~1e-300, same forces
    c.set_eigen(o["Ha"], o["Ei"], fe2)
    c.density()
    assert c.counter(7) == len(fe)
    np.testing.assert_allclose(c.forces(), F1, rtol=0, atol=1e-12)
    c.close()


def test_partition_matches_full(case, api):
    """two contexts owning half the atoms each reproduce the single-context H/S panels and forces
    (the data path of the 2-GPU run, emulated on one device; owner blocks exchanged by hand)."""
    wl, o, c = case
    na = wl["na"]
    c.get_near_neighbours()
    H, S = c.assemble_HS()
    half = na // 2
    Hp = np.zeros_like(H); Sp = np.zeros_like(S)
    for (a, b) in ((1, half), (half + 1, na)):
        cp = _ctx(api, wl)
        cp.get_near_neighbours()
        cp.set_partition(a, b)
        N = na * 9
        Hh = np.zeros((N, N), order="F"); Sh = np.zeros((N, N), order="F")
        cp._ck(cp.lib.xtb_assemble_HS(cp.h, api._dp(Hh), api._dp(Sh)))
        Hp += Hh; Sp += Sh
        cp.close()
    assert np.array_equal(Hp, H) and np.array_equal(Sp, S)


def test_error_paths(api, wk):
    wl = wk.au_fcc((2, 2, 2))
    c = api.TBContext(0)
    with pytest.raises(api.XtbError):
        c.set_system(wl["koa"], wl["S"], wl["supce"])   # parameters first
    c.set_workload(wl)
    with pytest.raises(api.XtbError):
        c.assemble_HS()                                  # neighbours first
    c.get_near_neighbours()
    with pytest.raises(api.XtbError):
        c.density()                                      # eigenvectors first
    bad = wl["koa"].copy(); bad[0] = 7
    with pytest.raises(api.XtbError):
        c.set_system(bad, wl["S"], wl["supce"])
    c.close()


def test_al500_triclinic_HS_and_forces(wk, orc, api):
    """config 3 cell (triclinic, 176 neighbours/atom): lists bit-exact, H/S parity on the full cell;
    density + forces against the sparse oracle on a sample of atoms with synthetic eigenvectors."""
    wl = wk.al500()
    nb = orc.neighbours(wl)
    c = _ctx(api, wl)
    g = c.get_near_neighbours()
    np.testing.assert_array_equal(g["size"], nb.size)
    m = nb.max_nn
    mask = np.arange(m)[None, :] < nb.size[:, None]
    np.testing.assert_array_equal(g["list"][mask], nb.list[:, :m][mask])
    assert np.array_equal(g["dist"][mask], nb.dist[:, :m][mask])
    lmn, _, _ = orc.geometry(wl, nb)
    rad = orc.nrl_radial(wl, nb)
    Ho, So = orc.nrl_assemble_HS(wl, nb, lmn, rad)
    H, S = c.assemble_HS()
    _assert_rel(H, Ho, 1e-12, "H al500"); _assert_rel(S, So, 1e-12, "S al500")
    N = wl["na"] * 9
    C_, E, fe = wk.synthetic_eigen(N, filling=3.0 / 18.0, Te_eV=0.1, seed=7)
    c.set_eigen(C_, E, fe)
    c.density()
    F = c.forces()
    # oracle needs rho blocks of the sampled atoms AND of their neighbours' on-site blocks: compute all on-site
    # rows cheaply by restricting to atoms 0..11 and comparing the pair part through the full sparse formula
    # on a tiny sub-sample is not separable, so evaluate the oracle for all atoms' blocks of the sample's
    # neighbourhood: here simply all atoms but with K limited by the occupations (f = 0 states drop out).
    keep = fe > 0
    rho, rhoE = orc.density_blocks(nb, 9, np.asfortranarray(C_[:, keep]), fe[keep], E[keep])
    Fo = orc.nrl_forces_sparse(wl, nb, lmn, rad, rho, rhoE, 0, 24)
    dF = np.abs(orc.forces_cartesian(F[:24], wl["supce"]) - orc.forces_cartesian(Fo[:24], wl["supce"])).max()
    assert dF <= 1e-9, dF
    c.close()
