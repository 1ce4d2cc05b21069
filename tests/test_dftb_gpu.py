"""GPU parity tests of the DFTB branch (xtant-3_b200/csrc/dftb.cu) through the C ABI against oracle/dftb_oracle.c.

Tolerances as in test_parity_gpu.py: lists bit-exact; radial tables, H, S 1e-12 relative; forces 1e-9 eV/A;
pressure 1e-8 GPa.  Cases: Si/Al zincblende (two kinds, sp3d5, hetero-nuclear tables), C/H (two kinds, sp3),
H simple cubic (s basis: assembly and density only -- the reference's s-basis force formula is not a derivative).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _assert_rel(got, ref, rel, what):
    scale = np.abs(ref).max()
    diff = np.abs(got - ref)
    big = np.abs(ref) >= 1e-3 * scale
    if big.any():
        r = (diff[big] / np.abs(ref[big])).max()
        assert r <= rel, f"{what}: elementwise relative error {r:.3e} > {rel}"
    assert diff.max() <= rel * scale, f"{what}: abs error {diff.max():.3e} > {rel}*{scale:.3e}"


@pytest.fixture(scope="module", params=["sial", "ch", "sial_3x2x2"])
def case(request, wk, orc, api):
    wl = {"sial": wk.sial_zb, "ch": wk.ch_zb, "sial_3x2x2": lambda: wk.sial_zb((3, 2, 2), sigma=0.1, seed=7)}[request.param]()
    o = orc.dftb_step(wl, with_pressure=True)
    c = api.TBContext(0)
    c.set_workload(wl)
    yield wl, o, c
    c.close()


def test_neighbours_default_cutoff_bit_exact(case):
    wl, o, c = case
    nb = o["nb"]
    g = c.get_near_neighbours()                 # rm <= 0: rcut + 10 d derived by the library (Atomic_tools.f90:588-591)
    np.testing.assert_array_equal(g["size"], nb.size)
    m = nb.max_nn
    mask = np.arange(m)[None, :] < nb.size[:, None]
    np.testing.assert_array_equal(g["list"][mask], nb.list[:, :m][mask])
    assert np.array_equal(g["dist"][mask], nb.dist[:, :m][mask])


def test_radial_tables(case):
    wl, o, c = case
    nb = o["nb"]
    c.get_near_neighbours()
    t = c.radial_tables()
    nv = {1: 1, 4: 4, 9: 10}[wl["n_orb"]]
    worst = 0.0
    for i in range(0, wl["na"], 5):
        for k in range(nb.size[i]):
            j = nb.list[i, k]
            for key, off in (("M_Vij", 0), ("M_dVij", 10), ("M_SVij", 20), ("M_dSVij", 30)):
                ref = o["rad"][i, k, off:off + nv]
                got = t[key][i, j, :]
                worst = max(worst, np.abs(got - ref).max() / max(1e-300, np.abs(ref).max()))
            np.testing.assert_allclose(t["M_lmn"][:, i, j], o["lmn"][i, k], rtol=1e-15, atol=0)
    assert worst <= 1e-12, worst


def test_HS_parity(case):
    wl, o, c = case
    c.get_near_neighbours()
    H, S = c.assemble_HS()
    _assert_rel(H, o["H"], 1e-12, "H_non")
    _assert_rel(S, o["S"], 1e-12, "Sij")
    assert np.array_equal(H, H.T) and np.array_equal(S, S.T)
    assert ((S == 0) == (o["S"] == 0)).all()


def test_density_forces_pressure_parity(case, orc):
    wl, o, c = case
    no = wl["n_orb"]
    c.get_near_neighbours()
    c.set_eigen(o["Ha"], o["Ei"], o["fe"])
    A, AE = c.density(dense=True)
    nb = o["nb"]
    gA, gAE = orc.gather_blocks(nb, no, A), orc.gather_blocks(nb, no, AE)
    mask = (np.arange(nb.cap + 1)[None, :] <= nb.size[:, None])
    for got, ref, nm in ((gA, o["rho"], "rho"), (gAE, o["rhoE"], "rho_E")):
        d = np.abs(got[mask] - ref[mask]).max()
        assert d <= 1e-11 * np.abs(ref[mask]).max(), f"{nm}: {d:.3e}"
    F = c.forces()
    dF = np.abs(orc.forces_cartesian(F, wl["supce"]) - orc.forces_cartesian(o["F"], wl["supce"])).max()
    assert dF <= 1e-9, f"forces differ by {dF:.3e} eV/A"
    P = c.pressure()
    p_gpu, s_gpu = orc.pressure_GPa(P, wl["supce"])
    p_ref, s_ref = orc.pressure_GPa(o["P"], wl["supce"])
    assert abs(p_gpu - p_ref) <= 1e-8 and np.abs(s_gpu - s_ref).max() <= 1e-8, (p_gpu, p_ref)
    assert np.array_equal(c.forces(), F) and np.array_equal(c.pressure(), P)      # deterministic


def test_s_basis(wk, orc, api):
    wl = wk.h_sc()
    o = orc.dftb_step(wl, with_pressure=False)
    c = api.TBContext(0)
    c.set_workload(wl)
    c.get_near_neighbours()
    H, S = c.assemble_HS()
    _assert_rel(H, o["H"], 1e-12, "H s-basis"); _assert_rel(S, o["S"], 1e-12, "S s-basis")
    c.set_eigen(o["Ha"], o["Ei"], o["fe"])
    A, AE = c.density(dense=True)
    g = orc.gather_blocks(o["nb"], 1, A)
    mask = (np.arange(o["nb"].cap + 1)[None, :] <= o["nb"].size[:, None])
    assert np.abs(g[mask] - o["rho"][mask]).max() <= 1e-11 * np.abs(o["rho"][mask]).max()
    # forces of the s basis AS WRITTEN in the reference (get_forces_DFTB / d_Hamilton_one_DFTB hand d_KS_s the value vectors,
    # TB_DFTB.f90:1047-1048): one number for every atom and direction; the oracle's verbatim loop nest gives the same
    Fv = orc.dftb_forces_verbatim(wl, o["nb"], o["lmn"], o["rad"], orc.density_dense(o["Ha"], o["fe"]), orc.density_dense(o["Ha"], o["fe"], o["Ei"]))
    F = c.forces()
    assert np.abs(Fv).max() > 1e-3 and np.abs(Fv - Fv[0, 0]).max() <= 1e-9 * np.abs(Fv).max()
    assert np.abs(F - Fv).max() <= 1e-10 * np.abs(Fv).max(), (F[0], Fv[0])
    assert np.array_equal(F, np.full_like(F, F[0, 0]))
    with pytest.raises(api.XtbError):
        c.pressure()                            # the supercell-force arm of the s basis stays refused (TB_DFTB.f90:1393-1394)
    c.close()


def test_partition_matches_full(case, api):
    """two contexts owning half the atoms each reproduce the single-context H/S column panels
    (the data path of the 2-GPU run, emulated on one device)"""
    wl, o, c = case
    na, no = wl["na"], wl["n_orb"]
    c.get_near_neighbours()
    H, S = c.assemble_HS()
    half = na // 2
    Hp = np.zeros_like(H); Sp = np.zeros_like(S)
    for (a, b) in ((1, half), (half + 1, na)):
        cp = api.TBContext(0)
        cp.set_workload(wl)
        cp.get_near_neighbours()
        cp.set_partition(a, b)
        N = na * no
        Hh = np.zeros((N, N), order="F"); Sh = np.zeros((N, N), order="F")
        cp._ck(cp.lib.xtb_assemble_HS(cp.h, api._dp(Hh), api._dp(Sh)))
        Hp += Hh; Sp += Sh
        cp.close()
    assert np.array_equal(Hp, H) and np.array_equal(Sp, S)


def test_error_paths(api, wk):
    import ctypes as C
    abi = __import__("importlib").import_module("xtant-3_b200._abi")
    wl = wk.ch_zb()
    c = api.TBContext(0)
    arr, keep = abi.dftb_pair_array(wl["params"])
    assert c.lib.xtb_set_params_dftb(c.h, 2, 5, arr) == abi.XTB_EINVAL           # n_orb must be 1, 4 or 9
    bad, keep2 = abi.dftb_pair_array(wl["params"])
    bad[1].n_r = 0
    assert c.lib.xtb_set_params_dftb(c.h, 2, 4, bad) == abi.XTB_EINVAL
    assert c.lib.xtb_set_params_dftb(c.h, 2, 4, arr) == 0
    c.close()
