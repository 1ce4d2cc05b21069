"""GPU tests of the on-device eigensolve glue (xtant-3_b200/csrc/eigen.cu): a complete
device-resident step  positions -> neighbours -> H/S -> Loewdin + Dsyevd -> occupations -> rho/rho_E -> forces
against the oracle, whose eigensolve is the reference's LAPACK route (oracle.eigensolve_nonorthogonal).
Eigenvalues 1e-9 eV; forces 1e-8 eV/A (two different LAPACK implementations sit between H and the forces)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["au32", "si64", "sial", "ch", "diamond64"])
def test_device_resident_step(name, wk, orc, api):
    wl = {"au32": lambda: wk.au_fcc((2, 2, 2), sigma=0.12, seed=11), "si64": lambda: wk.si_dia((2, 2, 2), sigma=0.05, seed=2),
          "sial": wk.sial_zb, "ch": wk.ch_zb, "diamond64": wk.diamond64}[name]()
    o = {"nrl": orc.nrl_step, "dftb": orc.dftb_step, "pettifor": orc.pettifor_step}[wl["tb"]](wl, with_pressure=False)
    c = api.TBContext(0)
    c.set_workload(wl)
    c.get_near_neighbours_device()
    c.assemble_HS(to_host=False)
    Ei, wS = c.eigensolve()
    assert np.abs(Ei - o["Ei"]).max() <= 1e-9 * max(1.0, np.abs(o["Ei"]).max())
    if wl["tb"] != "pettifor":
        assert np.abs(wS - o["eigen_S"]).max() <= 1e-10
    assert np.array_equal(c.get_eigenvalues(), Ei)
    fe, mu = wk.occupations(Ei, wl["n_electrons"], wl["Te_eV"])
    c.set_occupations(fe)
    c.density()
    F = c.forces()
    dF = np.abs(orc.forces_cartesian(F, wl["supce"]) - orc.forces_cartesian(o["F"], wl["supce"])).max()
    assert dF <= 1e-8, f"{name}: forces differ by {dF:.3e} eV/A"
    assert c.stage_ms(8) > 0.0
    c.close()


@pytest.mark.parametrize("name", ["si64", "diamond64"])
def test_64bit_solver_interface(name, wk, orc, api, monkeypatch):
    """cusolverDnXsyevd (workspace sizes as size_t: what the north-star N = 36 864 needs, the legacy interface counts its
    ~2 N^2 doubles in an int) forced at a small size with XTB_EIGEN_X64=1: eigenvalues against the oracle and against the
    legacy interface, forces against the oracle."""
    wl = {"si64": lambda: wk.si_dia((2, 2, 2), sigma=0.05, seed=2), "diamond64": wk.diamond64}[name]()
    o = {"nrl": orc.nrl_step, "pettifor": orc.pettifor_step}[wl["tb"]](wl, with_pressure=False)
    res = []
    for x64 in ("0", "1"):
        monkeypatch.setenv("XTB_EIGEN_X64", x64)
        c = api.TBContext(0)
        c.set_workload(wl)
        c.get_near_neighbours_device()
        c.assemble_HS(to_host=False)
        Ei, wS = c.eigensolve()
        assert np.abs(Ei - o["Ei"]).max() <= 1e-9 * max(1.0, np.abs(o["Ei"]).max())
        fe, mu = wk.occupations(Ei, wl["n_electrons"], wl["Te_eV"])
        c.set_occupations(fe)
        c.density()
        F = c.forces()
        dF = np.abs(orc.forces_cartesian(F, wl["supce"]) - orc.forces_cartesian(o["F"], wl["supce"])).max()
        assert dF <= 1e-8, f"{name} x64={x64}: forces differ by {dF:.3e} eV/A"
        res.append(Ei)
        c.close()
    assert np.abs(res[0] - res[1]).max() <= 1e-10 * max(1.0, np.abs(res[0]).max())


def test_eigenvectors_are_S_orthonormal(wk, orc, api):
    wl = wk.au_fcc((2, 2, 2), sigma=0.12, seed=11)
    c = api.TBContext(0)
    c.set_workload(wl)
    c.get_near_neighbours_device()
    H, S = c.assemble_HS()
    Ei, wS, Ha = c.eigensolve(want_Ha=True)
    assert np.abs(Ha.T @ S @ Ha - np.eye(len(Ei))).max() <= 1e-10          # C^T S C = I
    assert np.abs(H @ Ha - S @ Ha * Ei[None, :]).max() <= 1e-9 * np.abs(H).max()   # H C = S C E
    c.close()


def test_call_order_errors(wk, api):
    wl = wk.au_fcc((2, 2, 2))
    c = api.TBContext(0)
    c.set_workload(wl)
    c.get_near_neighbours_device()
    with pytest.raises(api.XtbError):
        c.eigensolve()                      # no dense H / S yet
    with pytest.raises(api.XtbError):
        c.set_occupations(np.zeros(wl["na"] * 9))   # no eigenvectors yet
    c.close()
