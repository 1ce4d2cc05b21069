"""GPU parity of xtb_coupling (coupling.cu: block-sparse S*Ha kernel + library DGEMM + in-place antisymmetrise kernel)
against the oracle's restatement of Electron_ion_coupling_Mij (Nonadiabatic.f90:782-945).  Tolerance 1e-12 absolute on
matrix elements of order 0.1-1 (different summation order; the reference normalises the vectors before the dot products,
the kernel scales G afterwards)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _two(wk, orc, wl, step, seed=0):
    o0 = step(wl, with_pressure=False)
    w1 = dict(wl); w1["S"] = np.mod(wl["S"] + np.random.default_rng(seed).normal(0, 1e-3, wl["S"].shape), 1.0)
    return o0, w1, step(w1, with_pressure=False)


@pytest.mark.parametrize("name", ["au32", "si64", "au108"])
@pytest.mark.parametrize("kind_M", [0, -1])
def test_coupling_nrl(wk, orc, api, name, kind_M):
    wl = {"au32": lambda: wk.au_fcc((2, 2, 2), sigma=0.1, seed=21), "si64": lambda: wk.si_dia((2, 2, 2), sigma=0.05, seed=2),
          "au108": wk.au108}[name]()
    o0, w1, o1 = _two(wk, orc, wl, orc.nrl_step)
    ref = orc.coupling_Mij(o1["Ha"], o0["Ha"], o1["S"], kind_M)
    c = api.TBContext(0)
    c.set_workload(w1)
    c.get_near_neighbours_device()
    c.assemble_HS(to_host=False)
    c.set_eigen(o0["Ha"], o0["Ei"], o0["fe"])
    c.save_eigen_prev()                                     # Ha0 = Ha, device to device
    c.set_eigen(o1["Ha"], o1["Ei"], o1["fe"])
    M = c.Electron_ion_coupling_Mij(kind_M)
    assert np.abs(M - ref).max() <= 1e-12, np.abs(M - ref).max()
    assert np.array_equal(M, M.T) and not np.diag(M).any()
    assert np.array_equal(M, c.Electron_ion_coupling_Mij(kind_M))     # deterministic
    c.close()


def test_coupling_orthogonal(wk, orc, api):
    wl = wk.diamond64()
    o0, w1, o1 = _two(wk, orc, wl, orc.pettifor_step)
    ref = orc.coupling_Mij(o1["Ha"], o0["Ha"], None, 0)
    c = api.TBContext(0)
    c.set_workload(w1)
    c.get_near_neighbours_device()
    c.set_eigen_prev(o0["Ha"])
    c.set_eigen(o1["Ha"], o1["Ei"], o1["fe"])
    M = c.Electron_ion_coupling_Mij(0)
    assert np.abs(M - ref).max() <= 1e-12
    c.close()


def test_coupling_errors(wk, api):
    c = api.TBContext(0)
    c.set_workload(wk.au_fcc((2, 2, 2)))
    with pytest.raises(api.XtbError) as e:
        c.Electron_ion_coupling_Mij(0)
    assert e.value.code == 6
    c.close()
