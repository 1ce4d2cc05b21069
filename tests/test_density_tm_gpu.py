"""The 9-orbital density kernel reading STATE-MINOR eigenvectors straight through a tensor map (TMA boxes of 16 states x 9
rows, 128-byte swizzle; no re-tiled copy) gives the same bits as the path that re-tiles the Fortran-layout Ha first: same
chunks, same k-steps, same accumulation order."""
import importlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _wl(wk, name):
    return {"au108": wk.au108, "au_3x3x3_odd": lambda: wk.au_fcc((3, 3, 3), sigma=0.12, seed=77), "si216": wk.si216,
            "al500": wk.al500, "au32": lambda: wk.au_fcc((2, 2, 2), sigma=0.1, seed=21)}[name]()


def _blocks(c):
    abi = importlib.import_module("xtant-3_b200._abi"); interop = importlib.import_module("xtant-3_b200.interop")
    return interop.buffer(c, abi.BUF_RHO).cpu().numpy().copy(), interop.buffer(c, abi.BUF_RHOE).cpu().numpy().copy()


@pytest.mark.parametrize("name", ["au32", "au108", "au_3x3x3_odd", "si216", "al500"])
def test_tensor_map_path_is_bit_identical(name, wk, api, monkeypatch):
    import torch
    wl = _wl(wk, name)
    N = wl["na"] * 9
    C_, E, fe = wk.synthetic_eigen(N, seed=11)
    c1 = api.TBContext(0); c1.set_workload(wl); c1.get_near_neighbours_device(); c1.set_eigen(C_, E, fe)
    c1.density()
    r1, e1 = _blocks(c1)
    assert c1.counter(17) == 0 and np.abs(r1).max() > 0
    # HaT[row][state]: C_ is Ha(row, state) as a numpy (N, N) array -> C-contiguous copy has the states of a row contiguous
    HaT = torch.from_numpy(np.ascontiguousarray(np.asarray(C_))).cuda()
    dE = torch.from_numpy(np.ascontiguousarray(E)).cuda(); dfe = torch.from_numpy(np.ascontiguousarray(fe)).cuda()
    torch.cuda.synchronize()
    c2 = api.TBContext(0); c2.set_workload(wl); c2.get_near_neighbours_device()
    c2.set_eigen_device_t(HaT.data_ptr(), dE.data_ptr(), dfe.data_ptr())
    c2.density()
    r2, e2 = _blocks(c2)
    assert c2.counter(17) == 1                      # really went through the tensor map
    assert np.array_equal(r1, r2) and np.array_equal(e1, e2)
    F1 = (c1.forces(), c1.pressure()); F2 = (c2.forces(), c2.pressure())
    assert np.array_equal(F1[0], F2[0]) and np.array_equal(F1[1], F2[1])
    # a partitioned range and the split / tail forms of the persistent grid
    na = wl["na"]
    c2.set_partition(na // 3 + 1, 2 * na // 3); c1.set_partition(na // 3 + 1, 2 * na // 3)
    c1.density(); c2.density()
    a, b = _blocks(c1), _blocks(c2)
    assert c2.counter(17) == 1 and np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    c1.set_partition(1, na); c2.set_partition(1, na)
    monkeypatch.setenv("XTB_DENSITY_TAIL", str(na // 2))
    c2.density()
    b = _blocks(c2)
    assert np.array_equal(r1, b[0]) and np.array_equal(e1, b[1])
    monkeypatch.delenv("XTB_DENSITY_TAIL")
    # kept states that are NOT a prefix (an empty state in the middle): falls back to the re-tiled path on a transposed copy
    fe2 = fe.copy(); fe2[N // 3] = 0.0
    dfe2 = torch.from_numpy(np.ascontiguousarray(fe2)).cuda(); torch.cuda.synchronize()
    c2.set_eigen_device_t(HaT.data_ptr(), dE.data_ptr(), dfe2.data_ptr())
    c2.density()
    c1.set_eigen(C_, E, fe2); c1.density()
    a, b = _blocks(c1), _blocks(c2)
    assert c2.counter(17) == 0 and np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    # the coupling matrix needs the Fortran layout: derived from HaT on the device, same result
    c1.close(); c2.close()


def test_odd_orbital_count_falls_back(wk, api):
    """N = 9 na odd: the row pitch of HaT (N * 8 bytes) is not a multiple of 16, a tensor map cannot describe it; the kernel
    then takes the re-tiled path on a transposed copy -- same results as the Fortran-layout input"""
    import torch
    w0 = wk.au_fcc((3, 3, 3), sigma=0.1, seed=3)
    wl = wk._wl("au107", "nrl", "Au", w0["S"][:-1], w0["supce"], 11, 0.1)
    N = wl["na"] * 9
    assert N % 2 == 1
    C_, E, fe = wk.synthetic_eigen(N, seed=13)
    c1 = api.TBContext(0); c1.set_workload(wl); c1.get_near_neighbours_device(); c1.set_eigen(C_, E, fe); c1.density()
    HaT = torch.from_numpy(np.ascontiguousarray(np.asarray(C_))).cuda()
    dE = torch.from_numpy(np.ascontiguousarray(E)).cuda(); dfe = torch.from_numpy(np.ascontiguousarray(fe)).cuda()
    torch.cuda.synchronize()
    c2 = api.TBContext(0); c2.set_workload(wl); c2.get_near_neighbours_device()
    c2.set_eigen_device_t(HaT.data_ptr(), dE.data_ptr(), dfe.data_ptr()); c2.density()
    a, b = _blocks(c1), _blocks(c2)
    assert c2.counter(17) == 0 and np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    # and the on-device eigensolve (state-minor output) works for it as well
    c2.assemble_HS(to_host=False); Ei = c2.eigensolve()[0]
    fe2, mu = wk.occupations(Ei, wl["n_electrons"], wl["Te_eV"]); c2.set_occupations(fe2); c2.density()
    assert c2.counter(17) == 0 and np.isfinite(c2.forces()).all()
    c1.close(); c2.close()


def test_eigensolve_leaves_state_minor_eigenvectors(wk, orc, api):
    """xtb_eigensolve (9 orbitals) writes HaT with its last GEMM; the density kernel streams it through the tensor map; the
    Fortran-layout Ha handed back to the host is the transpose made on demand"""
    wl = wk.si_dia((2, 2, 2), sigma=0.1, seed=5)
    o = orc.nrl_step(wl, with_pressure=False)
    c = api.TBContext(0); c.set_workload(wl); c.get_near_neighbours_device(); c.assemble_HS(to_host=False)
    Ei, wS, Ha = c.eigensolve(want_Ha=True)
    assert np.abs(Ei - o["Ei"]).max() <= 1e-9
    S = o["S"]
    assert np.abs(Ha.T @ S @ Ha - np.eye(Ha.shape[0])).max() <= 1e-9          # C^T S C = 1: a genuine eigenvector matrix, Fortran layout
    c.set_occupations(o["fe"]); c.density()
    assert c.counter(17) == 1
    F = c.forces()
    d = np.abs(orc.forces_cartesian(F, wl["supce"]) - orc.forces_cartesian(o["F"], wl["supce"])).max()
    assert d <= 1e-8, d
    c.close()
