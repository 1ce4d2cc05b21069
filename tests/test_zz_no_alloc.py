"""GPU test of xtb_vdw_pressure_noalloc, the allocation-free variant of xtb_vdw_pressure (vdw.cu; no counterpart in the reference,
added for callers that cannot call cudaMalloc).  Collected last on purpose (file name):
a failure here must not hide any other test.

Checks, through the C ABI: the result equals the existing entry point's on two calls in a row (tolerance of the vdW tests: 1e-11
of the largest element; in fact the same kernels run, so the bits agree), with a host and with a device output pointer; the
device's free memory does not shrink across a call; a partition is honoured; sizes beyond XTB_NOALLOC_MAX_ATOMS are refused
with XTB_ECAPACITY and a message."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _noalloc(c, out_ptr):
    return c.lib.xtb_vdw_pressure_noalloc(c.h, C.c_void_p(out_ptr))


@pytest.mark.parametrize("name", ["ac216", "ac1000"])
def test_noalloc_matches_existing_entry_point_twice(wk, api, name):
    import torch
    wl = getattr(wk, name)()
    c = api.TBContext(0)
    c.set_workload(wl)
    ref = c.vdW_Pressure()                                   # existing entry point (context buffers), (3,3) column-major
    assert np.abs(ref).max() > 1e-3
    tol = 1e-11 * np.abs(ref).max()
    for call in range(2):                                    # two calls in a row share the fixed scratch array
        out = np.full(9, np.nan)
        assert _noalloc(c, out.ctypes.data) == 0, c.lib.xtb_last_error(c.h)
        got = out.reshape(3, 3, order="F")
        assert np.abs(got - ref).max() <= tol, (call, got, ref)
        assert np.array_equal(got, ref), call                # same kernels, same order of summation
    # device output pointer, and no allocation on the device across the call
    dev = torch.full((9,), float("nan"), dtype=torch.float64, device="cuda:0")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    assert _noalloc(c, dev.data_ptr()) == 0
    free1 = torch.cuda.mem_get_info(0)[0]
    assert free1 >= free0, (free0, free1)
    assert np.array_equal(dev.cpu().numpy().reshape(3, 3, order="F"), ref)
    # an unrelated call in between (it overwrites nothing the variant relies on), then again
    c.get_vdW_s(); c.vdW_forces()
    out = np.zeros(9)
    assert _noalloc(c, out.ctypes.data) == 0
    assert np.array_equal(out.reshape(3, 3, order="F"), ref)
    # partial sums of a partition, as the existing entry point returns them
    na = wl["na"]
    c.set_partition(na // 3 + 1, 2 * na // 3)
    part_ref = c.vdW_Pressure()
    assert _noalloc(c, out.ctypes.data) == 0
    assert np.array_equal(out.reshape(3, 3, order="F"), part_ref)
    c.close()


def test_noalloc_on_a_fresh_context_and_state_errors(wk, api):
    """first use of the context's vdW term is the allocation-free call itself; missing parameters are a state error"""
    abi = api._abi
    wl = wk.ac216()
    c = api.TBContext(0)
    c.set_workload(wl)
    out = np.zeros(9)
    assert _noalloc(c, out.ctypes.data) == 0                 # nothing of the vdW path has run on this context yet
    ref = c.vdW_Pressure()
    assert np.array_equal(out.reshape(3, 3, order="F"), ref)
    c.set_params_vdw(None, None)
    assert _noalloc(c, out.ctypes.data) == abi.XTB_ESTATE
    assert b"xtb_set_params_vdw" in c.lib.xtb_last_error(c.h)
    assert _noalloc(c, 0) == abi.XTB_EINVAL                  # a null output is refused before anything else
    c.close()


def test_noalloc_refuses_sizes_beyond_the_fixed_maximum(wk, api):
    abi = api._abi
    S, h = wk.lattice(wk._FCC, 4.0782, (13, 13, 13), 0.02, 5)          # 8788 atoms > XTB_NOALLOC_MAX_ATOMS
    wl = wk._wl("au_big", "nrl", "Au", S, h, 11, 0.1)
    wl["vdw_kind"] = "LJ"
    wl["vdw_params"] = [{"eps": 0.05, "r0": 3.2, "n": 6.0, "d0_short": 2.0, "dd_short": 0.03, "d0_cut": 5.0, "dd_cut": 0.1}]
    assert wl["na"] > abi.NOALLOC_MAX_ATOMS
    c = api.TBContext(0)
    c.set_workload(wl)
    out = np.zeros(9)
    assert _noalloc(c, out.ctypes.data) == abi.XTB_ECAPACITY
    msg = c.lib.xtb_last_error(c.h)
    assert b"XTB_NOALLOC_MAX_ATOMS" in msg and b"8788" in msg, msg
    # inside the limit once the rank owns a slice (the limit is on the OWNED atoms), and equal to the existing entry point there
    c.set_partition(1, 4000)
    assert _noalloc(c, out.ctypes.data) == 0
    assert np.array_equal(out.reshape(3, 3, order="F"), c.vdW_Pressure())
    c.close()
