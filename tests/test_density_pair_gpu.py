"""rho / rho_E contraction with two column atoms per CTA (csrc/density_pair.cuh, XTB_DENSITY_IMPL=2) against the one-atom
DMMA kernel (XTB_DENSITY_IMPL=1) and the CUDA-core cross-check kernel (XTB_DENSITY_IMPL=0).

Both DMMA kernels add the states of a block element in the same order (chunks of 16 states, 4 k-steps, the ninth orbital as
four k-slices summed by the same butterfly), so their blocks must be BIT-identical whatever the pairing, the windows
(sub-tasks) and the deal of the halves to the warps; the oracle parity of the blocks themselves is tests/test_fullsize_gpu.py
and tests/test_parity_gpu.py (which run whatever kernel is the default)."""
import importlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _blocks(api, wl, eig, impl, monkeypatch, part=None, layout=None):
    abi = importlib.import_module("xtant-3_b200._abi"); interop = importlib.import_module("xtant-3_b200.interop")
    monkeypatch.setenv("XTB_DENSITY_IMPL", str(impl))
    c = api.TBContext(0)
    c.set_workload(wl)
    if layout:
        c.set_block_layout(layout)
    if part:
        c.set_partition(*part)
    c.get_near_neighbours_device()
    c.set_eigen(*eig)
    c.density()
    out = (interop.buffer(c, abi.BUF_RHO).cpu().numpy().copy(), interop.buffer(c, abi.BUF_RHOE).cpu().numpy().copy(),
           c.counter(9), c.counter(10))
    c.close()
    return out


CASES = {
    "au108": lambda wk: wk.au108(),                       # 108 atoms from the reference's SAVE files
    "au_3x3x3_odd": lambda wk: _drop_last(wk.au_fcc((3, 3, 3))),   # odd atom count: the last unit is a single atom
    "si216": lambda wk: wk.si216(),                       # 46 neighbours per atom
    "al500": lambda wk: wk.al500(),                       # 176 neighbours per atom: several windows per unit
    "au_2x2x2": lambda wk: wk.au_fcc((2, 2, 2)),          # tiny cell: every atom is everybody's neighbour (images)
}


def _drop_last(wl):
    wl = dict(wl)
    wl["na"] -= 1
    wl["koa"] = wl["koa"][:-1].copy(); wl["S"] = wl["S"][:-1].copy()
    wl["n_electrons"] = wl["n_electrons"] * wl["na"] / (wl["na"] + 1)
    return wl


@pytest.mark.parametrize("name", list(CASES))
def test_pair_kernel_is_bit_identical_to_the_one_atom_kernel(name, wk, api, monkeypatch):
    wl = CASES[name](wk)
    eig = wk.synthetic_eigen(wl["na"] * 9, seed=11)
    r1, e1, _, _ = _blocks(api, wl, eig, 1, monkeypatch)
    r2, e2, slots, stages = _blocks(api, wl, eig, 2, monkeypatch)
    assert np.abs(r1).max() > 0 and slots >= 4 and stages >= 3
    assert np.array_equal(r1, r2) and np.array_equal(e1, e2), (np.abs(r1 - r2).max(), np.abs(e1 - e2).max())
    r0, e0, _, _ = _blocks(api, wl, eig, 0, monkeypatch)
    assert np.abs(r0 - r2).max() <= 1e-11 * np.abs(r0).max() and np.abs(e0 - e2).max() <= 1e-11 * np.abs(e0).max()


@pytest.mark.parametrize("part", [(1, 54), (55, 108), (20, 20), (3, 101)])
def test_pair_kernel_on_a_partition(part, wk, api, monkeypatch):
    """owned atom ranges (multi-GPU): units are formed inside the range; blocks of other atoms stay zero"""
    wl = wk.au108()
    eig = wk.synthetic_eigen(wl["na"] * 9, seed=12)
    r1, e1, _, _ = _blocks(api, wl, eig, 1, monkeypatch, part=part, layout=2)
    r2, e2, _, _ = _blocks(api, wl, eig, 2, monkeypatch, part=part, layout=2)
    assert np.abs(r1).max() > 0
    assert np.array_equal(r1, r2) and np.array_equal(e1, e2)
