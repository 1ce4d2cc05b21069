"""The plain-C caller of the ABI (tests/abi_driver.c, built by __graft_entry__.build()): the shim's call sequence for one
NRL step, then the same step as one xtb_step call, compared bit for bit inside the C program; with >= 2 GPUs also the
partitioned step of n processes that exchange the NCCL id through a file (as MPI ranks would through MPI_Bcast)."""
import ctypes as C
import importlib
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DRIVER = os.path.join(ROOT, "tests", "abi_driver")


def _write_problem(tmp_path, wl):
    abi = importlib.import_module("xtant-3_b200._abi")
    p = abi.nrl_pair_from_dict(wl["params"][0])
    fp = tmp_path / "params.bin"; fs = tmp_path / "system.bin"
    fp.write_bytes(bytes(p))
    with open(fs, "wb") as f:
        f.write(np.int32(wl["na"]).tobytes())
        f.write(np.asfortranarray(wl["supce"], dtype=np.float64).tobytes(order="F"))
        f.write(np.ascontiguousarray(wl["S"], dtype=np.float64).tobytes())
        f.write(np.float64(wl["n_electrons"]).tobytes()); f.write(np.float64(wl["Te_eV"]).tobytes())
    return str(fp), str(fs)


def _run(args):
    if not os.path.exists(DRIVER):
        pytest.fail("tests/abi_driver is not built: python -c 'import __graft_entry__ as g; g.build()'")
    r = subprocess.run([DRIVER] + args, capture_output=True, text=True, timeout=900)
    print(r.stdout); print(r.stderr)
    assert r.returncode == 0, (r.returncode, r.stdout[-2000:], r.stderr[-2000:])
    assert "OK all" in r.stdout
    return r.stdout


def test_c_driver_single_gpu(tmp_path, wk):
    wl = wk.au_fcc((3, 3, 3), sigma=0.08, seed=33)
    out = _run(list(_write_problem(tmp_path, wl)))
    assert "OK staged na=108" in out and out.count("OK step") == 2


def test_c_driver_two_ranks(tmp_path, wk):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    wl = wk.au_fcc((3, 3, 6), sigma=0.08, seed=34)           # 216 atoms, two z-slabs
    out = _run(list(_write_problem(tmp_path, wl)) + ["2"])
    assert out.count("OK rank") == 4
