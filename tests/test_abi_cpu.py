"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol the
header declares, and refuses to work without a GPU (no CPU fallback)."""
import ctypes as C
import importlib
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    txt = open(os.path.join(ROOT, "include", "xtb_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(xtb_[A-Za-z_0-9]+)\s*\(", txt)))


def test_header_declares_expected_entry_points():
    syms = _header_symbols()
    for s in ("xtb_create", "xtb_neighbours", "xtb_radial", "xtb_assemble_HS", "xtb_set_eigen", "xtb_density",
              "xtb_forces", "xtb_pressure", "xtb_set_params_nrl", "xtb_set_params_pettifor"):
        assert s in syms


def test_library_exports_every_declared_symbol():
    abi = importlib.import_module("xtant-3_b200._abi")
    lib = abi.load()            # raises if the .so is missing: the CUDA extension is mandatory
    for s in _header_symbols():
        assert hasattr(lib, s), f"{s} declared in include/xtb_b200.h but not exported"
    assert set(abi.SIGNATURES) == set(_header_symbols())
    assert b"sm_100a" in lib.xtb_version()


def test_struct_layout_matches_header():
    abi = importlib.import_module("xtant-3_b200._abi")
    assert C.sizeof(abi.NrlPair) == 8 + 3 * 8 + 16 * 8 + 80 * 8
    assert C.sizeof(abi.PettiforPair) == (2 + 28 + 4) * 8
    assert C.sizeof(abi.PettiforRepPair) == (1 + 5 + 7 + 4) * 8


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    abi = importlib.import_module("xtant-3_b200._abi")
    lib = abi.load()
    h = C.c_void_p()
    assert lib.xtb_create(C.byref(h), 0) == abi.XTB_ENODEV
    api = importlib.import_module("xtant-3_b200.api")
    with pytest.raises(api.XtbError):
        api.TBContext(0)


def test_product_never_imports_oracle():
    pk = os.path.join(ROOT, "xtant-3_b200")
    for dp, _, fs in os.walk(pk):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".f90")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                for line in txt.splitlines():
                    if re.search(r"^\s*(import|from|#include)\b.*oracle", line) or "liboracle" in line:
                        raise AssertionError(f"{f} reaches into oracle/: {line.strip()}")
