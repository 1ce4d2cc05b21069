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
    assert C.sizeof(abi.WallPair) == 4 * 8
    assert C.sizeof(abi.VdwPair) == (2 + 16 + 4 + 2) * 8
    hdr = open(os.path.join(ROOT, "include", "xtb_b200.h")).read()      # member order of the header struct == ctypes mirror
    body = re.search(r"typedef struct xtb_vdw_pair \{(.*?)\} xtb_vdw_pair;", hdr, flags=re.S).group(1)
    names = [n for line in body.splitlines() for n in re.findall(r"\b(\w+)\s*[,;]", line.split("/*")[0])]
    assert tuple(names) == abi.VDW_FIELDS


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


def test_fortran_shim_interfaces_match_the_header():
    """the ISO_C_BINDING interface block of the shim (no Fortran compiler here: checked by parsing) binds only symbols the
    header declares, with the same number of arguments, and covers every entry point a Fortran caller of the path needs"""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "xtb_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", " ", hdr, flags=re.S)
    protos = {}
    for m in re.finditer(r"\b(xtb_[A-Za-z_0-9]+)\s*\(([^;{]*?)\)\s*;", hdr, flags=re.S):
        args = m.group(2).strip()
        protos[m.group(1)] = 0 if args in ("", "void") else len(args.split(","))
    shim = open(os.path.join(root, "xtant-3_b200", "fortran", "xtant_b200_shim.f90")).read()
    bound = {}
    for m in re.finditer(r"function\s+(xtb_[A-Za-z_0-9]+)\s*\(([^)]*)\)\s*bind\(C,\s*name=\"(xtb_[A-Za-z_0-9]+)\"\)", shim, flags=re.I):
        assert m.group(1).lower() == m.group(3).lower()
        bound[m.group(3)] = len([a for a in m.group(2).split(",") if a.strip()])
    assert len(bound) >= 30
    for name, n in bound.items():
        assert name in protos, f"shim binds {name}, which the header does not declare"
        assert protos[name] == n, f"{name}: {n} arguments in the shim, {protos[name]} in the header"
    # introspection / zero-copy / stream entry points are for the Python and C++ hosts; everything else must be bound
    host_only = {"xtb_version", "xtb_set_stream", "xtb_counter", "xtb_device_buffer", "xtb_neighbours_compact", "xtb_set_eigen_device", "xtb_set_eigen_device_t"}
    # and the shim's get_near_neighbours keeps the reference's `dm` contract: a present dm returns the radius, no list
    body = shim[shim.index("subroutine get_near_neighbours"):shim.index("end subroutine get_near_neighbours")]
    assert re.search(r"if \(present\(dm\)\) then\s*!.*\n\s*dm = rm\s*\n\s*else\s*\n\s*call push_system", body)
    assert "xtb_cutoff" in body and "TB_Rep_DFTB" in body and "include_vdW" in body
    assert set(protos) - set(bound) <= host_only, sorted(set(protos) - set(bound) - host_only)


def test_fortran_shim_struct_mirrors_match_ctypes_layout():
    """the bind(C) derived types of the shim have the field order, counts and sizes of the ctypes mirrors of the header
    structs (which test_struct_layout_matches_header ties to the header itself)"""
    import re
    import ctypes as CT
    import importlib
    abi = importlib.import_module("xtant-3_b200._abi")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    shim = open(os.path.join(root, "xtant-3_b200", "fortran", "xtant_b200_shim.f90")).read()
    mirrors = {"xtb_nrl_pair": abi.NrlPair, "xtb_pettifor_pair": abi.PettiforPair,
               "xtb_pettifor_rep_pair": abi.PettiforRepPair, "xtb_dftb_pair": abi.DftbPair,
               "xtb_wall_pair": abi.WallPair, "xtb_vdw_pair": abi.VdwPair, "xtb_coulomb_pair": abi.CoulombPair, "xtb_short_pair": abi.ShortPair,
               "xtb_step_io": abi.StepIO}
    size = {"c_double": 8, "c_int": 4, "c_int32_t": 4, "c_ptr": 8}
    for name, cls in mirrors.items():
        m = re.search(r"type,\s*bind\(C\)\s*::\s*" + name + r"\b(.*?)end type", shim, flags=re.S | re.I)
        assert m, name
        fields = []
        for line in m.group(1).splitlines():
            line = line.split("!")[0].strip()
            mm = re.match(r"(?:real|integer|type)\((\w+)\)\s*::\s*(.*)", line)
            if not mm:
                continue
            for item in re.findall(r"(\w+)(?:\((\d+)\))?", mm.group(2)):
                fields.append((item[0], size[mm.group(1)], int(item[1]) if item[1] else 1))
        ct = [(n, CT.sizeof(t)) for n, t in cls._fields_]
        assert [f[0] for f in fields] == [n for n, _ in ct], (name, [f[0] for f in fields], [n for n, _ in ct])
        assert [f[1] * f[2] for f in fields] == [s for _, s in ct], name
        off = 0                                          # natural alignment, as bind(C) and the C compiler lay it out
        for _, sz, cnt in fields:
            off = (off + sz - 1) // sz * sz + sz * cnt
        assert (off + 7) // 8 * 8 == CT.sizeof(cls), name


def test_fortran_shim_wrappers_keep_the_reference_argument_lists():
    """every reference procedure the shim re-exposes has the reference's own dummy-argument list (names and order), as
    recorded from the reference sources by tests/golden/make_signatures.py"""
    import json
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ref = json.load(open(os.path.join(root, "tests", "golden", "reference_signatures.json")))
    shim = open(os.path.join(root, "xtant-3_b200", "fortran", "xtant_b200_shim.f90")).read()
    shim = re.sub(r"&\s*\n\s*&?", " ", shim)
    mine = {}
    for m in re.finditer(r"^\s*subroutine\s+(\w+)\s*\(([^)]*)\)", shim, flags=re.M | re.I):
        mine.setdefault(m.group(1).lower(), [a.strip().lower() for a in m.group(2).split(",") if a.strip()])
    pub = re.search(r"public ::(.*?)\n\s*\n", shim, flags=re.S).group(1).lower()
    assert len(ref) == 29
    for name, sig in ref.items():
        assert name.lower() in mine, name
        assert re.search(r"\b" + name.lower() + r"\b", pub), f"{name} is not public in the shim"
        assert mine[name.lower()] == sig["args"], (name, sig["file"], mine[name.lower()], sig["args"])


def test_python_host_mirror_has_the_reference_procedure_names():
    """the host-side mirror (api.TBContext) exposes every re-exposed reference procedure under its own name"""
    import importlib
    import json
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    api = importlib.import_module("xtant-3_b200.api")
    ref = json.load(open(os.path.join(root, "tests", "golden", "reference_signatures.json")))
    missing = [n for n in ref if not callable(getattr(api.TBContext, n, None))]
    assert not missing, missing
