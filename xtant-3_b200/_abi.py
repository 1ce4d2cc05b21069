"""ctypes view of include/xtb_b200.h: structs, constants and the library loader.

The CUDA extension is mandatory: `load()` raises if libxtb_b200.so is missing or cannot be
loaded -- there is no CPU fallback anywhere in the product path.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# XTB_LIB_VARIANT=<tag> loads libxtb_b200_<tag>.so (kernel experiments built with tools/r02/build_variants.sh); default: the product library
LIB_PATH = os.path.join(_HERE, "libxtb_b200%s.so" % (("_" + os.environ["XTB_LIB_VARIANT"]) if os.environ.get("XTB_LIB_VARIANT") else ""))

XTB_OK, XTB_EINVAL, XTB_ENODEV, XTB_ECUDA, XTB_ENOMEM, XTB_ECAPACITY, XTB_ESTATE, XTB_ECOMM = range(8)
XTB_TB_NONE, XTB_TB_NRL, XTB_TB_PETTIFOR, XTB_TB_DFTB = range(4)

(BUF_H, BUF_S, BUF_RHO, BUF_RHOE, BUF_FORCES, BUF_ROWPTR, BUF_NNLIST, BUF_NNSIZE, BUF_SCFORCE, BUF_HBLK, BUF_SBLK, BUF_FREP, BUF_MIJ,
 BUF_FWALL, BUF_POS, BUF_VEL, BUF_EI, BUF_FE, BUF_FVDW, BUF_FCOUL) = range(1, 21)
CNT_NA, CNT_NORB, CNT_N, CNT_MAXNN, CNT_NBLOCKS, CNT_NPAIRS, CNT_LAUNCHES, CNT_KKEPT, CNT_CAP = range(9)
STAGE_NEIGH, STAGE_ASSEMBLE, STAGE_DENSITY, STAGE_FORCES, STAGE_PRESSURE = range(1, 6)
STAGE_REP_FORCES, STAGE_REP_PRESSURE, STAGE_COUPLING = 9, 10, 11
STAGE_XCHG_HS, STAGE_XCHG_RHO = 14, 15
CNT_XCHG_HS_BYTES, CNT_XCHG_RHO_BYTES, CNT_HALO_SEND, CNT_HALO_RECV = 12, 13, 14, 15
XCHG_HS, XCHG_RHO = 1, 2
FORCE_ATT, FORCE_REP, FORCE_WALL, FORCE_VDW, FORCE_COUL = 1, 2, 4, 8, 16
VDW_GIRIFALCO, VDW_LJ_CUT, VDW_ILJ_CUT = 1, 2, 3
STAGE_VDW, STAGE_COULOMB = 17, 18
NOALLOC_MAX_ATOMS = 8192
COULOMB_K = 1.602176634e-19 / (4.0 * 3.1415926535897932384626433832795 * 8.854187817620e-12) * 1.0e10     # Read_input_data.f90:1907 [eV A]
STEP_NEIGHBOURS, STEP_HS, STEP_EIGEN, STEP_DENSITY, STEP_FORCES, STEP_PRESSURE, STEP_PAIR_TERMS, STEP_NO_DENSE, STEP_GRAPH = 1, 2, 4, 8, 16, 32, 64, 128, 256
STEP_ALL = STEP_NEIGHBOURS | STEP_HS | STEP_DENSITY | STEP_FORCES
UNIQUE_ID_BYTES = 128

_d4 = C.c_double * 4
_d10 = C.c_double * 10


class NrlPair(C.Structure):
    """struct xtb_nrl_pair == type(TB_H_NRL), Objects.f90:190-200"""
    _fields_ = [("ind_split", C.c_int32), ("ind_overlap", C.c_int32),
                ("lambd", C.c_double), ("Rc", C.c_double), ("lden", C.c_double),
                ("al", _d4), ("bl", _d4), ("cl", _d4), ("dl", _d4),
                ("ellm", _d10), ("fllm", _d10), ("gllm", _d10), ("hllm", _d10),
                ("pllm", _d10), ("qllm", _d10), ("rllm", _d10), ("sllm", _d10)]


class PettiforPair(C.Structure):
    """struct xtb_pettifor_pair == type(TB_H_Pettifor), Objects.f90:170-174"""
    _fields_ = [("Es", C.c_double), ("Ep", C.c_double),
                ("nc", _d4), ("rc", _d4), ("c0", _d4), ("c1", _d4), ("c2", _d4), ("c3", _d4), ("V0", _d4),
                ("r0", C.c_double), ("n", C.c_double), ("r1", C.c_double), ("rm", C.c_double)]


class PettiforRepPair(C.Structure):
    """struct xtb_pettifor_rep_pair == type(TB_Rep_Pettifor), Objects.f90:95-100"""
    _fields_ = [("E0_TB", C.c_double), ("a0", C.c_double * 5),
                ("phi0", C.c_double), ("d0", C.c_double), ("m", C.c_double), ("mc", C.c_double), ("dc", C.c_double),
                ("d1", C.c_double), ("dm", C.c_double), ("c0", _d4)]


class WallPair(C.Structure):
    """struct xtb_wall_pair == type(TB_Exp_wall_simple), Objects.f90:458-466"""
    _fields_ = [("C", C.c_double), ("r0", C.c_double), ("d0", C.c_double), ("dd", C.c_double)]


VDW_FIELDS = ("d0_cut", "dd_cut", "C12", "C6", "dm", "d_cut", "a", "b", "c", "d", "dsm", "ds_cut", "as", "bs", "cs", "ds", "es", "fs",
              "eps", "r0", "n", "m", "d0_short", "dd_short")


class VdwPair(C.Structure):
    """struct xtb_vdw_pair == class(TB_vdW) Objects.f90:43-46 + TB_vdW_Girifalco :302-328 / TB_vdW_LJ_cut :336-343 / TB_vdW_ILJ_cut :346-354"""
    _fields_ = [(k, C.c_double) for k in VDW_FIELDS]


def vdw_pair_from_dict(d):
    """missing members are zero (each class reads only its own)"""
    p = VdwPair()
    for k in VDW_FIELDS:
        setattr(p, k, float(d.get(k, 0.0)))
    return p


SHORT_MAX_POW = 8


class ShortPair(C.Structure):
    """struct xtb_short_pair == type(TB_Short_Rep), Objects.f90:437-454 (tables by address)"""
    _fields_ = [("d0", C.c_double), ("dd", C.c_double),
                ("use_exp", C.c_int32), ("use_inv_exp", C.c_int32), ("use_zbl", C.c_int32), ("use_nlh", C.c_int32),
                ("exp_Phi", C.c_double), ("exp_r0", C.c_double), ("exp_a", C.c_double),
                ("inv_C", C.c_double), ("inv_r0", C.c_double),
                ("n_pow", C.c_int32), ("n_tab", C.c_int32),
                ("pow_Phi", C.c_double * SHORT_MAX_POW), ("pow_r0", C.c_double * SHORT_MAX_POW), ("pow_m", C.c_double * SHORT_MAX_POW),
                ("nlh_a", C.c_double * 3), ("nlh_b", C.c_double * 3),
                ("tab_spline", C.c_int32), ("pad_", C.c_int32),
                ("tab_R", C.POINTER(C.c_double)), ("tab_E", C.POINTER(C.c_double)), ("tab_a", C.POINTER(C.c_double)),
                ("tab_b", C.POINTER(C.c_double)), ("tab_c", C.POINTER(C.c_double)), ("tab_d", C.POINTER(C.c_double))]


def short_pair_array(params):
    """list of dicts as readers.read_Short_Rep_TB returns them -> (ShortPair array, objects to keep alive during the call).
    dict keys: d0, dd; exp = (Phi, r0, a); inv_exp = (C, r0); pow = [(Phi, r0, m), ...]; zbl, nlh = bool; nlh_a, nlh_b;
    tab = dict(R, E, spline[, a, b, c, d])"""
    import numpy as np
    arr = (ShortPair * len(params))()
    keep = []
    for q, d in enumerate(params):
        s = arr[q]
        s.d0 = float(d.get("d0", 0.0)); s.dd = float(d.get("dd", 0.01))                  # defaults of read_Short_Rep_TB
        if d.get("exp") is not None:
            s.use_exp = 1; s.exp_Phi, s.exp_r0, s.exp_a = (float(x) for x in d["exp"])
        if d.get("inv_exp") is not None:
            s.use_inv_exp = 1; s.inv_C, s.inv_r0 = (float(x) for x in d["inv_exp"])
        pw = d.get("pow") or []
        if len(pw) > SHORT_MAX_POW:
            raise ValueError(f"{len(pw)} power terms (max {SHORT_MAX_POW})")
        s.n_pow = len(pw)
        for k, (Phi, r0, m) in enumerate(pw):
            s.pow_Phi[k] = float(Phi); s.pow_r0[k] = float(r0); s.pow_m[k] = float(m)
        s.use_zbl = 1 if d.get("zbl") else 0
        s.use_nlh = 1 if d.get("nlh") else 0
        for k in range(3):
            s.nlh_a[k] = float(d.get("nlh_a", (0, 0, 0))[k]); s.nlh_b[k] = float(d.get("nlh_b", (0, 0, 0))[k])
        t = d.get("tab")
        if t is not None:
            s.n_tab = len(t["R"]); s.tab_spline = 1 if t.get("spline") else 0
            for name in ("R", "E", "a", "b", "c", "d"):
                if t.get(name) is not None:
                    a = np.ascontiguousarray(t[name], dtype=np.float64)
                    keep.append(a)
                    setattr(s, "tab_" + name, a.ctypes.data_as(C.POINTER(C.c_double)))
    return arr, keep


class CoulombPair(C.Structure):
    """struct xtb_coulomb_pair == type(TB_Coulomb_cut), Objects.f90:359-366 (k, dm, dd)"""
    _fields_ = [("k", C.c_double), ("dm", C.c_double), ("dd", C.c_double)]


class StepIO(C.Structure):
    """struct xtb_step_io"""
    _fields_ = [("S", C.POINTER(C.c_double)), ("supce", C.POINTER(C.c_double)), ("fe", C.POINTER(C.c_double)),
                ("n_electrons", C.c_double), ("Te_eV", C.c_double),
                ("Ei", C.POINTER(C.c_double)), ("fe_out", C.POINTER(C.c_double)), ("mu", C.POINTER(C.c_double)),
                ("E_band", C.POINTER(C.c_double)), ("F_att", C.POINTER(C.c_double)), ("SCforce_att", C.POINTER(C.c_double)),
                ("E_rep", C.POINTER(C.c_double)), ("F_rep", C.POINTER(C.c_double)), ("SCforce_rep", C.POINTER(C.c_double)),
                ("E_wall", C.POINTER(C.c_double)), ("F_wall", C.POINTER(C.c_double)), ("SCforce_wall", C.POINTER(C.c_double)),
                ("E_vdw", C.POINTER(C.c_double)), ("F_vdw", C.POINTER(C.c_double)), ("SCforce_vdw", C.POINTER(C.c_double)),
                ("E_coul", C.POINTER(C.c_double)), ("F_coul", C.POINTER(C.c_double)), ("SCforce_coul", C.POINTER(C.c_double))]


class DftbPair(C.Structure):
    """struct xtb_dftb_pair == type(TB_H_DFTB), Objects.f90:202-222 (tables flattened)"""
    _fields_ = [("rcut", C.c_double), ("d", C.c_double),
                ("Ed", C.c_double), ("Ep", C.c_double), ("Es", C.c_double),
                ("n_r", C.c_int32), ("pad_", C.c_int32),
                ("Rr", C.POINTER(C.c_double)), ("Vr", C.POINTER(C.c_double)), ("Sr", C.POINTER(C.c_double))]


def nrl_pair_from_dict(d):
    p = NrlPair()
    p.ind_split, p.ind_overlap = int(d["ind_split"]), int(d["ind_overlap"])
    p.lambd, p.Rc, p.lden = float(d["lambd"]), float(d["Rc"]), float(d["lden"])
    for k in ("al", "bl", "cl", "dl"):
        setattr(p, k, _d4(*d[k]))
    for k in ("ellm", "fllm", "gllm", "hllm", "pllm", "qllm", "rllm", "sllm"):
        setattr(p, k, _d10(*d[k]))
    return p


def pettifor_pair_from_dict(d):
    p = PettiforPair()
    p.Es, p.Ep = float(d["Es"]), float(d["Ep"])
    for k in ("nc", "rc", "c0", "c1", "c2", "c3", "V0"):
        setattr(p, k, _d4(*d[k]))
    p.r0, p.n, p.r1, p.rm = float(d["r0"]), float(d["n"]), float(d["r1"]), float(d["rm"])
    return p


def pettifor_rep_pair_from_dict(d):
    p = PettiforRepPair()
    p.E0_TB = float(d["E0_TB"])
    p.a0 = (C.c_double * 5)(*d["a0"])
    for k in ("phi0", "d0", "m", "mc", "dc", "d1", "dm"):
        setattr(p, k, float(d[k]))
    p.c0 = _d4(*d["c0"])
    return p


def dftb_pair_array(params):
    """-> (ctypes array of DftbPair, list of the numpy arrays the pointers refer to: keep it alive during the call)"""
    import numpy as np
    arr = (DftbPair * len(params))()
    keep = []
    for i, d in enumerate(params):
        Rr = np.ascontiguousarray(d["Rr"], dtype=np.float64)
        Vr = np.asfortranarray(d["Vr"], dtype=np.float64)      # (ir, channel) column-major
        Sr = np.asfortranarray(d["Sr"], dtype=np.float64)
        assert Vr.shape == (Rr.shape[0], 10) and Sr.shape == Vr.shape
        keep += [Rr, Vr, Sr]
        p = arr[i]
        p.rcut, p.d = float(d["rcut"]), float(d["d"])
        p.Ed, p.Ep, p.Es = float(d["Ed"]), float(d["Ep"]), float(d["Es"])
        p.n_r, p.pad_ = int(Rr.shape[0]), 0
        p.Rr = Rr.ctypes.data_as(C.POINTER(C.c_double))
        p.Vr = Vr.ctypes.data_as(C.POINTER(C.c_double))
        p.Sr = Sr.ctypes.data_as(C.POINTER(C.c_double))
    return arr, keep


_pd = C.POINTER(C.c_double)
_pi = C.POINTER(C.c_int32)

# name -> (restype, argtypes); every symbol include/xtb_b200.h declares
SIGNATURES = {
    "xtb_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int]),
    "xtb_destroy": (C.c_int, [C.c_void_p]),
    "xtb_last_error": (C.c_char_p, [C.c_void_p]),
    "xtb_version": (C.c_char_p, []),
    "xtb_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "xtb_sync": (C.c_int, [C.c_void_p]),
    "xtb_set_params_nrl": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(NrlPair)]),
    "xtb_set_params_pettifor": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(PettiforPair)]),
    "xtb_set_params_dftb": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(DftbPair)]),
    "xtb_set_params_pettifor_rep": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(PettiforRepPair)]),
    "xtb_set_system": (C.c_int, [C.c_void_p, C.c_int, _pi, _pd, _pd]),
    "xtb_set_partition": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "xtb_set_block_layout": (C.c_int, [C.c_void_p, C.c_int]),
    "xtb_neighbours": (C.c_int, [C.c_void_p, C.c_double, _pi, _pi, _pd, _pd]),
    "xtb_cutoff": (C.c_int, [C.c_void_p, _pd]),
    "xtb_neighbours_compact": (C.c_int, [C.c_void_p, C.c_int, _pi, _pi, _pd, _pd, C.POINTER(C.c_int)]),
    "xtb_radial": (C.c_int, [C.c_void_p, _pd, _pd, _pd, _pd, _pd, _pd, _pd]),
    "xtb_assemble_HS": (C.c_int, [C.c_void_p, _pd, _pd]),
    "xtb_assemble_blocks": (C.c_int, [C.c_void_p]),
    "xtb_expand_dense": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _pd, _pd]),
    "xtb_set_eigen": (C.c_int, [C.c_void_p, _pd, _pd, _pd]),
    "xtb_set_eigen_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "xtb_set_eigen_device_t": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "xtb_eigensolve": (C.c_int, [C.c_void_p, _pd, _pd, _pd]),
    "xtb_get_eigenvalues": (C.c_int, [C.c_void_p, _pd]),
    "xtb_set_occupations": (C.c_int, [C.c_void_p, _pd]),
    "xtb_density": (C.c_int, [C.c_void_p, _pd, _pd]),
    "xtb_density_get": (C.c_int, [C.c_void_p, _pd, _pd]),
    "xtb_set_density_cutoff": (C.c_int, [C.c_void_p, C.c_double]),
    "xtb_density_mirror": (C.c_int, [C.c_void_p]),
    "xtb_forces": (C.c_int, [C.c_void_p, _pd]),
    "xtb_pressure": (C.c_int, [C.c_void_p, _pd]),
    "xtb_repulsive_energy": (C.c_int, [C.c_void_p, _pd, _pd]),
    "xtb_repulsive_forces": (C.c_int, [C.c_void_p, _pd]),
    "xtb_repulsive_pressure": (C.c_int, [C.c_void_p, _pd]),
    "xtb_set_params_wall": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(WallPair)]),
    "xtb_wall_energy": (C.c_int, [C.c_void_p, _pd, _pd]),
    "xtb_wall_forces": (C.c_int, [C.c_void_p, _pd]),
    "xtb_wall_pressure": (C.c_int, [C.c_void_p, _pd]),
    "xtb_set_params_short": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(ShortPair), _pd]),
    "xtb_short_energy": (C.c_int, [C.c_void_p, _pd, _pd]),
    "xtb_short_forces": (C.c_int, [C.c_void_p, _pd]),
    "xtb_short_pressure": (C.c_int, [C.c_void_p, _pd]),
    "xtb_set_params_vdw": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(VdwPair)]),
    "xtb_vdw_image_cells": (C.c_int, [C.c_void_p, C.POINTER(C.c_int), _pd]),
    "xtb_vdw_energy": (C.c_int, [C.c_void_p, _pd, _pd]),
    "xtb_vdw_forces": (C.c_int, [C.c_void_p, _pd]),
    "xtb_vdw_pressure": (C.c_int, [C.c_void_p, _pd]),
    "xtb_vdw_pressure_noalloc": (C.c_int, [C.c_void_p, C.c_void_p]),
    "xtb_set_params_coulomb": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(CoulombPair)]),
    "xtb_set_charge": (C.c_int, [C.c_void_p, C.c_double]),
    "xtb_coulomb_image_cells": (C.c_int, [C.c_void_p, C.POINTER(C.c_int)]),
    "xtb_coulomb_energy": (C.c_int, [C.c_void_p, _pd, _pd]),
    "xtb_coulomb_forces": (C.c_int, [C.c_void_p, _pd]),
    "xtb_coulomb_pressure": (C.c_int, [C.c_void_p, _pd]),
    "xtb_fermi_occupations": (C.c_int, [C.c_void_p, C.c_double, C.c_double, _pd, _pd, _pd]),
    "xtb_md_set_velocities": (C.c_int, [C.c_void_p, _pd]),
    "xtb_md_half_step": (C.c_int, [C.c_void_p, C.c_int, C.c_double, _pd, _pd, C.c_uint]),
    "xtb_md_get_state": (C.c_int, [C.c_void_p, _pd, _pd, _pd, _pd]),
    "xtb_comm_unique_id": (C.c_int, [C.c_void_p]),
    "xtb_comm_init": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "xtb_comm_destroy": (C.c_int, [C.c_void_p]),
    "xtb_exchange_blocks": (C.c_int, [C.c_void_p, C.c_int]),
    "xtb_exchange_forces": (C.c_int, [C.c_void_p, _pd]),
    "xtb_allreduce_scforce": (C.c_int, [C.c_void_p, _pd]),
    "xtb_step": (C.c_int, [C.c_void_p, C.c_uint, C.POINTER(StepIO)]),
    "xtb_save_eigen_prev": (C.c_int, [C.c_void_p]),
    "xtb_set_eigen_prev": (C.c_int, [C.c_void_p, _pd]),
    "xtb_coupling": (C.c_int, [C.c_void_p, C.c_int, _pd]),
    "xtb_device_buffer": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]),
    "xtb_counter": (C.c_longlong, [C.c_void_p, C.c_int]),
    "xtb_stage_ms": (C.c_double, [C.c_void_p, C.c_int]),
}

_lib = None


def load():
    """dlopen libxtb_b200.so (built in-tree by __graft_entry__.build()).  Fails loudly."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"CUDA extension {LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`. "
            "There is no CPU fallback for the tight-binding hot path.")
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if a declared symbol is not exported
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib
