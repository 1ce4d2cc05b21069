"""Host-side binding of the C ABI (include/xtb_b200.h) used by tests, smoke() and bench.py.

`TBContext` is a thin object over an `xtb_ctx*`; its methods are named after the reference
procedures they stand in for (same argument meaning, results in the reference's array shapes):

    get_near_neighbours            Source_files/Atomic_tools.f90:562
    Construct_M_x1 / Construct_Vij_NRL   Source_files/TB.f90:3430, TB_NRL.f90:1956   (-> radial_tables)
    construct_TB_H_NRL / construct_TB_H_Pettifor / construct_TB_H_DFTB   TB_NRL.f90:74, TB_Pettifor.f90:1427, TB_DFTB.f90:284  (-> assemble_HS)
    Construct_Aij / Construct_Aij_x_En   TB.f90:4328,4401                             (-> density)
    get_dHij_drij_NRL / dHij_s     TB_NRL.f90:2412, TB_Pettifor.f90:141               (-> forces)
    Attract_TB_Forces_Press_NRL / Attract_TB_Forces_Press  TB_NRL.f90:2742, TB_Pettifor.f90:1134 (-> pressure)
    get_Erep_s / dErdr_s / dErdr_Pressure_s   TB_Pettifor.f90:2074,2162,2634        (same names)
    Electron_ion_coupling_Mij      Nonadiabatic.f90:782                               (same name)

Errors: every non-zero ABI status raises XtbError(code, text) -- the reference records
(Save_error_details) and continues; a Python caller gets an exception instead.
Nothing here computes: all arithmetic happens in libxtb_b200.so on the GPU.
"""
import ctypes as C

import numpy as np

from . import _abi


class XtbError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"xtb error {code}: {text}")
        self.code = code


def _dp(a):
    return None if a is None else a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return None if a is None else a.ctypes.data_as(C.POINTER(C.c_int32))


class TBContext:
    def __init__(self, device=0):
        self.lib = _abi.load()
        h = C.c_void_p()
        rc = self.lib.xtb_create(C.byref(h), int(device))
        if rc != 0:
            raise XtbError(rc, "xtb_create failed (no CUDA device? there is no CPU fallback)")
        self.h = h
        self.na = 0
        self.n_orb = 0
        self.tb = None

    # -- plumbing ---------------------------------------------------------------------------
    def _ck(self, rc):
        if rc != 0:
            raise XtbError(rc, self.lib.xtb_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            self.lib.xtb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def counter(self, which):
        return int(self.lib.xtb_counter(self.h, which))

    def stage_ms(self, stage):
        return float(self.lib.xtb_stage_ms(self.h, stage))

    def set_stream(self, cuda_stream_ptr):
        self._ck(self.lib.xtb_set_stream(self.h, C.c_void_p(cuda_stream_ptr)))

    def sync(self):
        self._ck(self.lib.xtb_sync(self.h))

    def device_buffer(self, which):
        p = C.c_void_p(); n = C.c_size_t()
        self._ck(self.lib.xtb_device_buffer(self.h, which, C.byref(p), C.byref(n)))
        return p.value, n.value

    # -- inputs -----------------------------------------------------------------------------
    def set_workload(self, wl):
        """parameters + system of a workloads.py dict"""
        nk = wl["n_kinds"]
        if wl["tb"] == "nrl":
            arr = (_abi.NrlPair * (nk * nk))(*[_abi.nrl_pair_from_dict(d) for d in wl["params"]])
            self._ck(self.lib.xtb_set_params_nrl(self.h, nk, arr))
            self.n_orb = 9
        elif wl["tb"] == "pettifor":
            arr = (_abi.PettiforPair * (nk * nk))(*[_abi.pettifor_pair_from_dict(d) for d in wl["params"]])
            self._ck(self.lib.xtb_set_params_pettifor(self.h, nk, arr))
            self.n_orb = 4
            if wl.get("rep_params"):
                rep = (_abi.PettiforRepPair * (nk * nk))(*[_abi.pettifor_rep_pair_from_dict(d) for d in wl["rep_params"]])
                self._ck(self.lib.xtb_set_params_pettifor_rep(self.h, nk, rep))
        elif wl["tb"] == "dftb":
            arr, keep = _abi.dftb_pair_array(wl["params"])     # tables are copied by the library during the call
            self._ck(self.lib.xtb_set_params_dftb(self.h, nk, int(wl["n_orb"]), arr))
            self.n_orb = int(wl["n_orb"])
        else:
            raise ValueError(wl["tb"])
        self.tb = wl["tb"]
        if wl.get("wall_params"):
            self.set_params_wall(wl["wall_params"])
        if wl.get("short_params"):
            self.set_params_short(wl["short_params"], wl["Z"])
        if wl.get("vdw_params"):
            self.set_params_vdw(wl["vdw_kind"], wl["vdw_params"])
        if wl.get("coulomb_params"):
            self.set_params_coulomb(wl["coulomb_params"])
            self.set_charge(wl.get("Q", 0.0))
        self.set_system(wl["koa"], wl["S"], wl["supce"])

    def set_params_coulomb(self, params):
        """Scell%TB_Coul(:,:) of class TB_Coulomb_cut: list of dict(dm, dd[, k]) (k defaults to the reader's constant,
        Read_input_data.f90:1907), (koa1, koa2) column-major; None / [] removes the term"""
        if not params:
            self._ck(self.lib.xtb_set_params_coulomb(self.h, 0, None))
            return
        nk = int(round(len(params) ** 0.5))
        arr = (_abi.CoulombPair * len(params))(*[_abi.CoulombPair(float(d.get("k", _abi.COULOMB_K)), float(d["dm"]), float(d["dd"])) for d in params])
        self._ck(self.lib.xtb_set_params_coulomb(self.h, nk, arr))

    def set_charge(self, Q):
        """Scell%Q: mean unbalanced charge per atom [e]"""
        self._ck(self.lib.xtb_set_charge(self.h, float(Q)))

    def set_params_vdw(self, kind, params):
        """Scell%TB_Waals(:,:): kind 'Girifalco' | 'LJ' | 'ILJ' (or XTB_VDW_*), list of dicts with the members of that class
        (readers.read_vdW_TB), (koa1, koa2) column-major; None / [] removes the term"""
        if not params:
            self._ck(self.lib.xtb_set_params_vdw(self.h, 0, 0, None))
            return
        kind = {"Girifalco": _abi.VDW_GIRIFALCO, "LJ": _abi.VDW_LJ_CUT, "ILJ": _abi.VDW_ILJ_CUT}.get(kind, kind)
        nk = int(round(len(params) ** 0.5))
        arr = (_abi.VdwPair * len(params))(*[_abi.vdw_pair_from_dict(d) for d in params])
        self._ck(self.lib.xtb_set_params_vdw(self.h, int(kind), nk, arr))

    def set_params_short(self, params, Z):
        """Scell%TB_Expwall(:,:) of class TB_Short_Rep ('General'): list of dicts (readers.read_Short_Rep_TB), (koa1, koa2)
        column-major, Z = atomic numbers per kind; replaces a simple wall; None / [] removes it"""
        if not params:
            self._ck(self.lib.xtb_set_params_short(self.h, 0, None, None))
            return
        nk = int(round(len(params) ** 0.5))
        arr, keep = _abi.short_pair_array(params)
        Zs = np.ascontiguousarray(Z, dtype=np.float64)
        self._ck(self.lib.xtb_set_params_short(self.h, nk, arr, _dp(Zs)))

    def set_params_wall(self, params):
        """Scell%TB_Expwall(:,:) of class TB_Exp_wall_simple: list of dict(C, r0, d0, dd), (koa1, koa2) column-major; None / [] removes it"""
        if not params:
            self._ck(self.lib.xtb_set_params_wall(self.h, 0, None))
            return
        nk = int(round(len(params) ** 0.5))
        arr = (_abi.WallPair * len(params))(*[_abi.WallPair(float(d["C"]), float(d["r0"]), float(d["d0"]), float(d["dd"])) for d in params])
        self._ck(self.lib.xtb_set_params_wall(self.h, nk, arr))

    def set_system(self, koa, S, supce):
        koa = np.ascontiguousarray(koa, dtype=np.int32)
        S = np.ascontiguousarray(S, dtype=np.float64)
        h = np.asfortranarray(supce, dtype=np.float64)
        self.na = int(koa.shape[0])
        self._ck(self.lib.xtb_set_system(self.h, self.na, _ip(koa), _dp(S), _dp(h)))

    def set_partition(self, first, last):
        self._ck(self.lib.xtb_set_partition(self.h, int(first), int(last)))

    def set_block_layout(self, parts):
        """block stores in `parts` equal-capacity segments (one per contiguous atom range of the row partition): the
        ranges are then exchanged with one in-place all-gather (multigpu.allgather_blocks).  Before the neighbour list."""
        self._ck(self.lib.xtb_set_block_layout(self.h, int(parts)))

    # -- S1 ---------------------------------------------------------------------------------
    def get_near_neighbours(self, rm=0.0, dense=False):
        """returns dict(size, list, dist, dist_s); dense=True gives the reference shapes
        Near_neighbor_size(na), _list(na,na) (1-based, 0 = empty), _dist(na,na,4), _dist_s(na,na,3);
        dense=False the compact (na,max_nn[,c]) arrays with 0-based atom indices."""
        na = self.na
        if dense:
            size = np.zeros(na, dtype=np.int32)
            lst = np.zeros((na, na), dtype=np.int32, order="F")
            dist = np.zeros((na, na, 4), order="F")
            dist_s = np.zeros((na, na, 3), order="F")
            self._ck(self.lib.xtb_neighbours(self.h, float(rm), _ip(size), _ip(lst), _dp(dist), _dp(dist_s)))
            return dict(size=size, list=lst, dist=dist, dist_s=dist_s)
        self._ck(self.lib.xtb_neighbours(self.h, float(rm), None, None, None, None))
        mx = max(self.counter(_abi.CNT_MAXNN), 1)
        size = np.zeros(na, dtype=np.int32)
        lst = np.zeros((na, mx), dtype=np.int32)
        dist = np.zeros((na, mx, 4)); dist_s = np.zeros((na, mx, 3))
        m = C.c_int()
        self._ck(self.lib.xtb_neighbours_compact(self.h, mx, _ip(size), _ip(lst), _dp(dist), _dp(dist_s), C.byref(m)))
        return dict(size=size, list=lst, dist=dist, dist_s=dist_s, max_nn=m.value)

    def get_near_neighbours_device(self, rm=0.0):
        """S1 with every output kept on the device (NULL host pointers)."""
        self._ck(self.lib.xtb_neighbours(self.h, float(rm), None, None, None, None))

    # -- S2 ---------------------------------------------------------------------------------
    def radial_tables(self):
        """Construct_M_x1 + Construct_Vij_NRL / Construct_Vij_DFTB in the reference's dense shapes (small systems)."""
        na = self.na
        nv = 10 if self.tb == "nrl" else {1: 1, 4: 4, 9: 10}[self.n_orb]
        out = {k: np.zeros((na, na, nv), order="F") for k in ("M_Vij", "M_dVij", "M_SVij", "M_dSVij")}
        out.update({k: np.zeros((3, na, na), order="F") for k in ("M_lmn", "M_x1", "M_xrr")})
        self._ck(self.lib.xtb_radial(self.h, _dp(out["M_Vij"]), _dp(out["M_dVij"]), _dp(out["M_SVij"]), _dp(out["M_dSVij"]),
                                     _dp(out["M_lmn"]), _dp(out["M_x1"]), _dp(out["M_xrr"])))
        return out

    def assemble_HS(self, to_host=True):
        N = self.na * self.n_orb
        if not to_host:
            self._ck(self.lib.xtb_assemble_HS(self.h, None, None))
            return None, None
        H = np.zeros((N, N), order="F")
        S = np.zeros((N, N), order="F") if self.tb != "pettifor" else None
        self._ck(self.lib.xtb_assemble_HS(self.h, _dp(H), _dp(S)))
        return H, S

    def assemble_blocks(self):
        """S2 for the owned atoms, block stores only (multi-GPU: allgather the blocks, then expand_dense)."""
        self._ck(self.lib.xtb_assemble_blocks(self.h))

    def expand_dense(self, first=1, last=None, to_host=False):
        last = self.na if last is None else last
        if not to_host:
            self._ck(self.lib.xtb_expand_dense(self.h, int(first), int(last), None, None))
            return None, None
        N = self.na * self.n_orb
        H = np.zeros((N, N), order="F")
        S = np.zeros((N, N), order="F") if self.tb != "pettifor" else None
        self._ck(self.lib.xtb_expand_dense(self.h, int(first), int(last), _dp(H), _dp(S)))
        return H, S

    # -- eigensolve boundary ------------------------------------------------------------------
    def set_eigen(self, Ha, Ei, fe):
        Ha = np.asfortranarray(Ha, dtype=np.float64)
        self._keep = (Ha, np.ascontiguousarray(Ei, dtype=np.float64), np.ascontiguousarray(fe, dtype=np.float64))
        self._ck(self.lib.xtb_set_eigen(self.h, _dp(self._keep[0]), _dp(self._keep[1]), _dp(self._keep[2])))

    def set_eigen_device(self, dHa, dEi, dfe):
        """device pointers (ints), e.g. torch tensors' data_ptr(); zero copy"""
        self._ck(self.lib.xtb_set_eigen_device(self.h, C.c_void_p(dHa), C.c_void_p(dEi), C.c_void_p(dfe)))

    def set_eigen_device_t(self, dHaT, dEi, dfe):
        """state-minor device eigenvectors HaT[row][state] (device pointers as ints); zero copy, no re-tiling"""
        self._ck(self.lib.xtb_set_eigen_device_t(self.h, C.c_void_p(dHaT), C.c_void_p(dEi), C.c_void_p(dfe)))

    def eigensolve(self, want_Ha=False):
        """on-device Loewdin + Dsyevd + back-transform (TB_NRL.f90:985-1060,402-421); returns (Ei, eigen_S[, Ha])"""
        N = self.na * self.n_orb
        Ei = np.zeros(N); wS = np.zeros(N)
        Ha = np.zeros((N, N), order="F") if want_Ha else None
        self._ck(self.lib.xtb_eigensolve(self.h, _dp(Ei), _dp(wS), _dp(Ha)))
        return (Ei, wS, Ha) if want_Ha else (Ei, wS)

    def get_eigenvalues(self):
        Ei = np.zeros(self.na * self.n_orb)
        self._ck(self.lib.xtb_get_eigenvalues(self.h, _dp(Ei)))
        return Ei

    def set_occupations(self, fe):
        fe = np.ascontiguousarray(fe, dtype=np.float64)
        self._ck(self.lib.xtb_set_occupations(self.h, _dp(fe)))

    def set_density_cutoff(self, f_cut):
        self._ck(self.lib.xtb_set_density_cutoff(self.h, float(f_cut)))

    # -- S3 ---------------------------------------------------------------------------------
    def density(self, dense=False):
        if not dense:
            self._ck(self.lib.xtb_density(self.h, None, None))
            return None, None
        N = self.na * self.n_orb
        A = np.zeros((N, N), order="F"); AE = np.zeros((N, N), order="F")
        self._ck(self.lib.xtb_density(self.h, _dp(A), _dp(AE)))
        return A, AE

    def density_mirror(self):
        self._ck(self.lib.xtb_density_mirror(self.h))

    # -- S4 / S5 ------------------------------------------------------------------------------
    def forces(self):
        F = np.zeros((self.na, 3))
        self._ck(self.lib.xtb_forces(self.h, _dp(F)))
        return F

    def forces_device(self):
        """S4 with the result kept on the device (XTB_BUF_FORCES)."""
        self._ck(self.lib.xtb_forces(self.h, None))

    def pressure_device(self):
        self._ck(self.lib.xtb_pressure(self.h, None))

    def pressure(self):
        att = np.zeros(9)
        self._ck(self.lib.xtb_pressure(self.h, _dp(att)))
        return att.reshape(3, 3, order="F")

    # -- repulsive pair terms of the Pettifor branch ----------------------
    def get_Erep_s(self):
        """get_Erep_s, TB_Pettifor.f90:2074 -> (E_rep [eV], MDAtoms(:)%Epot)"""
        E = np.zeros(1); Epot = np.zeros(self.na)
        self._ck(self.lib.xtb_repulsive_energy(self.h, _dp(E), _dp(Epot)))
        return float(E[0]), Epot

    def dErdr_s(self):
        """dErdr_s, TB_Pettifor.f90:2162 -> forces%rep (na,3), d/ds units [eV]"""
        F = np.zeros((self.na, 3))
        self._ck(self.lib.xtb_repulsive_forces(self.h, _dp(F)))
        return F

    def dErdr_Pressure_s(self):
        """dErdr_Pressure_s, TB_Pettifor.f90:2634 -> SCforce%rep(3,3)"""
        rep = np.zeros(9)
        self._ck(self.lib.xtb_repulsive_pressure(self.h, _dp(rep)))
        return rep.reshape(3, 3, order="F")

    # -- short-range wall (second half) ------------------------------------
    def get_Exp_wall_s(self):
        """get_Exp_wall_s, Exponential_wall.f90:990 -> (E_wall [eV], per-atom halves added to MDAtoms(:)%Epot)"""
        E = np.zeros(1); Epot = np.zeros(self.na)
        self._ck(self.lib.xtb_wall_energy(self.h, _dp(E), _dp(Epot)))
        return float(E[0]), Epot

    def d_Exp_wall_pot_s(self):
        """d_Exp_wall_pot_s, Exponential_wall.f90:1200 -> the term added to forces%rep, (na,3), d/ds units [eV]"""
        F = np.zeros((self.na, 3))
        self._ck(self.lib.xtb_wall_forces(self.h, _dp(F)))
        return F

    def d_Exp_wall_Pressure_s(self):
        """d_Exp_wall_Pressure_s, Exponential_wall.f90:1334 -> the term added to SCforce%rep(3,3)"""
        P = np.zeros(9)
        self._ck(self.lib.xtb_wall_pressure(self.h, _dp(P)))
        return P.reshape(3, 3, order="F")

    # -- the 'General' short-range class in the slot of the wall ----------------------------------
    def get_short_range_rep_s(self):
        """get_short_range_rep_s, Exponential_wall.f90:48 -> (E [eV], per-atom halves added to MDAtoms(:)%Epot)"""
        E = np.zeros(1); Epot = np.zeros(self.na)
        self._ck(self.lib.xtb_short_energy(self.h, _dp(E), _dp(Epot)))
        return float(E[0]), Epot

    def d_Short_range_pot_s(self):
        """d_Short_range_pot_s, Exponential_wall.f90:519 -> the term added to forces%rep, (na,3)"""
        F = np.zeros((self.na, 3))
        self._ck(self.lib.xtb_short_forces(self.h, _dp(F)))
        return F

    def d_Short_range_Pressure_s(self):
        """d_Short_range_Pressure_s, Exponential_wall.f90:664 -> the term added to SCforce%rep(3,3)"""
        P = np.zeros(9)
        self._ck(self.lib.xtb_short_pressure(self.h, _dp(P)))
        return P.reshape(3, 3, order="F")

    # -- van der Waals term (third part) -------------------------------------
    def get_mirror_cell_num(self):
        """get_mirror_cell_num, Van_der_Waals.f90:132 -> ((Nx, Ny, Nz), R_cut [A])"""
        N = (C.c_int * 3)(); R = np.zeros(1)
        self._ck(self.lib.xtb_vdw_image_cells(self.h, N, _dp(R)))
        return (int(N[0]), int(N[1]), int(N[2])), float(R[0])

    def get_vdW_s(self):
        """get_vdW_s, Van_der_Waals.f90:158 -> (E_vdW [eV], per-atom halves added to MDAtoms(:)%Epot)"""
        E = np.zeros(1); Epot = np.zeros(self.na)
        self._ck(self.lib.xtb_vdw_energy(self.h, _dp(E), _dp(Epot)))
        return float(E[0]), Epot

    def vdW_forces(self):
        """vdW_forces -> Construct_B + d_Forces_s, TB.f90:3599,3638 -> the term added to forces%rep, (na,3), d/ds units [eV]"""
        F = np.zeros((self.na, 3))
        self._ck(self.lib.xtb_vdw_forces(self.h, _dp(F)))
        return F

    def vdW_Pressure(self):
        """vdW_forces -> d_Forces_Pressure, TB.f90:3785 -> the term added to SCforce%rep(3,3)"""
        P = np.zeros(9)
        self._ck(self.lib.xtb_vdw_pressure(self.h, _dp(P)))
        return P.reshape(3, 3, order="F")

    # -- Coulomb term of a charged supercell (fourth part) -------------------
    def get_mirror_cell_num_C(self):
        """get_mirror_cell_num_C, Coulomb.f90:435 -> (Nx, Ny, Nz)"""
        N = (C.c_int * 3)()
        self._ck(self.lib.xtb_coulomb_image_cells(self.h, N))
        return int(N[0]), int(N[1]), int(N[2])

    def get_Coulomb_s(self):
        """get_Coulomb_s, Coulomb.f90:454 -> (E_coul [eV] (not halved, as written), what is added to MDAtoms(:)%Epot)"""
        E = np.zeros(1); Epot = np.zeros(self.na)
        self._ck(self.lib.xtb_coulomb_energy(self.h, _dp(E), _dp(Epot)))
        return float(E[0]), Epot

    def Coulomb_forces(self):
        """Coulomb_forces -> Construct_B_C + d_Forces_s, TB.f90:3560,3638 -> the term added to forces%rep, (na,3)"""
        F = np.zeros((self.na, 3))
        self._ck(self.lib.xtb_coulomb_forces(self.h, _dp(F)))
        return F

    def Coulomb_Pressure(self):
        """Coulomb_forces -> d_Forces_Pressure, TB.f90:3785 -> the term added to SCforce%rep(3,3)"""
        P = np.zeros(9)
        self._ck(self.lib.xtb_coulomb_pressure(self.h, _dp(P)))
        return P.reshape(3, 3, order="F")

    # -- Fermi occupations, velocity-Verlet halves, the fused step (step.cu) ----------------------
    def fermi_occupations(self, n_electrons, Te_eV):
        """Electron_Fixed_Te + set_Fermi (Electron_tools.f90:2660,1913) on the device -> (fe, mu, sum f E)"""
        fe = np.zeros(self.na * self.n_orb); mu = np.zeros(1); eb = np.zeros(1)
        self._ck(self.lib.xtb_fermi_occupations(self.h, float(n_electrons), float(Te_eV), _dp(fe), _dp(mu), _dp(eb)))
        return fe, float(mu[0]), float(eb[0])

    def md_set_velocities(self, SV=None):
        SV = None if SV is None else np.ascontiguousarray(SV, dtype=np.float64)
        self._ck(self.lib.xtb_md_set_velocities(self.h, _dp(SV)))

    def make_time_step_atoms(self, ind, dt_fs, mass_kg, Vsupce=None, force_terms=_abi.FORCE_ATT):
        """save_last_timestep + make_time_step_atoms(ind) (Atomic_tools.f90:1177,1873,1897), velocity-Verlet half on the device"""
        m = np.ascontiguousarray(mass_kg, dtype=np.float64)
        V = None if Vsupce is None else np.asfortranarray(Vsupce, dtype=np.float64)
        self._ck(self.lib.xtb_md_half_step(self.h, int(ind), float(dt_fs), _dp(m), _dp(V), int(force_terms)))

    def md_get_state(self, mass_kg=None):
        """-> (S (na,3), SV (na,3), E_kin [eV] or None)"""
        S = np.zeros((self.na, 3)); SV = np.zeros((self.na, 3))
        ek = np.zeros(1) if mass_kg is not None else None
        m = None if mass_kg is None else np.ascontiguousarray(mass_kg, dtype=np.float64)
        self._ck(self.lib.xtb_md_get_state(self.h, _dp(S), _dp(SV), _dp(ek), _dp(m)))
        return S, SV, (None if ek is None else float(ek[0]))

    def step(self, flags=_abi.STEP_ALL, S=None, supce=None, fe=None, n_electrons=0.0, Te_eV=0.0, want=(), graph=False):
        """xtb_step: the selected stages as one call.  want: names of host outputs to fetch, any of
        Ei, fe_out, mu, E_band, F_att, SCforce_att, E_rep, F_rep, SCforce_rep, E_wall, F_wall, SCforce_wall,
        E_vdw, F_vdw, SCforce_vdw, E_coul, F_coul, SCforce_coul -> dict"""
        N = self.na * self.n_orb
        io = _abi.StepIO()
        keep = []
        def inp(a, order="C"):
            if a is None:
                return None
            a = np.asfortranarray(a, dtype=np.float64) if order == "F" else np.ascontiguousarray(a, dtype=np.float64)
            keep.append(a)
            return _dp(a)
        io.S = inp(S); io.supce = inp(supce, "F"); io.fe = inp(fe)
        io.n_electrons = float(n_electrons); io.Te_eV = float(Te_eV)
        shapes = {"Ei": (N,), "fe_out": (N,), "mu": (1,), "E_band": (1,), "F_att": (self.na, 3), "SCforce_att": (9,),
                  "E_rep": (1,), "F_rep": (self.na, 3), "SCforce_rep": (9,), "E_wall": (1,), "F_wall": (self.na, 3), "SCforce_wall": (9,),
                  "E_vdw": (1,), "F_vdw": (self.na, 3), "SCforce_vdw": (9,),
                  "E_coul": (1,), "F_coul": (self.na, 3), "SCforce_coul": (9,)}
        out = {}
        for k in want:
            out[k] = np.zeros(shapes[k])
            setattr(io, k, _dp(out[k]))
        self._ck(self.lib.xtb_step(self.h, int(flags) | (_abi.STEP_GRAPH if graph else 0), C.byref(io)))
        for k in out:
            if k.startswith("SCforce"):
                out[k] = out[k].reshape(3, 3, order="F")
            elif shapes[k] == (1,):
                out[k] = float(out[k][0])
        return out

    # -- multi-GPU inside the library (comm.cu) -------------------------------------------------
    @staticmethod
    def comm_unique_id():
        buf = C.create_string_buffer(_abi.UNIQUE_ID_BYTES)
        rc = _abi.load().xtb_comm_unique_id(buf)
        if rc != 0:
            raise XtbError(rc, "xtb_comm_unique_id failed (NCCL not found?)")
        return buf.raw

    def comm_init(self, unique_id, rank, nranks):
        self._ck(self.lib.xtb_comm_init(self.h, C.c_char_p(unique_id), int(rank), int(nranks)))

    def comm_destroy(self):
        self._ck(self.lib.xtb_comm_destroy(self.h))

    def exchange_blocks(self, which):
        self._ck(self.lib.xtb_exchange_blocks(self.h, int(which)))

    def exchange_forces(self, to_host=False):
        F = np.zeros((self.na, 3)) if to_host else None
        self._ck(self.lib.xtb_exchange_forces(self.h, _dp(F)))
        return F

    def allreduce_scforce(self):
        P = np.zeros(9)
        self._ck(self.lib.xtb_allreduce_scforce(self.h, _dp(P)))
        return P.reshape(3, 3, order="F")

    # -- nonadiabatic coupling matrix -------------------------------------
    def save_eigen_prev(self):
        """Scell%Ha0 = Scell%Ha (TB_NRL.f90:88), device to device"""
        self._ck(self.lib.xtb_save_eigen_prev(self.h))

    def set_eigen_prev(self, Ha0):
        Ha0 = np.asfortranarray(Ha0, dtype=np.float64)
        self._ck(self.lib.xtb_set_eigen_prev(self.h, _dp(Ha0)))

    def Electron_ion_coupling_Mij(self, kind_M=0, to_host=True):
        """Electron_ion_coupling_Mij(wr, Ha, Ha0, Mij, kind_M, Sij), Nonadiabatic.f90:782: Ha = current eigenvectors of the
        context, Ha0 = save_eigen_prev / set_eigen_prev, Sij = the assembled overlap (absent for the orthogonal branch)."""
        if not to_host:
            self._ck(self.lib.xtb_coupling(self.h, int(kind_M), None))
            return None
        N = self.na * self.n_orb
        M = np.zeros((N, N), order="F")
        self._ck(self.lib.xtb_coupling(self.h, int(kind_M), _dp(M)))
        return M

    # -- the reference's own procedure names for the same calls (the parametrisation arms of TB.f90:247-317,459-487) ----
    def _arm(self, *kinds):
        if self.tb not in kinds:
            raise XtbError(-1, f"this context holds a {self.tb!r} parametrisation; the procedure belongs to {kinds}")

    def Construct_Vij_NRL(self):
        """Construct_Vij_NRL, TB_NRL.f90:1956 -> dict(M_Vij, M_dVij, M_SVij, M_dSVij, M_lmn, M_x1, M_xrr), dense shapes"""
        self._arm("nrl")
        return self.radial_tables()

    def Construct_Vij_DFTB(self):
        """Construct_Vij_DFTB, TB_DFTB.f90:56"""
        self._arm("dftb")
        return self.radial_tables()

    def construct_TB_H_NRL(self):
        """construct_TB_H_NRL -> Hamil_tot_NRL assembly part, TB_NRL.f90:74,226-397 -> (H_non, Sij); the eigensolve stays outside"""
        self._arm("nrl")
        return self.assemble_HS()

    def construct_TB_H_Pettifor(self):
        """construct_TB_H_Pettifor -> Hamil_tot, TB_Pettifor.f90:1427,1658 -> H (orthogonal basis: no overlap matrix)"""
        self._arm("pettifor")
        return self.assemble_HS()[0]

    def construct_TB_H_DFTB(self):
        """construct_TB_H_DFTB -> Hamil_tot_DFTB, TB_DFTB.f90:284,319 (non-SCC arm) -> (H_non, Sij)"""
        self._arm("dftb")
        return self.assemble_HS()

    def Construct_Aij(self):
        """Construct_Aij, TB.f90:4328 -> Aij (N,N), zero outside the neighbour pattern"""
        return self.density(dense=True)[0]

    def Construct_Aij_x_En(self):
        """Construct_Aij_x_En, TB.f90:4401 -> Aij_x_Ei (N,N)"""
        return self.density(dense=True)[1]

    def get_dHij_drij_NRL(self):
        """get_dHij_drij_NRL, TB_NRL.f90:2412 -> MDatoms(:)%forces%att as (na, 3), dE/ds [eV]"""
        self._arm("nrl")
        return self.forces()

    def dHij_s(self):
        """dHij_s, TB_Pettifor.f90:141"""
        self._arm("pettifor")
        return self.forces()

    def get_dHij_drij_DFTB(self):
        """get_dHij_drij_DFTB, TB_DFTB.f90:781"""
        self._arm("dftb")
        return self.forces()

    def Attract_TB_Forces_Press_NRL(self):
        """Attract_TB_Forces_Press_NRL, TB_NRL.f90:2742 -> SCforce%att (3,3)"""
        self._arm("nrl")
        return self.pressure()

    def Attract_TB_Forces_Press(self):
        """Attract_TB_Forces_Press, TB_Pettifor.f90:1134"""
        self._arm("pettifor")
        return self.pressure()

    def Attract_TB_Forces_Press_DFTB(self):
        """Attract_TB_Forces_Press_DFTB, TB_DFTB.f90:1135"""
        self._arm("dftb")
        return self.pressure()
