/*
 * xtb_b200.h -- C ABI of the B200-native tight-binding hot path (XTANT-3 graft).
 *
 * This is the drop-in boundary for the per-MD-step path
 *     neighbour list -> radial tables -> H/S block assembly -> [eigensolve, outside]
 *     -> rho = C f C^T, rho_E = C (f E) C^T -> Hellmann-Feynman forces (+ supercell force)
 * of the reference Fortran code.  The reference has no FFI of its own (plain module
 * procedures on derived types), so each entry point below names the reference
 * procedure it replaces (file:line under Source_files/ of N-Medvedev/XTANT-3).
 * The Fortran-side ISO_C_BINDING stub that forwards to these symbols is
 * xtant-3_b200/fortran/xtant_b200_shim.f90 and is described in INTEGRATION.md.
 *
 * Conventions
 *  - every array is Fortran column-major, FP64 (real(8)) or int32 (default integer);
 *  - atom / kind indices in arguments and list outputs are 1-based, as in the reference;
 *  - a NULL output pointer means "keep the result on the device only";
 *  - host arrays are owned by the caller; nothing is retained past the call;
 *  - return value 0 = ok, otherwise an XTB_E* code; text via xtb_last_error()
 *    (the Fortran shim maps non-zero to Save_error_details(Err, 6, text), Objects.f90:1309);
 *  - one caller thread per context (the reference calls this path from the master
 *    thread only, TB.f90:227-272); contexts are independent.
 *  - there is NO CPU fallback: without a CUDA device xtb_create fails with XTB_ENODEV.
 */
#ifndef XTB_B200_H
#define XTB_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define XTB_OK        0
#define XTB_EINVAL    1   /* bad argument / call order */
#define XTB_ENODEV    2   /* no usable CUDA device */
#define XTB_ECUDA     3   /* CUDA runtime error (text has the detail) */
#define XTB_ENOMEM    4   /* device allocation failed */
#define XTB_ECAPACITY 5   /* neighbour capacity exceeded */
#define XTB_ESTATE    6   /* stage called before its inputs exist */
#define XTB_ECOMM     7   /* NCCL not available / collective failed */

/* TB parametrisation kinds; mirrors the `select type` arms of TB.f90:247-317,459-487 */
#define XTB_TB_NONE     0
#define XTB_TB_NRL      1   /* type is (TB_H_NRL)      spd, non-orthogonal, 9 orbitals/atom */
#define XTB_TB_PETTIFOR 2   /* type is (TB_H_Pettifor) sp3, orthogonal,     4 orbitals/atom */
#define XTB_TB_DFTB     3   /* type is (TB_H_DFTB)     s/sp3/sp3d5 tables,  1/4/9 orbitals/atom */

/* type(TB_H_NRL), Objects.f90:190-200; file order of the 10 channels is
 * ss,sp,pp_sigma,pp_pi,sd,pd_sigma,pd_pi,dd_sigma,dd_pi,dd_delta (Read_input_data.f90:4075-4150) */
typedef struct xtb_nrl_pair {
    int32_t ind_split;     /* 1: eg on-site uses set 4, else set 3 */
    int32_t ind_overlap;   /* 1: overlap radial form Eq.(23) with delta(l,l') */
    double  lambd;         /* embedded-density exponent */
    double  Rc, lden;      /* cut-off radius [bohr], screening length */
    double  al[4], bl[4], cl[4], dl[4];
    double  ellm[10], fllm[10], gllm[10], hllm[10];
    double  pllm[10], qllm[10], rllm[10], sllm[10];
} xtb_nrl_pair;

/* type(TB_H_Pettifor), Objects.f90:170-174 */
typedef struct xtb_pettifor_pair {
    double Es, Ep;
    double nc[4], rc[4], c0[4], c1[4], c2[4], c3[4], V0[4];
    double r0, n, r1, rm;
} xtb_pettifor_pair;

/* type(TB_Rep_Pettifor), Objects.f90:95-100 (repulsive part of the Pettifor-type parametrisation;
 * file order read_Pettifor_TB_repulsive, Read_input_data.f90:2908-3056) */
typedef struct xtb_pettifor_rep_pair {
    double E0_TB;            /* [eV] constant shift per atom */
    double a0[5];            /* polynomial f(x) = a0 + a1 x + ... + a4 x^4 */
    double phi0, d0, m, mc, dc, d1, dm;
    double c0[4];            /* cubic tail of phi(r) on d1 < r <= dm */
} xtb_pettifor_rep_pair;

/* type(TB_H_DFTB), Objects.f90:202-222 -- tables flattened, one entry per (KOA1,KOA2) */
typedef struct xtb_dftb_pair {
    double  rcut, d;          /* Fermi-like cut-off [A] */
    double  Ed, Ep, Es;       /* on-site energies [eV] (of kind KOA1) */
    int32_t n_r;              /* radial grid points */
    int32_t pad_;
    const double* Rr;         /* [n_r]     radial grid [A], uniform, Rr[0] = dr */
    const double* Vr;         /* [n_r*10]  column-major (ir, channel) [eV] */
    const double* Sr;         /* [n_r*10]  overlaps */
} xtb_dftb_pair;

/* type(TB_Exp_wall_simple), Objects.f90:458-466 (file order read_Exponential_wall_TB, Read_input_data.f90:2313-2356):
 * Pot(r) = C exp(1 / (r - r0)) / (1 + exp((r - d0) / dd)), zero beyond d0 + 10 dd */
typedef struct xtb_wall_pair {
    double C;     /* [eV] */
    double r0;    /* [A] wall position */
    double d0;    /* [A] cut-off radius */
    double dd;    /* [A] cut-off smoothing width */
} xtb_wall_pair;

/* type(TB_Short_Rep), Objects.f90:437-454, the 'General' class of <El>_<El>_TB_short.txt / _TB_wall.txt (reader read_Short_Rep_TB,
 * Read_input_data.f90:2110-2312): the sum of up to six functions of r under one Fermi cut-off,
 *   Pot(r) = [ C exp(1/(r - r0))  +  Phi exp(-(r - r0)/a)  +  sum_i Phi_i (r/r0_i)^m_i  +  ZBL(Z1,Z2,r)  +  NLH(Z1,Z2,r)  +  Tab(r) ] / (1 + exp((r - d0)/dd)),
 * zero for r >= d0 + 10 dd (Shortrange_pot, Exponential_wall.f90:130-157).  (f_cut_inv is read by the reference but never used.)
 * Tab: grid tab_R[n_tab] (increasing), tab_E[n_tab]; tab_spline = 0: linear interpolation (linear_interpolation, Little_subroutines.f90:1031),
 * tab_spline = 1: cubic a + b x + c x^2 + d x^3 with the coefficients of make_cubic_splines (Algebra_tools.f90:115-190), all [n_tab];
 * the arrays are copied by xtb_set_params_short during the call.  NLH: a[3], b[3] of the pair from nlh_coeffs.dat (get_NLH_coefficients). */
#define XTB_SHORT_MAX_POW 8
typedef struct xtb_short_pair {
    double  d0, dd;                                   /* f_cut [A] */
    int32_t use_exp, use_inv_exp, use_zbl, use_nlh;
    double  exp_Phi, exp_r0, exp_a;                   /* f_exp */
    double  inv_C, inv_r0;                            /* f_inv_exp */
    int32_t n_pow, n_tab;
    double  pow_Phi[XTB_SHORT_MAX_POW], pow_r0[XTB_SHORT_MAX_POW], pow_m[XTB_SHORT_MAX_POW];   /* f_pow(1:n_pow) */
    double  nlh_a[3], nlh_b[3];                       /* f_NLH */
    int32_t tab_spline, pad_;
    const double *tab_R, *tab_E, *tab_a, *tab_b, *tab_c, *tab_d;      /* f_tab; a..d may be NULL when tab_spline = 0 */
} xtb_short_pair;

/* class(TB_vdW), Objects.f90:43-46 (parent: d0_cut, dd_cut) with its three usable extensions in one record:
 *   XTB_VDW_GIRIFALCO  type(TB_vdW_Girifalco) :302-328, file order read_vdW_Girifalco_TB (Read_input_data.f90:2392-2574):
 *                      C12 / r^12 - C6 / r^6 for dsm <= r <= dm, cubic a r^3 + b r^2 + c r + d on (dm, d_cut], quintic
 *                      as r^5 + bs r^4 + cs r^3 + ds r^2 + es r + fs on [ds_cut, dsm), zero outside [ds_cut, d_cut];
 *                      the reader sets d0_cut = d_cut, dd_cut = 0 (:1799-1800)
 *   XTB_VDW_LJ_CUT     type(TB_vdW_LJ_cut) :336-343: eps ((r0/r)^n - 2) (r0/r)^n times the short- and long-range Fermi cut-offs
 *   XTB_VDW_ILJ_CUT    type(TB_vdW_ILJ_cut) :346-354: eps (m/(n-m) (r0/r)^n - n/(n-m) (r0/r)^m) times the same cut-offs
 * (TB_vdW_Dumitrica is marked "UNFINISHED, DO NOT USE" in the reference and returns zero: not offered.) */
#define XTB_VDW_GIRIFALCO 1
#define XTB_VDW_LJ_CUT    2
#define XTB_VDW_ILJ_CUT   3
typedef struct xtb_vdw_pair {
    double d0_cut, dd_cut;                       /* [A] long-range cut-off radius and width (parent type) */
    double C12, C6, dm, d_cut, a, b, c, d;       /* Girifalco */
    double dsm, ds_cut, as, bs, cs, ds, es, fs;
    double eps, r0, n, m;                        /* LJ_cut (m unused) / ILJ_cut */
    double d0_short, dd_short;                   /* [A] short-range cut-off radius and width */
} xtb_vdw_pair;

/* type(TB_Coulomb_cut), Objects.f90:359-366 (reader Read_input_data.f90:1874-1910, 2360-2386): V(r) = k Q^2 / r / (1 + exp((r - dm) / dd)),
 * zero beyond dm + 10 dd; k = g_e / (4 pi g_e0) * 1e10 [eV A] is set by the reader.  (`ind`, `alpha` select the Wolf / Ewald forms of
 * the SCC part, Coulomb.f90:69-164, which this path does not reach.) */
typedef struct xtb_coulomb_pair {
    double k;     /* [eV A] coupling constant */
    double dm;    /* [A] cut-off radius */
    double dd;    /* [A] cut-off smoothing width */
} xtb_coulomb_pair;

typedef struct xtb_ctx xtb_ctx;

/* ---- life cycle ------------------------------------------------------------------- */
int  xtb_create (xtb_ctx** out, int device_ordinal);
int  xtb_destroy(xtb_ctx* ctx);
const char* xtb_last_error(const xtb_ctx* ctx);
const char* xtb_version(void);
/* run all kernels of this context on `cuda_stream` (a cudaStream_t); default: own stream */
int  xtb_set_stream(xtb_ctx* ctx, void* cuda_stream);
int  xtb_sync(xtb_ctx* ctx);

/* ---- parameters: flattening of Scell%TB_Hamil(:,:) (TB.f90:459-487) ---------------- */
/* p[(koa1-1) + n_kinds*(koa2-1)] = TB_Hamil(koa1,koa2) */
int  xtb_set_params_nrl     (xtb_ctx*, int n_kinds, const xtb_nrl_pair* p);
int  xtb_set_params_pettifor(xtb_ctx*, int n_kinds, const xtb_pettifor_pair* p);
int  xtb_set_params_dftb    (xtb_ctx*, int n_kinds, int n_orb /*1,4,9*/, const xtb_dftb_pair* p);
/* Scell%TB_Repuls(:,:) of type TB_Rep_Pettifor; call after xtb_set_params_pettifor (n_kinds must match).
 * Raises the neighbour cut-off to max(dm) when that is longer than max(rm) (Atomic_tools.f90:609-613). */
int  xtb_set_params_pettifor_rep(xtb_ctx*, int n_kinds, const xtb_pettifor_rep_pair* p);

/* ---- system: MDAtoms(:)%KOA, %S and Scell%supce (Objects.f90:552-581,692-845) ------ */
int  xtb_set_system(xtb_ctx*, int na, const int32_t* koa /*[na], 1-based*/,
                    const double* S /*[3*na] relative coords*/, const double supce[9] /*(3,3) col-major*/);
/* rows (atoms, 1-based inclusive range) this context owns in a multi-GPU run; default all.
 * Replaces the round-robin split `do i = 1+rank, N, nranks` of the MPI build (TB_NRL.f90:267-327). */
int  xtb_set_partition(xtb_ctx*, int atom_first, int atom_last);
/* Multi-GPU block layout (optional; call before xtb_neighbours, same value on every rank): the block stores (H, S, rho,
 * rho_E) are laid out in `parts` equal-capacity segments, one per contiguous range of ceil(na/parts) atoms, so that the
 * owners' ranges are exchanged by ONE in-place all-gather (no packing).  rowptr then has gaps; every access is
 * rowptr[atom] + q, so nothing else changes.  parts = 1 (default): compact.  No reference counterpart (the MPI build
 * all-reduces whole dense arrays, MPI_subroutines.f90:3965-3982). */
int  xtb_set_block_layout(xtb_ctx*, int parts);

/* ---- S1: get_near_neighbours -> Find_nearest_neighbours (Atomic_tools.f90:562,4230;
 *      minimum image = shortest_distance_OLD :4028-4131).  rm_angstrom <= 0: derive the
 *      cut-off from the parameters exactly as get_near_neighbours does (:574-586).
 *      Dense outputs have the reference shapes:
 *        nn_size(na), nn_list(na,na), nn_dist(na,na,4), nn_dist_s(na,na,3); unused = 0 / 1d8 / untouched. */
int  xtb_neighbours(xtb_ctx*, double rm_angstrom, int32_t* nn_size, int32_t* nn_list,
                    double* nn_dist, double* nn_dist_s);
/* the cut-off [A] get_near_neighbours derives from the TB parameters (Atomic_tools.f90:570-613: attractive part, raised to
 * the Pettifor repulsive dm); a caller that also has vdW / DFTB-repulsive terms enlarges it on the host (:615-640) and passes
 * the result to xtb_neighbours.  This is also what the reference returns through its optional `dm` argument (:645-649). */
int  xtb_cutoff(xtb_ctx*, double* rm_angstrom);
/* compact (ELL) copy: list[k + cap*(i-1)], dist[c + 4*(k + cap*(i-1))], dist_s[c + 3*(...)] */
int  xtb_neighbours_compact(xtb_ctx*, int cap, int32_t* nn_size, int32_t* nn_list,
                            double* nn_dist, double* nn_dist_s, int* max_nn_out);

/* ---- S2a: Construct_M_x1 (TB.f90:3430) and Construct_Vij_NRL (TB_NRL.f90:1956,2116) /
 *      Construct_M_Vs (TB_Pettifor.f90:227) / Construct_Vij_DFTB (TB_DFTB.f90:56).
 *      Dense outputs M_V*(na,na,nv), M_lmn/M_x1/M_xrr(3,na,na); any may be NULL. */
int  xtb_radial(xtb_ctx*, double* M_Vij, double* M_dVij, double* M_SVij, double* M_dSVij,
                double* M_lmn, double* M_x1, double* M_xrr);

/* ---- S2b: Hamil_tot_NRL assembly part (TB_NRL.f90:226-397) / Hamil_tot (TB_Pettifor.f90:1658)
 *      / Hamil_tot_DFTB (TB_DFTB.f90:319): dense H_non [eV] and Sij, (N,N), N = na*n_orb.
 *      S may be NULL for orthogonal parametrisations.  In a partitioned run only the column
 *      panels of the owned atoms are written (host copies cover the same panels). */
int  xtb_assemble_HS(xtb_ctx*, double* H_non, double* Sij);

/* Multi-GPU form of S2b (replaces the MPI build's `MPI_Allreduce` of the dense Hij, Sij, TB_NRL.f90:326-327):
 *   xtb_assemble_blocks  computes only the 4x4 / 9x9 blocks of the OWNED atoms into the block stores
 *                        XTB_BUF_HBLK / XTB_BUF_SBLK ([rowptr[j]+q][mu + n_orb*nu], rows = neighbour q-1 or j itself, cols = j);
 *   the caller allgathers the contiguous block ranges of all ranks (0.4 GB instead of 21.7 GB of dense panels at the
 *   north-star size), then
 *   xtb_expand_dense     writes the dense column panels of atoms first..last (1-based, inclusive) from the block stores
 *                        (every element exactly once, zeros included); host pointers may be NULL. */
int  xtb_assemble_blocks(xtb_ctx*);
int  xtb_expand_dense(xtb_ctx*, int atom_first, int atom_last, double* H_non, double* Sij);

/* ---- eigensolve boundary (outside the graft): Scell%Ha(N,N) (column n = state n),
 *      Scell%Ei(N) [eV], Scell%fe(N) (occupations incl. spin factor).  *_device: pointers
 *      are device memory that stays valid until the next xtb_set_eigen* call (zero copy). */
int  xtb_set_eigen       (xtb_ctx*, const double* Ha, const double* Ei, const double* fe /*NULL: occupations follow through xtb_set_occupations*/);
int  xtb_set_eigen_device(xtb_ctx*, const double* dHa, const double* dEi, const double* dfe);
/* State-minor device eigenvectors: dHaT[n + K*r] = Ha(r, n) (row r of state n; the K = N states of a row are contiguous),
 * 16-byte aligned.  This is the layout xtb_eigensolve leaves for 9-orbital sets (its last GEMM writes it at no extra cost) and
 * the one the rho / rho_E kernel streams straight through a tensor map -- with it no re-tiled copy of the eigenvectors is made.
 * The Fortran-layout Ha is derived on the device only when something asks for it (xtb_coupling, xtb_save_eigen_prev, a host copy). */
int  xtb_set_eigen_device_t(xtb_ctx*, const double* dHaT, const double* dEi, const double* dfe);
/* (xtb_set_eigen_device reads dfe once, to select the occupied states: call it again after changing the buffers' contents) */

/* Optional on-device eigensolve (the north star feeds end-to-end steps this way so that H, S and Ha
 *      never cross PCIe).  Same algebra as Loewdin_Orthogonalization (TB_NRL.f90:985-1060) + sym_diagonalize (Algebra_tools.f90:
 *      882-966) + the back-transform Ha = X V (TB_NRL.f90:421), on the device-resident H_non / Sij of the last xtb_assemble_HS,
 *      with cuSOLVER Dsyevd / cuBLAS Dgemm bound at run time (XTB_ESTATE if the libraries are absent).  Leaves Ha and Ei on the
 *      device; host copies of Ei(N), eigen_S(N) (eigenvalues of S, Scell%eigen_S) and Ha(N,N) are optional (NULL = skip).
 *      xtb_set_occupations then supplies Scell%fe(N) (the host computes it from Ei as Electron_tools does) and arms xtb_density. */
int  xtb_eigensolve(xtb_ctx*, double* Ei, double* eigen_S, double* Ha);
int  xtb_get_eigenvalues(xtb_ctx*, double* Ei /*N*/);       /* Scell%Ei from the device (any eigen source) */
int  xtb_set_occupations(xtb_ctx*, const double* fe /*N*/);   /* Scell%fe; eigenvectors / eigenvalues stay as they are */

/* ---- S3: Construct_Aij (TB.f90:4328) and Construct_Aij_x_En (TB.f90:4401).
 *      Computed on the neighbour pattern only (every consumer reads only those blocks,
 *      TB.f90:250-299,5544-5573); dense outputs (N,N) are zero outside the pattern.
 *      f_cut: states with fe(n) <= f_cut are skipped (0 = bit-safe: only exact zeros). */
int  xtb_density(xtb_ctx*, double* Aij, double* Aij_x_Ei);
/* dense copies of the LAST pass without contracting again (the pass makes rho and rho_E together: the Fortran shim runs it
 * once per step from Construct_Aij and serves Construct_Aij_x_En from it) */
int  xtb_density_get(xtb_ctx*, double* Aij, double* Aij_x_Ei);
int  xtb_set_density_cutoff(xtb_ctx*, double f_cut);
/* multi-GPU only: after the owner blocks of all ranks have been exchanged into XTB_BUF_RHO/RHOE
 * (one allgather, replaces the MPI_Allreduce of the dense Aij, TB.f90:4374), fill the mirrored blocks. */
int  xtb_density_mirror(xtb_ctx*);

/* ---- S4: get_dHij_drij_NRL (TB_NRL.f90:2412) / dHij_s (TB_Pettifor.f90:141) /
 *      get_dHij_drij_DFTB (TB_DFTB.f90:781): MDatoms(k)%forces%att(1:3) = dE/ds_k [eV].
 *      F_att(3,na); in a partitioned run only owned atoms are written. */
int  xtb_forces(xtb_ctx*, double* F_att);

/* ---- S5: Attract_TB_Forces_Press_NRL (TB_NRL.f90:2742) / Attract_TB_Forces_Press
 *      (TB_Pettifor.f90:1134) / _DFTB (TB_DFTB.f90:1135): Scell%SCforce%att(3,3) col-major.
 *      In a partitioned run this is the partial sum over owned atoms. */
int  xtb_pressure(xtb_ctx*, double SCforce_att[9]);

/* ---- repulsive pair terms of the Pettifor branch on the same neighbour stream.
 *      xtb_repulsive_energy    get_Erep_s        (TB_Pettifor.f90:2074): E_rep = sum_i [f(sum_j phi(r_ij)) + E0_TB] [eV],
 *                              Epot(na) = MDAtoms(i)%Epot (may be NULL)
 *      xtb_repulsive_forces    dErdr_s           (TB_Pettifor.f90:2162): MDatoms(k)%forces%rep(1:3) = dE_rep/ds_k, F_rep(3,na)
 *      xtb_repulsive_pressure  dErdr_Pressure_s  (TB_Pettifor.f90:2634): Scell%SCforce%rep(3,3) col-major
 *      Need only the neighbour list (no eigenvectors).  Partitioned run: owned atoms / partial sums, as S4 / S5. */
int  xtb_repulsive_energy  (xtb_ctx*, double* E_rep, double* Epot);
int  xtb_repulsive_forces  (xtb_ctx*, double* F_rep);
int  xtb_repulsive_pressure(xtb_ctx*, double SCforce_rep[9]);

/* ---- SURVEY.md 8f item 2, second half: the short-range "exponential wall" (class TB_Exp_wall_simple) on the same stream,
 *      for every TB branch (dispatch Exponential_wall_forces TB.f90:3533-3556; energy TB.f90:3983).  p[(koa1-1) + n_kinds*(koa2-1)].
 *      n_kinds = 0 / p = NULL removes the wall (the reference proceeds "without additional short-range forces" when the file is
 *      absent or its class line is not recognised, Read_input_data.f90:2013-2034).
 *      xtb_wall_energy    get_Exp_wall_s        (Exponential_wall.f90:990-1067): E_wall [eV]; Epot(na) = the per-atom halves
 *      xtb_wall_forces    d_Exp_wall_pot_s      (:1200-1329): the term ADDED to MDatoms(k)%forces%rep(1:3), F_wall(3,na)
 *      xtb_wall_pressure  d_Exp_wall_Pressure_s (:1334-1385): the term ADDED to Scell%SCforce%rep(3,3) */
int  xtb_set_params_wall(xtb_ctx*, int n_kinds, const xtb_wall_pair* p);
int  xtb_wall_energy  (xtb_ctx*, double* E_wall, double* Epot);
int  xtb_wall_forces  (xtb_ctx*, double* F_wall);
int  xtb_wall_pressure(xtb_ctx*, double SCforce_wall[9]);

/* ---- The 'General' short-range class (TB_Short_Rep) in the same slot as the simple wall (Scell%TB_Expwall holds ONE class; setting
 *      one here replaces the other; dispatch Exponential_wall_forces TB.f90:3533-3556, Exponential_wall_s TB.f90:3975-3993).
 *      Z[n_kinds] = matter%Atoms(:)%Z, the atomic numbers (ZBL / NLH).  Results land in the wall buffers: xtb_step reports them as
 *      E_wall / F_wall / SCforce_wall, XTB_FORCE_WALL selects them in the integrator.
 *      xtb_short_energy    get_short_range_rep_s    (Exponential_wall.f90:48-127)
 *      xtb_short_forces    d_Short_range_pot_s      (:519-659): the term ADDED to forces%rep
 *      xtb_short_pressure  d_Short_range_Pressure_s (:664-719): the term ADDED to SCforce%rep */
int  xtb_set_params_short(xtb_ctx*, int n_kinds, const xtb_short_pair* p, const double* Z);
int  xtb_short_energy  (xtb_ctx*, double* E_short, double* Epot);
int  xtb_short_forces  (xtb_ctx*, double* F_short);
int  xtb_short_pressure(xtb_ctx*, double SCforce_short[9]);

/* ---- SURVEY.md 8f item 2, third part: the van der Waals pair term (class TB_vdW; Van_der_Waals.f90, dispatch vdW_forces
 *      TB.f90:3599-3633 / vdW_s TB.f90:4017-4035).  Unlike every other term it does NOT use the neighbour list: it sums over
 *      ALL atom pairs in (2Nx+1)(2Ny+1)(2Nz+1) images of the supercell, the image counts from get_mirror_cell_num
 *      (Van_der_Waals.f90:132-157: R_cut = max(TB cut-off, max d0_cut + 10 max dd_cut), images until n |supce row| >= 2 R_cut,
 *      at least one per axis).  p[(koa1-1) + n_kinds*(koa2-1)] = TB_Waals(koa1, koa2); n_kinds = 0 / p = NULL removes the term
 *      (a missing file or an unrecognised class line -- the shipped 'Girifalco:exclude' -- "proceeds without van der Waals
 *      forces", Read_input_data.f90:1777-1784).  Needs only xtb_set_system.
 *      xtb_vdw_image_cells  get_mirror_cell_num: N[3], R_cut [A]
 *      xtb_vdw_energy       get_vdW_s     (Van_der_Waals.f90:158-250): E_vdW [eV]; Epot(na) = the per-atom halves
 *      xtb_vdw_forces       Construct_B + d_Forces_s (:251-421, TB.f90:3638-3781): the term ADDED to forces%rep, F_vdw(3,na)
 *      xtb_vdw_pressure     d_Forces_Pressure (TB.f90:3785-3848): the term ADDED to Scell%SCforce%rep(3,3) */
int  xtb_set_params_vdw(xtb_ctx*, int kind /*XTB_VDW_**/, int n_kinds, const xtb_vdw_pair* p);
int  xtb_vdw_image_cells(xtb_ctx*, int N[3], double* R_cut);
int  xtb_vdw_energy  (xtb_ctx*, double* E_vdw, double* Epot);
int  xtb_vdw_forces  (xtb_ctx*, double* F_vdw);
int  xtb_vdw_pressure(xtb_ctx*, double SCforce_vdw[9]);

/* ---- Allocation-free variant of xtb_vdw_pressure.  It has NO counterpart in the reference: it was added
 *      for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 *      forbid it).  Same inputs (the context's system and vdW table, both already resident) and the same result as
 *      xtb_vdw_pressure, but the per-atom partial sums and the 9 sums live in a fixed file-scope device array instead of the
 *      context's grow-only buffers: the call performs no cudaMalloc / cudaFree / event creation and uses no allocating library.
 *      Limits and rules: at most XTB_NOALLOC_MAX_ATOMS owned atoms (more: XTB_ECAPACITY with a message; use xtb_vdw_pressure);
 *      every GPU that loads the library holds the array (590 KB) until exit; all calls on one GPU share it, so calls that overlap
 *      on a GPU (streams, contexts, threads) race -- serialise them.  SCforce_vdw[9]: host OR device pointer. */
#define XTB_NOALLOC_MAX_ATOMS 8192
int  xtb_vdw_pressure_noalloc(xtb_ctx*, double SCforce_vdw[9]);

/* ---- SURVEY.md 8f item 2, fourth part: the Coulomb term of a charged supercell (class TB_Coulomb_cut; Coulomb.f90, dispatch
 *      Coulomb_forces TB.f90:3560-3596 / Coulomb_s TB.f90:3995-4012).  Same all-pairs-in-image-cells structure as the van der Waals
 *      term and the same d_Forces_s / d_Forces_Pressure, with the reference's differences kept: image counts from
 *      R_cut = TB_Coul(1,1)%dm alone (get_mirror_cell_num_C :435-452); the energy is NOT halved (get_Coulomb_s :549, Epot(i) gets the
 *      full sum); Construct_B_C contracts the distance with the TRANSPOSE of supce (:864-866) and returns zero forces when
 *      Q <= 1e-12 (:762).  Q = Scell%Q, the mean unbalanced charge per atom [e], changes from step to step: xtb_set_charge.
 *      xtb_coulomb_energy    get_Coulomb_s   (Coulomb.f90:454-550): E_coul [eV]; Epot(na) = what is added to MDAtoms(:)%Epot
 *      xtb_coulomb_forces    Construct_B_C + d_Forces_s (:715-880, TB.f90:3638): the term ADDED to forces%rep, F_coul(3,na)
 *      xtb_coulomb_pressure  d_Forces_Pressure (TB.f90:3785): the term ADDED to Scell%SCforce%rep(3,3) */
int  xtb_set_params_coulomb(xtb_ctx*, int n_kinds, const xtb_coulomb_pair* p);
int  xtb_set_charge(xtb_ctx*, double Q);
int  xtb_coulomb_image_cells(xtb_ctx*, int N[3]);
int  xtb_coulomb_energy  (xtb_ctx*, double* E_coul, double* Epot);
int  xtb_coulomb_forces  (xtb_ctx*, double* F_coul);
int  xtb_coulomb_pressure(xtb_ctx*, double SCforce_coul[9]);

/* ---- Fermi occupations on the device (between the eigensolve and S3; lets a step stay on the device):
 *      Electron_Fixed_Te (Electron_tools.f90:2660-2706: bisection of mu at fixed Te until |N(mu) - Ne| <= 1e-12 Ne) +
 *      set_Fermi (:1913-1969), spin-degenerate (f_norm = 2).  Uses the eigenvalues on the device (xtb_eigensolve or
 *      xtb_set_eigen*), arms xtb_density.  fe(N), mu [eV], E_band = sum f_n E_n [eV] optional host outputs. */
int  xtb_fermi_occupations(xtb_ctx*, double n_electrons, double Te_eV, double* fe, double* mu, double* E_band);

/* ---- velocity-Verlet halves on the device, make_time_step_atoms_SC (Atomic_tools.f90:1897-1974)
 *      preceded by save_last_timestep (:1177-1192) as in the main loop (XTANT_MAIN_FILE.f90:338-341,380-383):
 *        ind = 2: S = S0 + dt SV0 + a dt^2/2 (periodic wrap, :2654,:2692), SV = SV0 + a dt/2     [MD_step, TB.f90:113]
 *        ind = 1: SV = SV0 + a dt/2 with the new forces                                        [AT_MOVE_2]
 *      a = -g_e 1e-10 / M * g^-1 F_total - g^-1 (dg/dt) SV0; F_total = sum of the selected device-resident force terms
 *      (XTB_FORCE_*; each [eV], d/ds as returned by xtb_forces).  Vsupce(3,3) = NULL for a fixed cell.  mass_kg[n_kinds].
 *      Positions are those of xtb_set_system and are updated in place (the next xtb_neighbours / xtb_step sees them). */
#define XTB_FORCE_ATT   1u   /* XTB_BUF_FORCES  (xtb_forces) */
#define XTB_FORCE_REP   2u   /* XTB_BUF_FREP    (xtb_repulsive_forces) */
#define XTB_FORCE_WALL  4u   /* XTB_BUF_FWALL   (xtb_wall_forces) */
#define XTB_FORCE_VDW   8u   /* XTB_BUF_FVDW    (xtb_vdw_forces) */
#define XTB_FORCE_COUL 16u   /* XTB_BUF_FCOUL   (xtb_coulomb_forces) */
int  xtb_md_set_velocities(xtb_ctx*, const double* SV /*[3*na] relative velocities [1/fs]; NULL = zero*/);
int  xtb_md_half_step(xtb_ctx*, int ind, double dt_fs, const double* mass_kg, const double* Vsupce, unsigned force_terms);
int  xtb_md_get_state(xtb_ctx*, double* S, double* SV, double* E_kin /*[eV]*/, const double* mass_kg /*for E_kin*/);

/* ---- multi-GPU inside the library (one process / MPI rank per GPU; NCCL bound at run time with dlopen).
 *      Replaces the MPI_Allreduce of whole dense arrays of the MPI build (TB_NRL.f90:326-327, TB.f90:4374,
 *      MPI_subroutines.f90:3965-3982):
 *        rank 0:  xtb_comm_unique_id(id)  ->  MPI_Bcast(id, 128 bytes)  ->  every rank: xtb_comm_init(ctx, id, rank, nranks)
 *      xtb_comm_init gives the context the atom range of its rank (ceil(na/nranks) atoms each, contiguous) and lays the block
 *      stores out in one segment per rank (xtb_set_block_layout).  Then, per step, either xtb_step (below) or by hand:
 *        xtb_neighbours -> xtb_assemble_blocks -> xtb_exchange_blocks(XTB_XCHG_HS) -> xtb_expand_dense(1, na) -> [eigensolve]
 *        -> xtb_density -> xtb_exchange_blocks(XTB_XCHG_RHO) -> xtb_density_mirror -> xtb_forces -> xtb_exchange_forces
 *        [-> xtb_pressure -> xtb_allreduce_scforce]
 *      XTB_XCHG_HS:  in-place all-gather of the ranks' H / S block segments (every rank expands the replicated dense matrices);
 *      XTB_XCHG_RHO: halo only -- point-to-point transfers of the owner blocks a rank mirrors into its own rows and of the
 *                    on-site blocks of the neighbours of its atoms (XTB_RHO_EXCHANGE=allgather: everything to everyone). */
#define XTB_UNIQUE_ID_BYTES 128
#define XTB_XCHG_HS   1
#define XTB_XCHG_RHO  2
int  xtb_comm_unique_id(void* id_out /*XTB_UNIQUE_ID_BYTES*/);
int  xtb_comm_init   (xtb_ctx*, const void* unique_id, int rank, int nranks);
int  xtb_comm_destroy(xtb_ctx*);
int  xtb_exchange_blocks  (xtb_ctx*, int which);
int  xtb_exchange_forces  (xtb_ctx*, double* F_att /*(3,na) complete, or NULL*/);
int  xtb_allreduce_scforce(xtb_ctx*, double SCforce_att[9] /*or NULL*/);

/* ---- the whole step as ONE call: the stages selected in `flags`, in the order of get_Hamilonian_and_E (TB.f90:196-330),
 *      with the exchanges of a communicating context in between, everything enqueued on the context stream and ONE host
 *      synchronisation at the end.  On a single GPU the neighbour list is rebuilt without a size read-back from the second
 *      step on (capacity, neighbour bound and block capacity of the previous step, checked on the device; a step that
 *      outgrows them is re-run through the sizing path, transparently).  Exceptions that do wait in the middle: the first
 *      step / a communicating context (list sizes size the launches and the NCCL counts), XTB_STEP_EIGEN (cuSOLVER reports
 *      its status through the host) and new occupations (the number of occupied states sizes the density launches).
 *      io: host arrays, every pointer optional (NULL = keep on the device / not given). */
#define XTB_STEP_NEIGHBOURS   1u
#define XTB_STEP_HS           2u    /* blocks [+ exchange] + dense H_non, Sij */
#define XTB_STEP_EIGEN        4u    /* on-device eigensolve (xtb_eigensolve) + Fermi occupations (io->n_electrons, io->Te_eV) */
#define XTB_STEP_DENSITY      8u    /* rho, rho_E [+ halo exchange + mirror] */
#define XTB_STEP_FORCES      16u    /* attractive forces [+ all-gather] */
#define XTB_STEP_PRESSURE    32u    /* supercell force [+ all-reduce] (also of the pair terms when XTB_STEP_PAIR_TERMS) */
#define XTB_STEP_PAIR_TERMS  64u    /* repulsive part of the Pettifor branch, the short-range wall and the van der Waals term, and the Coulomb term, whichever are set */
#define XTB_STEP_NO_DENSE   128u    /* with XTB_STEP_HS: block stores only */
#define XTB_STEP_GRAPH      256u    /* single GPU, no XTB_STEP_EIGEN / new occupations: once a steady step with the same flags has
                                       run, replay its launch sequence as ONE CUDA graph (launch-bound small systems); host
                                       inputs / outputs travel through pinned staging; falls back to the eager path otherwise */
#define XTB_STEP_ALL (XTB_STEP_NEIGHBOURS | XTB_STEP_HS | XTB_STEP_DENSITY | XTB_STEP_FORCES)
typedef struct xtb_step_io {
    const double* S;          /* in  [3*na] new MDAtoms%S */
    const double* supce;      /* in  [9]    new Scell%supce */
    const double* fe;         /* in  [N]    Scell%fe for this step (ignored with XTB_STEP_EIGEN) */
    double  n_electrons;      /* in  Scell%Ne_low, for XTB_STEP_EIGEN */
    double  Te_eV;            /* in  Scell%TeeV,   for XTB_STEP_EIGEN */
    double* Ei;               /* out [N] */
    double* fe_out;           /* out [N]   (XTB_STEP_EIGEN) */
    double* mu;               /* out [1]   (XTB_STEP_EIGEN) */
    double* E_band;           /* out [1]   sum f_n E_n (XTB_STEP_EIGEN) */
    double* F_att;            /* out [3*na] */
    double* SCforce_att;      /* out [9] */
    double* E_rep;  double* F_rep;  double* SCforce_rep;      /* out: Pettifor repulsive part */
    double* E_wall; double* F_wall; double* SCforce_wall;     /* out: short-range wall */
    double* E_vdw;  double* F_vdw;  double* SCforce_vdw;      /* out: van der Waals term */
    double* E_coul; double* F_coul; double* SCforce_coul;     /* out: Coulomb term */
} xtb_step_io;
int  xtb_step(xtb_ctx*, unsigned flags, const xtb_step_io* io);

/* ---- Electron_ion_coupling_Mij (Nonadiabatic.f90:782-945), the two more N^3 contractions of an
 *      "electron-ion coupling" run.  Ha0 = eigenvectors of the previous step (Scell%Ha0, saved by construct_TB_H_* as
 *      `Ha0 = Ha`, TB_NRL.f90:88): xtb_save_eigen_prev copies the context's current eigenvectors device-to-device,
 *      xtb_set_eigen_prev uploads a host Ha0(N,N).  xtb_coupling: kind_M = -1 -> the `case(-1)` arm, anything else ->
 *      Tully's expression (default arm); the overlap matrix is the block-sparse S of the last xtb_assemble_HS /
 *      xtb_assemble_blocks (all atoms), absent for the orthogonal branch.  Mij(N,N) host copy optional (NULL = device only).
 *      Not partitioned: in a multi-GPU run every rank that calls it computes the whole matrix. */
int  xtb_save_eigen_prev(xtb_ctx*);
int  xtb_set_eigen_prev (xtb_ctx*, const double* Ha0);
int  xtb_coupling       (xtb_ctx*, int kind_M, double* Mij);

/* ---- introspection / zero-copy interop (device pointers valid until the producing stage
 *      is re-run or the context destroyed) ------------------------------------------- */
#define XTB_BUF_H        1   /* dense H_non (N,N) */
#define XTB_BUF_S        2   /* dense Sij (N,N) */
#define XTB_BUF_RHO      3   /* block-sparse rho   [n_blocks][n_orb*n_orb] */
#define XTB_BUF_RHOE     4   /* block-sparse rho_E */
#define XTB_BUF_FORCES   5   /* (3,na) */
#define XTB_BUF_ROWPTR   6   /* int32 [na+1]: first block of atom i (block 0 = on-site) */
#define XTB_BUF_NNLIST   7   /* int32 ELL list, 0-based, [cap*na] */
#define XTB_BUF_NNSIZE   8   /* int32 [na] */
#define XTB_BUF_SCFORCE  9   /* 9 doubles */
#define XTB_BUF_HBLK    10   /* block-sparse H [n_blocks][n_orb*n_orb] (eV) */
#define XTB_BUF_SBLK    11   /* block-sparse S */
#define XTB_BUF_FREP    12   /* (3,na) repulsive forces */
#define XTB_BUF_MIJ     13   /* (N,N) coupling matrix of the last xtb_coupling */
#define XTB_BUF_FWALL   14   /* (3,na) wall forces */
#define XTB_BUF_POS     15   /* (3,na) relative coordinates S */
#define XTB_BUF_VEL     16   /* (3,na) relative velocities SV */
#define XTB_BUF_EI      17   /* (N) eigenvalues */
#define XTB_BUF_FE      18   /* (N) occupations */
#define XTB_BUF_FVDW    19   /* (3,na) van der Waals forces */
#define XTB_BUF_FCOUL   20   /* (3,na) Coulomb forces */
int  xtb_device_buffer(xtb_ctx*, int which, void** dptr, size_t* bytes);
/* counters: 0 na, 1 n_orb, 2 N, 3 max_nn, 4 n_blocks (incl. on-site), 5 n_pairs (directed),
 *           6 kernels launched since create, 7 states kept by the last xtb_density, 8 cap,
 *           9 / 10 shared-memory slots / pipeline stages of the last density kernel,
 *           11 work units (atom pairs + single atoms) of the last two-column-atom density launch,
 *           12 / 13 bytes received per rank by the last H/S / rho exchange, 14 / 15 halo blocks sent / received, 16 CUDA graphs captured,
 *           17 = 1 if the last density launch read the state-minor eigenvectors through the tensor map (no re-tiled copy) */
long long xtb_counter(const xtb_ctx*, int which);
/* device time [ms] of the last run of a stage (CUDA events on the context stream):
 * 1 neighbours, 2 radial+assembly, 3 density, 4 forces, 5 pressure,
 * 6 dense-panel kernel alone (inside 2), 7 density contraction kernel alone (inside 3), 8 xtb_eigensolve,
 * 9 pair-term energy + forces (repulsive / wall), 10 pair-term pressure, 11 xtb_coupling (12 / 13: its norms + S*Ha part / its DGEMM),
 * 14 H/S block exchange, 15 rho block exchange (xtb_counter 12 / 13: bytes received per rank by each),
 * 0 the last graph replay of xtb_step as a whole (xtb_counter 16: graphs captured so far),
 * 16 the eigenvector re-tiling kernel alone (inside 3), 17 van der Waals term, 18 Coulomb term (energy / forces / supercell force of one call).  Stage 6 is the dense H_non / Sij update: the matrices stay resident and
 * only the blocks of the previous and of the new pattern are touched (XTB_DENSE_FULL=1: every element, every step).  In the NRL
 * branch on one rank the values are written by the block kernel of stage 2 in the same pass as the block stores, and stage 6 is
 * the preparation alone (dropped blocks cleared, new pattern recorded); XTB_DENSE_FUSED=0: separate scatter kernel, inside 6 */
double xtb_stage_ms(const xtb_ctx*, int stage);

#ifdef __cplusplus
}
#endif
#endif /* XTB_B200_H */
