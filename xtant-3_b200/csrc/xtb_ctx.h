// Internal context of libxtb_b200 (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>
#include "../../include/xtb_b200.h"

// Universal_constants.f90:67,103,104 -- copied verbatim; g_A2au is NOT 1/g_au2A in the reference.
#define XTB_G_RY   13.6056981
#define XTB_G_AU2A 0.529177249
#define XTB_G_A2AU 1.8897259885789

struct DevBuf {
    void*  p = nullptr;
    size_t bytes = 0;
};

// device record of one (KOA1, KOA2) entry of class TB_Short_Rep (wall.cu); tab -> R | E | a | b | c | d, each [n_tab]
struct ShortDev {
    double d0, dd;
    int use_exp, use_inv_exp, use_zbl, use_nlh;
    double exp_Phi, exp_r0, exp_a, inv_C, inv_r0;
    int n_pow, n_tab, tab_spline, pad_;
    double pow_Phi[XTB_SHORT_MAX_POW], pow_r0[XTB_SHORT_MAX_POW], pow_m[XTB_SHORT_MAX_POW];
    double nlh_a[3], nlh_b[3];
    double Z1, Z2;              // atomic numbers of KOA1, KOA2
    const double* tab;
};

struct xtb_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::string err;
    long long launches = 0;

    // parameters
    int tb = XTB_TB_NONE;
    int n_kinds = 0;
    int n_orb = 0;
    std::vector<xtb_nrl_pair> h_nrl;
    std::vector<xtb_pettifor_pair> h_pet;
    std::vector<xtb_pettifor_rep_pair> h_pet_rep;   // empty: no repulsive part set
    DevBuf d_rep_params;
    DevBuf d_params;            // array of the active pair struct (DFTB: array of DftbDev, dftb_device.cuh)
    DevBuf d_dftb_tab;          // DFTB: all radial grids and tables, one buffer of doubles
    double dftb_rm = 0.0;       // DFTB: max(rcut) + 10*max(d)  (get_near_neighbours, Atomic_tools.f90:588-591)

    // system
    int na = 0;
    int a0 = 0, a1 = 0;         // owned atom range [a0, a1) (0-based)
    int layout_parts = 1;       // block stores in this many equal-capacity segments (xtb_set_block_layout)
    double h_supce[9] = {0};
    DevBuf d_koa, d_S;          // int32 [na] (0-based kinds), double [3*na]
    // neighbour list (ELL, capacity cap) for ALL atoms (positions are replicated; the list is
    // cheap and every later stage needs the lists of neighbours of owned atoms as well)
    int cap = 0, max_nn = 0;    // max_nn: the neighbour count every later stage is sized with (steady-state form: a bound with slack)
    int max_nn_seen = 0;        // largest neighbour count of the last validated list
    long long n_pairs = 0, n_blocks = 0;
    long long blk_cap = 0;      // blocks the block stores are sized for (>= n_blocks; the steady-state form adds slack)
    bool have_shape = false;    // cap / max_nn / blk_cap are armed for xtb_impl_neighbours_steady
    int* h_meta = nullptr;      // pinned: {max count, real blocks, overflow flag, rowptr[na]} of the last steady-state list
    // CUDA-graph replay of a steady step (step.cu)
    void* graph_exec = nullptr;
    std::vector<char> graph_key;
    double* h_stage = nullptr; size_t h_stage_bytes = 0;     // pinned staging of the step's host inputs / outputs
    unsigned steady_flags_done = 0;                           // flag set of the last eager steady step (its buffers are sized)
    long long graph_captures = 0;
    long long alloc_epoch = 0;                                // bumped by every (re)allocation of a device buffer
    double rm = 0.0;
    DevBuf d_nn_size, d_nn_list, d_nn_dist, d_nn_dist_s, d_rev, d_own, d_rowptr, d_maxnn;
    bool have_neigh = false, have_dens = false;
    // NRL per-atom quantities
    DevBuf d_dens;              // embedded density rho_j
    DevBuf d_onsite_w;          // sum_l dh_l/drho * tr_l(rho_jj) * g_Ry  (force on-site weight)
    // block-sparse H, S ([n_blocks][n_orb^2], block (j,q): rows = neighbour / j itself, cols = j) and dense H, S
    DevBuf d_Hblk, d_Sblk;
    DevBuf d_H, d_S2;
    // the dense matrices stay resident; the pattern they were written with (xtb_dense_update, nrl_assemble.cu)
    DevBuf d_prev_size, d_prev_list, d_dense_out;
    int dense_state = 0, dense_na = 0, dense_norb = 0, dense_ortho = -1, dense_pcap = 0;
    bool have_HS = false, have_blocks = false;
    // eigen
    const double* d_Ha = nullptr; const double* d_Ei = nullptr; const double* d_fe = nullptr;
    // state-minor copy of the eigenvectors HaT[row r][state n] (states contiguous): what the 9-orbital density kernel streams
    // straight through tensor-map TMA.  Either layout may be the primary one; the other is made on demand (eigen.cu).
    const double* d_HaT = nullptr;
    DevBuf own_HaT, d_tmap;
    const void* tmap_for = nullptr; long long tmap_rows = 0, tmap_states = 0;
    bool keep_prefix = false;   // the kept states are 0 .. k_kept-1
    int density_tm = 0;         // the last density launch read HaT through the tensor map
    DevBuf own_Ha, own_Ei, own_fe;
    long long K = 0;            // number of states (columns of Ha)
    bool have_eigen = false;
    bool have_eigvec = false;   // Ha, Ei produced on the device by xtb_eigensolve (occupations pending)
    void* h_cusolver = nullptr; void* h_cublas = nullptr;      // library handles of this context (eigen.cu)
    DevBuf d_eig_work, d_eig_X, d_eig_T, d_eigS;   // cuSOLVER workspace, Loewdin X, scratch, eigenvalues of S
    // density
    double f_cut = 0.0;
    DevBuf d_keep, d_wf, d_wfe, d_Ct, d_nkeep;   // kept state indices, weights, tiled transposed C
    long long k_kept = 0;
    bool have_keep = false;     // d_keep / d_wf / d_wfe / k_kept match the current fe, Ei, f_cut
    DevBuf d_rho, d_rhoE;       // block-sparse [n_blocks][n_orb^2]
    DevBuf d_ctneed;            // partitioned run: atom groups whose eigenvector tiles are re-tiled
    DevBuf d_prog;              // density kernel: per-CTA progress counters (lock-step throttle)
    DevBuf d_units, d_pairw;    // pair density kernel: (i1, i2) per work unit + matching scratch; candidate-pair weights
    int density_pairing = 1;    // 0: no matching rounds (every atom a single unit; XTB_DENSITY_PAIRING, experiments)
    bool have_density = false;
    // forces / pressure
    DevBuf d_F, d_SC, d_scratch, d_hdev, d_ppart, d_tab;
    DevBuf d_fpair;             // NRL forces: per list entry the pair term (H - S, three coordinates) and the on-site term (nrl_forces.cu)
    // repulsive pair terms (pettifor_rep.cu)
    DevBuf d_Epot, d_dfpsi, d_Erep, d_Frep, d_SCrep;
    // short-range wall (wall.cu)
    std::vector<xtb_wall_pair> h_wall;     // empty: no wall set
    std::vector<ShortDev> h_short;         // class TB_Short_Rep in the same slot (at most one of h_wall / h_short is non-empty)
    DevBuf d_short_params, d_short_tab;
    DevBuf d_wall_params, d_Ewall_at, d_Ewall, d_Fwall, d_SCwall;
    // van der Waals term (vdw.cu)
    std::vector<xtb_vdw_pair> h_vdw;       // empty: no vdW term set
    int vdw_kind = 0;                      // XTB_VDW_*
    bool vdw_sym = true;                   // TB_Waals(k1,k2) == TB_Waals(k2,k1) bit for bit
    int vdw_cells[3] = {0, 0, 0};          // image counts of the last launch
    DevBuf d_vdw_params, d_Evdw_at, d_Evdw, d_Fvdw, d_SCvdw;
    // Coulomb term of a charged cell (vdw.cu): the device table holds k Q Q per pair, rebuilt when Q or the parameters change
    std::vector<xtb_coulomb_pair> h_coul;  // empty: no Coulomb term set
    double coul_Q = 0.0;
    bool coul_dirty = true, coul_sym = true;
    int coul_cells[3] = {0, 0, 0};
    DevBuf d_coul_params, d_Ecoul_at, d_Ecoul, d_Fcoul, d_SCcoul;
    // Fermi occupations and velocity-Verlet state (step.cu)
    DevBuf d_fermi;                        // {mu, N(mu), sum f E, iterations}
    DevBuf d_S0, d_SV, d_SV0, d_mdpar;
    bool have_md = false;
    // nonadiabatic coupling (coupling.cu): eigenvectors of the previous step, T = S Ha, G -> Mij, inverse column norms
    DevBuf d_Ha0, d_cpl_T, d_cpl_M, d_cpl_inv, d_cpl_union;
    bool have_Ha0 = false;
    // multi-GPU (comm.cu): NCCL communicator bound at run time, halo transfer lists of the current neighbour list
    void* comm = nullptr;
    int rank = 0, nranks = 1;
    cudaStream_t comm_stream = nullptr;    // the H/S block all-gather runs here while the step's stream expands the own panels
    cudaEvent_t ev_comm0 = nullptr, ev_comm1 = nullptr;
    bool rho_allgather = false;            // XTB_RHO_EXCHANGE=allgather: every rho block to every rank (comparison runs)
    DevBuf d_halo_tr, d_halo_cnt, d_halo_slots, d_halo_send, d_halo_recv;
    std::vector<int> halo_count;           // blocks per transfer: sends to the other ranks (ascending), then receives
    int halo_listcap = 0;
    long long halo_send_blocks = 0, halo_recv_blocks = 0;
    bool have_halo = false;
    double xchg_bytes[3] = {0, 0, 0};      // bytes received per rank by the last exchange [XTB_XCHG_HS | XTB_XCHG_RHO]
    // timing: one CUDA event pair per stage index of xtb_stage_ms (created on first use).  A stage records its pair on the
    // context stream; xtb_stage_finish() waits for the stream and reads the pending pairs -- except inside xtb_step
    // (deferred), where nothing waits and the pairs are read after the step's single synchronisation, and inside a
    // CUDA-graph capture (capturing), where no events are recorded at all.
    cudaEvent_t tev[24][2] = {};
    unsigned t_pending = 0, t_started = 0;
    bool deferred = false, capturing = false;
    int density_slotcap = 0, density_stages = 0;
    double stage_ms[24] = {0};   // see xtb_stage_ms
    int sm_count = 148;
    int density_dbg = 0;        // XTB_DENSITY_DBG (timing experiments only, separate kernel instantiation): 1 no math, 2 quarter of the tile copies, 4 helper warps nap while polling, 8 no DFMA
    int density_ncw = 8;        // consumer warps of the density kernel (8 or 12); XTB_DENSITY_NCW
    int density_lead = 64;      // chunks a CTA of the density kernel may run ahead of the slowest one (0 = no lock-step); XTB_DENSITY_LEAD
    int density_impl = 1;       // 1 = DMMA + bulk-copy pipeline, one column atom per CTA; 2 = the same with two column atoms per CTA (density_pair.cuh); 0 = simple CUDA-core kernel (debug / cross-check)
};

#define XTB_FAIL(ctx, code, ...)                                   \
    do {                                                           \
        char _b[512];                                              \
        snprintf(_b, sizeof _b, __VA_ARGS__);                      \
        (ctx)->err = _b;                                           \
        return (code);                                             \
    } while (0)

#define XTB_CUDA(ctx, call)                                                                   \
    do {                                                                                      \
        cudaError_t _e = (call);                                                              \
        if (_e != cudaSuccess) {                                                              \
            cudaGetLastError();     /* not sticky: do not let it surface at the next launch check */ \
            XTB_FAIL(ctx, XTB_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e),  \
                     __FILE__, __LINE__);                                                     \
        }                                                                                     \
    } while (0)

#define XTB_LAUNCH_CHECK(ctx)                                                                 \
    do {                                                                                      \
        (ctx)->launches++;                                                                    \
        cudaError_t _e = cudaGetLastError();                                                  \
        if (_e != cudaSuccess)                                                                \
            XTB_FAIL(ctx, XTB_ECUDA, "kernel launch failed: %s (%s:%d)",                      \
                     cudaGetErrorString(_e), __FILE__, __LINE__);                             \
    } while (0)

// blocks every block store is allocated for
static inline size_t xtb_blocks_alloc(const xtb_ctx* c) { return (size_t)(c->blk_cap > c->n_blocks ? c->blk_cap : c->n_blocks); }

#define XTB_T0(c, s) do { int _r = xtb_t0((c), (s)); if (_r) return _r; } while (0)
#define XTB_T1(c, s) do { int _r = xtb_t1((c), (s)); if (_r) return _r; } while (0)

int xtb_reserve(xtb_ctx* c, DevBuf& b, size_t bytes);   // grow-only device allocation
int xtb_t0(xtb_ctx* c, int stage);          // record the start / end event of a stage on the context stream
int xtb_t1(xtb_ctx* c, int stage);
int xtb_stage_finish(xtb_ctx* c);           // synchronous mode: wait for the stream and read the pending pairs into stage_ms
int xtb_collect_timing(xtb_ctx* c);         // read the pending pairs (the stream must be idle)
void xtb_free(DevBuf& b);

// stage implementations (one translation unit each)
int xtb_impl_neighbours(xtb_ctx* c, double rm);
int xtb_impl_expand_dense(xtb_ctx* c, int first, int last, bool timed = true);   // xtb_api.cu: dense panels from the block stores, no wait
int xtb_impl_neighbours_steady(xtb_ctx* c, double rm);   // no host round trip; validated by xtb_step after its synchronisation
int xtb_impl_prepare_occupations(xtb_ctx* c);            // density.cu: kept states + weights from fe, Ei (at occupation-set time)
int xtb_impl_nrl_assemble(xtb_ctx* c, bool dense);     // dense = false: block stores only
int xtb_impl_nrl_dens(xtb_ctx* c);          // embedded densities rho_j (needed by S2 and by the on-site force terms)
int xtb_impl_pettifor_assemble(xtb_ctx* c, bool dense);
int xtb_impl_density(xtb_ctx* c);
int xtb_impl_nrl_forces(xtb_ctx* c);
int xtb_impl_nrl_pressure(xtb_ctx* c);
int xtb_impl_pettifor_forces(xtb_ctx* c);
int xtb_impl_pettifor_pressure(xtb_ctx* c);
int xtb_impl_pettifor_rep_energy(xtb_ctx* c);      // pettifor_rep.cu
int xtb_impl_pettifor_rep_forces(xtb_ctx* c);
int xtb_impl_pettifor_rep_pressure(xtb_ctx* c);
int xtb_impl_mirror_density(xtb_ctx* c);             // rows of the owned atoms when the context communicates, else all rows
static inline bool xtb_has_wall(const xtb_ctx* c) { return !c->h_wall.empty() || !c->h_short.empty(); }
int xtb_impl_wall_energy(xtb_ctx* c);                // wall.cu (simple wall or the General class, whichever is set)
int xtb_impl_wall_forces(xtb_ctx* c);
int xtb_impl_wall_pressure(xtb_ctx* c);
int xtb_impl_vdw_cells(xtb_ctx* c, int N[3], double* R_cut);   // vdw.cu: get_mirror_cell_num on the host
int xtb_impl_vdw(xtb_ctx* c, unsigned what);         // what: 1 energy | 2 forces | 4 supercell force, one pass
int xtb_impl_coulomb_cells(xtb_ctx* c, int N[3]);     // vdw.cu: get_mirror_cell_num_C
int xtb_impl_coulomb(xtb_ctx* c, unsigned what);
int xtb_impl_fermi(xtb_ctx* c, double n_electrons, double Te_eV);     // step.cu
int xtb_impl_assemble(xtb_ctx* c, bool dense);       // xtb_api.cu: dispatch on the TB branch (the `select type` arms)
int xtb_impl_forces(xtb_ctx* c);
int xtb_impl_pressure(xtb_ctx* c);
int xtb_impl_pair_terms(xtb_ctx* c, const xtb_step_io* io, bool pressure);
double xtb_default_cutoff(const xtb_ctx* c);
int xtb_impl_comm_lists(xtb_ctx* c);                 // comm.cu
int xtb_impl_exchange_blocks(xtb_ctx* c, int which, bool side_stream = false);   // side_stream: enqueue on comm_stream, fenced by ev_comm0 / ev_comm1
int xtb_impl_exchange_forces(xtb_ctx* c, DevBuf& F);
int xtb_impl_allreduce9(xtb_ctx* c, double* d9);
int xtb_impl_coupling(xtb_ctx* c, int kind_M);      // coupling.cu
int xtb_rows_by_state(xtb_ctx* c, const double* Ha, long long K, double* X);      // density.cu
int xtb_blas_dgemm(xtb_ctx* c, bool ta, bool tb, int m, int n, int k, const double* A, int lda, const double* B, int ldb, double* Cm, int ldc);
void xtb_release_lib_handles(xtb_ctx* c);   // eigen.cu
int xtb_ensure_Ha_std(xtb_ctx* c);          // eigen.cu: the Fortran-layout Ha(N,N) on the device (transposed from HaT if that is all there is)
int xtb_impl_eigensolve(xtb_ctx* c);        // eigen.cu: Loewdin + Dsyevd + back-transform on the device-resident H, S
int xtb_impl_dftb_assemble(xtb_ctx* c, bool dense);
int xtb_impl_dftb_forces(xtb_ctx* c);
int xtb_impl_dftb_pressure(xtb_ctx* c);
int xtb_impl_dftb_tables(xtb_ctx* c, double* d_rad, double* d_geo);  // compact tables, nv channels per list entry
int xtb_impl_nrl_tables(xtb_ctx* c, double* d_rad, double* d_geo);   // compact M_V*/M_lmn/M_xrr/M_x1 per list entry
// dense column panels of the owned atoms from a block store; tr: block holds (rows j, cols neighbour)
int xtb_dense_update(xtb_ctx* c, int f0, int f1, bool scatter = true);      // resident dense H_non / Sij: clear the previous pattern, write the new one (scatter = false: the caller's kernel writes it)
int xtb_dense_from_blocks(xtb_ctx* c, const double* Ablk, const double* Bblk, double* A, double* B, bool tr, int f0 = -1, int f1 = -1);
