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
    int cap = 0, max_nn = 0;
    long long n_pairs = 0, n_blocks = 0;
    double rm = 0.0;
    DevBuf d_nn_size, d_nn_list, d_nn_dist, d_nn_dist_s, d_rev, d_own, d_rowptr, d_maxnn;
    bool have_neigh = false, have_dens = false;
    // NRL per-atom quantities
    DevBuf d_dens;              // embedded density rho_j
    DevBuf d_onsite_w;          // sum_l dh_l/drho * tr_l(rho_jj) * g_Ry  (force on-site weight)
    // block-sparse H, S ([n_blocks][n_orb^2], block (j,q): rows = neighbour / j itself, cols = j) and dense H, S
    DevBuf d_Hblk, d_Sblk;
    DevBuf d_H, d_S2;
    bool have_HS = false, have_blocks = false;
    // eigen
    const double* d_Ha = nullptr; const double* d_Ei = nullptr; const double* d_fe = nullptr;
    DevBuf own_Ha, own_Ei, own_fe;
    long long K = 0;            // number of states (columns of Ha)
    bool have_eigen = false;
    bool have_eigvec = false;   // Ha, Ei produced on the device by xtb_eigensolve (occupations pending)
    DevBuf d_eig_work, d_eig_X, d_eig_T, d_eigS;   // cuSOLVER workspace, Loewdin X, scratch, eigenvalues of S
    double eigen_ms = 0.0;
    // density
    double f_cut = 0.0;
    DevBuf d_keep, d_wf, d_wfe, d_Ct, d_nkeep;   // kept state indices, weights, tiled transposed C
    long long k_kept = 0;
    DevBuf d_rho, d_rhoE;       // block-sparse [n_blocks][n_orb^2]
    DevBuf d_ctneed;            // partitioned run: atom groups whose eigenvector tiles are re-tiled
    DevBuf d_prog;              // density kernel: per-CTA progress counters (lock-step throttle)
    DevBuf d_units, d_pairw;    // pair density kernel: (i1, i2) per work unit + matching scratch; candidate-pair weights
    int density_pairing = 1;    // 0: no matching rounds (every atom a single unit; XTB_DENSITY_PAIRING, experiments)
    bool have_density = false;
    // forces / pressure
    DevBuf d_F, d_SC, d_scratch, d_hdev, d_ppart, d_tab;
    // repulsive pair terms (pettifor_rep.cu)
    DevBuf d_Epot, d_dfpsi, d_Erep, d_Frep, d_SCrep;
    // nonadiabatic coupling (coupling.cu): eigenvectors of the previous step, T = S Ha, G -> Mij, inverse column norms
    DevBuf d_Ha0, d_cpl_T, d_cpl_M, d_cpl_inv;
    bool have_Ha0 = false;
    double coupling_spmm_ms = 0.0, coupling_gemm_ms = 0.0;
    // timing
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_mid = nullptr, ev_mid2 = nullptr;
    int density_slotcap = 0, density_stages = 0;
    double stage_ms[12] = {0};   // 1..7 and 9..11 (8 = eigen_ms), see xtb_stage_ms
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
        if (_e != cudaSuccess)                                                                \
            XTB_FAIL(ctx, XTB_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e),  \
                     __FILE__, __LINE__);                                                     \
    } while (0)

#define XTB_LAUNCH_CHECK(ctx)                                                                 \
    do {                                                                                      \
        (ctx)->launches++;                                                                    \
        cudaError_t _e = cudaGetLastError();                                                  \
        if (_e != cudaSuccess)                                                                \
            XTB_FAIL(ctx, XTB_ECUDA, "kernel launch failed: %s (%s:%d)",                      \
                     cudaGetErrorString(_e), __FILE__, __LINE__);                             \
    } while (0)

int xtb_reserve(xtb_ctx* c, DevBuf& b, size_t bytes);   // grow-only device allocation
void xtb_free(DevBuf& b);

// stage implementations (one translation unit each)
int xtb_impl_neighbours(xtb_ctx* c, double rm);
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
int xtb_impl_mirror_density(xtb_ctx* c);
int xtb_impl_coupling(xtb_ctx* c, int kind_M);      // coupling.cu
int xtb_blas_dgemm(xtb_ctx* c, bool ta, bool tb, int m, int n, int k, const double* A, int lda, const double* B, int ldb, double* Cm, int ldc);
int xtb_impl_eigensolve(xtb_ctx* c);        // eigen.cu: Loewdin + Dsyevd + back-transform on the device-resident H, S
int xtb_impl_dftb_assemble(xtb_ctx* c, bool dense);
int xtb_impl_dftb_forces(xtb_ctx* c);
int xtb_impl_dftb_pressure(xtb_ctx* c);
int xtb_impl_dftb_tables(xtb_ctx* c, double* d_rad, double* d_geo);  // compact tables, nv channels per list entry
int xtb_impl_nrl_tables(xtb_ctx* c, double* d_rad, double* d_geo);   // compact M_V*/M_lmn/M_xrr/M_x1 per list entry
// dense column panels of the owned atoms from a block store; tr: block holds (rows j, cols neighbour)
int xtb_dense_from_blocks(xtb_ctx* c, const double* Ablk, const double* Bblk, double* A, double* B, bool tr, int f0 = -1, int f1 = -1);
