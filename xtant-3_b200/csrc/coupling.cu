// Nonadiabatic electron-ion coupling matrix.
//
// Replaces Electron_ion_coupling_Mij (Source_files/Nonadiabatic.f90:782-945):
//   psi0_i = Ha0(:,i)/|Ha0(:,i)|, psi_j = Ha(:,j)/|Ha(:,j)|  (Euclidean norms, 0 when < 1e-13, :812-826)
//   tij(j,:) = S psi_j                                          (:829-842, only with an overlap matrix)
//   Tully (default arm, :876-938):  Mij(i<j) = 1/2 [ psi0_i . S psi_j - psi0_j . S psi_i ],  Mij(j,i) = Mij(i,j), Mij(i,i) = 0
//   kind_M = -1 (:816-874):         Mij(i<j) = psi0_i . S psi_j,                              Mij(j,i) = Mij(i,j)
// Reference data flow: tij is N^2 dot products of length N against the DENSE Sij (an un-blocked N^3 loop nest), then
// N^2/2 more dot products of length N -- 2 N^3 multiply-adds in scalar loops.
// Here:  T = S Ha with S taken BLOCK-SPARSE from the S2 block store (2*81*B*N flop instead of 2 N^3: 50x fewer at the
//        north-star size), G = Ha0^T T as one library DGEMM (a plain dense GEMM: cuBLAS, bound at run time in eigen.cu),
//        and one in-place tile-pair kernel that applies the norms, antisymmetrises and mirrors (HBM-bound, N^2 doubles
//        read and written once).  Normalisation is applied to G(i,j) instead of to the vectors (same value up to rounding).
#include "xtb_ctx.h"
#include <cstdlib>

namespace {

constexpr int CT = 256;

// inv[n] = 1/|A(:,n)| (0 when the norm is below eps); one CTA per column, fixed-order tree
__global__ void __launch_bounds__(CT) col_invnorm_kernel(long long N, const double* __restrict__ A, double eps, double* __restrict__ inv) {
    __shared__ double buf[CT];
    const double* col = A + (size_t)N * blockIdx.x;
    double s = 0.0;
    for (long long r = threadIdx.x; r < N; r += CT) { const double v = col[r]; s += v * v; }
    buf[threadIdx.x] = s;
    __syncthreads();
    for (int off = CT / 2; off > 0; off >>= 1) {
        if (threadIdx.x < off) buf[threadIdx.x] += buf[threadIdx.x + off];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double nr = sqrt(buf[0]);
        inv[blockIdx.x] = nr < eps ? 0.0 : 1.0 / nr;
    }
}

// T(rows of atom j, n) = sum_{q} blk(j,q)^T C(rows of atom i_q, n);  blk(j,q)(mu,nu) = S(i_q mu, j nu)  (block store of S2).
// CTA = (atom j, 2*ST consecutive states); the atom's blocks are staged once in shared memory (stride 82: 16-byte aligned
// rows), each thread carries two states: 18 accumulators, 81 broadcast LDS per 162 DFMA.
// XT = true: the eigenvector rows come from X[atom][state][NO] (xtb_rows_by_state), so the 32 lanes of a load cover 32
// consecutive states of one atom = 32 * 72 contiguous bytes (8 sectors per load instruction) instead of 32 columns of Ha
// (32 sectors per load instruction, each fetched for 8 of its 32 bytes at a time): the kernel was bound by those wavefronts.
constexpr int ST = 128;
template <int NO, bool XT>
__global__ void __launch_bounds__(ST) spmm_blocks_kernel(int na, long long N, long long K, int cap, const int* __restrict__ nn_size,
                                                        const int* __restrict__ nn_list, const int* __restrict__ rowptr,
                                                        const double* __restrict__ Sblk, const double* __restrict__ Cm,
                                                        double* __restrict__ T) {
    extern __shared__ __align__(16) double sb[];
    constexpr int NB = NO * NO, LD = NB + (NB & 1);
    const int j = blockIdx.x;
    const int m = nn_size[j];
    const double* g = Sblk + (size_t)rowptr[j] * NB;
    for (int e = threadIdx.x; e < (m + 1) * NB; e += ST) sb[(e / NB) * LD + (e % NB)] = g[e];
    __syncthreads();
    const long long n0 = (long long)blockIdx.y * (2 * ST) + threadIdx.x, n1 = n0 + ST;
    const bool ok0 = n0 < K, ok1 = n1 < K;
    if (!ok0) return;
    double a0[NO], a1[NO];
#pragma unroll
    for (int nu = 0; nu < NO; ++nu) { a0[nu] = 0.0; a1[nu] = 0.0; }
    for (int q = 0; q <= m; ++q) {
        const int i = q == 0 ? j : nn_list[(size_t)j * cap + q - 1];
        const double* c0 = XT ? Cm + ((size_t)i * K + n0) * NO : Cm + (size_t)i * NO + (size_t)N * n0;
        const double* c1 = XT ? Cm + ((size_t)i * K + (ok1 ? n1 : n0)) * NO : Cm + (size_t)i * NO + (size_t)N * (ok1 ? n1 : n0);
        double x0[NO], x1[NO];
#pragma unroll
        for (int mu = 0; mu < NO; ++mu) { x0[mu] = __ldg(c0 + mu); x1[mu] = __ldg(c1 + mu); }
        const double* b = sb + q * LD;
#pragma unroll
        for (int nu = 0; nu < NO; ++nu)
#pragma unroll
            for (int mu = 0; mu < NO; ++mu) {
                const double s = b[mu + NO * nu];
                a0[nu] = fma(s, x0[mu], a0[nu]);
                a1[nu] = fma(s, x1[mu], a1[nu]);
            }
    }
    double* t0 = T + (size_t)j * NO + (size_t)N * n0;
#pragma unroll
    for (int nu = 0; nu < NO; ++nu) t0[nu] = a0[nu];
    if (ok1) {
        double* t1 = T + (size_t)j * NO + (size_t)N * n1;
#pragma unroll
        for (int nu = 0; nu < NO; ++nu) t1[nu] = a1[nu];
    }
}

// ---- tiled form of T = S Ha (9 orbitals, rows-by-state eigenvectors) -------------------------------------------------------------
// spmm_blocks_kernel fetches the 72 bytes of (neighbour atom, state) once per column atom j, i.e. ~79 times: 2.25 flop per byte
// from L2, which bounds it at 11 TF.  Here a CTA owns TJ = 4 consecutive atoms (spatial neighbours in any sensible numbering: their
// lists overlap) and walks the UNION of their lists once: a fetched row vector feeds every tile atom that has the union atom as a
// neighbour (2.6 of 4 on the fcc metal).  Measured: no faster than the kernel above (see xtb_impl_coupling); kept as an opt-in variant.
//   tile_union_kernel   one warp per tile: the merged, index-ordered union of {j} U NN(j) of its atoms, and for every union atom
//                       the block of S it forms with each tile atom (or -1)
//   spmm_tile_kernel    CTA = (tile, 256 states), thread = 2 states, 2 x 4 x 9 accumulators; the <= 4 blocks of a union atom are
//                       staged in shared memory (double buffer, rows padded to 10 doubles: 128-bit broadcast loads)
constexpr int TJ = 4;
__global__ void __launch_bounds__(32) tile_union_kernel(int na, int cap, int umax, const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                       const int* __restrict__ rowptr, int* __restrict__ ucount, int* __restrict__ uat,
                                                       int* __restrict__ ublk /*[tile][umax][TJ]*/) {
    const int tile = blockIdx.x, j0 = tile * TJ;
    if (threadIdx.x != 0) return;      // a serial 4-way merge of short sorted lists (once per call and tile: ~1 k tiles)
    int pos[TJ], len[TJ];
    bool self_done[TJ];
    for (int t = 0; t < TJ; ++t) { pos[t] = 0; len[t] = (j0 + t < na) ? nn_size[j0 + t] : 0; self_done[t] = !(j0 + t < na); }
    int* ua = uat + (size_t)tile * umax;
    int* ub = ublk + (size_t)tile * umax * TJ;
    int n = 0;
    for (;;) {
        // smallest pending atom index over the four sequences (list of j with j itself inserted in order)
        int best = 0x7fffffff;
        for (int t = 0; t < TJ; ++t) {
            const int j = j0 + t;
            if (j >= na) continue;
            int head = pos[t] < len[t] ? nn_list[(size_t)j * cap + pos[t]] : 0x7fffffff;
            if (!self_done[t] && j < head) head = j;
            if (head < best) best = head;
        }
        if (best == 0x7fffffff) break;
        ua[n] = best;
        for (int t = 0; t < TJ; ++t) {
            const int j = j0 + t;
            int b = -1;
            if (j < na) {
                if (!self_done[t] && j == best) { b = rowptr[j]; self_done[t] = true; }
                else if (pos[t] < len[t] && nn_list[(size_t)j * cap + pos[t]] == best) { b = rowptr[j] + 1 + pos[t]; ++pos[t]; }
            }
            ub[n * TJ + t] = b;
        }
        ++n;
    }
    ucount[tile] = n;
}

constexpr int TS = 128;               // threads per CTA, two states each
__global__ void __launch_bounds__(TS, 2) spmm_tile_kernel(int na, long long N, long long K, int umax, const int* __restrict__ ucount,
                                                         const int* __restrict__ uat, const int* __restrict__ ublk,
                                                         const double* __restrict__ Sblk, const double* __restrict__ X,
                                                         double* __restrict__ T) {
    constexpr int NO = 9, LDR = 10, BLK = NO * LDR;            // staged block: [nu][10]
    extern __shared__ __align__(16) double sm[];
    double* sS = sm;                                           // [2][TJ][BLK]
    int* sU = reinterpret_cast<int*>(sS + 2 * TJ * BLK);       // [umax] union atoms, then [umax][TJ] block indices
    const int tile = blockIdx.x, j0 = tile * TJ;
    const int nu_ = ucount[tile];
    int* sB = sU + umax;
    for (int e = threadIdx.x; e < nu_; e += TS) sU[e] = uat[(size_t)tile * umax + e];
    for (int e = threadIdx.x; e < nu_ * TJ; e += TS) sB[e] = ublk[(size_t)tile * umax * TJ + e];
    __syncthreads();
    const long long n0 = (long long)blockIdx.y * (2 * TS) + threadIdx.x, n1 = n0 + TS;
    const bool ok0 = n0 < K, ok1 = n1 < K;
    const long long m0 = ok0 ? n0 : 0, m1 = ok1 ? n1 : m0;      // out-of-range threads read a valid row (results dropped) and keep the barriers
    double a0[TJ][NO], a1[TJ][NO];
#pragma unroll
    for (int t = 0; t < TJ; ++t)
#pragma unroll
        for (int v = 0; v < NO; ++v) { a0[t][v] = 0.0; a1[t][v] = 0.0; }
    // stage the blocks of union atom `it` into buffer `buf`: element (mu, nu) of block t -> sS[buf][t][nu * 10 + mu]
    auto stage = [&](int it, int buf) {
        for (int e = threadIdx.x; e < TJ * NO * NO; e += TS) {
            const int t = e / (NO * NO), r = e - t * (NO * NO);
            const int b = sB[it * TJ + t];
            if (b >= 0) { const int v = r / NO, mu = r - v * NO; sS[(buf * TJ + t) * BLK + v * LDR + mu] = __ldg(Sblk + (size_t)b * (NO * NO) + r); }
        }
    };
    if (nu_ > 0) stage(0, 0);
    __syncthreads();
    for (int it = 0; it < nu_; ++it) {
        const int buf = it & 1;
        const int u = sU[it];
        const double* c0 = X + ((size_t)u * K + m0) * NO;
        const double* c1 = X + ((size_t)u * K + m1) * NO;
        double x0[NO], x1[NO];
#pragma unroll
        for (int mu = 0; mu < NO; ++mu) { x0[mu] = __ldg(c0 + mu); x1[mu] = __ldg(c1 + mu); }
        if (it + 1 < nu_) stage(it + 1, buf ^ 1);
#pragma unroll
        for (int t = 0; t < TJ; ++t) {
            if (sB[it * TJ + t] < 0) continue;                   // uniform over the CTA
            const double* b = sS + (buf * TJ + t) * BLK;
#pragma unroll
            for (int v = 0; v < NO; ++v) {
                const double2 s01 = *reinterpret_cast<const double2*>(b + v * LDR), s23 = *reinterpret_cast<const double2*>(b + v * LDR + 2);
                const double2 s45 = *reinterpret_cast<const double2*>(b + v * LDR + 4), s67 = *reinterpret_cast<const double2*>(b + v * LDR + 6);
                const double s8 = b[v * LDR + 8];
                double p = a0[t][v], q = a1[t][v];
                p = fma(s01.x, x0[0], p); q = fma(s01.x, x1[0], q);
                p = fma(s01.y, x0[1], p); q = fma(s01.y, x1[1], q);
                p = fma(s23.x, x0[2], p); q = fma(s23.x, x1[2], q);
                p = fma(s23.y, x0[3], p); q = fma(s23.y, x1[3], q);
                p = fma(s45.x, x0[4], p); q = fma(s45.x, x1[4], q);
                p = fma(s45.y, x0[5], p); q = fma(s45.y, x1[5], q);
                p = fma(s67.x, x0[6], p); q = fma(s67.x, x1[6], q);
                p = fma(s67.y, x0[7], p); q = fma(s67.y, x1[7], q);
                p = fma(s8, x0[8], p);    q = fma(s8, x1[8], q);
                a0[t][v] = p; a1[t][v] = q;
            }
        }
        __syncthreads();                                          // buffer buf^1 is complete, buffer buf may be overwritten next
    }
#pragma unroll
    for (int t = 0; t < TJ; ++t) {
        const int j = j0 + t;
        if (j >= na) continue;
        if (ok0) { double* t0 = T + (size_t)j * NO + (size_t)N * n0;
#pragma unroll
            for (int v = 0; v < NO; ++v) t0[v] = a0[t][v]; }
        if (ok1) { double* t1 = T + (size_t)j * NO + (size_t)N * n1;
#pragma unroll
            for (int v = 0; v < NO; ++v) t1[v] = a1[t][v]; }
    }
}

// in place on G (N x N, column-major): tile pair (a, b), a <= b, 32 x 32 tiles through shared memory.
//   g(i,j) = G(i,j) inv0[i] inv1[j];  Tully: M(i,j) = M(j,i) = (g(i,j) - g(j,i))/2 for i < j;  kind -1: M(i,j) = M(j,i) = g(i,j);  M(i,i) = 0
__global__ void __launch_bounds__(256) antisym_kernel(long long N, int kind, const double* __restrict__ inv0,
                                                     const double* __restrict__ inv1, double* __restrict__ G) {
    __shared__ double U[32][33], L[32][33];
    const int a = blockIdx.x, b = blockIdx.y;                     // tile pair, a <= b
    if (a > b) return;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;       // 32 x 8
    const long long r0 = 32LL * a, c0 = 32LL * b;
    for (int yy = ty; yy < 32; yy += 8) {
        const long long iu = r0 + tx, ju = c0 + yy;                // upper tile element G(iu, ju)
        U[yy][tx] = (iu < N && ju < N) ? G[iu + N * ju] * inv0[iu] * inv1[ju] : 0.0;
        const long long il = c0 + tx, jl = r0 + yy;                // lower tile element G(il, jl)
        L[yy][tx] = (il < N && jl < N) ? G[il + N * jl] * inv0[il] * inv1[jl] : 0.0;
    }
    __syncthreads();
    for (int yy = ty; yy < 32; yy += 8) {
        // upper tile position (i = r0 + tx, j = c0 + yy): g(i,j) = U[yy][tx], g(j,i) = L[tx][yy]
        {
            const long long i = r0 + tx, j = c0 + yy;
            if (i < N && j < N) {
                double v;
                if (i == j) v = 0.0;
                else if (i < j) v = kind == -1 ? U[yy][tx] : 0.5 * (U[yy][tx] - L[tx][yy]);
                else v = kind == -1 ? L[tx][yy] : 0.5 * (L[tx][yy] - U[yy][tx]);      // diagonal tile, below the diagonal
                G[i + N * j] = v;
            }
        }
        if (a != b) {   // lower tile position (i = c0 + tx, j = r0 + yy), always i > j: M(i,j) = M(j,i), pair (j,i) sits in U[tx][yy]
            const long long i = c0 + tx, j = r0 + yy;
            if (i < N && j < N) G[i + N * j] = kind == -1 ? U[tx][yy] : 0.5 * (U[tx][yy] - L[yy][tx]);
        }
    }
}

}  // namespace

int xtb_impl_coupling(xtb_ctx* c, int kind_M) {
    { int rc0 = xtb_ensure_Ha_std(c); if (rc0) return rc0; }      // the coupling reads the Fortran-layout Ha
    const long long N = (long long)c->na * c->n_orb, K = c->K;
    if (K != N) XTB_FAIL(c, XTB_EINVAL, "xtb_coupling: needs all N = %lld states (have %lld)", N, K);
    if (N > 46000) XTB_FAIL(c, XTB_EINVAL, "xtb_coupling: N = %lld exceeds the 32-bit leading dimension of cublasDgemm", N);
    const bool ortho = c->tb == XTB_TB_PETTIFOR;
    const size_t nn = (size_t)N * N;
    int rc;
    if ((rc = xtb_reserve(c, c->d_cpl_M, sizeof(double) * nn))) return rc;
    if ((rc = xtb_reserve(c, c->d_cpl_inv, sizeof(double) * 2 * N))) return rc;
    if (!ortho && (rc = xtb_reserve(c, c->d_cpl_T, sizeof(double) * nn))) return rc;
    double* inv0 = (double*)c->d_cpl_inv.p; double* inv1 = inv0 + N;
    const double* Ha0 = (const double*)c->d_Ha0.p;
    XTB_T0(c, 11); XTB_T0(c, 12);
    col_invnorm_kernel<<<(unsigned)N, CT, 0, c->stream>>>(N, Ha0, 1.0e-13, inv0);        // Norm0, eps  :812,:883
    XTB_LAUNCH_CHECK(c);
    col_invnorm_kernel<<<(unsigned)N, CT, 0, c->stream>>>(N, c->d_Ha, 1.0e-13, inv1);    // Norm1       :884
    XTB_LAUNCH_CHECK(c);
    const double* Tm = c->d_Ha;                                    // orthogonal basis: S = 1
    if (!ortho) {
        const dim3 grid((unsigned)c->na, (unsigned)((K + 2 * ST - 1) / (2 * ST)));
        const size_t smem = sizeof(double) * (size_t)(c->max_nn + 1) * 82;
        // eigenvector rows by state (one HBM-bound transposition, 2 * 8 * N * K bytes) unless XTB_COUPLING_XT=0
        const char* ev = getenv("XTB_COUPLING_XT");
        const bool xt = ev ? atoi(ev) != 0 : true;
        const double* Cm = c->d_Ha;
        if (xt) {
            if ((rc = xtb_reserve(c, c->d_Ct, sizeof(double) * (size_t)N * K + 1024))) return rc;      // shares the density stage's scratch
            if ((rc = xtb_rows_by_state(c, c->d_Ha, K, (double*)c->d_Ct.p))) return rc;
            Cm = (const double*)c->d_Ct.p;
        }
        const char* evt = getenv("XTB_COUPLING_TILE");
        // XTB_COUPLING_TILE=1: four column atoms per CTA on the union of their lists (spmm_tile_kernel).  Parity-green and NOT
        // faster at au4096 (172.8 vs 169.5 ms for norms + S Ha, profiles/r02as_coupling_tiled_ab.json), so one column atom per CTA
        // (spmm_blocks_kernel) stays the default: the stage is not bound by the L2 fetches the tiling removes.
        const bool tiled = xt && c->n_orb == 9 && evt && atoi(evt) == 1;
        if (tiled) {
            const int ntile = (c->na + TJ - 1) / TJ, umax = TJ * (c->max_nn + 1);
            if ((rc = xtb_reserve(c, c->d_cpl_union, sizeof(int) * ((size_t)ntile * (1 + (size_t)umax * (1 + TJ)))))) return rc;
            int* ucount = (int*)c->d_cpl_union.p; int* uat = ucount + ntile; int* ublk = uat + (size_t)ntile * umax;
            tile_union_kernel<<<ntile, 32, 0, c->stream>>>(c->na, c->cap, umax, (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p,
                                                           (const int*)c->d_rowptr.p, ucount, uat, ublk);
            XTB_LAUNCH_CHECK(c);
            const size_t smt = sizeof(double) * 2 * TJ * 90 + sizeof(int) * (size_t)umax * (1 + TJ);
            XTB_CUDA(c, cudaFuncSetAttribute(spmm_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smt));
            spmm_tile_kernel<<<dim3((unsigned)ntile, (unsigned)((K + 2 * TS - 1) / (2 * TS))), TS, smt, c->stream>>>(
                c->na, N, K, umax, ucount, uat, ublk, (const double*)c->d_Sblk.p, Cm, (double*)c->d_cpl_T.p);
        } else
#define XTB_SPMM(NO, XT) { XTB_CUDA(c, cudaFuncSetAttribute(spmm_blocks_kernel<NO, XT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        spmm_blocks_kernel<NO, XT><<<grid, ST, smem, c->stream>>>(c->na, N, K, c->cap, (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p, \
                                                                  (const int*)c->d_rowptr.p, (const double*)c->d_Sblk.p, Cm, (double*)c->d_cpl_T.p); }
        {
        if (c->n_orb == 9) { if (xt) XTB_SPMM(9, true) else XTB_SPMM(9, false) }
        else if (c->n_orb == 4) { if (xt) XTB_SPMM(4, true) else XTB_SPMM(4, false) }
        else { if (xt) XTB_SPMM(1, true) else XTB_SPMM(1, false) }
        }
#undef XTB_SPMM
        XTB_LAUNCH_CHECK(c);
        Tm = (const double*)c->d_cpl_T.p;
    }
    XTB_T1(c, 12); XTB_T0(c, 13);
    if ((rc = xtb_blas_dgemm(c, true, false, (int)N, (int)N, (int)N, Ha0, (int)N, Tm, (int)N, (double*)c->d_cpl_M.p, (int)N))) return rc;   // G = Ha0^T T
    XTB_T1(c, 13);
    const unsigned nt = (unsigned)((N + 31) / 32);
    antisym_kernel<<<dim3(nt, nt), 256, 0, c->stream>>>(N, kind_M == -1 ? -1 : 0, inv0, inv1, (double*)c->d_cpl_M.p);
    XTB_LAUNCH_CHECK(c);
    XTB_T1(c, 11);
    return xtb_stage_finish(c);
}
