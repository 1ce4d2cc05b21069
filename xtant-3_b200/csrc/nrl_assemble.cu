// S2 -- NRL radial functions + Slater-Koster H/S block assembly + dense panels.
//
// Replaces Construct_Vij_NRL_OMP (Source_files/TB_NRL.f90:2116-2244), Hamil_tot_NRL's assembly part
// (:226-397) with Hamilton_one_NRL (:1798-1856), rho_NRL (:2248-2287), Get_overlap_S_matrix (:1724-1744)
// and KS_sp3d5 (TB_Koster_Slater.f90:5234-5348).
//
// Reference data flow: dense M_Vij/M_SVij (Na,Na,10) tables -> 9x9 blocks scattered into zero-filled
// dense Hij,Sij (N,N) -> mirror -> *g_Ry.  Here:
//   k1  nrl_dens_kernel    one warp per atom: embedded density rho_j (list order, as the reference sums)
//   k2  nrl_blocks_kernel  one thread per (unordered pair, matrix): radial functions in registers, SK rotation in
//                          registers, 81 entries to the block-sparse store Hblk/Sblk (block (j,q) holds
//                          rows of neighbour i, columns of outer atom j -- the reference placement) and their
//                          transpose to the partner's block; on one rank also the same values into the
//                          RESIDENT dense H_non / Sij (xtb_dense_update prepares: dense_zero_prev_kernel clears
//                          the blocks the new pattern drops, dense_remember_kernel records the pattern)
//   k3  dense_scatter_kernel  the dense values from the block stores (other branches, panels of other ranks,
//                          XTB_DENSE_FUSED=0); dense_panel_kernel (XTB_DENSE_FULL=1): one CTA per owned atom j
//                          writes the 9 dense columns of j for H and S, every element exactly once (zeros
//                          included), streaming stores: 2*8*N*(9*n_owned) bytes at 0.87 of the HBM roofline.
// Semantics kept: pair (a>b) evaluated once with TB(kind(b),kind(a)) and r from the list of b (:2151-2159);
// block (rows a, cols b) uses the direction cosines of a's list entry for b (M_lmn(:,a,b), :1849);
// the other triangle is the element transpose (:369-390); |S| <= 1e-12 -> 0 (:359); H *= g_Ry (:394).
#include "nrl_device.cuh"
#include "sk_device.cuh"

namespace {

// one warp per atom: the lanes evaluate the envelope of the list entries (two exp each), lane 0 adds them in list order
// (the reference's summation order; one thread per atom with the exp chain inside the loop took 77 us at 4096 atoms)
constexpr int ND_WARPS = 4;
__global__ void __launch_bounds__(32 * ND_WARPS) nrl_dens_kernel(int na, int cap, int mx, int n_kinds, const int* __restrict__ koa,
                                                                const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                                const double* __restrict__ nn_dist,
                                                                const xtb_nrl_pair* __restrict__ TB, double* __restrict__ dens) {
    extern __shared__ double xs[];                  // [ND_WARPS][mx]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int j = blockIdx.x * ND_WARPS + warp;
    if (j >= na) return;
    double* x = xs + (size_t)warp * mx;
    const int m = nn_size[j];
    for (int q = lane; q < m; q += 32) {
        const size_t e = (size_t)j * cap + q;
        const int i = nn_list[e];
        const xtb_nrl_pair& P = TB[koa[j] + n_kinds * koa[i]];
        const double r = nn_dist[4 * e + 3] * XTB_G_A2AU;
        double X, dX;
        nrl::envelope(P, r, X, dX);
        x[q] = X;
    }
    __syncwarp();
    if (lane == 0) {
        double rho = 0.0;
        for (int q = 0; q < m; ++q) rho += x[q];
        dens[j] = rho;
    }
}

// one thread per (pair, mat) does the arithmetic (radial functions, SK rotation: all in registers).
// The two blocks of an unordered pair -- (rows i, cols j) in the panel of j and (rows j, cols i) in the panel of i -- are
// element-wise transposes of ONE computed block C (rows = larger index, cols = smaller; same radial values, same direction
// cosines), so the pair is evaluated once -- by the atom that owns it in the density pass (own[e]: half of the neighbours of
// every atom, equal work per CTA) when the partner belongs to this rank's range too, else by the atom of the range -- and
// written to both places.  One CTA per atom: warp 0 compacts the entries the atom evaluates (list order), then every thread
// takes one (entry, matrix) item -- item 0/1 = the on-site block -- and leaves the packed upper triangle of C (45 values; the
// lower triangle is +-upper, sk::lower_sign) in its row of a shared-memory staging area; each warp then writes the blocks of
// its 32 threads out cooperatively, expanding the triangle on the way: every store instruction covers 256 contiguous bytes
// of a block (a thread writing its own 648-byte block touched 32 sectors per store).
constexpr int NB_THREADS = 96;        // ~ (39 owned pairs + on-site) x 2 matrices on an fcc metal: one pass
constexpr int NB_UP = 45;
enum { NBK_NONE = 0, NBK_DIRECT = 1, NBK_TRANSP = 2, NBK_PLAIN = 3 };   // stored block = C | C^T | symmetric without signs (on-site, zero)

// packed upper triangle of the block of entry (j, q-1) / of the on-site block (q == 0); returns the kind of the stored block (rows i, cols j)
__device__ __forceinline__ int nrl_block_upper(int j, int q, int mat, int i, int rv, size_t e, int cap, int n_kinds, const int* __restrict__ koa,
                                               const double* __restrict__ nn_dist, const xtb_nrl_pair* __restrict__ TB,
                                               const double* __restrict__ dens, double* __restrict__ out /*[45]*/) {
    if (q == 0) {   // on-site: diag(h_s, h_p x3, h_t2g x3, h_eg x2) * g_Ry ; S = I
        const xtb_nrl_pair& P = TB[koa[j] + n_kinds * koa[j]];
        const double rho = dens[j];
        double hv[4];
#pragma unroll
        for (int s = 0; s < 4; ++s) hv[s] = nrl::onsite(rho, P.al[s], P.bl[s], P.cl[s], P.dl[s]) * XTB_G_RY;
        for (int x = 0; x < NB_UP; ++x) out[x] = 0.0;
        for (int o = 0; o < 9; ++o) out[sk::ut(o, o)] = mat == 0 ? hv[nrl::onsite_set(o, P.ind_split)] : 1.0;
        return NBK_PLAIN;
    }
    if (rv < 0) {   // asymmetric list: the reference's dense tables hold zeros there
        for (int x = 0; x < NB_UP; ++x) out[x] = 0.0;
        return NBK_PLAIN;
    }
    const size_t ei = (size_t)i * cap + rv;      // entry of j in the list of i
    // radial: unique pair evaluated from the smaller index' entry with TB(kind(small), kind(large))
    const int small = i < j ? i : j, large = i < j ? j : i;
    const double r = nn_dist[4 * (i < j ? ei : e) + 3] * XTB_G_A2AU;
    const xtb_nrl_pair& P = TB[koa[small] + n_kinds * koa[large]];
    double V[10], dV[10];
    nrl::radial10(P, mat, r, V, dV);
    // direction cosines: the entry of the LARGER index atom (rows) for the smaller (cols)
    const size_t el = (i > j) ? ei : e;
    const double4 d4 = *reinterpret_cast<const double4*>(nn_dist + 4 * el);
    const double l = d4.x / d4.w, m = d4.y / d4.w, n = d4.z / d4.w;
    double up[NB_UP];
    sk::block_sp3d5_upper<double>(V, l, m, n, up);
    const double scale = mat == 0 ? XTB_G_RY : 1.0;
#pragma unroll
    for (int x = 0; x < NB_UP; ++x) {
        double u = up[x];
        if (mat == 1 && fabs(u) <= 1e-12) u = 0.0;
        out[x] = u * scale;
    }
    // computed block C has rows = larger atom, cols = smaller atom.  Stored block = (rows i, cols j): i > j: C ; i < j: C^T.
    return i > j ? NBK_DIRECT : NBK_TRANSP;
}

__global__ void __launch_bounds__(NB_THREADS, 6) nrl_blocks_kernel(int na, int a0, int a1, int cap, int mx, int n_kinds,
                                                               const int* __restrict__ koa, const int* __restrict__ nn_size,
                                                               const int* __restrict__ nn_list, const int* __restrict__ rev,
                                                               const signed char* __restrict__ own,
                                                               const double* __restrict__ nn_dist, const int* __restrict__ rowptr,
                                                               const xtb_nrl_pair* __restrict__ TB, const double* __restrict__ dens,
                                                               double* __restrict__ Hblk, double* __restrict__ Sblk,
                                                               size_t N, double* __restrict__ DH, double* __restrict__ DS) {
    // DH / DS != null: the resident dense H_non / Sij take the same values in the same pass (element (row atom, mu; col atom, nu)
    // at (9 row + mu) + N (9 col + nu)): the separate scatter kernel would read all blocks back (0.42 GB at 4096 atoms)
    extern __shared__ double stage_out[];          // [NB_THREADS][45], then int actq[mx]
    int* actq = reinterpret_cast<int*>(stage_out + NB_THREADS * NB_UP);
    // element e = a + 9 b of C: index of its value in the packed upper triangle | 0x40 if C(a,b) is the negated value
    // (a > b, odd l_a + l_b) | 0x80 if C(b,a) is
    __shared__ unsigned char tab[81];
    __shared__ int nact_sh;
    const int j = a0 + blockIdx.x;
    const int m = nn_size[j];
    if (threadIdx.x >= 32 && threadIdx.x - 32 < 81) {
        const int x = threadIdx.x - 32;
        const int a = x % 9, b = x / 9;
        const int lo = a < b ? a : b, hi = a < b ? b : a;
        const bool odd = sk::lower_sign(lo, hi) < 0.0;
        tab[x] = (unsigned char)(sk::ut(lo, hi) | ((a > b && odd) ? 0x40 : 0) | ((b > a && odd) ? 0x80 : 0));
    }
    if (NB_THREADS - 32 < 81 && threadIdx.x < 81 - (NB_THREADS - 32)) {      // the entries the 64 upper threads did not reach
        const int x = threadIdx.x + (NB_THREADS - 32);
        const int a = x % 9, b = x / 9;
        const int lo = a < b ? a : b, hi = a < b ? b : a;
        const bool odd = sk::lower_sign(lo, hi) < 0.0;
        tab[x] = (unsigned char)(sk::ut(lo, hi) | ((a > b && odd) ? 0x40 : 0) | ((b > a && odd) ? 0x80 : 0));
    }
    if (threadIdx.x < 32) {      // ordered compaction (warp 0) of the entries whose pair this atom evaluates
        int cnt = 0;
        for (int base = 0; base < m; base += 32) {
            const int q = base + threadIdx.x;
            bool f = false;
            if (q < m) {
                const size_t e = (size_t)j * cap + q;
                const int i = nn_list[e];
                f = !(own[e] == 0 && rev[e] >= 0 && i >= a0 && i < a1);      // else the partner (owner of the pair, in range) writes both blocks
            }
            const unsigned mask = __ballot_sync(0xffffffffu, f);
            if (f) actq[cnt + __popc(mask & ((1u << threadIdx.x) - 1u))] = q;
            cnt += __popc(mask);
        }
        if (threadIdx.x == 0) nact_sh = cnt;
    }
    __syncthreads();
    const int nitems = 2 * (nact_sh + 1);          // (on-site | evaluated entry) x (H | S)
    const int lane = threadIdx.x & 31;
    const double* wsrc = stage_out + (threadIdx.x - lane) * NB_UP;
    unsigned tb[3];
    tb[0] = tab[lane]; tb[1] = tab[lane + 32]; tb[2] = lane < 17 ? tab[lane + 64] : 0;
    size_t doff[3];                  // dense offset of this lane's elements e = lane + 32 p = mu + 9 nu inside a block
#pragma unroll
    for (int p = 0; p < 3; ++p) { const int e = lane + 32 * p; doff[p] = (size_t)(e % 9) + N * (size_t)(e / 9); }
    for (int w0 = 0; w0 < nitems; w0 += NB_THREADS) {      // (uniform per CTA; the staging rows are per warp)
        const int w = w0 + threadIdx.x;
        double* dst = nullptr;       // the block (j, q)
        double* dst2 = nullptr;      // the partner's block of the same pair (the transpose), when the partner is in range
        int kind = NBK_NONE, irow = j;
        if (w < nitems) {
            const int mat = w & 1, it = w >> 1;
            double* base = mat == 0 ? Hblk : Sblk;
            int q = 0, i = j, rv = 0;
            size_t e = 0;
            if (it > 0) {
                q = actq[it - 1] + 1;
                e = (size_t)j * cap + (q - 1);
                i = nn_list[e];
                rv = rev[e];
                if (rv >= 0 && i >= a0 && i < a1) dst2 = base + ((size_t)rowptr[i] + rv + 1) * 81;
            }
            dst = base + ((size_t)rowptr[j] + q) * 81;
            irow = i;
            kind = nrl_block_upper(j, q, mat, i, rv, e, cap, n_kinds, koa, nn_dist, TB, dens, stage_out + threadIdx.x * NB_UP);
        }
        __syncwarp();
#pragma unroll 2
        for (int s = 0; s < 32; ++s) {
            const int kd = __shfl_sync(0xffffffffu, kind, s);
            if (kd == NBK_NONE) continue;
            double* d1 = reinterpret_cast<double*>(__shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(dst), s));
            double* d2 = reinterpret_cast<double*>(__shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(dst2), s));
            const double* src = wsrc + s * NB_UP;
            const unsigned m1 = kd == NBK_DIRECT ? 0x40u : (kd == NBK_TRANSP ? 0x80u : 0u);    // sign bit of the stored block (j, q)
            const unsigned m2 = kd == NBK_DIRECT ? 0x80u : 0x40u;                              // and of the partner's (d2 != null only for pair blocks)
            // dense destinations: the staged items of a warp alternate H, S (item = 2 * entry + mat; passes start at even items)
            double* D = (s & 1) ? DS : DH;
            double* D1 = nullptr; double* D2 = nullptr;
            if (D) {
                const int ir = __shfl_sync(0xffffffffu, irow, s);
                D1 = D + (size_t)ir * 9 + N * ((size_t)j * 9);
                D2 = D + (size_t)j * 9 + N * ((size_t)ir * 9);
            }
#pragma unroll
            for (int p = 0; p < 3; ++p) {
                if (p == 2 && lane >= 17) break;
                const double v = src[tb[p] & 63u];
                const double v1 = (tb[p] & m1) ? v * -1.0 : v;
                d1[lane + 32 * p] = v1;
                if (D1) D1[doff[p]] = v1;
                if (d2) {
                    const double v2 = (tb[p] & m2) ? v * -1.0 : v;
                    d2[lane + 32 * p] = v2;
                    if (D2) D2[doff[p]] = v2;
                }
            }
        }
        __syncwarp();                // the staging rows are rewritten by the next pass
    }
}

}  // namespace

// Generic: dense column panels of the owned atoms from a block-sparse store.
// blk(b)[mu + no*nu]: mu = orbital of the ROW atom (neighbour i, or j itself for the first block), nu = orbital of j.
template <int NO, bool TR>
__global__ void __launch_bounds__(256) dense_panel_kernel(int na, int a0, int cap, const int* __restrict__ nn_size,
                                                         const int* __restrict__ nn_list, const int* __restrict__ rowptr,
                                                         const double* __restrict__ Ablk, const double* __restrict__ Bblk,
                                                         double* __restrict__ A, double* __restrict__ B) {
    extern __shared__ int sm_i[];
    int* nbl = sm_i;             // [cap]
    int* slot = sm_i + cap;      // [256]
    const int j = a0 + blockIdx.x;
    const int m = nn_size[j];
    for (int q = threadIdx.x; q < m; q += blockDim.x) nbl[q] = nn_list[(size_t)j * cap + q];
    __syncthreads();
    const size_t N = (size_t)na * NO;
    const size_t b0 = (size_t)rowptr[j];
    constexpr int TA = 256;      // row atoms per tile
    for (int tile = 0; tile < na; tile += TA) {
        {
            const int ia = tile + threadIdx.x;
            int s = -1;
            if (ia == j) s = 0;
            else if (ia < na) {
                int lo = 0, hi = m - 1;
                while (lo <= hi) {
                    const int mid = (lo + hi) >> 1;
                    const int v = nbl[mid];
                    if (v == ia) { s = mid + 1; break; }
                    if (v < ia) lo = mid + 1; else hi = mid - 1;
                }
            }
            slot[threadIdx.x] = s;
        }
        __syncthreads();
        const int nat = min(TA, na - tile);
        const int nrow = nat * NO;
        for (int c = 0; c < NO; ++c) {
            const size_t col = ((size_t)j * NO + c) * N + (size_t)tile * NO;
            for (int e = threadIdx.x; e < nrow; e += blockDim.x) {
                const int il = e / NO, orb = e - il * NO;
                const int s = slot[il];
                double va = 0.0, vb = 0.0;
                if (s >= 0) {
                    const size_t off = (b0 + s) * (NO * NO) + (TR ? c + NO * orb : orb + NO * c);
                    va = Ablk[off];
                    if (Bblk) vb = Bblk[off];
                }
                __stcs(A + col + e, va);
                if (B) __stcs(B + col + e, vb);
            }
        }
        __syncthreads();
    }
}

template <int NO>
static void launch_panel(xtb_ctx* c, int f0, int nown, size_t smem, const double* Ablk, const double* Bblk, double* A, double* B, bool tr) {
    if (tr)
        dense_panel_kernel<NO, true><<<nown, 256, smem, c->stream>>>(c->na, f0, c->cap, (const int*)c->d_nn_size.p,
                                                                     (const int*)c->d_nn_list.p, (const int*)c->d_rowptr.p, Ablk, Bblk, A, B);
    else
        dense_panel_kernel<NO, false><<<nown, 256, smem, c->stream>>>(c->na, f0, c->cap, (const int*)c->d_nn_size.p,
                                                                      (const int*)c->d_nn_list.p, (const int*)c->d_rowptr.p, Ablk, Bblk, A, B);
}

int xtb_dense_from_blocks(xtb_ctx* c, const double* Ablk, const double* Bblk, double* A, double* B, bool tr, int f0, int f1) {
    if (f0 < 0) { f0 = c->a0; f1 = c->a1; }      // default: the owned atoms
    const int nown = f1 - f0;
    if (nown <= 0) return XTB_OK;
    const size_t smem = sizeof(int) * (c->cap + 256);
    if (c->n_orb == 9) launch_panel<9>(c, f0, nown, smem, Ablk, Bblk, A, B, tr);
    else if (c->n_orb == 4) launch_panel<4>(c, f0, nown, smem, Ablk, Bblk, A, B, tr);
    else if (c->n_orb == 1) launch_panel<1>(c, f0, nown, smem, Ablk, Bblk, A, B, tr);
    else XTB_FAIL(c, XTB_EINVAL, "unsupported n_orb %d", c->n_orb);
    XTB_LAUNCH_CHECK(c);
    return XTB_OK;
}

// ---- dense H_non / Sij kept RESIDENT and updated in place -----------------------------------------------------------------
// The dense matrices are 2 * 8 * N^2 bytes (21.7 GB at the north-star size) of which 2 % are non-zero (the blocks of the
// neighbour pattern).  Writing every element every step costs 3.8 ms at the HBM roofline -- on every rank of a multi-GPU run
// with a replicated eigensolve.  Instead the matrices stay in HBM from step to step together with the pattern they were
// written with; a step clears the blocks of the PREVIOUS pattern and writes the blocks of the new one (2 * 2 * 8 * n_orb^2
// bytes per block: 0.84 GB), so every element still has exactly the value a full write would give it.  A full clear happens
// only when the buffers are new, were used as scratch (xtb_density_get), or the system changed shape.  XTB_DENSE_FULL=1
// restores the full column-panel write of every step (dense_panel_kernel; the HBM-roofline measurement of bench.py).
// nsize / nlist (stride ncap): the pattern the scatter that follows will write.  A block of the previous pattern that is also in the
// new one is overwritten in full by that scatter, so it is not cleared here (in a steady MD run that is every block: the kernel
// then only compares lists).  Membership is a binary search in the index-ordered list of the column atom; a miss on an unordered
// list merely clears a block that is rewritten anyway.  nsize = nullptr: clear every block of the previous pattern.
template <int NO>
__global__ void __launch_bounds__(256) dense_zero_prev_kernel(int f0, int pcap, const int* __restrict__ psize, const int* __restrict__ plist,
                                                             size_t N, double* __restrict__ A, double* __restrict__ B,
                                                             int ncap, const int* __restrict__ nsize, const int* __restrict__ nlist) {
    extern __shared__ unsigned char keep[];              // [pcap + 1]: 1 = the block stays (the scatter rewrites it)
    const int j = f0 + blockIdx.x;
    const int m = psize[j];
    if (m < 0) return;                                   // no previous content in this column panel
    int any = 1;
    if (nsize) {
        const int mn = nsize[j];
        const int* row = nlist + (size_t)j * ncap;
        int cleared = 0;
        for (int q = threadIdx.x; q <= m; q += 256) {
            unsigned char k = 1;                         // q == 0: the on-site block, part of every pattern
            if (q > 0) {
                const int i = plist[(size_t)j * pcap + q - 1];
                int lo = 0, hi = mn - 1;
                k = 0;
                while (lo <= hi) {
                    const int mid = (lo + hi) >> 1, v = row[mid];
                    if (v == i) { k = 1; break; }
                    if (v < i) lo = mid + 1; else hi = mid - 1;
                }
            }
            keep[q] = k;
            cleared |= !k;
        }
        any = __syncthreads_or(cleared);
    }
    if (!any) return;
    const int tot = (m + 1) * NO * NO;
    for (int idx = threadIdx.x; idx < tot; idx += 256) {
        const int q = idx / (NO * NO), e = idx - q * (NO * NO);
        if (nsize && keep[q]) continue;
        const int nu = e / NO, mu = e - nu * NO;
        const int i = q == 0 ? j : plist[(size_t)j * pcap + q - 1];
        const size_t d = ((size_t)i * NO + mu) + N * ((size_t)j * NO + nu);
        A[d] = 0.0;
        if (B) B[d] = 0.0;
    }
}
template <int NO>
__global__ void __launch_bounds__(256) dense_scatter_kernel(int f0, int cap, const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                           const int* __restrict__ rowptr, const double* __restrict__ Ablk,
                                                           const double* __restrict__ Bblk, size_t N, double* __restrict__ A, double* __restrict__ B,
                                                           int pcap, int* __restrict__ psize, int* __restrict__ plist) {
    extern __shared__ int rowat[];                       // [m + 1] row atom of block q (q = 0: j itself)
    const int j = f0 + blockIdx.x;
    const int m = nn_size[j];
    const int tot = (m + 1) * NO * NO;
    const size_t b0 = (size_t)rowptr[j] * (NO * NO);
    for (int q = threadIdx.x; q <= m; q += 256) rowat[q] = q == 0 ? j : nn_list[(size_t)j * cap + q - 1];
    __syncthreads();
    // four independent (load, load, store, store) groups per thread in flight: the kernel is bound by load latency otherwise
#pragma unroll 4
    for (int idx = threadIdx.x; idx < tot; idx += 256) {
        const int q = idx / (NO * NO), e = idx - q * (NO * NO);
        const int nu = e / NO, mu = e - nu * NO;
        const size_t d = ((size_t)rowat[q] * NO + mu) + N * ((size_t)j * NO + nu);
        const double va = __ldcs(Ablk + b0 + idx);
        A[d] = va;
        if (B) B[d] = __ldcs(Bblk + b0 + idx);
    }
    // remember the pattern this panel now holds
    for (int q = threadIdx.x; q < m; q += 256) plist[(size_t)j * pcap + q] = nn_list[(size_t)j * cap + q];
    if (threadIdx.x == 0) psize[j] = m;
}
// the pattern a panel holds after a pass that wrote its blocks elsewhere (nrl_blocks_kernel with dense destinations)
__global__ void dense_remember_kernel(int f0, int cap, const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                      int pcap, int* __restrict__ psize, int* __restrict__ plist) {
    const int j = f0 + blockIdx.x;
    const int m = nn_size[j];
    for (int q = threadIdx.x; q < m; q += blockDim.x) plist[(size_t)j * pcap + q] = nn_list[(size_t)j * cap + q];
    if (threadIdx.x == 0) psize[j] = m;
}
__global__ void fill_int_kernel(int n, int v, int* __restrict__ p) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// dense column panels of atoms [f0, f1) (0-based) of H_non (and Sij unless orthogonal) from the block stores
int xtb_dense_update(xtb_ctx* c, int f0, int f1, bool scatter) {
    const bool ortho = c->tb == XTB_TB_PETTIFOR;
    const int na = c->na, NO = c->n_orb;
    const size_t N = (size_t)na * NO;
    int rc;
    const void* pH = c->d_H.p; const void* pS = c->d_S2.p;
    if ((rc = xtb_reserve(c, c->d_H, sizeof(double) * N * N))) return rc;
    if (!ortho && (rc = xtb_reserve(c, c->d_S2, sizeof(double) * N * N))) return rc;
    if (f1 <= f0) return XTB_OK;
    double* A = (double*)c->d_H.p; double* B = ortho ? nullptr : (double*)c->d_S2.p;
    const double* Ablk = (const double*)c->d_Hblk.p; const double* Bblk = ortho ? nullptr : (const double*)c->d_Sblk.p;
    const char* ev = getenv("XTB_DENSE_FULL");
    if (ev && atoi(ev) != 0) {       // every element of the panels, every step
        c->dense_state = 0;
        return xtb_dense_from_blocks(c, Ablk, Bblk, A, B, false, f0, f1);
    }
    // (xtb_reserve zero-fills what it allocates: a new buffer is clean and empty)
    const bool fresh = c->d_H.p != pH || (!ortho && c->d_S2.p != pS);
    const bool shape = c->dense_na != na || c->dense_norb != NO || c->dense_ortho != (int)ortho;
    if ((rc = xtb_reserve(c, c->d_prev_size, sizeof(int) * na))) return rc;
    if (fresh || shape || c->dense_state == 0) {
        if (!(fresh && c->d_H.p != pH && (ortho || c->d_S2.p != pS))) {          // not both brand new: clear what is there
            XTB_CUDA(c, cudaMemsetAsync(A, 0, sizeof(double) * N * N, c->stream));
            if (B) XTB_CUDA(c, cudaMemsetAsync(B, 0, sizeof(double) * N * N, c->stream));
        }
        fill_int_kernel<<<(na + 255) / 256, 256, 0, c->stream>>>(na, -1, (int*)c->d_prev_size.p);
        XTB_LAUNCH_CHECK(c);
        c->dense_pcap = 0;
        c->dense_na = na; c->dense_norb = NO; c->dense_ortho = (int)ortho; c->dense_state = 1;
    }
    if (c->dense_pcap != c->cap) {
        if (c->dense_pcap > 0) {      // the list stride changed: clear every panel with the old layout, then adopt the new one
#define XTB_ZERO(NOV) dense_zero_prev_kernel<NOV><<<na, 256, 0, c->stream>>>(0, c->dense_pcap, (const int*)c->d_prev_size.p, (const int*)c->d_prev_list.p, N, A, B, 0, nullptr, nullptr)
            if (NO == 9) XTB_ZERO(9); else if (NO == 4) XTB_ZERO(4); else XTB_ZERO(1);
            XTB_LAUNCH_CHECK(c);
            fill_int_kernel<<<(na + 255) / 256, 256, 0, c->stream>>>(na, -1, (int*)c->d_prev_size.p);
            XTB_LAUNCH_CHECK(c);
        }
        if ((rc = xtb_reserve(c, c->d_prev_list, sizeof(int) * (size_t)na * c->cap))) return rc;
        c->dense_pcap = c->cap;
    }
    const int n = f1 - f0;
    // blocks of the previous pattern that the new pattern drops (XTB_DENSE_ZERO_ALL=1: every previous block, the earlier form)
    const bool diff = !(getenv("XTB_DENSE_ZERO_ALL") && atoi(getenv("XTB_DENSE_ZERO_ALL")) != 0);
    const int* nsz = diff ? (const int*)c->d_nn_size.p : nullptr;
    const int* nls = diff ? (const int*)c->d_nn_list.p : nullptr;
    const size_t zsm = (size_t)c->dense_pcap + 1;
#define XTB_ZERO2(NOV) dense_zero_prev_kernel<NOV><<<n, 256, zsm, c->stream>>>(f0, c->dense_pcap, (const int*)c->d_prev_size.p, (const int*)c->d_prev_list.p, N, A, B, c->cap, nsz, nls)
    if (NO == 9) XTB_ZERO2(9); else if (NO == 4) XTB_ZERO2(4); else XTB_ZERO2(1);
#undef XTB_ZERO2
    XTB_LAUNCH_CHECK(c);
    if (!scatter) {      // the caller's own kernel writes the blocks of the new pattern into A, B (after this call, same stream)
        dense_remember_kernel<<<n, 128, 0, c->stream>>>(f0, c->cap, (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p,
                                                       c->dense_pcap, (int*)c->d_prev_size.p, (int*)c->d_prev_list.p);
        XTB_LAUNCH_CHECK(c);
        return XTB_OK;
    }
#define XTB_SCATTER(NOV) dense_scatter_kernel<NOV><<<n, 256, sizeof(int) * (size_t)(c->cap + 1), c->stream>>>(f0, c->cap, (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p, \
        (const int*)c->d_rowptr.p, Ablk, Bblk, N, A, B, c->dense_pcap, (int*)c->d_prev_size.p, (int*)c->d_prev_list.p)
    if (NO == 9) XTB_SCATTER(9); else if (NO == 4) XTB_SCATTER(4); else XTB_SCATTER(1);
#undef XTB_SCATTER
#undef XTB_ZERO
    XTB_LAUNCH_CHECK(c);
    return XTB_OK;
}

// compact per-entry tables for the drop-in outputs of xtb_radial (never used by the device hot path)
__global__ void nrl_tables_kernel(int na, int cap, int n_kinds, const int* __restrict__ koa, const int* __restrict__ nn_size,
                                  const int* __restrict__ nn_list, const int* __restrict__ rev, const double* __restrict__ nn_dist,
                                  const xtb_nrl_pair* __restrict__ TB, const double* __restrict__ hmat,
                                  double* __restrict__ rad, double* __restrict__ geo) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (long long)na * cap) return;
    const int j = (int)(t / cap), q = (int)(t % cap);
    if (q >= nn_size[j]) return;
    const int i = nn_list[t];
    const double4 d4 = *reinterpret_cast<const double4*>(nn_dist + 4 * t);
    double* g = geo + 9 * t;
    const double l = d4.x / d4.w, m = d4.y / d4.w, n = d4.z / d4.w;
    g[0] = l; g[1] = m; g[2] = n;                               // M_lmn   TB.f90:3512-3514
    g[3] = l / d4.w; g[4] = m / d4.w; g[5] = n / d4.w;          // M_xrr   :3516-3518
    for (int a = 0; a < 3; ++a) g[6 + a] = d4.x * hmat[a] + d4.y * hmat[a + 3] + d4.z * hmat[a + 6];   // M_x1 :3505-3507
    double* o = rad + 40 * t;
    const int rv = rev[t];
    if (rv < 0) { for (int c = 0; c < 40; ++c) o[c] = 0.0; return; }
    const int small = i < j ? i : j, large = i < j ? j : i;
    const double r = (j < i ? d4.w : nn_dist[4 * ((size_t)i * cap + rv) + 3]) * XTB_G_A2AU;
    const xtb_nrl_pair& P = TB[koa[small] + n_kinds * koa[large]];
    nrl::radial10(P, 0, r, o, o + 10);
    nrl::radial10(P, 1, r, o + 20, o + 30);
}

int xtb_impl_nrl_tables(xtb_ctx* c, double* d_rad, double* d_geo) {
    const long long tot = (long long)c->na * c->cap;
    nrl_tables_kernel<<<(unsigned)((tot + 127) / 128), 128, 0, c->stream>>>(c->na, c->cap, c->n_kinds, (const int*)c->d_koa.p,
                                                                           (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p,
                                                                           (const int*)c->d_rev.p, (const double*)c->d_nn_dist.p,
                                                                           (const xtb_nrl_pair*)c->d_params.p, (const double*)c->d_hdev.p,
                                                                           d_rad, d_geo);
    XTB_LAUNCH_CHECK(c);
    return XTB_OK;
}

int xtb_impl_nrl_dens(xtb_ctx* c) {
    const int na = c->na;
    int rc;
    if ((rc = xtb_reserve(c, c->d_dens, sizeof(double) * na))) return rc;
    const int mx = c->max_nn > 0 ? c->max_nn : 1;
    const size_t smem = sizeof(double) * (size_t)mx * ND_WARPS;
    XTB_CUDA(c, cudaFuncSetAttribute(nrl_dens_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    nrl_dens_kernel<<<(na + ND_WARPS - 1) / ND_WARPS, 32 * ND_WARPS, smem, c->stream>>>(na, c->cap, mx, c->n_kinds, (const int*)c->d_koa.p,
                                                             (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p,
                                                             (const double*)c->d_nn_dist.p, (const xtb_nrl_pair*)c->d_params.p,
                                                             (double*)c->d_dens.p);
    XTB_LAUNCH_CHECK(c);
    c->have_dens = true;
    return XTB_OK;
}

int xtb_impl_nrl_assemble(xtb_ctx* c, bool dense) {
    const int na = c->na;
    int rc;
    if ((rc = xtb_reserve(c, c->d_Hblk, sizeof(double) * 81 * xtb_blocks_alloc(c)))) return rc;
    if ((rc = xtb_reserve(c, c->d_Sblk, sizeof(double) * 81 * xtb_blocks_alloc(c)))) return rc;
    XTB_T0(c, 2);
    if ((rc = xtb_impl_nrl_dens(c))) return rc;
    const int nown = c->a1 - c->a0;
    // dense H_non / Sij: written by the block kernel itself (XTB_DENSE_FUSED=0 or XTB_DENSE_FULL=1: by the separate update afterwards)
    const char* evf = getenv("XTB_DENSE_FUSED"); const char* evd = getenv("XTB_DENSE_FULL");
    const bool fused = dense && nown > 0 && !(evf && atoi(evf) == 0) && !(evd && atoi(evd) != 0);
    if (fused) {      // clear what the new pattern drops, remember the new one
        XTB_T0(c, 6);
        if ((rc = xtb_dense_update(c, c->a0, c->a1, false))) return rc;
        XTB_T1(c, 6);
    }
    if (nown > 0) {
        const int mx = c->max_nn > 0 ? c->max_nn : 1;
        const size_t stage_bytes = sizeof(double) * NB_UP * NB_THREADS + sizeof(int) * (size_t)mx;      // 34 KB: six CTAs per SM
        XTB_CUDA(c, cudaFuncSetAttribute(nrl_blocks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stage_bytes));
        nrl_blocks_kernel<<<nown, NB_THREADS, stage_bytes, c->stream>>>(
            na, c->a0, c->a1, c->cap, mx, c->n_kinds, (const int*)c->d_koa.p, (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p,
            (const int*)c->d_rev.p, (const signed char*)c->d_own.p, (const double*)c->d_nn_dist.p, (const int*)c->d_rowptr.p,
            (const xtb_nrl_pair*)c->d_params.p, (const double*)c->d_dens.p, (double*)c->d_Hblk.p, (double*)c->d_Sblk.p,
            (size_t)na * 9, fused ? (double*)c->d_H.p : nullptr, fused ? (double*)c->d_S2.p : nullptr);
        XTB_LAUNCH_CHECK(c);
    }
    if (!fused) {     // dense update alone (fused form: stage 6 is the preparation only, the values are written by nrl_blocks_kernel)
        XTB_T0(c, 6);
        if (dense && (rc = xtb_dense_update(c, c->a0, c->a1))) return rc;
        XTB_T1(c, 6);
    }
    XTB_T1(c, 2);
    if ((rc = xtb_stage_finish(c))) return rc;
    c->have_blocks = true;
    c->have_HS = dense;
    return XTB_OK;
}
