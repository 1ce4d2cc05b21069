// S4 / S5 -- Hellmann-Feynman attractive forces and supercell ("pressure") force, NRL branch.
//
// Replaces get_dHij_drij_NRL -> get_forces -> d_Hamilton_one_NRL -> Attract_TB_forces_NRL
// (Source_files/TB_NRL.f90:2412-2736) and Attract_TB_Forces_Press_NRL -> dHamil_tot_Press_NRL ->
// dHamilton_one_Press_NRL -> dHopping_Press_NRL (TB_NRL.f90:2742-3046), with the geometric derivatives
// drij_dska / ddija_dskb_kd / drij_dhab / dda_dhgd (TB_Koster_Slater.f90:59-161) and the on-site terms
// d_On_site_NRL / d_rho_NRL (TB_NRL.f90:3053-3118), d_On_site_NRL_Press / d_rho_NRL_Press (:3121-3198).
//
// Reference algorithm: for EACH atom k it rebuilds dense dH,dS(3,N,N) from ALL pair blocks (non-zero only
// when k is one of the two atoms, or for the on-site block when k is the atom or one of its neighbours)
// and contracts with dense rho, rho_E: O(Na * N^2) memory traffic.  Here each atom k gathers only its
// own non-zero terms, in a fixed order (deterministic, no atomics):
//   F_k = sum_{b in NN(k)} 2 * < dBlock(min(k,b) -> max(k,b); delta_k) , rho(min,max) > - same with dS, rho_E
//       + w_k * sum_{b in NN(k)} grad rho_k[b]  -  sum_{b in NN(k)} w_b * grad rho_b[k]
// where the pair block derivative is evaluated, as in the reference, from the list entry of the SMALLER
// index atom (rows) for the larger (cols) (TB_NRL.f90:2503-2513), the transposed block contributes the
// same amount because rho is symmetric (factor 2), and w_j = g_Ry * sum_orb dh_l/drho(rho_j) * rho_jj(orb,orb).
// Thread mapping: one CTA per atom; the rho / rho_E blocks of the atom (contiguous in the block store) are read once
// through L1; one thread per (neighbour, matrix, direction) differentiates the Slater-Koster block in registers
// (forward-mode dual numbers) and contracts it.  Each unordered pair is differentiated once (the partner's share is the
// exact negative, see nrl_force_pairs_kernel); a second small kernel adds the terms of every atom in list order.
#include <cstdlib>
#include "nrl_device.cuh"
#include "sk_device.cuh"

namespace {

// per-atom on-site weight w_j (forces and pressure share it)
__global__ void nrl_onsite_w_kernel(int na, int n_kinds, const int* __restrict__ koa, const int* __restrict__ rowptr,
                                    const xtb_nrl_pair* __restrict__ TB, const double* __restrict__ dens,
                                    const double* __restrict__ rho, double* __restrict__ w) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= na) return;
    const xtb_nrl_pair& P = TB[koa[j] + n_kinds * koa[j]];
    const double rj = dens[j];
    double dh[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) dh[s] = nrl::d_onsite(rj, P.bl[s], P.cl[s], P.dl[s]);
    const double* blk = rho + (size_t)rowptr[j] * 81;
    double acc = 0.0;
    for (int o = 0; o < 9; ++o) acc += dh[nrl::onsite_set(o, P.ind_split)] * blk[o * 10];
    w[j] = acc * XTB_G_RY;
}

// symmetrised contraction coefficients: for the packed upper triangle (a<=b) of a block derivative D with
// D(b,a) = sgn(a,b) D(a,b):   <D, A> = sum_{a<=b} D(a,b) * coef(a,b),  coef = A(a,b) + sgn*A(b,a) (diag: A(a,a)).
// `A` is the rho block whose rows correspond to the block's row atom; tr = read it transposed.
__device__ __forceinline__ double contract_upper(const double* up /*45, derivative values*/, const double* A, bool tr) {
    double s = 0.0;
#pragma unroll
    for (int a = 0; a < 9; ++a)
#pragma unroll
        for (int b = a; b < 9; ++b) {
            const double x = tr ? A[b + 9 * a] : A[a + 9 * b];
            const double y = tr ? A[a + 9 * b] : A[b + 9 * a];
            const double cf = (a == b) ? x : x + sk::lower_sign(a, b) * y;
            s += up[sk::ut(a, b)] * cf;
        }
    return s;
}

// derivative (w.r.t. one scalar parameter) of the reference-signed SK block, packed upper triangle,
// given V, dV/dparam, (l,m,n), d(l,m,n)/dparam.  Reproduces the reference quirk in (dxz,dx2-y2).
__device__ __forceinline__ void dblock_upper(const double* V, const double* dV, double l, double m, double n,
                                             double dl, double dm, double dn, double* up) {
    sk::Dual Vd[10], o[45];
#pragma unroll
    for (int c = 0; c < 10; ++c) Vd[c] = sk::mk(V[c], dV[c]);
    sk::block_sp3d5_upper<sk::Dual>(Vd, sk::mk(l, dl), sk::mk(m, dm), sk::mk(n, dn), o);
#pragma unroll
    for (int e = 0; e < 45; ++e) up[e] = o[e].d;
    up[sk::ut(5, 7)] = sk::ref_d_dxz_dx2y2(l, dl, m, dm, n, dn, V[7], dV[7], V[8], dV[8], V[9], dV[9]);
}

constexpr int FT = 256;

// one (atom k, list entry q, matrix, coordinate alpha) pair term: 2 < dBlock(min(k,b) -> max(k,b); delta_k), rho(min,max) > [* g_Ry for H]
__device__ __forceinline__ double pair_term(int k, int b, int rv, size_t e, int cap, int n_kinds, int mat, int al,
                                            const int* __restrict__ koa, const double* __restrict__ nn_dist,
                                            const xtb_nrl_pair* __restrict__ TB, const double* h, const double* A /* block (k, q) */) {
    // block rows = smaller index atom i, cols = larger j; geometry from i's entry (TB_NRL.f90:2503-2513)
    const bool ksmall = k < b;
    const size_t es = ksmall ? e : (size_t)b * cap + rv;
    const double4 d4 = *reinterpret_cast<const double4*>(nn_dist + 4 * es);
    const double x = d4.x, y = d4.y, z = d4.z, r = d4.w;
    const double delt = ksmall ? 1.0 : -1.0;              // delta_ik - delta_jk
    const int ia = ksmall ? k : b, ja = ksmall ? b : k;
    const xtb_nrl_pair& P = TB[koa[ia] + n_kinds * koa[ja]];
    double V[10], dV[10];
    nrl::radial10(P, mat, r * XTB_G_A2AU, V, dV);
    const double l = x / r, mm = y / r, n = z / r;
    // drij_dska (TB_Koster_Slater.f90:75) and ddija_dskb_kd (:102) for coordinate alpha = al
    const double drds = delt / r * (x * h[al] + y * h[al + 3] + z * h[al + 6]);     // supce(al,1..3)
    const double dl = (delt * h[al + 0] - x / r * drds) / r;     // supce(beta=al, alpha=1)
    const double dm = (delt * h[al + 3] - y / r * drds) / r;
    const double dn = (delt * h[al + 6] - z / r * drds) / r;
    const double drB = drds * XTB_G_A2AU;
#pragma unroll
    for (int c = 0; c < 10; ++c) dV[c] *= drB;
    double up[45];
    dblock_upper(V, dV, l, mm, n, dl, dm, dn, up);
    // rho(i,j) with rows i = smaller: block (k,q) read directly if k is the smaller, else transposed
    double res = 2.0 * contract_upper(up, A, !ksmall);
    if (mat == 0) res *= XTB_G_RY;
    return res;
}

// the two on-site terms of list entry q of atom k (centre k / neighbour b and centre b / neighbour k), three coordinates
__device__ __forceinline__ void onsite_terms(int k, int b, int rv, size_t e, int cap, int n_kinds, const int* __restrict__ koa,
                                             const double* __restrict__ nn_dist, const xtb_nrl_pair* __restrict__ TB,
                                             const double* __restrict__ wons, const double* h, double* o3) {
    o3[0] = o3[1] = o3[2] = 0.0;
    {   // centre k, neighbour b: + w_k * X'(r_kb; TB(kind k, kind b)) * g_A2au * dr/ds_k
        const double4 d4 = *reinterpret_cast<const double4*>(nn_dist + 4 * e);
        const xtb_nrl_pair& P = TB[koa[k] + n_kinds * koa[b]];
        double X, dX;
        nrl::envelope(P, d4.w * XTB_G_A2AU, X, dX);
        const double f = wons[k] * dX * XTB_G_A2AU / d4.w;
#pragma unroll
        for (int al = 0; al < 3; ++al) o3[al] += f * (d4.x * h[al] + d4.y * h[al + 3] + d4.z * h[al + 6]);
    }
    if (rv >= 0) {   // centre b, neighbour k: delta = -1
        const size_t eb = (size_t)b * cap + rv;
        const double4 d4 = *reinterpret_cast<const double4*>(nn_dist + 4 * eb);
        const xtb_nrl_pair& P = TB[koa[b] + n_kinds * koa[k]];
        double X, dX;
        nrl::envelope(P, d4.w * XTB_G_A2AU, X, dX);
        const double f = -wons[b] * dX * XTB_G_A2AU / d4.w;
#pragma unroll
        for (int al = 0; al < 3; ++al) o3[al] += f * (d4.x * h[al] + d4.y * h[al + 3] + d4.z * h[al + 6]);
    }
}

// ---- forces, single-kernel form (XTB_FORCES_MODE=0): one CTA per owned atom k evaluates every one of its list entries ----
__global__ void __launch_bounds__(FT, 2) nrl_forces_kernel(int na, int a0, int cap, int mx, int n_kinds, const int* __restrict__ koa,
                                                       const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                       const int* __restrict__ rev, const double* __restrict__ nn_dist,
                                                       const int* __restrict__ rowptr, const xtb_nrl_pair* __restrict__ TB,
                                                       const double* __restrict__ hmat, const double* __restrict__ wons,
                                                       const double* __restrict__ rho, const double* __restrict__ rhoE,
                                                       double* __restrict__ F) {
    extern __shared__ __align__(16) double smf[];
    const int k = a0 + blockIdx.x;
    const int m = nn_size[k];
    // mx = largest neighbour count (sizes the shared arrays); cap = ELL stride of the lists
    double* part = smf;                             // [mx][8]: 6 pair partials (mat*3+alpha) + (unused 2)
    double* ons = part + (size_t)mx * 8;            // [mx][3] on-site partials
    const double* blkR = rho + (size_t)rowptr[k] * 81;      // the blocks of atom k: contiguous, read once through L1
    const double* blkE = rhoE + (size_t)rowptr[k] * 81;
    __shared__ double h[9];
    if (threadIdx.x < 9) h[threadIdx.x] = hmat[threadIdx.x];
    __syncthreads();
    // ---- pair terms: thread -> (q, mat, alpha) ----
    for (int w = threadIdx.x; w < m * 6; w += FT) {
        const int q = w / 6, rem = w - q * 6;
        const int mat = rem / 3, al = rem - mat * 3;
        const size_t e = (size_t)k * cap + q;
        const int b = nn_list[e];
        const int rv = rev[e];
        double res = 0.0;
        if (rv >= 0) res = pair_term(k, b, rv, e, cap, n_kinds, mat, al, koa, nn_dist, TB, h, (mat == 0 ? blkR : blkE) + (size_t)(q + 1) * 81);
        part[q * 8 + rem] = res;
    }
    // ---- on-site terms: thread -> q ----
    for (int q = threadIdx.x; q < m; q += FT) {
        const size_t e = (size_t)k * cap + q;
        double o3[3];
        onsite_terms(k, nn_list[e], rev[e], e, cap, n_kinds, koa, nn_dist, TB, wons, h, o3);
        ons[q * 3] = o3[0]; ons[q * 3 + 1] = o3[1]; ons[q * 3 + 2] = o3[2];
    }
    __syncthreads();
    if (threadIdx.x < 3) {   // fixed-order reduction over the list
        const int al = threadIdx.x;
        double s = 0.0;
        for (int q = 0; q < m; ++q) s += (part[q * 8 + al] - part[q * 8 + 3 + al]) + ons[q * 3 + al];
        F[3 * k + al] = s;
    }
}

// who differentiates the pair of list entry e = (k, b): the atom that owns it in the density pass (own[e]: positive half-space
// rule of nn_rev_kernel, half of the neighbours of every atom -- equal work per CTA) when that atom belongs to this rank's
// range, else the atom of the range.  Entries the partner does not list (rev < 0) are owned by construction.
__device__ __forceinline__ bool force_theirs(signed char own_e, int rv, int b, int a0, int a1) {
    return own_e == 0 && rv >= 0 && b >= a0 && b < a1;
}

// ---- forces, pair-once form (default) ---------------------------------------------------------------------------------------
// The pair term of entry (k, b) is linear in delta = delta_ik - delta_jk and reads the same geometry (the entry of the smaller
// index) and the same numbers of rho (the two mirrored blocks are element-wise transposes), so what atom b gets from the pair
// is the exact negative, bit for bit, of what atom k gets: every floating-point operation on the path is odd in delta.  Each
// unordered pair is therefore differentiated ONCE -- by its owner of the density pass (force_theirs) -- and
// the (H - S) difference of its three coordinates is kept per list entry; the second kernel adds the entries of atom k in list
// order, taking the partner's value with the opposite sign.  Same bits as the single-kernel form at half the FP64 work.
__global__ void __launch_bounds__(FT, 2) nrl_force_pairs_kernel(int a0, int a1, int cap, int mx, int n_kinds, const int* __restrict__ koa,
                                                              const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                              const int* __restrict__ rev, const signed char* __restrict__ own,
                                                              const double* __restrict__ nn_dist,
                                                              const int* __restrict__ rowptr, const xtb_nrl_pair* __restrict__ TB,
                                                              const double* __restrict__ hmat, const double* __restrict__ wons,
                                                              const double* __restrict__ rho, const double* __restrict__ rhoE,
                                                              double* __restrict__ fpair /*[na*cap][6]: pair (H - S) x 3, on-site x 3*/) {
    extern __shared__ __align__(16) double smf[];
    const int k = a0 + blockIdx.x;
    const int m = nn_size[k];
    double* part = smf;                             // [mx][8]
    int* actq = reinterpret_cast<int*>(part + (size_t)mx * 8);      // [mx] list ordinals of the entries this atom evaluates
    __shared__ int nact_sh;
    const double* blkR = rho + (size_t)rowptr[k] * 81;
    const double* blkE = rhoE + (size_t)rowptr[k] * 81;
    __shared__ double h[9];
    if (threadIdx.x < 9) h[threadIdx.x] = hmat[threadIdx.x];
    if (threadIdx.x < 32) {      // ordered compaction (warp 0): the evaluated entries fill the first pass of the CTA
        int cnt = 0;
        for (int base = 0; base < m; base += 32) {
            const int q = base + threadIdx.x;
            bool f = false;
            if (q < m) {
                const size_t e = (size_t)k * cap + q;
                const int b = nn_list[e];
                f = !force_theirs(own[e], rev[e], b, a0, a1);
            }
            const unsigned mask = __ballot_sync(0xffffffffu, f);
            if (f) actq[cnt + __popc(mask & ((1u << threadIdx.x) - 1u))] = q;
            cnt += __popc(mask);
        }
        if (threadIdx.x == 0) nact_sh = cnt;
    }
    __syncthreads();
    const int nact = nact_sh;
    for (int w = threadIdx.x; w < nact * 6; w += FT) {
        const int qi = w / 6, rem = w - qi * 6;
        const int q = actq[qi];
        const int mat = rem / 3, al = rem - mat * 3;
        const size_t e = (size_t)k * cap + q;
        const int b = nn_list[e];
        const int rv = rev[e];
        double res = 0.0;
        if (rv >= 0) res = pair_term(k, b, rv, e, cap, n_kinds, mat, al, koa, nn_dist, TB, h, (mat == 0 ? blkR : blkE) + (size_t)(q + 1) * 81);
        part[qi * 8 + rem] = res;
    }
    for (int q = threadIdx.x; q < m; q += FT) {
        const size_t e = (size_t)k * cap + q;
        double o3[3];
        onsite_terms(k, nn_list[e], rev[e], e, cap, n_kinds, koa, nn_dist, TB, wons, h, o3);
        fpair[e * 6 + 3] = o3[0]; fpair[e * 6 + 4] = o3[1]; fpair[e * 6 + 5] = o3[2];
    }
    __syncthreads();
    for (int w = threadIdx.x; w < nact * 3; w += FT) {
        const int qi = w / 3, al = w - qi * 3;
        fpair[((size_t)k * cap + actq[qi]) * 6 + al] = part[qi * 8 + al] - part[qi * 8 + 3 + al];
    }
}

// fixed-order sum over the list of atom k; one warp per atom stages the terms, three lanes add them in list order
constexpr int FS_WARPS = 4;
__global__ void __launch_bounds__(32 * FS_WARPS) nrl_force_sum_kernel(int a0, int a1, int cap, int mx, int wpc, const int* __restrict__ nn_size,
                                                                     const int* __restrict__ nn_list, const int* __restrict__ rev,
                                                                     const signed char* __restrict__ own,
                                                                     const double* __restrict__ fpair, double* __restrict__ F) {
    extern __shared__ double sms[];                 // [wpc][mx][3]  (wpc = warps per CTA, <= FS_WARPS: fewer when mx is large)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int k = a0 + blockIdx.x * wpc + warp;
    if (k >= a1) return;
    double* t = sms + (size_t)warp * mx * 3;
    const int m = nn_size[k];
    for (int w = lane; w < m * 3; w += 32) {
        const int q = w / 3, al = w - q * 3;
        const size_t e = (size_t)k * cap + q;
        const int b = nn_list[e];
        const int rv = rev[e];
        const bool theirs = force_theirs(own[e], rv, b, a0, a1);
        const double d = theirs ? -fpair[((size_t)b * cap + rv) * 6 + al] : fpair[e * 6 + al];
        t[w] = d + fpair[e * 6 + 3 + al];
    }
    __syncwarp();
    if (lane < 3) {
        double s = 0.0;
        for (int q = 0; q < m; ++q) s += t[q * 3 + lane];
        F[3 * k + lane] = s;
    }
}

// ---- pressure: one CTA per owned atom i, the directed pairs (i, q) + on-site ----
// The term of the directed pair (j, i) equals the term of (i, j) bit for bit when the two list entries hold negated geometry
// (always, except when two periodic images tie and the scans of i and j keep different ones): the block derivative picks up
// (-1)^(l_a + l_b) from (l, m, n) -> -(l, m, n), the mirrored rho block is the transpose, dr/dh = r_delta s_gamma / r is even.
// PAIR_ONCE: such a pair is evaluated by its owner of the density pass (force_theirs) and its nine (H - S) values are kept per
// list entry (fp: [na * cap][19] = pair values, on-site values, index of the entry that holds the pair values); nrl_pressure_sum_kernel then adds the entries of every atom
// in list order, taking the partner's values for the skipped ones -- the same bits as evaluating every directed pair
// (XTB_FORCES_MODE=0: !PAIR_ONCE, everything in one kernel).
__device__ __forceinline__ bool pressure_theirs(signed char own_e, int j, int rv, size_t e, int cap, int a0, int a1,
                                                const double* __restrict__ nn_dist, const double* __restrict__ nn_dist_s) {
    if (!force_theirs(own_e, rv, j, a0, a1)) return false;
    const size_t eb = (size_t)j * cap + rv;
    const double4 u = *reinterpret_cast<const double4*>(nn_dist + 4 * e);
    const double4 v = *reinterpret_cast<const double4*>(nn_dist + 4 * eb);
    bool anti = u.x == -v.x && u.y == -v.y && u.z == -v.z && u.w == v.w;
#pragma unroll
    for (int g = 0; g < 3; ++g) anti = anti && nn_dist_s[3 * e + g] == -nn_dist_s[3 * eb + g];
    return anti;
}

constexpr int PF = 19;      // doubles per list entry in the pressure scratch: 9 pair values, 9 on-site values, source entry
template <bool PAIR_ONCE>
__global__ void __launch_bounds__(FT) nrl_pressure_kernel(int na, int a0, int a1, int cap, int mx, int n_kinds, const int* __restrict__ koa,
                                                         const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                         const int* __restrict__ rev, const signed char* __restrict__ own,
                                                         const double* __restrict__ nn_dist,
                                                         const double* __restrict__ nn_dist_s, const int* __restrict__ rowptr,
                                                         const xtb_nrl_pair* __restrict__ TB, const double* __restrict__ wons,
                                                         const double* __restrict__ rho, const double* __restrict__ rhoE,
                                                         double* __restrict__ partial /*[n_owned][9]*/, double* __restrict__ fp) {
    extern __shared__ __align__(16) double smf[];
    const int i = a0 + blockIdx.x;
    const int m = nn_size[i];
    double* part = smf;                        // [mx][18]: (mat, k1)
    double* ons = part + (size_t)mx * 18;      // [mx][9]
    int* actq = reinterpret_cast<int*>(ons + (size_t)mx * 9);      // [mx] PAIR_ONCE: list ordinals of the entries this atom evaluates
    __shared__ int nact_sh;
    const size_t b0 = (size_t)rowptr[i] * 81;
    if (PAIR_ONCE) {
        if (threadIdx.x < 32) {      // ordered compaction (warp 0)
            int cnt = 0;
            for (int base = 0; base < m; base += 32) {
                const int q = base + threadIdx.x;
                bool f = false;
                if (q < m) {
                    const size_t e = (size_t)i * cap + q;
                    const int j = nn_list[e], rv = rev[e];
                    f = !pressure_theirs(own[e], j, rv, e, cap, a0, a1, nn_dist, nn_dist_s);
                    fp[e * PF + 18] = (double)(f ? e : (size_t)j * cap + rv);      // the entry that holds the pair values (exact: < 2^53)
                }
                const unsigned mask = __ballot_sync(0xffffffffu, f);
                if (f) actq[cnt + __popc(mask & ((1u << threadIdx.x) - 1u))] = q;
                cnt += __popc(mask);
            }
            if (threadIdx.x == 0) nact_sh = cnt;
        }
        __syncthreads();
    }
    const int nact = PAIR_ONCE ? nact_sh : m;
    // thread -> (evaluated entry, mat, gamma): three delta components each
    for (int w = threadIdx.x; w < nact * 6; w += FT) {
        const int qi = w / 6, rem = w - qi * 6;
        const int q = PAIR_ONCE ? actq[qi] : qi;
        const int mat = rem / 3, ga = rem - mat * 3;
        const size_t e = (size_t)i * cap + q;
        const int j = nn_list[e];
        const int rv = rev[e];
        const double4 d4 = *reinterpret_cast<const double4*>(nn_dist + 4 * e);
        const double rij[3] = {d4.x, d4.y, d4.z};
        const double r = d4.w;
        const double sg = nn_dist_s[3 * e + ga];
        // radial table entry M_Vij(i,j,:) = pair value from the smaller index' entry (mirrored, TB_NRL.f90:2228-2239)
        double V[10], dV[10];
        bool ok = rv >= 0;
        if (ok) {
            const int sm_ = i < j ? i : j, lg_ = i < j ? j : i;
            const double rr = (i < j) ? r : nn_dist[4 * ((size_t)j * cap + rv) + 3];
            nrl::radial10(TB[koa[sm_] + n_kinds * koa[lg_]], mat, rr * XTB_G_A2AU, V, dV);
        }
        const double l = rij[0] / r, mm = rij[1] / r, n = rij[2] / r;
        const double* A = (mat == 0 ? rho : rhoE) + b0 + (size_t)(q + 1) * 81;
        const double r2 = r * r;
#pragma unroll
        for (int de = 0; de < 3; ++de) {
            double res = 0.0;
            if (ok) {
                // dda_dhgd (TB_Koster_Slater.f90:153-161) * g_au2A ; drij_dhab (:146-150)
                const double pre = sg / (r2 * r);
                const double dl = pre * (r2 * (de == 0 ? 1.0 : 0.0) - rij[0] * rij[de]) * XTB_G_AU2A;
                const double dm = pre * (r2 * (de == 1 ? 1.0 : 0.0) - rij[1] * rij[de]) * XTB_G_AU2A;
                const double dn = pre * (r2 * (de == 2 ? 1.0 : 0.0) - rij[2] * rij[de]) * XTB_G_AU2A;
                const double drdh = rij[de] * sg / r;
                double dVs[10];
#pragma unroll
                for (int c = 0; c < 10; ++c) dVs[c] = dV[c] * drdh;
                double up[45];
                dblock_upper(V, dVs, l, mm, n, dl, dm, dn, up);
                // block (rows i, cols j) = d_KS_sp3d5_TEST^T = d_KS_sp3d5 convention; contracted with rho(i,j) = block (i,q)
                res = contract_upper(up, A, false) * XTB_G_A2AU;
                if (mat == 0) res *= XTB_G_RY;
            }
            part[qi * 18 + mat * 9 + ga * 3 + de] = res;
        }
    }
    // on-site: thread -> (q, k1): w_i * X'(r) * (r_delta s_gamma / r) * g_A2au
    for (int w = threadIdx.x; w < m * 9; w += FT) {
        const int q = w / 9, k1 = w - q * 9;
        const int ga = k1 / 3, de = k1 - ga * 3;
        const size_t e = (size_t)i * cap + q;
        const int j = nn_list[e];
        const xtb_nrl_pair& P = TB[koa[i] + n_kinds * koa[j]];
        const double r = nn_dist[4 * e + 3];
        double X, dX;
        nrl::envelope(P, r * XTB_G_A2AU, X, dX);
        ons[q * 9 + k1] = wons[i] * dX * (nn_dist[4 * e + de] * nn_dist_s[3 * e + ga] / r) * XTB_G_A2AU;
    }
    __syncthreads();
    if (PAIR_ONCE) {
        for (int w = threadIdx.x; w < m * 9; w += FT) {
            const int q = w / 9, k1 = w - q * 9;
            fp[((size_t)i * cap + q) * PF + 9 + k1] = ons[q * 9 + k1];
        }
        for (int w = threadIdx.x; w < nact * 9; w += FT) {
            const int qi = w / 9, k1 = w - qi * 9;
            fp[((size_t)i * cap + actq[qi]) * PF + k1] = part[qi * 18 + k1] - part[qi * 18 + 9 + k1];
        }
    } else if (threadIdx.x < 9) {
        const int k1 = threadIdx.x;
        double s = 0.0;
        for (int q = 0; q < m; ++q) s += (part[q * 18 + k1] - part[q * 18 + 9 + k1]) + ons[q * 9 + k1];
        partial[(size_t)blockIdx.x * 9 + k1] = s;
    }
}

// fixed-order sum over the list of atom i (one warp per atom stages the terms, nine lanes add them in list order)
__global__ void __launch_bounds__(32 * FS_WARPS) nrl_pressure_sum_kernel(int a0, int a1, int cap, int mx, int wpc, const int* __restrict__ nn_size,
                                                                        const int* __restrict__ nn_list, const int* __restrict__ rev,
                                                                        const signed char* __restrict__ own,
                                                                        const double* __restrict__ nn_dist, const double* __restrict__ nn_dist_s,
                                                                        const double* __restrict__ fp, double* __restrict__ partial) {
    extern __shared__ double sms[];                 // [wpc][mx][9]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int i = a0 + blockIdx.x * wpc + warp;
    if (i >= a1) return;
    double* t = sms + (size_t)warp * mx * 9;
    const int m = nn_size[i];
    for (int w = lane; w < m * 9; w += 32) {
        const int q = w / 9, k1 = w - q * 9;
        const size_t e = (size_t)i * cap + q;
        const size_t se = (size_t)fp[e * PF + 18];      // own entry or the partner's (recorded by nrl_pressure_kernel)
        t[w] = fp[se * PF + k1] + fp[e * PF + 9 + k1];
    }
    __syncwarp();
    if (lane < 9) {
        double s = 0.0;
        for (int q = 0; q < m; ++q) s += t[q * 9 + lane];
        partial[(size_t)(i - a0) * 9 + lane] = s;
    }
}

// fixed-order final reduction of the per-atom partials (deterministic)
__global__ void sum9_kernel(int n, const double* __restrict__ partial, double* __restrict__ out) {
    __shared__ double buf[9][128];
    const int k1 = blockIdx.x;
    double s = 0.0;
    for (int a = threadIdx.x; a < n; a += 128) s += partial[(size_t)a * 9 + k1];
    buf[k1][threadIdx.x] = s;
    __syncthreads();
    for (int off = 64; off > 0; off >>= 1) {
        if (threadIdx.x < off) buf[k1][threadIdx.x] += buf[k1][threadIdx.x + off];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[k1] = buf[k1][0];
}

int launch_onsite_w(xtb_ctx* c) {
    int rc;
    if (!c->have_dens && (rc = xtb_impl_nrl_dens(c))) return rc;
    if ((rc = xtb_reserve(c, c->d_onsite_w, sizeof(double) * c->na))) return rc;
    nrl_onsite_w_kernel<<<(c->na + 127) / 128, 128, 0, c->stream>>>(c->na, c->n_kinds, (const int*)c->d_koa.p,
                                                                   (const int*)c->d_rowptr.p, (const xtb_nrl_pair*)c->d_params.p,
                                                                   (const double*)c->d_dens.p, (const double*)c->d_rho.p,
                                                                   (double*)c->d_onsite_w.p);
    XTB_LAUNCH_CHECK(c);
    return XTB_OK;
}

}  // namespace

int xtb_impl_nrl_forces(xtb_ctx* c) {
    const int nown = c->a1 - c->a0;
    int rc;
    if ((rc = xtb_reserve(c, c->d_F, sizeof(double) * 3 * (size_t)(c->na + c->nranks)))) return rc;
    const char* ev = getenv("XTB_FORCES_MODE");              // 0: every directed pair evaluated by its atom, one kernel (cross-check of the default)
    const int mode = ev ? atoi(ev) : 1;
    if (mode != 0 && (rc = xtb_reserve(c, c->d_fpair, sizeof(double) * 6 * (size_t)c->na * c->cap))) return rc;
    XTB_T0(c, 4);
    if ((rc = launch_onsite_w(c))) return rc;
    if (nown > 0) {
        const int mx = c->max_nn > 0 ? c->max_nn : 1;
        const int* koa = (const int*)c->d_koa.p; const int* nsz = (const int*)c->d_nn_size.p; const int* nls = (const int*)c->d_nn_list.p;
        const int* rv = (const int*)c->d_rev.p; const double* nd = (const double*)c->d_nn_dist.p; const int* rp = (const int*)c->d_rowptr.p;
        const xtb_nrl_pair* TB = (const xtb_nrl_pair*)c->d_params.p;
        if (mode == 0) {
            const size_t smem = sizeof(double) * ((size_t)mx * 11);
            XTB_CUDA(c, cudaFuncSetAttribute(nrl_forces_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            nrl_forces_kernel<<<nown, FT, smem, c->stream>>>(c->na, c->a0, c->cap, mx, c->n_kinds, koa, nsz, nls, rv, nd, rp, TB,
                                                             (const double*)c->d_hdev.p, (const double*)c->d_onsite_w.p,
                                                             (const double*)c->d_rho.p, (const double*)c->d_rhoE.p, (double*)c->d_F.p);
            XTB_LAUNCH_CHECK(c);
        } else {
            const size_t smem = sizeof(double) * ((size_t)mx * 8) + sizeof(int) * (size_t)mx;
            XTB_CUDA(c, cudaFuncSetAttribute(nrl_force_pairs_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            nrl_force_pairs_kernel<<<nown, FT, smem, c->stream>>>(c->a0, c->a1, c->cap, mx, c->n_kinds, koa, nsz, nls, rv, (const signed char*)c->d_own.p, nd, rp, TB,
                                                                  (const double*)c->d_hdev.p, (const double*)c->d_onsite_w.p,
                                                                  (const double*)c->d_rho.p, (const double*)c->d_rhoE.p, (double*)c->d_fpair.p);
            XTB_LAUNCH_CHECK(c);
            int wpc = FS_WARPS;                                     // warps (atoms) per CTA: as many as the staged terms leave room for
            while (wpc > 1 && sizeof(double) * (size_t)mx * 3 * wpc > 160 * 1024) wpc >>= 1;
            const size_t smem2 = sizeof(double) * ((size_t)mx * 3 * wpc);
            XTB_CUDA(c, cudaFuncSetAttribute(nrl_force_sum_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
            nrl_force_sum_kernel<<<(nown + wpc - 1) / wpc, 32 * wpc, smem2, c->stream>>>(c->a0, c->a1, c->cap, mx, wpc, nsz, nls, rv,
                                                                                                    (const signed char*)c->d_own.p, (const double*)c->d_fpair.p, (double*)c->d_F.p);
            XTB_LAUNCH_CHECK(c);
        }
    }
    XTB_T1(c, 4);
    if ((rc = xtb_stage_finish(c))) return rc;
    return XTB_OK;
}

int xtb_impl_nrl_pressure(xtb_ctx* c) {
    const int nown = c->a1 - c->a0;
    int rc;
    if ((rc = xtb_reserve(c, c->d_SC, sizeof(double) * 9))) return rc;
    if ((rc = xtb_reserve(c, c->d_ppart, sizeof(double) * 9 * (size_t)(nown > 0 ? nown : 1)))) return rc;
    const char* ev = getenv("XTB_FORCES_MODE");              // 0: every directed pair evaluated by its atom (cross-check of the default)
    const int mode = ev ? atoi(ev) : 1;
    if (mode != 0 && (rc = xtb_reserve(c, c->d_fpair, sizeof(double) * PF * (size_t)c->na * c->cap))) return rc;
    XTB_T0(c, 5);
    if ((rc = launch_onsite_w(c))) return rc;
    if (nown > 0) {
        const int mx = c->max_nn > 0 ? c->max_nn : 1;
        const size_t smem = sizeof(double) * ((size_t)mx * 27) + sizeof(int) * (size_t)mx;
        auto kern = mode == 0 ? nrl_pressure_kernel<false> : nrl_pressure_kernel<true>;
        XTB_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<nown, FT, smem, c->stream>>>(c->na, c->a0, c->a1, c->cap, mx, c->n_kinds, (const int*)c->d_koa.p,
                                            (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p, (const int*)c->d_rev.p,
                                            (const signed char*)c->d_own.p, (const double*)c->d_nn_dist.p, (const double*)c->d_nn_dist_s.p,
                                            (const int*)c->d_rowptr.p, (const xtb_nrl_pair*)c->d_params.p,
                                            (const double*)c->d_onsite_w.p, (const double*)c->d_rho.p,
                                            (const double*)c->d_rhoE.p, (double*)c->d_ppart.p, (double*)c->d_fpair.p);
        XTB_LAUNCH_CHECK(c);
        if (mode != 0) {
            int wpc = FS_WARPS;
            while (wpc > 1 && sizeof(double) * (size_t)mx * 9 * wpc > 160 * 1024) wpc >>= 1;
            const size_t smem2 = sizeof(double) * ((size_t)mx * 9 * wpc);
            XTB_CUDA(c, cudaFuncSetAttribute(nrl_pressure_sum_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
            nrl_pressure_sum_kernel<<<(nown + wpc - 1) / wpc, 32 * wpc, smem2, c->stream>>>(
                c->a0, c->a1, c->cap, mx, wpc, (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p, (const int*)c->d_rev.p,
                (const signed char*)c->d_own.p, (const double*)c->d_nn_dist.p, (const double*)c->d_nn_dist_s.p, (const double*)c->d_fpair.p, (double*)c->d_ppart.p);
            XTB_LAUNCH_CHECK(c);
        }
        sum9_kernel<<<9, 128, 0, c->stream>>>(nown, (const double*)c->d_ppart.p, (double*)c->d_SC.p);
        XTB_LAUNCH_CHECK(c);
    } else {
        XTB_CUDA(c, cudaMemsetAsync(c->d_SC.p, 0, sizeof(double) * 9, c->stream));
    }
    XTB_T1(c, 5);
    if ((rc = xtb_stage_finish(c))) return rc;
    return XTB_OK;
}
