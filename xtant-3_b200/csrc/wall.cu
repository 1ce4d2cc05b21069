// Short-range "exponential wall" pair term on the neighbour stream (second half).
//
// Replaces, in Source_files/Exponential_wall.f90 (class TB_Exp_wall_simple, Objects.f90:458-466; dispatch TB.f90:3533-3556
// for the forces and TB.f90:3983 for the energy):
//   get_Exp_wall_s :990-1067            E_wall = 1/2 sum_i sum_{j in NN(i)} Pot(r_ij);  MDAtoms(i)%Epot += 1/2 sum_j Pot
//   Exp_wall_pot :1070-1092             Pot = C exp(1 / (r - r0)) / (1 + exp((r - d0) / dd)),  0 for r > d0 + 10 dd
//   Wall_potential / d_Wall_potential :1095-1122   (the "atoms too close" arm tests |r - r0| <= log(tiny) < 0: never taken)
//   d_Exp_wall_pot :1140-1166           dPot/dr  (f_cut_L_C, d_f_cut_L_C: Coulomb.f90:610-629)
//   d_Exp_wall_pot_s :1200-1329         forces%rep(k) += 1/2 sum_{i1} sum_{j1} dPot (d_ik - d_jk) / r * x1,  x1 = h . (x, y, z)
//   d_Exp_wall_Pressure_s :1334-1385    SCforce%rep(l,k) += 1/2 sum_i sum_j dPot r_k s_l / r
// The reference loops, for every atom ian, over every atom i1 and its whole list (O(Na^2 nn)); only i1 = ian and the i1 that
// list ian contribute.  Here: one warp per atom, lanes stride its list; the mirrored term of a neighbour is read from the
// NEIGHBOUR's own record (located with `rev`), as the reference does; fixed-order reductions.  Works with every TB branch
// (the wall does not change the neighbour cut-off, Atomic_tools.f90:562-651).  HBM-bound, tiny: 36 bytes per list entry.
//
// The 'General' class of the same slot (type TB_Short_Rep, Objects.f90:437-454) runs through the same three kernels with another
// pair function (template parameter):
//   get_short_range_rep_s :48-127, Shortrange_pot :130-157, d_Short_range_pot :160-203, d_Short_range_pot_s :519-659,
//   d_Short_range_Pressure_s :664-719; exp_function :935 / inv_exp_function :894 / power_function :850 / tabulated_potential :726
//   and their derivatives; ZBL_pot, d_ZBL_pot, NLH_pot, d_NLH_pot (ZBL_potential.f90:150-178,366-452).
#include "xtb_ctx.h"

namespace {

constexpr int WW = 4;
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ double wall_pot(const xtb_wall_pair& W, double r) {          // Exp_wall_pot :1070
    if (r > W.d0 + W.dd * 10.0) return 0.0;
    const double fcut = 1.0 / (1.0 + exp((r - W.d0) / W.dd));                           // f_cut_L_C
    return W.C * exp(1.0 / (r - W.r0)) * fcut;                                          // Wall_potential :1095
}
__device__ __forceinline__ double wall_dpot(const xtb_wall_pair& W, double r) {         // d_Exp_wall_pot :1140
    if (r > W.d0 + W.dd * 10.0) return 0.0;
    const double rr0 = 1.0 / (r - W.r0);
    const double E_C = W.C * exp(rr0);
    const double d_E_C = -W.C * exp(rr0) * rr0 * rr0;                                   // d_Wall_potential :1109
    const double arg = (r - W.d0) / W.dd;
    const double fcut = 1.0 / (1.0 + exp(arg));
    double dfcut = 0.0;                                                                 // d_f_cut_L_C
    if (arg < 709.782712893384) {
        const double er = exp(arg), e2 = 1.0 + er;
        dfcut = -er / (W.dd * e2 * e2);
    }
    return d_E_C * fcut + E_C * dfcut;
}
struct SimpleWall {
    typedef xtb_wall_pair Param;
    static __device__ __forceinline__ double pot(const Param& W, double r) { return wall_pot(W, r); }
    static __device__ __forceinline__ double dpot(const Param& W, double r) { return wall_dpot(W, r); }
};

// ---- class TB_Short_Rep ----
constexpr double ZBL_A_U = 0.8854, ZBL_PHI1 = 0.1818, ZBL_PHI2 = 0.5099, ZBL_PHI3 = 0.2802, ZBL_PHI4 = 0.02817;     // ZBL_potential.f90:43-51
constexpr double ZBL_EXP1 = -3.2, ZBL_EXP2 = -0.9423, ZBL_EXP3 = -0.4028, ZBL_EXP4 = -0.2016;
constexpr double SR_G_PI = 3.1415926535897932384626433832795, SR_G_E0 = 8.854187817620e-12, SR_G_E = 1.602176634e-19, SR_G_A0 = 0.5291772085936;
constexpr double SR_M_K = 1.0 / (4.0 * SR_G_PI * SR_G_E0);                                                           // m_k :52

__device__ __forceinline__ double sr_coul_part(double Z1, double Z2, double r) { return SR_M_K * Z1 * Z2 * SR_G_E / (1.0e-10 * r); }
__device__ __forceinline__ double zbl_a(double Z1, double Z2) { return ZBL_A_U * SR_G_A0 / (pow(Z1, 0.23) + pow(Z2, 0.23)); }      // ZBL_a :448
__device__ __forceinline__ double zbl_phi(double Z1, double Z2, double r) {                                          // ZBL_phi :434
    const double x = r / zbl_a(Z1, Z2);
    return ZBL_PHI1 * exp(ZBL_EXP1 * x) + ZBL_PHI2 * exp(ZBL_EXP2 * x) + ZBL_PHI3 * exp(ZBL_EXP3 * x) + ZBL_PHI4 * exp(ZBL_EXP4 * x);
}
__device__ __forceinline__ double zbl_dphi(double Z1, double Z2, double r) {                                         // d_ZBL_phi :406
    const double a = zbl_a(Z1, Z2), x = r / a;
    return (ZBL_PHI1 * ZBL_EXP1 * exp(ZBL_EXP1 * x) + ZBL_PHI2 * ZBL_EXP2 * exp(ZBL_EXP2 * x) + ZBL_PHI3 * ZBL_EXP3 * exp(ZBL_EXP3 * x) +
            ZBL_PHI4 * ZBL_EXP4 * exp(ZBL_EXP4 * x)) / a;
}
// Find_monotonous_LE (Little_subroutines.f90:1184): 0-based index i with R[i] <= r < R[i+1]; 0 below the grid, n-1 at / above its end
__device__ __forceinline__ int tab_find(const double* R, int n, double r) {
    if (r <= R[0]) return 0;
    if (r >= R[n - 1]) return n - 1;
    int lo = 0, hi = n - 1;
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (R[mid] <= r) lo = mid; else hi = mid; }
    return lo;
}
__device__ __forceinline__ double tab_pot(const ShortDev& W, double r) {                                             // tabulated_potential :726
    const int n = W.n_tab;
    if (n <= 0) return 0.0;
    const double *R = W.tab, *E = R + n;
    if (r > R[n - 1]) return 0.0;
    int ic = tab_find(R, n, r);
    if (W.tab_spline) {
        if (ic == n - 1) ic = n - 2;
        const double x = r - R[ic], x2 = x * x;
        return R[2 * n + ic] + R[3 * n + ic] * x + R[4 * n + ic] * x2 + R[5 * n + ic] * x * x2;                       // cubic_function
    }
    int i = ic + 1;                      // upper point; the reference would read past the grid at r == R(n): clamped
    if (i > n - 1) i = n - 1;
    // linear_interpolation, default arm (no x0, y0, replac), i > 1 in Fortran numbering
    if (fabs(E[i] - E[i - 1]) < 1.0e-6 * fabs(E[i])) return E[i];
    if (fabs(R[i] - R[i - 1]) < 1.0e-6 * fabs(R[i])) return (E[i - 1] + E[i]) * 0.5;
    if (fabs(r - R[i - 1]) >= 1.0e-6 * fabs(R[i - 1]) || fabs(R[i]) > 1.0e-8)
        return E[i - 1] + (E[i] - E[i - 1]) / (R[i] - R[i - 1]) * (r - R[i - 1]);
    return E[i];
}
__device__ __forceinline__ double tab_dpot(const ShortDev& W, double r) {                                            // d_tabulated_potential :766
    const int n = W.n_tab;
    if (n <= 0) return 0.0;
    const double *R = W.tab, *E = R + n;
    if (r > R[n - 1]) return 0.0;
    int ic = tab_find(R, n, r);
    if (ic == 0) return (E[1] - E[0]) / (R[1] - R[0]);
    if (W.tab_spline) {
        if (ic == n - 1) ic = n - 2;
        const double x = r - R[ic];
        return R[3 * n + ic] + 2.0 * R[4 * n + ic] * x + 3.0 * R[5 * n + ic] * (x * x);                              // d_cubic_function
    }
    int i = ic + 1;
    if (i > n - 1) i = n - 1;
    return (E[i] - E[i - 1]) / (R[i] - R[i - 1]);
}

struct ShortRep {
    typedef ShortDev Param;
    // the six functions and, when D, their derivatives; returns through f / df
    template <bool D> static __device__ __forceinline__ void terms(const Param& W, double r, double& f, double& df) {
        double f_exp = 0.0, d_exp = 0.0, f_inv = 0.0, d_inv = 0.0, f_pow = 0.0, d_pow = 0.0, f_zbl = 0.0, d_zbl = 0.0, f_nlh = 0.0, d_nlh = 0.0;
        if (W.use_exp && fabs(W.exp_a) > 1.0e-5) {                              // exp_function :935, d_exp_function :952
            f_exp = W.exp_Phi * exp(-(r - W.exp_r0) / W.exp_a);
            if (D) d_exp = -W.exp_Phi * exp(-(r - W.exp_r0) / W.exp_a) / W.exp_a;
        }
        if (W.use_inv_exp) {                                                    // Wall_potential :1095, d_Wall_potential :1109
            const double rr0 = 1.0 / (r - W.inv_r0);
            f_inv = W.inv_C * exp(rr0);
            if (D) d_inv = -W.inv_C * exp(rr0) * rr0 * rr0;
        }
        for (int i = 0; i < W.n_pow; ++i) {                                     // power_function :850, d_power_function :864
            const double t = W.pow_Phi[i] * pow(r / W.pow_r0[i], W.pow_m[i]);
            f_pow = f_pow + t;
            if (D) d_pow = d_pow + t * W.pow_m[i] / r;
        }
        if (W.use_zbl) {                                                        // ZBL_pot :366, d_ZBL_pot :376
            const double part = sr_coul_part(W.Z1, W.Z2, r), phi = zbl_phi(W.Z1, W.Z2, r);
            f_zbl = part * phi;
            if (D) d_zbl = part * (zbl_dphi(W.Z1, W.Z2, r) - phi / r);
        }
        if (W.use_nlh) {                                                        // NLH_pot :150, d_NLH_pot :164
            const double part = sr_coul_part(W.Z1, W.Z2, r);
            const double e0 = exp(-W.nlh_b[0] * r), e1 = exp(-W.nlh_b[1] * r), e2 = exp(-W.nlh_b[2] * r);
            const double phi = W.nlh_a[0] * e0 + W.nlh_a[1] * e1 + W.nlh_a[2] * e2;
            f_nlh = part * phi;
            if (D) d_nlh = part * ((-W.nlh_b[0] * W.nlh_a[0] * e0 + -W.nlh_b[1] * W.nlh_a[1] * e1 + -W.nlh_b[2] * W.nlh_a[2] * e2) - phi / r);
        }
        const double f_tab = tab_pot(W, r);
        f = f_inv + f_exp + f_pow + f_zbl + f_nlh + f_tab;                      // :154
        if (D) df = d_inv + d_exp + d_pow + d_zbl + d_nlh + tab_dpot(W, r);     // :199
    }
    static __device__ __forceinline__ double pot(const Param& W, double r) {    // Shortrange_pot :130-157
        if (!(r < W.d0 + W.dd * 10.0)) return 0.0;
        double f, df;
        terms<false>(W, r, f, df);
        return f * (1.0 / (1.0 + exp((r - W.d0) / W.dd)));
    }
    static __device__ __forceinline__ double dpot(const Param& W, double r) {   // d_Short_range_pot :160-203
        if (!(r < W.d0 + W.dd * 10.0)) return 0.0;
        double f, df;
        terms<true>(W, r, f, df);
        const double arg = (r - W.d0) / W.dd;
        const double fcut = 1.0 / (1.0 + exp(arg));
        double dfcut = 0.0;                                                     // d_f_cut_L_C
        if (arg < 709.782712893384) {
            const double er = exp(arg), e2 = 1.0 + er;
            dfcut = -er / (W.dd * e2 * e2);
        }
        return f * dfcut + df * fcut;                                           // :201
    }
};

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(FULL, v, off);
    return v;
}

template <class P>
__global__ void __launch_bounds__(32 * WW) wall_energy_kernel(int na, int cap, int n_kinds, const int* __restrict__ koa,
                                                              const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                              const double* __restrict__ nn_dist, const typename P::Param* __restrict__ WP,
                                                              double* __restrict__ Epot) {
    const int i = blockIdx.x * WW + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= na) return;
    const int m = nn_size[i], ki = koa[i];
    double s = 0.0;
    for (int q = lane; q < m; q += 32) {
        const size_t e = (size_t)i * cap + q;
        s += P::pot(WP[ki + n_kinds * koa[nn_list[e]]], nn_dist[4 * e + 3]);            // TB_Expwall(KOA1,KOA2) :1051
    }
    s = warp_sum(s);
    if (lane == 0) Epot[i] = 0.5 * s;                                                  // :1058
}

__global__ void __launch_bounds__(256) wall_esum_kernel(int na, const double* __restrict__ Epot, double* __restrict__ out) {
    __shared__ double buf[256];
    double s = 0.0;
    for (int i = threadIdx.x; i < na; i += 256) s += Epot[i];
    buf[threadIdx.x] = s;
    __syncthreads();
    for (int off = 128; off > 0; off >>= 1) {
        if (threadIdx.x < off) buf[threadIdx.x] += buf[threadIdx.x + off];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[0] = buf[0];
}

template <class P>
__global__ void __launch_bounds__(32 * WW) wall_forces_kernel(int a0, int a1, int cap, int n_kinds, const int* __restrict__ koa,
                                                              const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                              const int* __restrict__ rev, const double* __restrict__ nn_dist,
                                                              const typename P::Param* __restrict__ WP, const double* __restrict__ hdev,
                                                              double* __restrict__ F) {
    const int k = a0 + blockIdx.x * WW + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (k >= a1) return;
    double h[9];
#pragma unroll
    for (int e = 0; e < 9; ++e) h[e] = hdev[e];             // supce(a,b) = h[a + 3b]
    const int m = nn_size[k], kk = koa[k];
    double acc[3] = {0.0, 0.0, 0.0};
    for (int q = lane; q < m; q += 32) {
        const size_t e = (size_t)k * cap + q;
        const int j = nn_list[e], kj = koa[j];
        {   // i1 = k, j1 = j: dik - djk = +1
            const double4 d = *reinterpret_cast<const double4*>(nn_dist + 4 * e);
            const double bd = P::dpot(WP[kk + n_kinds * kj], d.w) * (1.0 / d.w);
#pragma unroll
            for (int al = 0; al < 3; ++al) acc[al] += bd * (d.x * h[al] + d.y * h[al + 3] + d.z * h[al + 6]);   // x1 :1291-1293
        }
        const int rv = rev[e];
        if (rv >= 0) {   // i1 = j, j1 = k: dik - djk = -1, geometry from j's own record
            const double4 d = *reinterpret_cast<const double4*>(nn_dist + 4 * ((size_t)j * cap + rv));
            const double bd = P::dpot(WP[kj + n_kinds * kk], d.w) * (-1.0 / d.w);
#pragma unroll
            for (int al = 0; al < 3; ++al) acc[al] += bd * (d.x * h[al] + d.y * h[al + 3] + d.z * h[al + 6]);
        }
    }
#pragma unroll
    for (int al = 0; al < 3; ++al) acc[al] = 0.5 * warp_sum(acc[al]);                   // :1309
    if (lane < 3) F[3 * (size_t)k + lane] = lane == 0 ? acc[0] : (lane == 1 ? acc[1] : acc[2]);
}

template <class P>
__global__ void __launch_bounds__(32 * WW) wall_pressure_kernel(int a0, int a1, int cap, int n_kinds, const int* __restrict__ koa,
                                                                const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                                const double* __restrict__ nn_dist, const double* __restrict__ nn_dist_s,
                                                                const typename P::Param* __restrict__ WP, double* __restrict__ partial) {
    const int i = a0 + blockIdx.x * WW + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= a1) return;
    const int m = nn_size[i], ki = koa[i];
    double acc[9];
#pragma unroll
    for (int e = 0; e < 9; ++e) acc[e] = 0.0;
    for (int q = lane; q < m; q += 32) {
        const size_t e = (size_t)i * cap + q;
        const double4 d = *reinterpret_cast<const double4*>(nn_dist + 4 * e);
        const double rc[3] = {d.x, d.y, d.z};
        const double dp = P::dpot(WP[ki + n_kinds * koa[nn_list[e]]], d.w);
#pragma unroll
        for (int k = 0; k < 3; ++k)
#pragma unroll
            for (int l = 0; l < 3; ++l) acc[l + 3 * k] += dp * rc[k] * nn_dist_s[3 * e + l] / d.w;     // :1369
    }
#pragma unroll
    for (int e = 0; e < 9; ++e) {
        const double s = warp_sum(acc[e]);
        if (lane == 0) partial[(size_t)(i - a0) * 9 + e] = 0.5 * s;                     // :1377
    }
}

__global__ void wall_sum9_kernel(int n, const double* __restrict__ partial, double* __restrict__ out) {
    __shared__ double buf[128];
    const int k1 = blockIdx.x;
    double s = 0.0;
    for (int a = threadIdx.x; a < n; a += 128) s += partial[(size_t)a * 9 + k1];
    buf[threadIdx.x] = s;
    __syncthreads();
    for (int off = 64; off > 0; off >>= 1) {
        if (threadIdx.x < off) buf[threadIdx.x] += buf[threadIdx.x + off];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[k1] = buf[0];
}

}  // namespace

int xtb_impl_wall_energy(xtb_ctx* c) {
    const int na = c->na;
    int rc;
    if ((rc = xtb_reserve(c, c->d_Ewall_at, sizeof(double) * na))) return rc;
    if ((rc = xtb_reserve(c, c->d_Ewall, sizeof(double)))) return rc;
    XTB_T0(c, 9);
    if (!c->h_short.empty())
        wall_energy_kernel<ShortRep><<<(na + WW - 1) / WW, 32 * WW, 0, c->stream>>>(na, c->cap, c->n_kinds, (const int*)c->d_koa.p, (const int*)c->d_nn_size.p,
                                                                                 (const int*)c->d_nn_list.p, (const double*)c->d_nn_dist.p,
                                                                                 (const ShortDev*)c->d_short_params.p, (double*)c->d_Ewall_at.p);
    else
        wall_energy_kernel<SimpleWall><<<(na + WW - 1) / WW, 32 * WW, 0, c->stream>>>(na, c->cap, c->n_kinds, (const int*)c->d_koa.p, (const int*)c->d_nn_size.p,
                                                                                   (const int*)c->d_nn_list.p, (const double*)c->d_nn_dist.p,
                                                                                   (const xtb_wall_pair*)c->d_wall_params.p, (double*)c->d_Ewall_at.p);
    XTB_LAUNCH_CHECK(c);
    wall_esum_kernel<<<1, 256, 0, c->stream>>>(na, (const double*)c->d_Ewall_at.p, (double*)c->d_Ewall.p);
    XTB_LAUNCH_CHECK(c);
    XTB_T1(c, 9);
    return xtb_stage_finish(c);
}

int xtb_impl_wall_forces(xtb_ctx* c) {
    const int nown = c->a1 - c->a0;
    int rc;
    if ((rc = xtb_reserve(c, c->d_Fwall, sizeof(double) * 3 * (size_t)(c->na + c->nranks)))) return rc;
    XTB_T0(c, 9);
    if (nown > 0) {
        if (!c->h_short.empty())
            wall_forces_kernel<ShortRep><<<(nown + WW - 1) / WW, 32 * WW, 0, c->stream>>>(c->a0, c->a1, c->cap, c->n_kinds, (const int*)c->d_koa.p,
                                                                                     (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p,
                                                                                     (const int*)c->d_rev.p, (const double*)c->d_nn_dist.p,
                                                                                     (const ShortDev*)c->d_short_params.p,
                                                                                     (const double*)c->d_hdev.p, (double*)c->d_Fwall.p);
        else
            wall_forces_kernel<SimpleWall><<<(nown + WW - 1) / WW, 32 * WW, 0, c->stream>>>(c->a0, c->a1, c->cap, c->n_kinds, (const int*)c->d_koa.p,
                                                                                       (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p,
                                                                                       (const int*)c->d_rev.p, (const double*)c->d_nn_dist.p,
                                                                                       (const xtb_wall_pair*)c->d_wall_params.p,
                                                                                       (const double*)c->d_hdev.p, (double*)c->d_Fwall.p);
        XTB_LAUNCH_CHECK(c);
    }
    XTB_T1(c, 9);
    return xtb_stage_finish(c);
}

int xtb_impl_wall_pressure(xtb_ctx* c) {
    const int nown = c->a1 - c->a0;
    int rc;
    if ((rc = xtb_reserve(c, c->d_SCwall, sizeof(double) * 9))) return rc;
    if ((rc = xtb_reserve(c, c->d_ppart, sizeof(double) * 9 * (size_t)(nown > 0 ? nown : 1)))) return rc;
    XTB_T0(c, 10);
    if (nown > 0) {
        if (!c->h_short.empty())
            wall_pressure_kernel<ShortRep><<<(nown + WW - 1) / WW, 32 * WW, 0, c->stream>>>(c->a0, c->a1, c->cap, c->n_kinds, (const int*)c->d_koa.p,
                                                                                       (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p,
                                                                                       (const double*)c->d_nn_dist.p, (const double*)c->d_nn_dist_s.p,
                                                                                       (const ShortDev*)c->d_short_params.p, (double*)c->d_ppart.p);
        else
            wall_pressure_kernel<SimpleWall><<<(nown + WW - 1) / WW, 32 * WW, 0, c->stream>>>(c->a0, c->a1, c->cap, c->n_kinds, (const int*)c->d_koa.p,
                                                                                         (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p,
                                                                                         (const double*)c->d_nn_dist.p, (const double*)c->d_nn_dist_s.p,
                                                                                         (const xtb_wall_pair*)c->d_wall_params.p, (double*)c->d_ppart.p);
        XTB_LAUNCH_CHECK(c);
        wall_sum9_kernel<<<9, 128, 0, c->stream>>>(nown, (const double*)c->d_ppart.p, (double*)c->d_SCwall.p);
        XTB_LAUNCH_CHECK(c);
    } else {
        XTB_CUDA(c, cudaMemsetAsync(c->d_SCwall.p, 0, sizeof(double) * 9, c->stream));
    }
    XTB_T1(c, 10);
    return xtb_stage_finish(c);
}
