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
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(FULL, v, off);
    return v;
}

__global__ void __launch_bounds__(32 * WW) wall_energy_kernel(int na, int cap, int n_kinds, const int* __restrict__ koa,
                                                              const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                              const double* __restrict__ nn_dist, const xtb_wall_pair* __restrict__ WP,
                                                              double* __restrict__ Epot) {
    const int i = blockIdx.x * WW + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= na) return;
    const int m = nn_size[i], ki = koa[i];
    double s = 0.0;
    for (int q = lane; q < m; q += 32) {
        const size_t e = (size_t)i * cap + q;
        s += wall_pot(WP[ki + n_kinds * koa[nn_list[e]]], nn_dist[4 * e + 3]);          // TB_Expwall(KOA1,KOA2) :1051
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

__global__ void __launch_bounds__(32 * WW) wall_forces_kernel(int a0, int a1, int cap, int n_kinds, const int* __restrict__ koa,
                                                              const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                              const int* __restrict__ rev, const double* __restrict__ nn_dist,
                                                              const xtb_wall_pair* __restrict__ WP, const double* __restrict__ hdev,
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
            const double bd = wall_dpot(WP[kk + n_kinds * kj], d.w) * (1.0 / d.w);
#pragma unroll
            for (int al = 0; al < 3; ++al) acc[al] += bd * (d.x * h[al] + d.y * h[al + 3] + d.z * h[al + 6]);   // x1 :1291-1293
        }
        const int rv = rev[e];
        if (rv >= 0) {   // i1 = j, j1 = k: dik - djk = -1, geometry from j's own record
            const double4 d = *reinterpret_cast<const double4*>(nn_dist + 4 * ((size_t)j * cap + rv));
            const double bd = wall_dpot(WP[kj + n_kinds * kk], d.w) * (-1.0 / d.w);
#pragma unroll
            for (int al = 0; al < 3; ++al) acc[al] += bd * (d.x * h[al] + d.y * h[al + 3] + d.z * h[al + 6]);
        }
    }
#pragma unroll
    for (int al = 0; al < 3; ++al) acc[al] = 0.5 * warp_sum(acc[al]);                   // :1309
    if (lane < 3) F[3 * (size_t)k + lane] = lane == 0 ? acc[0] : (lane == 1 ? acc[1] : acc[2]);
}

__global__ void __launch_bounds__(32 * WW) wall_pressure_kernel(int a0, int a1, int cap, int n_kinds, const int* __restrict__ koa,
                                                                const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                                const double* __restrict__ nn_dist, const double* __restrict__ nn_dist_s,
                                                                const xtb_wall_pair* __restrict__ WP, double* __restrict__ partial) {
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
        const double dp = wall_dpot(WP[ki + n_kinds * koa[nn_list[e]]], d.w);
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
    wall_energy_kernel<<<(na + WW - 1) / WW, 32 * WW, 0, c->stream>>>(na, c->cap, c->n_kinds, (const int*)c->d_koa.p, (const int*)c->d_nn_size.p,
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
        wall_forces_kernel<<<(nown + WW - 1) / WW, 32 * WW, 0, c->stream>>>(c->a0, c->a1, c->cap, c->n_kinds, (const int*)c->d_koa.p,
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
        wall_pressure_kernel<<<(nown + WW - 1) / WW, 32 * WW, 0, c->stream>>>(c->a0, c->a1, c->cap, c->n_kinds, (const int*)c->d_koa.p,
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
