// Repulsive pair terms of the Pettifor-type branch on the neighbour stream.
//
// Replaces, in Source_files/TB_Pettifor.f90:
//   get_Erep_s :2074-2105 (fx :2108, phi :2126)               E_rep = sum_i [ f(psi_i) + E0_TB ],  psi_i = sum_j phi(r_ij)
//   dErdr_s :2162-2306 (dphi :2695, dfx :2776)                forces%rep(k) = dE_rep/ds_k
//   dErdr_Pressure_s :2634-2690                               SCforce%rep(3,3)
// Reference data flow: dErdr_s loops, for EVERY atom ian, over EVERY atom i1 and its whole list (O(Na^2 nn)) although
// only i1 = ian and the i1 that list ian contribute.  Here: one warp per atom; pass 1 gives psi_i, f(psi_i), f'(psi_i)
// (lanes stride the list, fixed-order shuffle reduction); pass 2 gives the force of atom k from its own entries
// (d_ik = 1) and from the mirrored entries of its neighbours (d_jk = 1, located with the `rev` index and read from the
// NEIGHBOUR's record, as the reference does) ; the supercell force is one 9-vector per atom, summed in fixed order.
// HBM-bound, tiny: 36 bytes per list entry, ~3 transcendental calls per entry.
#include "xtb_ctx.h"

namespace {

constexpr int RW = 4;                    // warps (atoms) per CTA
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ double rep_phi(const xtb_pettifor_rep_pair& R, double r) {        // :2126
    if (r > R.dm) return 0.0;
    if (r > R.d1) {
        const double x1 = r - R.d1, x2 = x1 * x1;
        return R.c0[0] + R.c0[1] * x1 + R.c0[2] * x2 + R.c0[3] * x1 * x2;
    }
    if (r <= 0.0) return 0.0;
    return R.phi0 * pow(R.d0 / r, R.m) * exp(R.m * (-pow(r / R.dc, R.mc) + pow(R.d0 / R.dc, R.mc)));
}
__device__ __forceinline__ double rep_dphi(const xtb_pettifor_rep_pair& R, double r) {       // :2695
    if (r > R.dm) return 0.0;
    if (r > R.d1) {
        const double x1 = r - R.d1;
        return R.c0[1] + R.c0[2] * x1 * 2.0 + R.c0[3] * x1 * x1 * 3.0;
    }
    if (r <= 0.0) return 0.0;
    return -R.m * rep_phi(R, r) / r * (1.0 + R.mc * pow(r / R.dc, R.mc));
}
__device__ __forceinline__ double rep_fx(const xtb_pettifor_rep_pair& R, double x) {         // :2108
    const double x2 = x * x;
    return R.a0[0] + R.a0[1] * x + R.a0[2] * x2 + R.a0[3] * x2 * x + R.a0[4] * x2 * x2;
}
__device__ __forceinline__ double rep_dfx(const xtb_pettifor_rep_pair& R, double x) {        // :2776
    const double x2 = x * x;
    return R.a0[1] + R.a0[2] * x * 2.0 + R.a0[3] * x2 * 3.0 + R.a0[4] * x * x2 * 4.0;
}
__device__ __forceinline__ double warp_sum(double v) {      // fixed tree: same bits on every run
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(FULL, v, off);
    return v;
}

// pass 1: every atom (neighbours of owned atoms are needed in pass 2).  out: Epot[i], dfpsi[i]
__global__ void __launch_bounds__(32 * RW) rep_psi_kernel(int na, int cap, int n_kinds, const int* __restrict__ koa,
                                                          const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                          const double* __restrict__ nn_dist,
                                                          const xtb_pettifor_rep_pair* __restrict__ REP,
                                                          double* __restrict__ Epot, double* __restrict__ dfpsi) {
    const int i = blockIdx.x * RW + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= na) return;
    const int m = nn_size[i], ki = koa[i];
    double psi = 0.0;
    for (int q = lane; q < m; q += 32) {
        const size_t e = (size_t)i * cap + q;
        psi += rep_phi(REP[koa[nn_list[e]] + n_kinds * ki], nn_dist[4 * e + 3]);     // TB_Repuls(KOA(j1), KOA(i1))  :2094
    }
    psi = warp_sum(psi);
    if (lane == 0) {
        const xtb_pettifor_rep_pair& R = REP[ki + n_kinds * ki];
        Epot[i] = rep_fx(R, psi) + R.E0_TB;                                           // :2099
        dfpsi[i] = rep_dfx(R, psi);
    }
}

// total energy: one CTA, fixed order
__global__ void __launch_bounds__(256) rep_esum_kernel(int na, const double* __restrict__ Epot, double* __restrict__ out) {
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

// pass 2: force on owned atom k.  Own entries: +f'(psi_k) dphi(r)/r x1 ; mirrored entries of neighbour j: -f'(psi_j) dphi(r')/r' x1'
__global__ void __launch_bounds__(32 * RW) rep_forces_kernel(int a0, int a1, int cap, int n_kinds, const int* __restrict__ koa,
                                                             const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                             const int* __restrict__ rev, const double* __restrict__ nn_dist,
                                                             const xtb_pettifor_rep_pair* __restrict__ REP,
                                                             const double* __restrict__ hdev, const double* __restrict__ dfpsi,
                                                             double* __restrict__ F) {
    const int k = a0 + blockIdx.x * RW + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (k >= a1) return;
    double h[9];
#pragma unroll
    for (int e = 0; e < 9; ++e) h[e] = hdev[e];             // supce(a,b) = h[a + 3b]
    const int m = nn_size[k], kk = koa[k];
    const double dfk = dfpsi[k];
    double acc[3] = {0.0, 0.0, 0.0};
    for (int q = lane; q < m; q += 32) {
        const size_t e = (size_t)k * cap + q;
        const int j = nn_list[e], kj = koa[j];
        {   // i1 = k, j1 = j: dik - djk = +1
            const double4 d = *reinterpret_cast<const double4*>(nn_dist + 4 * e);
            const double bd = dfk * (rep_dphi(REP[kj + n_kinds * kk], d.w) * (1.0 / d.w));
#pragma unroll
            for (int al = 0; al < 3; ++al) acc[al] += bd * (d.x * h[al] + d.y * h[al + 3] + d.z * h[al + 6]);   // x1(al) :2273
        }
        const int rv = rev[e];
        if (rv >= 0) {   // i1 = j, j1 = k: dik - djk = -1, geometry from j's own record
            const double4 d = *reinterpret_cast<const double4*>(nn_dist + 4 * ((size_t)j * cap + rv));
            const double bd = dfpsi[j] * (rep_dphi(REP[kk + n_kinds * kj], d.w) * (-1.0 / d.w));
#pragma unroll
            for (int al = 0; al < 3; ++al) acc[al] += bd * (d.x * h[al] + d.y * h[al + 3] + d.z * h[al + 6]);
        }
    }
#pragma unroll
    for (int al = 0; al < 3; ++al) acc[al] = warp_sum(acc[al]);
    if (lane < 3) F[3 * (size_t)k + lane] = lane == 0 ? acc[0] : (lane == 1 ? acc[1] : acc[2]);
}

// supercell force: per owned atom i, Rep_Pr(l,k) = sum_j dphi(r) r_k s_l / r, times f'(psi_i)   (:2648-2676)
__global__ void __launch_bounds__(32 * RW) rep_pressure_kernel(int a0, int a1, int cap, int n_kinds, const int* __restrict__ koa,
                                                               const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                               const double* __restrict__ nn_dist, const double* __restrict__ nn_dist_s,
                                                               const xtb_pettifor_rep_pair* __restrict__ REP,
                                                               const double* __restrict__ dfpsi, double* __restrict__ partial) {
    const int i = a0 + blockIdx.x * RW + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= a1) return;
    const int m = nn_size[i], ki = koa[i];
    double acc[9];
#pragma unroll
    for (int e = 0; e < 9; ++e) acc[e] = 0.0;
    for (int q = lane; q < m; q += 32) {
        const size_t e = (size_t)i * cap + q;
        const double4 d = *reinterpret_cast<const double4*>(nn_dist + 4 * e);
        const double rc[3] = {d.x, d.y, d.z};
        const double dp = rep_dphi(REP[koa[nn_list[e]] + n_kinds * ki], d.w);
#pragma unroll
        for (int k = 0; k < 3; ++k)
#pragma unroll
            for (int l = 0; l < 3; ++l) acc[l + 3 * k] += dp * rc[k] * nn_dist_s[3 * e + l] / d.w;
    }
    const double df = dfpsi[i];
#pragma unroll
    for (int e = 0; e < 9; ++e) {
        const double s = warp_sum(acc[e]);
        if (lane == 0) partial[(size_t)(i - a0) * 9 + e] = df * s;
    }
}

__global__ void rep_sum9_kernel(int n, const double* __restrict__ partial, double* __restrict__ out) {
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

// psi / Epot / f'(psi) of every atom + the total energy; shared by the three entry points
static int rep_pass1(xtb_ctx* c) {
    const int na = c->na;
    int rc;
    if ((rc = xtb_reserve(c, c->d_Epot, sizeof(double) * na))) return rc;
    if ((rc = xtb_reserve(c, c->d_dfpsi, sizeof(double) * na))) return rc;
    if ((rc = xtb_reserve(c, c->d_Erep, sizeof(double)))) return rc;
    rep_psi_kernel<<<(na + RW - 1) / RW, 32 * RW, 0, c->stream>>>(na, c->cap, c->n_kinds, (const int*)c->d_koa.p, (const int*)c->d_nn_size.p,
                                                                   (const int*)c->d_nn_list.p, (const double*)c->d_nn_dist.p,
                                                                   (const xtb_pettifor_rep_pair*)c->d_rep_params.p,
                                                                   (double*)c->d_Epot.p, (double*)c->d_dfpsi.p);
    XTB_LAUNCH_CHECK(c);
    return XTB_OK;
}

int xtb_impl_pettifor_rep_energy(xtb_ctx* c) {
    XTB_T0(c, 9);
    int rc = rep_pass1(c);
    if (rc) return rc;
    rep_esum_kernel<<<1, 256, 0, c->stream>>>(c->na, (const double*)c->d_Epot.p, (double*)c->d_Erep.p);
    XTB_LAUNCH_CHECK(c);
    XTB_T1(c, 9);
    if ((rc = xtb_stage_finish(c))) return rc;
    return XTB_OK;
}

int xtb_impl_pettifor_rep_forces(xtb_ctx* c) {
    const int nown = c->a1 - c->a0;
    int rc;
    if ((rc = xtb_reserve(c, c->d_Frep, sizeof(double) * 3 * (size_t)(c->na + c->nranks)))) return rc;
    XTB_T0(c, 9);
    if ((rc = rep_pass1(c))) return rc;
    if (nown > 0) {
        rep_forces_kernel<<<(nown + RW - 1) / RW, 32 * RW, 0, c->stream>>>(c->a0, c->a1, c->cap, c->n_kinds, (const int*)c->d_koa.p,
                                                                         (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p,
                                                                         (const int*)c->d_rev.p, (const double*)c->d_nn_dist.p,
                                                                         (const xtb_pettifor_rep_pair*)c->d_rep_params.p,
                                                                         (const double*)c->d_hdev.p, (const double*)c->d_dfpsi.p,
                                                                         (double*)c->d_Frep.p);
        XTB_LAUNCH_CHECK(c);
    }
    XTB_T1(c, 9);
    if ((rc = xtb_stage_finish(c))) return rc;
    return XTB_OK;
}

int xtb_impl_pettifor_rep_pressure(xtb_ctx* c) {
    const int nown = c->a1 - c->a0;
    int rc;
    if ((rc = xtb_reserve(c, c->d_SCrep, sizeof(double) * 9))) return rc;
    if ((rc = xtb_reserve(c, c->d_ppart, sizeof(double) * 9 * (size_t)(nown > 0 ? nown : 1)))) return rc;
    XTB_T0(c, 10);
    if ((rc = rep_pass1(c))) return rc;
    if (nown > 0) {
        rep_pressure_kernel<<<(nown + RW - 1) / RW, 32 * RW, 0, c->stream>>>(c->a0, c->a1, c->cap, c->n_kinds, (const int*)c->d_koa.p,
                                                                           (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p,
                                                                           (const double*)c->d_nn_dist.p, (const double*)c->d_nn_dist_s.p,
                                                                           (const xtb_pettifor_rep_pair*)c->d_rep_params.p,
                                                                           (const double*)c->d_dfpsi.p, (double*)c->d_ppart.p);
        XTB_LAUNCH_CHECK(c);
        rep_sum9_kernel<<<9, 128, 0, c->stream>>>(nown, (const double*)c->d_ppart.p, (double*)c->d_SCrep.p);
        XTB_LAUNCH_CHECK(c);
    } else {
        XTB_CUDA(c, cudaMemsetAsync(c->d_SCrep.p, 0, sizeof(double) * 9, c->stream));
    }
    XTB_T1(c, 10);
    if ((rc = xtb_stage_finish(c))) return rc;
    return XTB_OK;
}
