// Pettifor-type orthogonal sp3 branch: H assembly, Hellmann-Feynman forces, supercell force.
//
// Replaces, in Source_files/TB_Pettifor.f90:
//   Hamil_tot :1658-1810 -> Hamilton_one :1813-1868 -> Hopping :1872-1930 -> Vs :1935-1976          (S2)
//   dHij_s :141-223 -> Construct_M_Vs :227-350, dHamil_tot_s :463-521, dHamilton_one_s :571-618,
//                      dHopping_s :909-1130, Attract_TB_H3_near :354-399                             (S4)
//   Attract_TB_Forces_Press :1134-1198 -> dHamil_tot_Press :1201, dHopping_Press_h :1283-1423        (S5)
// Reference data flow: Hamil_tot is serial and re-runs the 27-image minimum search per pair; dHij_s rebuilds three
// dense (4Na)^2 matrices for EVERY atom k although only blocks with k as one of the two atoms are non-zero.
// Here: one thread per 4x4 block for the assembly (radial functions and direction cosines in registers, block store
// -> dense panels by the shared dense_panel_kernel), one CTA per atom for the forces (the two non-zero blocks per
// neighbour, contracted with the atom's own contiguous rho blocks; fixed-order reduction, no atomics).
// Reference semantics kept: block (rows neighbour i, cols outer j) from the minimum image of (i, j) with
// TB(kind j, kind i) (:1769,1842); Vs takes the cubic tail for r >= r1 but dVs for r > r1 (:1967 vs :2010);
// force blocks are stored transposed (dHijx(:,ki) = dtsx(ki,:), :597); no on-site derivative terms.
#include "xtb_ctx.h"

namespace {

__device__ __forceinline__ double pet_Vs(const xtb_pettifor_pair& P, int i, double r) {
    if (r > P.rm) return 0.0;
    if (r >= P.r1) {
        const double d = r - P.r1, d2 = d * d;
        return P.V0[i] * (P.c0[i] + P.c1[i] * d + P.c2[i] * d2 + P.c3[i] * d2 * d);
    }
    return P.V0[i] * pow(P.r0 / r, P.n) * exp(P.n * (-pow(r / P.rc[i], P.nc[i]) + pow(P.r0 / P.rc[i], P.nc[i])));
}
__device__ __forceinline__ double pet_dVs(const xtb_pettifor_pair& P, int i, double r, double V) {
    if (r > P.rm) return 0.0;
    if (r > P.r1) {
        const double d = r - P.r1;
        return P.V0[i] * (P.c1[i] + P.c2[i] * d * 2.0 + P.c3[i] * d * d * 3.0);
    }
    return -P.n * V / r * (1.0 + P.nc[i] * pow(r / P.rc[i], P.nc[i]));
}

// ts(a,b) of Hopping (:1898-1913), o[a + 4*b]
__device__ __forceinline__ void hopping_block(const double* V, double xr, double yr, double zr, double* o) {
    const double w = V[2] - V[3];
    o[0] = V[0];        o[4] = -xr * V[1];            o[8] = -yr * V[1];             o[12] = -zr * V[1];
    o[1] = xr * V[1];   o[5] = xr * xr * w + V[3];    o[9] = xr * yr * w;            o[13] = xr * zr * w;
    o[2] = yr * V[1];   o[6] = xr * yr * w;           o[10] = yr * yr * w + V[3];    o[14] = yr * zr * w;
    o[3] = zr * V[1];   o[7] = xr * zr * w;           o[11] = yr * zr * w;           o[15] = zr * zr * w + V[3];
}

// one thread per block (atom j, slot q): q = 0 on-site, else list entry q-1 (neighbour i).  Block store layout
// [mu + 4*nu], mu = orbital of the row atom i, nu = orbital of j (Hij(k,l) = Hij1(i1,j1), :1776).
__global__ void __launch_bounds__(128) pet_blocks_kernel(int na, int a0, int a1, int cap, int n_kinds, const int* __restrict__ koa,
                                                        const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                        const int* __restrict__ rev, const double* __restrict__ nn_dist,
                                                        const int* __restrict__ rowptr, const xtb_pettifor_pair* __restrict__ TB,
                                                        double* __restrict__ Hblk) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int j = a0 + (int)(t / (cap + 1));
    if (j >= a1) return;
    const int q = (int)(t % (cap + 1));
    if (q > nn_size[j]) return;
    double* out = Hblk + ((size_t)rowptr[j] + q) * 16;
    double o[16];
    if (q == 0) {
        const xtb_pettifor_pair& P = TB[koa[j] + n_kinds * koa[j]];
#pragma unroll
        for (int e = 0; e < 16; ++e) o[e] = 0.0;
        o[0] = P.Es; o[5] = P.Ep; o[10] = P.Ep; o[15] = P.Ep;                       // :1831-1835
    } else {
        const size_t e = (size_t)j * cap + (q - 1);
        const int i = nn_list[e];
        const int rv = rev[e];
        // shortest_distance(i, j) (:1842) = the list entry of i for j; if j is missing from i's list (possible only
        // for a pair exactly at the cut-off) the mirrored entry of j is used.
        double x, y, z, r;
        if (rv >= 0) {
            const double4 d = *reinterpret_cast<const double4*>(nn_dist + 4 * ((size_t)i * cap + rv));
            x = d.x; y = d.y; z = d.z; r = d.w;
        } else {
            const double4 d = *reinterpret_cast<const double4*>(nn_dist + 4 * e);
            x = -d.x; y = -d.y; z = -d.z; r = d.w;
        }
        const xtb_pettifor_pair& P = TB[koa[j] + n_kinds * koa[i]];                 // TB_Hamil(KOA(j), KOA(i)) :1769
        if (r <= P.rm) {
            double V[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) V[c] = pet_Vs(P, c, r);
            hopping_block(V, x / r, y / r, z / r, o);
        } else {
#pragma unroll
            for (int c = 0; c < 16; ++c) o[c] = 0.0;
        }
    }
    double2* o2 = reinterpret_cast<double2*>(out);
#pragma unroll
    for (int c = 0; c < 8; ++c) o2[c] = make_double2(o[2 * c], o[2 * c + 1]);
}

// dts_al(a,b) of dHopping_s (:1035-1117) for the three directions, given V, dV, the direction cosines, dr/ds_al and
// d(l,m,n)/ds_al; returns the contraction sum_{a,b} dH(a,b) A(a,b) with dH(a,b) = dts(b,a) (:597) for each direction.
// A(a,b) is read as Ab[a*sa + b*sb].
__device__ __forceinline__ void pet_contract(const double* V, const double* dV, double xr, double yr, double zr,
                                             const double* drds, const double* dl, const double* dm, const double* dn,
                                             const double* Ab, int sa, int sb, int ncomp, double* res) {
    const double w = V[2] - V[3], dw = dV[2] - dV[3];
    for (int al = 0; al < ncomp; ++al) {
        double d[16];
        const double dr = drds[al];
        d[0] = dV[0] * dr;
        d[1] = -(dl[al] * V[1] + xr * dV[1] * dr);  d[4] = -d[1];
        d[2] = -(dm[al] * V[1] + yr * dV[1] * dr);  d[8] = -d[2];
        d[3] = -(dn[al] * V[1] + zr * dV[1] * dr);  d[12] = -d[3];
        d[5] = 2.0 * xr * w * dl[al] + (xr * xr * dw + dV[3]) * dr;
        d[6] = (xr * dm[al] + yr * dl[al]) * w + xr * yr * dw * dr;  d[9] = d[6];
        d[7] = (xr * dn[al] + zr * dl[al]) * w + xr * zr * dw * dr;  d[13] = d[7];
        d[10] = 2.0 * yr * w * dm[al] + (yr * yr * dw + dV[3]) * dr;
        d[11] = (yr * dn[al] + zr * dm[al]) * w + yr * zr * dw * dr; d[14] = d[11];
        d[15] = 2.0 * zr * w * dn[al] + (zr * zr * dw + dV[3]) * dr;
        double s = 0.0;
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) s += d[b + 4 * a] * Ab[a * sa + b * sb];     // dH(a,b) = dts(b,a)
        res[al] = s;
    }
}

constexpr int PT = 128;

// forces: one CTA per owned atom k; thread q handles neighbour q: the block (k outer, q) with delta = +1 and the
// block (neighbour outer, k) with delta = -1 (the only non-zero blocks of dHamil_tot_s for this k).
__global__ void __launch_bounds__(PT) pet_forces_kernel(int na, int a0, int cap, int n_kinds, const int* __restrict__ koa,
                                                       const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                       const int* __restrict__ rev, const double* __restrict__ nn_dist,
                                                       const int* __restrict__ rowptr, const xtb_pettifor_pair* __restrict__ TB,
                                                       const double* __restrict__ hmat, const double* __restrict__ rho,
                                                       double* __restrict__ F) {
    extern __shared__ double part[];      // [m][3]
    __shared__ double h[9];
    const int k = a0 + blockIdx.x;
    const int m = nn_size[k];
    if (threadIdx.x < 9) h[threadIdx.x] = hmat[threadIdx.x];
    __syncthreads();
    const double* blk = rho + (size_t)rowptr[k] * 16;
    for (int q = threadIdx.x; q < m; q += PT) {
        const size_t e = (size_t)k * cap + q;
        const int b = nn_list[e];
        const int rv = rev[e];
        const double* Ab = blk + (size_t)(q + 1) * 16;        // rho(k mu, b nu) at [mu + 4 nu]
        double acc[3] = {0.0, 0.0, 0.0};
#pragma unroll
        for (int side = 0; side < 2; ++side) {
            if (side == 1 && rv < 0) break;                    // b's list does not hold k: that block does not exist
            const size_t es = side == 0 ? e : (size_t)b * cap + rv;
            const double4 d4 = *reinterpret_cast<const double4*>(nn_dist + 4 * es);
            const double x = d4.x, y = d4.y, z = d4.z, r = d4.w;
            const xtb_pettifor_pair& P = side == 0 ? TB[koa[k] + n_kinds * koa[b]] : TB[koa[b] + n_kinds * koa[k]];
            const double bb = (side == 0 ? 1.0 : -1.0) / r;    // (dik - djk)/r  :940
            double V[4], dV[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) { V[c] = pet_Vs(P, c, r); dV[c] = pet_dVs(P, c, r, V[c]); }
            const double xrr = x / r / r, yrr = y / r / r, zrr = z / r / r;
            double drds[3], dl[3], dm[3], dn[3];
#pragma unroll
            for (int al = 0; al < 3; ++al) {
                drds[al] = bb * (x * h[al] + y * h[al + 3] + z * h[al + 6]);          // b * M_x1(al)
                dl[al] = bb * h[al] - xrr * drds[al];                                   // supce(al,1)
                dm[al] = bb * h[al + 3] - yrr * drds[al];
                dn[al] = bb * h[al + 6] - zrr * drds[al];
            }
            double res[3];
            // side 0: rows = k orbitals (a), cols = b orbitals: A(a,b) = Ab[a + 4b]
            // side 1: rows = b orbitals (a), cols = k orbitals: A(a,b) = rho(b a, k b') = rho(k b', b a) = Ab[b' + 4a]
            pet_contract(V, dV, x / r, y / r, z / r, drds, dl, dm, dn, Ab, side == 0 ? 1 : 4, side == 0 ? 4 : 1, 3, res);
            acc[0] += res[0]; acc[1] += res[1]; acc[2] += res[2];
        }
        part[q * 3] = acc[0]; part[q * 3 + 1] = acc[1]; part[q * 3 + 2] = acc[2];
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        double s = 0.0;
        for (int q = 0; q < m; ++q) s += part[q * 3 + threadIdx.x];
        F[3 * k + threadIdx.x] = s;
    }
}

// pressure: one CTA per owned atom i, thread q -> list entry q, 9 components (k1 = gamma*3 + delta)
__global__ void __launch_bounds__(PT) pet_pressure_kernel(int na, int a0, int cap, int n_kinds, const int* __restrict__ koa,
                                                         const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                         const double* __restrict__ nn_dist, const double* __restrict__ nn_dist_s,
                                                         const int* __restrict__ rowptr, const xtb_pettifor_pair* __restrict__ TB,
                                                         const double* __restrict__ rho, double* __restrict__ partial) {
    extern __shared__ double part[];      // [m][9]
    const int i = a0 + blockIdx.x;
    const int m = nn_size[i];
    const double* blk = rho + (size_t)rowptr[i] * 16;
    for (int q = threadIdx.x; q < m; q += PT) {
        const size_t e = (size_t)i * cap + q;
        const int j = nn_list[e];
        const double4 d4 = *reinterpret_cast<const double4*>(nn_dist + 4 * e);
        const double xyz[3] = {d4.x, d4.y, d4.z};
        const double r = d4.w;
        const xtb_pettifor_pair& P = TB[koa[i] + n_kinds * koa[j]];
        double V[4], dV[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) { V[c] = pet_Vs(P, c, r); dV[c] = pet_dVs(P, c, r, V[c]); }
        const double* Ab = blk + (size_t)(q + 1) * 16;
        for (int ga = 0; ga < 3; ++ga) {
            const double sr = nn_dist_s[3 * e + ga] / r;
            double drdh[3], dl[3], dm[3], dn[3];
#pragma unroll
            for (int de = 0; de < 3; ++de) {
                drdh[de] = xyz[de] * sr;                                              // drijdhgd  :1339-1365
                const double f = drdh[de] / (r * r);
                dl[de] = (de == 0 ? sr : 0.0) - xyz[0] * f;
                dm[de] = (de == 1 ? sr : 0.0) - xyz[1] * f;
                dn[de] = (de == 2 ? sr : 0.0) - xyz[2] * f;
            }
            double res[3];
            pet_contract(V, dV, xyz[0] / r, xyz[1] / r, xyz[2] / r, drdh, dl, dm, dn, Ab, 1, 4, 3, res);
            part[q * 9 + ga * 3] = res[0]; part[q * 9 + ga * 3 + 1] = res[1]; part[q * 9 + ga * 3 + 2] = res[2];
        }
    }
    __syncthreads();
    if (threadIdx.x < 9) {
        double s = 0.0;
        for (int q = 0; q < m; ++q) s += part[q * 9 + threadIdx.x];
        partial[(size_t)blockIdx.x * 9 + threadIdx.x] = s;
    }
}

__global__ void pet_sum9_kernel(int n, const double* __restrict__ partial, double* __restrict__ out) {
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

int xtb_impl_pettifor_assemble(xtb_ctx* c, bool dense) {
    const int na = c->na;
    int rc;
    if ((rc = xtb_reserve(c, c->d_Hblk, sizeof(double) * 16 * xtb_blocks_alloc(c)))) return rc;
    XTB_T0(c, 2);
    const int nown = c->a1 - c->a0;
    const long long nthreads = (long long)(c->cap + 1) * nown;
    if (nthreads > 0) {
        pet_blocks_kernel<<<(unsigned)((nthreads + 127) / 128), 128, 0, c->stream>>>(
            na, c->a0, c->a1, c->cap, c->n_kinds, (const int*)c->d_koa.p, (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p,
            (const int*)c->d_rev.p, (const double*)c->d_nn_dist.p, (const int*)c->d_rowptr.p,
            (const xtb_pettifor_pair*)c->d_params.p, (double*)c->d_Hblk.p);
        XTB_LAUNCH_CHECK(c);
    }
    XTB_T0(c, 6);      // dense panel kernel alone
    if (dense && (rc = xtb_dense_update(c, c->a0, c->a1))) return rc;
    XTB_T1(c, 6);
    XTB_T1(c, 2);
    if ((rc = xtb_stage_finish(c))) return rc;
    c->have_blocks = true;
    c->have_HS = dense;
    return XTB_OK;
}

int xtb_impl_pettifor_forces(xtb_ctx* c) {
    const int nown = c->a1 - c->a0;
    int rc;
    if ((rc = xtb_reserve(c, c->d_F, sizeof(double) * 3 * (size_t)(c->na + c->nranks)))) return rc;
    XTB_T0(c, 4);
    if (nown > 0) {
        const int mx = c->max_nn > 0 ? c->max_nn : 1;
        const size_t smem = sizeof(double) * 3 * (size_t)mx;
        XTB_CUDA(c, cudaFuncSetAttribute(pet_forces_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        pet_forces_kernel<<<nown, PT, smem, c->stream>>>(c->na, c->a0, c->cap, c->n_kinds, (const int*)c->d_koa.p,
                                                         (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p, (const int*)c->d_rev.p,
                                                         (const double*)c->d_nn_dist.p, (const int*)c->d_rowptr.p,
                                                         (const xtb_pettifor_pair*)c->d_params.p, (const double*)c->d_hdev.p,
                                                         (const double*)c->d_rho.p, (double*)c->d_F.p);
        XTB_LAUNCH_CHECK(c);
    }
    XTB_T1(c, 4);
    if ((rc = xtb_stage_finish(c))) return rc;
    return XTB_OK;
}

int xtb_impl_pettifor_pressure(xtb_ctx* c) {
    const int nown = c->a1 - c->a0;
    int rc;
    if ((rc = xtb_reserve(c, c->d_SC, sizeof(double) * 9))) return rc;
    if ((rc = xtb_reserve(c, c->d_ppart, sizeof(double) * 9 * (size_t)(nown > 0 ? nown : 1)))) return rc;
    XTB_T0(c, 5);
    if (nown > 0) {
        const int mx = c->max_nn > 0 ? c->max_nn : 1;
        const size_t smem = sizeof(double) * 9 * (size_t)mx;
        XTB_CUDA(c, cudaFuncSetAttribute(pet_pressure_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        pet_pressure_kernel<<<nown, PT, smem, c->stream>>>(c->na, c->a0, c->cap, c->n_kinds, (const int*)c->d_koa.p,
                                                           (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p,
                                                           (const double*)c->d_nn_dist.p, (const double*)c->d_nn_dist_s.p,
                                                           (const int*)c->d_rowptr.p, (const xtb_pettifor_pair*)c->d_params.p,
                                                           (const double*)c->d_rho.p, (double*)c->d_ppart.p);
        XTB_LAUNCH_CHECK(c);
        pet_sum9_kernel<<<9, 128, 0, c->stream>>>(nown, (const double*)c->d_ppart.p, (double*)c->d_SC.p);
        XTB_LAUNCH_CHECK(c);
    } else {
        XTB_CUDA(c, cudaMemsetAsync(c->d_SC.p, 0, sizeof(double) * 9, c->stream));
    }
    XTB_T1(c, 5);
    if ((rc = xtb_stage_finish(c))) return rc;
    return XTB_OK;
}
