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
// Thread mapping: one CTA per atom; rho / rho_E blocks of the atom (contiguous in the block store) are
// streamed with 128-bit loads; one thread per (neighbour, matrix, direction) differentiates the
// Slater-Koster block in registers (forward-mode dual numbers) and contracts it; partial sums are
// reduced in shared memory in list order.
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

// ---- forces: one CTA per owned atom k ----
template <int MINB>
__global__ void __launch_bounds__(FT, MINB) nrl_forces_kernel(int na, int a0, int cap, int mx, int n_kinds, const int* __restrict__ koa,
                                                       const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                       const int* __restrict__ rev, const double* __restrict__ nn_dist,
                                                       const int* __restrict__ rowptr, const xtb_nrl_pair* __restrict__ TB,
                                                       const double* __restrict__ hmat, const double* __restrict__ wons,
                                                       const double* __restrict__ rho, const double* __restrict__ rhoE,
                                                       int stage_blocks, double* __restrict__ F) {
    extern __shared__ __align__(16) double smf[];
    const int k = a0 + blockIdx.x;
    const int m = nn_size[k];
    // mx = largest neighbour count (sizes the shared arrays); cap = ELL stride of the lists
    double* part = smf;                             // [mx][8]: 6 pair partials (mat*3+alpha) + (unused 2)
    double* ons = part + (size_t)mx * 8;            // [mx][3] on-site partials
    double* sR = ons + (size_t)mx * 3 + ((mx & 1) ? 1 : 0);   // [(mx+1)*81] rho blocks of atom k (16 B aligned)
    double* sE = sR + (size_t)(mx + 1) * 81 + ((mx + 1) & 1);
    const double* blkR = stage_blocks ? sR : rho + (size_t)rowptr[k] * 81;
    const double* blkE = stage_blocks ? sE : rhoE + (size_t)rowptr[k] * 81;
    __shared__ double h[9];
    if (threadIdx.x < 9) h[threadIdx.x] = hmat[threadIdx.x];
    if (stage_blocks) {   // stream the blocks of atom k: contiguous (m+1)*81 doubles per matrix, 128-bit loads when aligned
        const size_t b0 = (size_t)rowptr[k] * 81;
        const int nd = (m + 1) * 81;
        const double* gR = rho + b0; const double* gE = rhoE + b0;
        if ((b0 & 1) == 0) {
            const int n2 = nd >> 1;
            const double2* gR2 = reinterpret_cast<const double2*>(gR);
            const double2* gE2 = reinterpret_cast<const double2*>(gE);
            double2* sR2 = reinterpret_cast<double2*>(sR);
            double2* sE2 = reinterpret_cast<double2*>(sE);
            for (int e = threadIdx.x; e < n2; e += FT) { sR2[e] = __ldg(gR2 + e); sE2[e] = __ldg(gE2 + e); }
            if ((nd & 1) && threadIdx.x == 0) { sR[nd - 1] = gR[nd - 1]; sE[nd - 1] = gE[nd - 1]; }
        } else {
            for (int e = threadIdx.x; e < nd; e += FT) { sR[e] = gR[e]; sE[e] = gE[e]; }
        }
    }
    __syncthreads();
    // ---- pair terms: thread -> (q, mat, alpha) ----
    for (int w = threadIdx.x; w < m * 6; w += FT) {
        const int q = w / 6, rem = w - q * 6;
        const int mat = rem / 3, al = rem - mat * 3;
        const size_t e = (size_t)k * cap + q;
        const int b = nn_list[e];
        const int rv = rev[e];
        double res = 0.0;
        if (rv >= 0) {
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
            const double* A = (mat == 0 ? blkR : blkE) + (size_t)(q + 1) * 81;
            res = 2.0 * contract_upper(up, A, !ksmall);
            if (mat == 0) res *= XTB_G_RY;
        }
        part[q * 8 + rem] = res;
    }
    // ---- on-site terms: thread -> q ----
    for (int q = threadIdx.x; q < m; q += FT) {
        const size_t e = (size_t)k * cap + q;
        const int b = nn_list[e];
        const int rv = rev[e];
        double o3[3] = {0.0, 0.0, 0.0};
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

// ---- pressure: one CTA per owned atom i, all directed pairs (i, q) + on-site; 9 partials per CTA ----
__global__ void __launch_bounds__(FT) nrl_pressure_kernel(int na, int a0, int cap, int mx, int n_kinds, const int* __restrict__ koa,
                                                         const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                         const int* __restrict__ rev, const double* __restrict__ nn_dist,
                                                         const double* __restrict__ nn_dist_s, const int* __restrict__ rowptr,
                                                         const xtb_nrl_pair* __restrict__ TB, const double* __restrict__ wons,
                                                         const double* __restrict__ rho, const double* __restrict__ rhoE,
                                                         double* __restrict__ partial /*[n_owned][9]*/) {
    extern __shared__ __align__(16) double smf[];
    const int i = a0 + blockIdx.x;
    const int m = nn_size[i];
    double* part = smf;                        // [mx][18]: (mat, k1)
    double* ons = part + (size_t)mx * 18;      // [mx][9]
    const size_t b0 = (size_t)rowptr[i] * 81;
    // thread -> (q, mat, gamma): three delta components each
    for (int w = threadIdx.x; w < m * 6; w += FT) {
        const int q = w / 6, rem = w - q * 6;
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
            part[q * 18 + mat * 9 + ga * 3 + de] = res;
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
    if (threadIdx.x < 9) {
        const int k1 = threadIdx.x;
        double s = 0.0;
        for (int q = 0; q < m; ++q) s += (part[q * 18 + k1] - part[q * 18 + 9 + k1]) + ons[q * 9 + k1];
        partial[(size_t)blockIdx.x * 9 + k1] = s;
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
    XTB_T0(c, 4);
    if ((rc = launch_onsite_w(c))) return rc;
    if (nown > 0) {
        const int mx = c->max_nn > 0 ? c->max_nn : 1;
        const size_t small = sizeof(double) * ((size_t)mx * 11 + 2);
        const size_t big = small + sizeof(double) * (((size_t)(mx + 1) * 81 + 1) * 2);
        const char* ev = getenv("XTB_FORCES_MODE");              // experiments: 0 = staged blocks, 1 CTA/SM; 1 = direct reads, 2 CTAs/SM
        const int mode = ev ? atoi(ev) : 1;
        const int stage_blocks = (mode == 0 && big <= 200 * 1024) ? 1 : 0;      // else read the blocks straight from L1 / L2
        const size_t smem = stage_blocks ? big : small;
        auto kern = mode == 0 ? nrl_forces_kernel<1> : nrl_forces_kernel<2>;
        XTB_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<nown, FT, smem, c->stream>>>(c->na, c->a0, c->cap, mx, c->n_kinds, (const int*)c->d_koa.p,
                                                         (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p, (const int*)c->d_rev.p,
                                                         (const double*)c->d_nn_dist.p, (const int*)c->d_rowptr.p,
                                                         (const xtb_nrl_pair*)c->d_params.p, (const double*)c->d_hdev.p,
                                                         (const double*)c->d_onsite_w.p, (const double*)c->d_rho.p,
                                                         (const double*)c->d_rhoE.p, stage_blocks, (double*)c->d_F.p);
        XTB_LAUNCH_CHECK(c);
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
    XTB_T0(c, 5);
    if ((rc = launch_onsite_w(c))) return rc;
    if (nown > 0) {
        const int mx = c->max_nn > 0 ? c->max_nn : 1;
        const size_t smem = sizeof(double) * ((size_t)mx * 27);
        XTB_CUDA(c, cudaFuncSetAttribute(nrl_pressure_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        nrl_pressure_kernel<<<nown, FT, smem, c->stream>>>(c->na, c->a0, c->cap, mx, c->n_kinds, (const int*)c->d_koa.p,
                                                           (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p, (const int*)c->d_rev.p,
                                                           (const double*)c->d_nn_dist.p, (const double*)c->d_nn_dist_s.p,
                                                           (const int*)c->d_rowptr.p, (const xtb_nrl_pair*)c->d_params.p,
                                                           (const double*)c->d_onsite_w.p, (const double*)c->d_rho.p,
                                                           (const double*)c->d_rhoE.p, (double*)c->d_ppart.p);
        XTB_LAUNCH_CHECK(c);
        sum9_kernel<<<9, 128, 0, c->stream>>>(nown, (const double*)c->d_ppart.p, (double*)c->d_SC.p);
        XTB_LAUNCH_CHECK(c);
    } else {
        XTB_CUDA(c, cudaMemsetAsync(c->d_SC.p, 0, sizeof(double) * 9, c->stream));
    }
    XTB_T1(c, 5);
    if ((rc = xtb_stage_finish(c))) return rc;
    return XTB_OK;
}
