// van der Waals and Coulomb pair terms (third part; classes TB_vdW and TB_Coulomb_cut): the two terms of
// the reference that sum over ALL atom pairs in the periodic images of the supercell instead of over the neighbour list.
//
// Replaces, in Source_files/Van_der_Waals.f90 and TB.f90 (dispatch vdW_forces TB.f90:3599-3633, vdW_s TB.f90:4017-4035):
//   get_mirror_cell_num :132-157        R_cut = max(TB cut-off, max d0_cut + 10 max dd_cut) (get_near_neighbours with include_vdW,
//                                       Atomic_tools.f90:636-641); get_number_of_image_cells (Atomic_tools.f90:654-702); >= 1 per axis
//   get_vdW_s :158-250                  E = 1/2 sum_cells sum_i sum_j V(r_ij,cell), (j /= i) or (cell /= 0); Epot(i) += 1/2 sum_j V
//   vdW_energy / dvdW :578-627          dispatch on the class: Girifalco :933-1057, LJ_cut :650-675, ILJ_cut :718-745
//   Construct_B / cell_cycle_B :251-421 Bij = dV/dr with TB_Waals(KOA(j1), KOA(i1)), geometry from distance_to_given_cell
//                                       (Atomic_tools.f90:4135-4166)
//   d_Forces_s TB.f90:3638-3781         forces%rep(ian) += 1/2 sum_{i1} sum_{j1,cell} Bij (d_{ian,i1} - d_{ian,j1}) / r * (h . r)
//   d_Forces_Pressure TB.f90:3785-3848  SCforce%rep(l,k) += 1/2 sum_cells sum_i sum_j Bij r_k s_l / r
// The Coulomb term (Coulomb.f90; dispatch Coulomb_forces TB.f90:3560-3596, Coulomb_s TB.f90:3995-4012) goes through the same
// d_Forces_s / d_Forces_Pressure with its own arrays:
//   get_mirror_cell_num_C :435-452      R_cut = TB_Coul(1,1)%dm alone
//   get_Coulomb_s :454-550              E = sum_cells sum_i sum_j V -- NO factor 1/2 (as written); Epot(i) += sum_j V
//   Coulomb_pot :552-587 / dCoulomb :641-672   V = k Q Q / r * f_cut_L_C, zero for r > dm + 10 dd; Q = Scell%Q, the mean unbalanced charge
//   Construct_B_C :715-880              all zero when Q <= 1e-12; XijSupce = x supce(1,1) + y supce(2,1) + z supce(3,1): the
//                                       TRANSPOSE of the contraction Construct_B uses (identical for a symmetric supce) -- kept
// The reference builds eleven (cells, Na, Na) arrays and then loops ian x i1 x j1 x cells (O(Na^3)); only i1 = ian and j1 = ian
// contribute, and the j1 = ian term of cell zb is the mirror image of the i1 = ian term of cell -zb (the displacement changes
// sign exactly in floating point), so one pass over (j, cell) per atom gives energy, force and supercell force together.  Here:
// one CTA per owned atom, threads stride the other atoms, every thread walks the image cells; fixed-order reductions, nothing
// stored per pair.  Every class is exactly zero (value and slope) beyond R_max = max d_cut (Girifalco) / max (d0_cut + 10 dd_cut)
// (Fermi_function returns 0.0 there), so image cells whose slab |s_a + c_a| * (spacing of the cell planes along a) > R_max are
// not visited: they would add +-0.0.  At config 4 (17.8 A cell, R_max 6 A) that leaves ~1.4 of the 27 images per pair.
// FP64 CUDA cores, compute-bound (no memory traffic to speak of: positions and kinds of Na atoms, L1-resident).
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>
#include "xtb_ctx.h"

namespace {

constexpr int VT = 128;             // threads per CTA (one CTA per atom)
constexpr unsigned FULL = 0xffffffffu;
enum : unsigned { DO_E = 1u, DO_F = 2u, DO_P = 4u };

__device__ __forceinline__ double fermi_f(double rcut, double d, double r) {        // Fermi_function, Little_subroutines.f90:349
    if (r >= rcut + 10.0 * d) return 0.0;
    return 1.0 / (1.0 + exp((r - rcut) / d));
}
__device__ __forceinline__ double d_fermi_f(double rcut, double d, double r) {      // d_Fermi_function :360 (cut-off on)
    if (r >= rcut + 10.0 * d) return 0.0;
    const double e = exp((r - rcut) / d), e1 = 1.0 + e;
    return -e / (d * e1 * e1);
}

template <int KIND> struct Vdw;
template <> struct Vdw<XTB_VDW_GIRIFALCO> {
    static __device__ __forceinline__ double V(const xtb_vdw_pair& W, double r) {   // vdW_Girifalco :933-991
        if (r > W.d_cut || r < W.ds_cut) return 0.0;
        if (r > W.dm) { const double r2 = r * r; return W.a * r2 * r + W.b * r2 + W.c * r + W.d; }
        if (r < W.dsm) {
            const double r2 = r * r, r4 = r2 * r2;
            return W.as * r4 * r + W.bs * r4 + W.cs * r2 * r + W.ds * r2 + W.es * r + W.fs;
        }
        const double r2 = r * r, r3 = r2 * r, r6 = r3 * r3, r12 = r6 * r6;          // LJ_potential :1116, fast_pow(r, 6)
        return W.C12 / r12 - W.C6 / r6;
    }
    static __device__ __forceinline__ double dV(const xtb_vdw_pair& W, double r) {  // d_vdW_Girifalco :995-1057
        if (r > W.d_cut || r < W.ds_cut) return 0.0;
        if (r > W.dm) return 3.0 * W.a * r * r + 2.0 * W.b * r + W.c;
        if (r < W.dsm) {
            const double r2 = r * r, r4 = r2 * r2;
            return 5.0 * W.as * r4 + 4.0 * W.bs * r2 * r + 3.0 * W.cs * r2 + 2.0 * W.ds * r + W.es;
        }
        const double r2 = r * r, r3 = r2 * r, r6 = r3 * r3, r7 = r6 * r, r13 = r6 * r7;   // d_LJ_potential :1126
        return -12.0 * W.C12 / r13 + 6.0 * W.C6 / r7;
    }
};
template <> struct Vdw<XTB_VDW_LJ_CUT> {
    static __device__ __forceinline__ double V(const xtb_vdw_pair& W, double r) {   // double_truncated_LJ :650-660
        double v = 0.0;                                                             // General_LJ_potential :839
        if (r > 0.0) { const double q = pow(W.r0 / r, W.n); v = W.eps * (q - 2.0) * q; }
        return v * (1.0 - fermi_f(W.d0_short, W.dd_short, r)) * fermi_f(W.d0_cut, W.dd_cut, r);
    }
    static __device__ __forceinline__ double dV(const xtb_vdw_pair& W, double r) {  // d_double_truncated_LJ :663-675
        double v = 0.0, dv = 0.0;
        if (r > 0.0) {
            const double q = pow(W.r0 / r, W.n);
            v = W.eps * (q - 2.0) * q;
            dv = -W.eps * 2.0 * W.n / r * q * (q - 1.0);                            // d_General_LJ_potential :853
        }
        const double fl = fermi_f(W.d0_cut, W.dd_cut, r), dfl = d_fermi_f(W.d0_cut, W.dd_cut, r);
        const double fs = 1.0 - fermi_f(W.d0_short, W.dd_short, r), dfs = -d_fermi_f(W.d0_short, W.dd_short, r);
        return v * (dfs * fl + fs * dfl) + dv * (fs * fl);
    }
};
template <> struct Vdw<XTB_VDW_ILJ_CUT> {
    static __device__ __forceinline__ double V(const xtb_vdw_pair& W, double r) {   // ILJ :718-727, Improved_LJ_potential :882
        double v = 0.0;
        if (r > 0.0) { const double q = W.r0 / r, nm = W.n - W.m; v = W.eps * (W.m / nm * pow(q, W.n) - W.n / nm * pow(q, W.m)); }
        return v * (1.0 - fermi_f(W.d0_short, W.dd_short, r)) * fermi_f(W.d0_cut, W.dd_cut, r);
    }
    static __device__ __forceinline__ double dV(const xtb_vdw_pair& W, double r) {  // d_ILJ :730-745
        double v = 0.0, dv = 0.0;
        if (r > 0.0) {
            const double q = W.r0 / r, nm = W.n - W.m, qn = pow(q, W.n), qm = pow(q, W.m);
            v = W.eps * (W.m / nm * qn - W.n / nm * qm);
            dv = -W.eps * (qn - qm) * W.n * W.m / (nm * r);                         // d_Improved_LJ_potential :896
        }
        const double fl = fermi_f(W.d0_cut, W.dd_cut, r), dfl = d_fermi_f(W.d0_cut, W.dd_cut, r);
        const double fs = 1.0 - fermi_f(W.d0_short, W.dd_short, r), dfs = -d_fermi_f(W.d0_short, W.dd_short, r);
        return v * (dfs * fl + fs * dfl) + dv * (fs * fl);
    }
};

constexpr int KIND_COULOMB = 4;      // internal: the table entry carries eps = k Q Q, d0_cut = dm, dd_cut = dd
template <> struct Vdw<KIND_COULOMB> {
    static __device__ __forceinline__ double V(const xtb_vdw_pair& W, double r) {   // Coulomb_pot :552-587
        if (r > W.d0_cut + W.dd_cut * 10.0) return 0.0;
        const double fcut = 1.0 / (1.0 + exp((r - W.d0_cut) / W.dd_cut));           // f_cut_L_C :610
        return W.eps / r * fcut;                                                    // Coulomb_potential :589
    }
    static __device__ __forceinline__ double dV(const xtb_vdw_pair& W, double r) {  // dCoulomb :641-672
        if (r > W.d0_cut + W.dd_cut * 10.0) return 0.0;
        const double arg = (r - W.d0_cut) / W.dd_cut;
        const double fcut = 1.0 / (1.0 + exp(arg));
        const double E_C = W.eps / r, d_E_C = -W.eps / (r * r);                     // d_Coulomb_potential :596
        double dfcut = 0.0;                                                         // d_f_cut_L_C :617-629
        if (arg < 709.782712893384) {
            const double er = exp(arg), e2 = 1.0 + er;
            dfcut = -er / (W.dd_cut * e2 * e2);
        }
        return d_E_C * fcut + E_C * dfcut;
    }
};

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(FULL, v, off);
    return v;
}

// atom i = a0 + blockIdx.x.  sym: TB_Waals(k1,k2) == TB_Waals(k2,k1) bit for bit (the mirrored force term then equals the direct one).
template <int KIND>
// reach[a] = R_max / (plane spacing along cell axis a), a little widened: the range of |s_a| inside which a pair can interact
__global__ void __launch_bounds__(VT) vdw_atom_kernel(int a0, int na, int n_kinds, int Nx, int Ny, int Nz, unsigned what, int sym,
                                                      double reach_x, double reach_y, double reach_z, double e_scale, int xs_t,
                                                      const int* __restrict__ koa, const double* __restrict__ S,
                                                      const xtb_vdw_pair* __restrict__ WP, const double* __restrict__ hdev,
                                                      double* __restrict__ Epot, double* __restrict__ F, double* __restrict__ ppart) {
    __shared__ double red[VT / 32][13];
    const int i = a0 + blockIdx.x, tid = threadIdx.x;
    double h[9];
#pragma unroll
    for (int e = 0; e < 9; ++e) h[e] = hdev[e];             // supce(a,b) = h[a + 3b]
    const double sxi = S[3 * i], syi = S[3 * i + 1], szi = S[3 * i + 2];
    const int ki = koa[i];
    double acc[13];                                          // 0 energy, 1-3 force, 4-12 supercell force (l + 3k)
#pragma unroll
    for (int e = 0; e < 13; ++e) acc[e] = 0.0;
    for (int j = tid; j < na; j += VT) {
        const int kj = koa[j];
        const xtb_vdw_pair& Wij = WP[ki + n_kinds * kj];    // TB_Waals(KOAi, KOAj): energy (:597-604); Bij of (i1 = j, j1 = i)
        const xtb_vdw_pair& Wji = WP[kj + n_kinds * ki];    // TB_Waals(KOA(j1), KOA(i1)): Bij of (i1 = i, j1 = j) (:353)
        const double dsx = sxi - S[3 * j], dsy = syi - S[3 * j + 1], dsz = szi - S[3 * j + 2];
        // image range per axis: |ds_a + c_a| <= reach_a, inside [-N_a, N_a] (compared as doubles: reach may be huge)
        const int x0 = (int)fmax(-(double)Nx, ceil(-reach_x - dsx)), x1 = (int)fmin((double)Nx, floor(reach_x - dsx));
        const int y0 = (int)fmax(-(double)Ny, ceil(-reach_y - dsy)), y1 = (int)fmin((double)Ny, floor(reach_y - dsy));
        const int z0 = (int)fmax(-(double)Nz, ceil(-reach_z - dsz)), z1 = (int)fmin((double)Nz, floor(reach_z - dsz));
        for (int cx = x0; cx <= x1; ++cx)
            for (int cy = y0; cy <= y1; ++cy)
                for (int cz = z0; cz <= z1; ++cz) {
                    if (j == i && cx == 0 && cy == 0 && cz == 0) continue;          // no self-interaction in the origin cell
                    const double s[3] = {dsx + (double)cx, dsy + (double)cy, dsz + (double)cz};     // distance_to_given_cell :4163-4165
                    double x[3];
#pragma unroll
                    for (int al = 0; al < 3; ++al) x[al] = s[0] * h[3 * al] + s[1] * h[1 + 3 * al] + s[2] * h[2 + 3 * al];   // :4153-4155
                    const double r = sqrt(x[0] * x[0] + x[1] * x[1] + x[2] * x[2]);
                    if (what & DO_E) acc[0] += Vdw<KIND>::V(Wij, r);
                    if (what & (DO_F | DO_P)) {
                        const double b = Vdw<KIND>::dV(Wji, r);
                        if ((what & DO_F) && j != i) {
                            const double bb = (sym ? b + b : b + Vdw<KIND>::dV(Wij, r)) / r;
#pragma unroll
                            for (int al = 0; al < 3; ++al)      // XijSupce :343-345; Coulomb.f90:864-866 contracts with the transpose
                                acc[1 + al] += bb * (xs_t ? x[0] * h[3 * al] + x[1] * h[3 * al + 1] + x[2] * h[3 * al + 2]
                                                          : x[0] * h[al] + x[1] * h[al + 3] + x[2] * h[al + 6]);
                        }
                        if (what & DO_P) {
                            const double br = b / r;
#pragma unroll
                            for (int k = 0; k < 3; ++k)
#pragma unroll
                                for (int l = 0; l < 3; ++l) acc[4 + l + 3 * k] += br * x[k] * s[l];                                   // TB.f90:3829
                        }
                    }
                }
    }
#pragma unroll
    for (int e = 0; e < 13; ++e) {
        const double v = warp_sum(acc[e]);
        if ((tid & 31) == 0) red[tid >> 5][e] = v;
    }
    __syncthreads();
    if (tid < 13) {
        double v = 0.0;
#pragma unroll
        for (int w = 0; w < VT / 32; ++w) v += red[w][tid];
        v *= tid == 0 ? e_scale : 0.5;                       // :246 (vdW 1/2; Coulomb.f90:549 none), TB.f90:3765, TB.f90:3838
        if (tid == 0) { if (what & DO_E) Epot[i] = v; }
        else if (tid < 4) { if (what & DO_F) F[3 * (size_t)i + tid - 1] = v; }
        else if (what & DO_P) ppart[(size_t)blockIdx.x * 9 + tid - 4] = v;
    }
}

__global__ void __launch_bounds__(256) vdw_esum_kernel(int na, const double* __restrict__ Epot, double* __restrict__ out) {
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

__global__ void vdw_sum9_kernel(int n, const double* __restrict__ partial, double* __restrict__ out) {
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

// one image-cell pair term of the context: its table, class and output buffers
struct PairSet {
    int kind;  bool sym;  double Rmax, e_scale;  int xs_t;
    DevBuf *params, *E_at, *E, *F, *SC;
    int stage;
};

// ppart: where the per-atom supercell-force partial sums go (default: the context's grow-only scratch buffer)
int launch_atoms(xtb_ctx* c, const PairSet& t, int first, int count, unsigned what, const int N[3], double* ppart = nullptr) {
    if (count <= 0) return XTB_OK;
    // reach of the pair table in units of the plane spacings of the cell (volume / face area)
    double reach[3];
    {
        const double* h = c->h_supce;               // row a of supce = cell vector a: (h[a], h[a+3], h[a+6])
        double v[3][3], vol;
        for (int a = 0; a < 3; ++a) for (int k = 0; k < 3; ++k) v[a][k] = h[a + 3 * k];
        auto cross = [](const double* p, const double* q, double* o) { o[0] = p[1] * q[2] - p[2] * q[1]; o[1] = p[2] * q[0] - p[0] * q[2]; o[2] = p[0] * q[1] - p[1] * q[0]; };
        double bc[3];
        cross(v[1], v[2], bc);
        vol = fabs(v[0][0] * bc[0] + v[0][1] * bc[1] + v[0][2] * bc[2]);
        for (int a = 0; a < 3; ++a) {
            double n[3];
            cross(v[(a + 1) % 3], v[(a + 2) % 3], n);
            const double area = sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
            const double spacing = area > 0.0 ? vol / area : 0.0;
            reach[a] = (spacing > 0.0 && !getenv("XTB_VDW_NO_PRUNE")) ? t.Rmax / spacing * (1.0 + 1e-9) + 1e-9 : 1e300;
        }
    }
#define XTB_VDW_LAUNCH(KIND)                                                                                                          \
    vdw_atom_kernel<KIND><<<count, VT, 0, c->stream>>>(first, c->na, c->n_kinds, N[0], N[1], N[2], what, t.sym ? 1 : 0,                \
                                                       reach[0], reach[1], reach[2], t.e_scale, t.xs_t,                               \
                                                       (const int*)c->d_koa.p, (const double*)c->d_S.p, (const xtb_vdw_pair*)t.params->p, \
                                                       (const double*)c->d_hdev.p, (double*)t.E_at->p, (double*)t.F->p,                \
                                                       ppart ? ppart : (double*)c->d_ppart.p)
    switch (t.kind) {
        case XTB_VDW_GIRIFALCO: XTB_VDW_LAUNCH(XTB_VDW_GIRIFALCO); break;
        case XTB_VDW_LJ_CUT:    XTB_VDW_LAUNCH(XTB_VDW_LJ_CUT); break;
        case XTB_VDW_ILJ_CUT:   XTB_VDW_LAUNCH(XTB_VDW_ILJ_CUT); break;
        case KIND_COULOMB:      XTB_VDW_LAUNCH(KIND_COULOMB); break;
        default: XTB_FAIL(c, XTB_ESTATE, "pair term: unknown class %d", t.kind);
    }
#undef XTB_VDW_LAUNCH
    XTB_LAUNCH_CHECK(c);
    return XTB_OK;
}

// get_number_of_image_cells (Atomic_tools.f90:654-702) for cut-off Rc, then "at least one image per axis"
int image_cells(xtb_ctx* c, double Rc, int N[3], const char* who) {
    const double* h = c->h_supce;
    long long cells = 1;
    for (int a = 0; a < 3; ++a) {
        auto dist = [&](double zb) { const double x0 = zb * h[a], y0 = zb * h[a + 3], z0 = zb * h[a + 6]; return sqrt(x0 * x0 + y0 * y0 + z0 * z0); };
        if (!(dist(1.0) > 0.0)) XTB_FAIL(c, XTB_EINVAL, "%s: supercell vector %d has no length", who, a + 1);
        int n = 0;
        double R = dist(1.0);
        while (R < 2.0 * Rc) {
            n++;
            R = dist((double)n);
            if (n > 64) XTB_FAIL(c, XTB_EINVAL, "%s: more than 64 images of the supercell along axis %d (R_cut = %g A)", who, a + 1, Rc);
        }
        if (n < 1) n = 1;
        N[a] = n;
        cells *= 2 * n + 1;
    }
    if (cells > 100000) XTB_FAIL(c, XTB_EINVAL, "%s: %lld image cells (R_cut = %g A)", who, cells, Rc);
    return XTB_OK;
}

// Energy of every atom on every rank (replicated, like the wall); forces and the supercell-force partial sums of the owned
// atoms (exchanged / all-reduced by the caller when the context communicates).
int run_set(xtb_ctx* c, const PairSet& t, unsigned what, const int N[3], bool forces_off) {
    const int na = c->na, nown = c->a1 - c->a0;
    int rc;
    if ((rc = xtb_reserve(c, *t.E_at, sizeof(double) * na))) return rc;
    if ((rc = xtb_reserve(c, *t.E, sizeof(double)))) return rc;
    if ((rc = xtb_reserve(c, *t.F, sizeof(double) * 3 * (size_t)(na + c->nranks)))) return rc;
    if ((rc = xtb_reserve(c, *t.SC, sizeof(double) * 9))) return rc;
    if ((rc = xtb_reserve(c, c->d_ppart, sizeof(double) * 9 * (size_t)(nown > 0 ? nown : 1)))) return rc;
    XTB_T0(c, t.stage);
    if (forces_off) {                       // Construct_B_C with Q <= 1e-12: every array zero
        if (what & DO_F) XTB_CUDA(c, cudaMemsetAsync(t.F->p, 0, sizeof(double) * 3 * (size_t)na, c->stream));
        if (what & DO_P) XTB_CUDA(c, cudaMemsetAsync(t.SC->p, 0, sizeof(double) * 9, c->stream));
        what &= DO_E;
    }
    if (what && (rc = launch_atoms(c, t, c->a0, nown, what, N))) return rc;
    if (what & DO_E) {
        if ((rc = launch_atoms(c, t, 0, c->a0, DO_E, N)) || (rc = launch_atoms(c, t, c->a1, na - c->a1, DO_E, N))) return rc;
        vdw_esum_kernel<<<1, 256, 0, c->stream>>>(na, (const double*)t.E_at->p, (double*)t.E->p);
        XTB_LAUNCH_CHECK(c);
    }
    if (what & DO_P) {
        if (nown > 0) {
            vdw_sum9_kernel<<<9, 128, 0, c->stream>>>(nown, (const double*)c->d_ppart.p, (double*)t.SC->p);
            XTB_LAUNCH_CHECK(c);
        } else XTB_CUDA(c, cudaMemsetAsync(t.SC->p, 0, sizeof(double) * 9, c->stream));
    }
    XTB_T1(c, t.stage);
    return xtb_stage_finish(c);
}

}  // namespace

// get_mirror_cell_num (Van_der_Waals.f90:132-157) on the host: supce and the parameters live there
int xtb_impl_vdw_cells(xtb_ctx* c, int N[3], double* R_cut) {
    double Rc = xtb_default_cutoff(c), r = 0.0, d = 0.0;
    bool first = true;
    for (auto& p : c->h_vdw) {                                // maxval over the pair table, Atomic_tools.f90:637-638
        if (first || p.d0_cut > r) r = p.d0_cut;
        if (first || p.dd_cut > d) d = p.dd_cut;
        first = false;
    }
    const double X = r + 10.0 * d;
    if (Rc < X) Rc = X;
    if (R_cut) *R_cut = Rc;
    return image_cells(c, Rc, N, "vdW");
}

// what: 1 energy | 2 forces | 4 supercell force, one pass
int xtb_impl_vdw(xtb_ctx* c, unsigned what) {
    int rc, N[3];
    if ((rc = xtb_impl_vdw_cells(c, N, nullptr))) return rc;
    c->vdw_cells[0] = N[0]; c->vdw_cells[1] = N[1]; c->vdw_cells[2] = N[2];
    PairSet t{c->vdw_kind, c->vdw_sym, 0.0, 0.5, 0, &c->d_vdw_params, &c->d_Evdw_at, &c->d_Evdw, &c->d_Fvdw, &c->d_SCvdw, 17};
    for (auto& p : c->h_vdw) {              // beyond this every class returns exactly zero, value and slope
        const double r = c->vdw_kind == XTB_VDW_GIRIFALCO ? p.d_cut : p.d0_cut + 10.0 * p.dd_cut;
        if (r > t.Rmax) t.Rmax = r;
    }
    return run_set(c, t, what, N, false);
}

// get_mirror_cell_num_C (Coulomb.f90:435-452): R_cut = TB_Coul(1,1)%dm
int xtb_impl_coulomb_cells(xtb_ctx* c, int N[3]) {
    if (c->h_coul.empty()) XTB_FAIL(c, XTB_ESTATE, "Coulomb: no parameters");
    return image_cells(c, c->h_coul[0].dm, N, "Coulomb");
}

int xtb_impl_coulomb(xtb_ctx* c, unsigned what) {
    int rc, N[3];
    if ((rc = xtb_impl_coulomb_cells(c, N))) return rc;
    c->coul_cells[0] = N[0]; c->coul_cells[1] = N[1]; c->coul_cells[2] = N[2];
    const size_t cnt = c->h_coul.size();
    if (c->coul_dirty) {                    // the device table carries k Q Q (Coulomb_potential: k*Q1*Q2, left to right)
        std::vector<xtb_vdw_pair> tab(cnt);
        memset(tab.data(), 0, sizeof(xtb_vdw_pair) * cnt);
        for (size_t q = 0; q < cnt; ++q) { tab[q].eps = c->h_coul[q].k * c->coul_Q * c->coul_Q; tab[q].d0_cut = c->h_coul[q].dm; tab[q].dd_cut = c->h_coul[q].dd; }
        if ((rc = xtb_reserve(c, c->d_coul_params, sizeof(xtb_vdw_pair) * cnt))) return rc;
        XTB_CUDA(c, cudaMemcpyAsync(c->d_coul_params.p, tab.data(), sizeof(xtb_vdw_pair) * cnt, cudaMemcpyHostToDevice, c->stream));
        XTB_CUDA(c, cudaStreamSynchronize(c->stream));        // `tab` is a host temporary
        c->coul_dirty = false;
    }
    PairSet t{KIND_COULOMB, c->coul_sym, 0.0, 1.0, 1, &c->d_coul_params, &c->d_Ecoul_at, &c->d_Ecoul, &c->d_Fcoul, &c->d_SCcoul, 18};
    for (auto& p : c->h_coul) { const double r = p.dm + p.dd * 10.0; if (r > t.Rmax) t.Rmax = r; }
    return run_set(c, t, what, N, c->coul_Q <= 1.0e-12);      // CHARGE: Coulomb.f90:762
}

// ---- allocation-free variant of the supercell-force pass (no counterpart in the reference; added
//      for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that forbid it) ----
// The per-atom partial sums (9 doubles per owned atom) and the 9 results live in ONE file-scope device array of fixed size
// instead of the context's grow-only buffers: every GPU that loads the library holds 9 * (XTB_NOALLOC_MAX_ATOMS + 1) * 8 bytes
// (= 590 KB) until the process exits.  In host code `g_noalloc_scratch` is only a stand-in for the symbol, NOT a device address:
// the address on a device comes from cudaGetSymbolAddress with that device current (cached per device below).  Every call on a
// device shares the array, which keeps the contents of the last call: each element read by a call is written by that same call
// first (nothing accumulates, so nothing has to start at zero; the empty-partition case zero-fills its 9 results), and calls that
// overlap on one GPU (two streams, two contexts, two threads) race on it.
static __device__ __align__(256) double g_noalloc_scratch[9 * (XTB_NOALLOC_MAX_ATOMS + 1)];

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
double* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr};      // device address of g_noalloc_scratch, per device ordinal
std::mutex g_noalloc_mutex;

int noalloc_scratch(xtb_ctx* c, double** out) {
    if (c->device < 0 || c->device >= NOALLOC_MAX_DEVICES) XTB_FAIL(c, XTB_EINVAL, "xtb_vdw_pressure_noalloc: device ordinal %d", c->device);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[c->device]) {        // once per device, with that device current (the caller has made it so)
        double* p = nullptr;
        XTB_CUDA(c, cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
        if (!p) XTB_FAIL(c, XTB_ECUDA, "xtb_vdw_pressure_noalloc: cudaGetSymbolAddress returned a null address");
        g_noalloc_ptr[c->device] = p;
    }
    *out = g_noalloc_ptr[c->device];
    return XTB_OK;
}
}  // namespace

int xtb_vdw_pressure_noalloc(xtb_ctx* c, double* SCforce_vdw) {
    if (!c) return XTB_EINVAL;
    if (!SCforce_vdw) XTB_FAIL(c, XTB_EINVAL, "xtb_vdw_pressure_noalloc: null output");
    if (c->h_vdw.empty()) XTB_FAIL(c, XTB_ESTATE, "xtb_vdw_pressure_noalloc: no vdW parameters (xtb_set_params_vdw)");
    if (c->na <= 0) XTB_FAIL(c, XTB_ESTATE, "xtb_vdw_pressure_noalloc: no system set");
    const int nown = c->a1 - c->a0;
    if (nown > XTB_NOALLOC_MAX_ATOMS)
        XTB_FAIL(c, XTB_ECAPACITY, "xtb_vdw_pressure_noalloc: %d owned atoms, the fixed scratch array holds XTB_NOALLOC_MAX_ATOMS = %d (use xtb_vdw_pressure)",
                 nown, XTB_NOALLOC_MAX_ATOMS);
    XTB_CUDA(c, cudaSetDevice(c->device));
    int rc, N[3];
    double* scratch = nullptr;
    if ((rc = noalloc_scratch(c, &scratch))) return rc;
    if ((rc = xtb_impl_vdw_cells(c, N, nullptr))) return rc;          // host arithmetic only
    PairSet t{c->vdw_kind, c->vdw_sym, 0.0, 0.5, 0, &c->d_vdw_params, &c->d_Evdw_at, &c->d_Evdw, &c->d_Fvdw, &c->d_SCvdw, 17};
    for (auto& p : c->h_vdw) {
        const double r = c->vdw_kind == XTB_VDW_GIRIFALCO ? p.d_cut : p.d0_cut + 10.0 * p.dd_cut;
        if (r > t.Rmax) t.Rmax = r;
    }
    double* result = scratch + 9 * (size_t)XTB_NOALLOC_MAX_ATOMS;     // the 9 sums, behind the partial sums
    if (nown > 0) {
        if ((rc = launch_atoms(c, t, c->a0, nown, DO_P, N, scratch))) return rc;      // what = DO_P: writes ppart only
        vdw_sum9_kernel<<<9, 128, 0, c->stream>>>(nown, scratch, result);
        XTB_LAUNCH_CHECK(c);
    } else XTB_CUDA(c, cudaMemsetAsync(result, 0, sizeof(double) * 9, c->stream));
    // the output may be a host or a device pointer (unified addressing decides); no staging buffer is allocated
    XTB_CUDA(c, cudaMemcpyAsync(SCforce_vdw, result, sizeof(double) * 9, cudaMemcpyDefault, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    return XTB_OK;
}
