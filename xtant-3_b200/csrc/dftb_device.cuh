// DFTB (Slater-Koster table) radial functions and hetero-nuclear two-centre blocks, device side.
// Reference: Source_files/TB_DFTB.f90:56-280 (Construct_Vij_DFTB, DFTB_radial_function, d_DFTB_radial_function),
//            Little_subroutines.f90:349-381 (Fermi_function, d_Fermi_function), :1031-1125 (linear_interpolation),
//            TB_Koster_Slater.f90:4125-4189,4406-4639 (KS_sp3_hetero, KS_sp3d5_hetero and their derivatives).
#pragma once
#include "xtb_ctx.h"
#include "sk_device.cuh"

// device image of one type(TB_H_DFTB) entry: scalars + offsets (in doubles) into the shared table buffer
struct DftbDev {
    double rcut, d, Ed, Ep, Es;
    int n_r, pad_;
    long long off_R, off_V, off_S;       // Rr[n_r], Vr[n_r*10], Sr[n_r*10] (column-major (ir, channel))
};

namespace dftb {

__host__ __device__ __forceinline__ int n_v(int n_orb) { return n_orb == 1 ? 1 : (n_orb == 4 ? 4 : 10); }
// compact channel c of the basis -> table column (sp3: ss, sp, pp_sigma, pp_pi = columns 0,1,3,4; TB_DFTB.f90:175-182);
// the same number is the position in the 10-channel vector the Slater-Koster routines take
__host__ __device__ __forceinline__ int col(int n_orb, int c) { return (n_orb == 4 && c >= 2) ? c + 1 : c; }

// linear_interpolation without optional arguments (Little_subroutines.f90:1031-1125); i is 1-based
__device__ __forceinline__ double lin_interp(const double* __restrict__ xa, const double* __restrict__ ya, double x, int i) {
    if (i > 1) {
        const double y1 = ya[i - 1], y0 = ya[i - 2], x1 = xa[i - 1], x0 = xa[i - 2];
        if (fabs(y1 - y0) < 1.0e-6 * fabs(y1)) return y1;
        if (fabs(x1 - x0) < 1.0e-6 * fabs(x1)) return (y0 + y1) * 0.5;
        if (fabs(x - x0) >= 1.0e-6 * fabs(x0) || fabs(x1) > 1.0e-8) return y0 + (y1 - y0) / (x1 - x0) * (x - x0);
        return y1;
    }
    if (fabs(xa[1] - xa[0]) > 1.0e-8) return ya[0] - (ya[1] - ya[0]) / (xa[1] - xa[0]) * (xa[0] - x);
    return ya[0];
}

// M_Vij(j,i,:) / M_SVij (mat = 1) and their r-derivatives for one list entry at distance r [A], Fermi cut-off included
// (Construct_Vij_DFTB, TB_DFTB.f90:170-207).  Results go to the 10-channel positions; unused channels are zero.
template <int NO>
__device__ __forceinline__ void radial(const DftbDev& P, const double* __restrict__ tab, int mat, double r, double* V, double* dV) {
#pragma unroll
    for (int c = 0; c < 10; ++c) { V[c] = 0.0; dV[c] = 0.0; }
    const double* Rr = tab + P.off_R;
    const double* T = tab + (mat == 0 ? P.off_V : P.off_S);
    const double dr = Rr[0];
    const int ia = (int)ceil(r / dr);
    double F = 0.0, dF = 0.0;
    if (!(r >= P.rcut + 10.0 * P.d)) {
        const double ex = exp((r - P.rcut) / P.d);
        const double e1 = 1.0 + ex;
        F = 1.0 / e1;
        dF = -ex / (P.d * e1 * e1);
    }
    constexpr int NV = NO == 1 ? 1 : (NO == 4 ? 4 : 10);
#pragma unroll
    for (int c = 0; c < NV; ++c) {
        const int k = col(NO, c);
        const double* y = T + (size_t)P.n_r * k;
        double v, dv;
        if (ia > P.n_r) { v = y[P.n_r - 1]; dv = 0.0; }
        else {
            v = lin_interp(Rr, y, r, ia);
            dv = ia == 1 ? (y[1] - y[0]) / dr : (y[ia - 1] - y[ia - 2]) / dr;
        }
        dV[k] = dv * F + v * dF;
        V[k] = v * F;
    }
}

// hetero-nuclear block entry (a,b) from the reference-signed upper triangles of KS_sp3d5(V12) and KS_sp3d5(V21):
//   a <= b : up12(a,b)            a > b : (-1)^(l_a+l_b) * up21(b,a)        (KS_sp3d5_hetero, TB_Koster_Slater.f90:4406-4520;
//   the sp3 table :4125-4155 is its leading 4x4 part)
__device__ __forceinline__ double hetero(const double* up12, const double* up21, int a, int b) {
    return a <= b ? up12[sk::ut(a, b)] : sk::lower_sign(a, b) * up21[sk::ut(b, a)];
}

// derivative block D(a,b) as d_KS_sp3d5_hetero defines it (:4524-4639): every entry carries the extra parity sign
// (-1)^(l_a+l_b) relative to the derivative of the value table, i.e.
//   a <= b : sgn(a,b) * d up12(a,b)        a > b : d up21(b,a)
// contracted with a density block A(a,b) (tr: A is stored transposed):  sum_ab D(a,b) A(a,b)
template <int NO>
__device__ __forceinline__ double contract_hetero(const double* d12, const double* d21, const double* A, bool tr) {
    double s = 0.0;
#pragma unroll
    for (int a = 0; a < NO; ++a)
#pragma unroll
        for (int b = a; b < NO; ++b) {
            const double x = tr ? A[b + NO * a] : A[a + NO * b];      // A(a,b)
            const double y = tr ? A[a + NO * b] : A[b + NO * a];      // A(b,a)
            s += sk::lower_sign(a, b) * d12[sk::ut(a, b)] * x;
            if (a != b) s += d21[sk::ut(a, b)] * y;
        }
    return s;
}

// d/dparam of the reference-signed upper triangle (with the reference's own d_t_dxz_dx2_y2, see sk_device.cuh)
__device__ __forceinline__ void dupper(const double* V, const double* dV, double l, double m, double n,
                                       double dl, double dm, double dn, double* up) {
    sk::Dual Vd[10], o[45];
#pragma unroll
    for (int c = 0; c < 10; ++c) Vd[c] = sk::mk(V[c], dV[c]);
    sk::block_sp3d5_upper<sk::Dual>(Vd, sk::mk(l, dl), sk::mk(m, dm), sk::mk(n, dn), o);
#pragma unroll
    for (int e = 0; e < 45; ++e) up[e] = o[e].d;
    up[sk::ut(5, 7)] = sk::ref_d_dxz_dx2y2(l, dl, m, dm, n, dn, V[7], dV[7], V[8], dV[8], V[9], dV[9]);
}

}  // namespace dftb
