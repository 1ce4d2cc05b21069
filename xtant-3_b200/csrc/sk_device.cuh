// Slater-Koster two-centre blocks for the sp3d5 (and sp3) basis, device side.
//
// Own formulation (Slater & Koster, Phys. Rev. 94, 1498 (1954), Table I) evaluated on a generic
// scalar type T: T = double gives the block, T = Dual gives block + directional derivative by
// forward-mode differentiation.  This is deliberately NOT the reference's hand-derived d_t_*
// routines (Source_files/TB_Koster_Slater.f90:633-967): the parity tests compare this
// independent path against the oracle's term-by-term restatement of those routines.
//
// The one place where the reference's derivative is not the derivative of its own value
// function -- d_t_dxz_dx2_y2, TB_Koster_Slater.f90:846,851 -- is reproduced explicitly
// (strict_reference) because forces must match the reference, not finite differences.
//
// Orbital order (reference): 0 s, 1 px, 2 py, 3 pz, 4 dxy, 5 dxz, 6 dyz, 7 dx2-y2, 8 d3z2-r2.
// Channels V[0..9]: ss_s, sp_s, sd_s, pp_s, pp_p, pd_s, pd_p, dd_s, dd_p, dd_d.
// Sign convention of KS_sp3d5 (TB_Koster_Slater.f90:5234-5348): upper-triangle s-p and p-d
// entries carry a minus; lower triangle = upper * (-1)^(l_a + l_b).
#pragma once
#include <cuda_runtime.h>

namespace sk {

struct Dual {
    double v, d;
};
__host__ __device__ __forceinline__ Dual mk(double v, double d) { Dual r; r.v = v; r.d = d; return r; }
__host__ __device__ __forceinline__ Dual operator+(Dual a, Dual b) { return mk(a.v + b.v, a.d + b.d); }
__host__ __device__ __forceinline__ Dual operator-(Dual a, Dual b) { return mk(a.v - b.v, a.d - b.d); }
__host__ __device__ __forceinline__ Dual operator*(Dual a, Dual b) { return mk(a.v * b.v, a.d * b.v + a.v * b.d); }
__host__ __device__ __forceinline__ Dual operator*(double a, Dual b) { return mk(a * b.v, a * b.d); }
__host__ __device__ __forceinline__ Dual operator*(Dual a, double b) { return mk(a.v * b, a.d * b); }
__host__ __device__ __forceinline__ Dual operator+(double a, Dual b) { return mk(a + b.v, b.d); }
__host__ __device__ __forceinline__ Dual operator+(Dual a, double b) { return mk(a.v + b, a.d); }
__host__ __device__ __forceinline__ Dual operator-(double a, Dual b) { return mk(a - b.v, -b.d); }
__host__ __device__ __forceinline__ Dual operator-(Dual a, double b) { return mk(a.v - b, a.d); }
__host__ __device__ __forceinline__ Dual operator-(Dual a) { return mk(-a.v, -a.d); }

__host__ __device__ __forceinline__ double val(double a) { return a; }
__host__ __device__ __forceinline__ double val(Dual a) { return a.v; }
__host__ __device__ __forceinline__ double der(Dual a) { return a.d; }

// index of (a,b), a <= b, in the packed upper triangle of a 9x9 block
__host__ __device__ __forceinline__ constexpr int ut(int a, int b) { return a * 9 - (a * (a - 1)) / 2 + (b - a); }

#define SK_SQRT3 1.7320508075688772

// Packed upper triangle (45 entries) of the reference-signed block.
template <class T>
__host__ __device__ __forceinline__ void block_sp3d5_upper(const T* V, T l, T m, T n, T* o) {
    const T ss = V[0], sp = V[1], sd = V[2], pps = V[3], ppp = V[4], pds = V[5], pdp = V[6], dds = V[7], ddp = V[8], ddd = V[9];
    const T l2 = l * l, m2 = m * m, n2 = n * n;
    const T lm = l * m, ln = l * n, mn = m * n;
    const T a = l2 + m2;          // l^2 + m^2
    const T b = l2 - m2;          // l^2 - m^2
    const T g = n2 - 0.5 * a;     // n^2 - (l^2+m^2)/2
    const double r3 = SK_SQRT3, hr3 = 0.5 * SK_SQRT3;
    // s-s
    o[ut(0, 0)] = ss;
    // s-p  (minus: KS_sp3d5 convention)
    o[ut(0, 1)] = -(l * sp);
    o[ut(0, 2)] = -(m * sp);
    o[ut(0, 3)] = -(n * sp);
    // s-d
    const T r3sd = r3 * sd;
    o[ut(0, 4)] = lm * r3sd;
    o[ut(0, 5)] = ln * r3sd;
    o[ut(0, 6)] = mn * r3sd;
    o[ut(0, 7)] = hr3 * (b * sd);
    o[ut(0, 8)] = g * sd;
    // p-p
    const T dpp = pps - ppp;
    o[ut(1, 1)] = dpp * l2 + ppp;
    o[ut(1, 2)] = dpp * lm;
    o[ut(1, 3)] = dpp * ln;
    o[ut(2, 2)] = dpp * m2 + ppp;
    o[ut(2, 3)] = dpp * mn;
    o[ut(3, 3)] = dpp * n2 + ppp;
    // p-d  (minus: KS_sp3d5 convention)
    const T q1 = r3 * pds - 2.0 * pdp;            // sqrt3 pds - 2 pdp
    const T q2 = hr3 * pds - pdp;                 // sqrt3/2 pds - pdp
    const T r3pds = r3 * pds, r3pdp = r3 * pdp;
    const T lmn = lm * n;
    o[ut(1, 4)] = -((r3pds * l2 + (1.0 - 2.0 * l2) * pdp) * m);   // x,xy
    o[ut(1, 5)] = -((r3pds * l2 + (1.0 - 2.0 * l2) * pdp) * n);   // x,xz
    o[ut(1, 6)] = -(q1 * lmn);                                    // x,yz
    o[ut(1, 7)] = -((q2 * b + pdp) * l);                          // x,x2-y2
    o[ut(1, 8)] = -((g * pds - r3pdp * n2) * l);                  // x,3z2-r2
    o[ut(2, 4)] = -((r3pds * m2 + (1.0 - 2.0 * m2) * pdp) * l);   // y,xy
    o[ut(2, 5)] = -(q1 * lmn);                                    // y,xz
    o[ut(2, 6)] = -((r3pds * m2 + (1.0 - 2.0 * m2) * pdp) * n);   // y,yz
    o[ut(2, 7)] = -((q2 * b - pdp) * m);                          // y,x2-y2
    o[ut(2, 8)] = -((g * pds - r3pdp * n2) * m);                  // y,3z2-r2
    o[ut(3, 4)] = -(q1 * lmn);                                    // z,xy
    o[ut(3, 5)] = -((r3pds * n2 + (1.0 - 2.0 * n2) * pdp) * l);   // z,xz
    o[ut(3, 6)] = -((r3pds * n2 + (1.0 - 2.0 * n2) * pdp) * m);   // z,yz
    o[ut(3, 7)] = -(q2 * b * n);                                  // z,x2-y2
    o[ut(3, 8)] = -((g * pds + r3pdp * a) * n);                   // z,3z2-r2
    // d-d
    const T l2m2 = l2 * m2, l2n2 = l2 * n2, m2n2 = m2 * n2;
    o[ut(4, 4)] = 3.0 * l2m2 * dds + (l2 + m2 - 4.0 * l2m2) * ddp + (n2 + l2m2) * ddd;      // xy,xy
    o[ut(5, 5)] = 3.0 * l2n2 * dds + (l2 + n2 - 4.0 * l2n2) * ddp + (m2 + l2n2) * ddd;      // xz,xz
    o[ut(6, 6)] = 3.0 * m2n2 * dds + (m2 + n2 - 4.0 * m2n2) * ddp + (l2 + m2n2) * ddd;      // yz,yz
    o[ut(4, 5)] = (3.0 * l2 * dds + (1.0 - 4.0 * l2) * ddp + (l2 - 1.0) * ddd) * mn;        // xy,xz
    o[ut(4, 6)] = (3.0 * m2 * dds + (1.0 - 4.0 * m2) * ddp + (m2 - 1.0) * ddd) * ln;        // xy,yz
    o[ut(5, 6)] = (3.0 * n2 * dds + (1.0 - 4.0 * n2) * ddp + (n2 - 1.0) * ddd) * lm;        // xz,yz
    o[ut(4, 7)] = (1.5 * dds - 2.0 * ddp + 0.5 * ddd) * lm * b;                              // xy,x2-y2
    o[ut(6, 7)] = (1.5 * b * dds - (1.0 + 2.0 * b) * ddp + (1.0 + 0.5 * b) * ddd) * mn;      // yz,x2-y2
    o[ut(5, 7)] = (1.5 * b * dds + (1.0 - 2.0 * b) * ddp - (1.0 - 0.5 * b) * ddd) * ln;      // xz,x2-y2
    const T e1 = r3 * (g * dds - 2.0 * n2 * ddp) + hr3 * ((1.0 + n2) * ddd);
    o[ut(4, 8)] = e1 * lm;                                                                   // xy,3z2-r2
    const T e2 = (g * dds + (a - n2) * ddp - 0.5 * a * ddd) * r3;
    o[ut(6, 8)] = e2 * mn;                                                                   // yz,3z2-r2
    o[ut(5, 8)] = e2 * ln;                                                                   // xz,3z2-r2
    const T b2 = b * b;
    o[ut(7, 7)] = 0.75 * b2 * dds + (a - b2) * ddp + (n2 + 0.25 * b2) * ddd;                 // x2-y2,x2-y2
    o[ut(7, 8)] = (0.5 * g * dds - n2 * ddp + 0.25 * (1.0 + n2) * ddd) * (r3 * b);           // x2-y2,3z2-r2
    o[ut(8, 8)] = g * g * dds + 3.0 * n2 * a * ddp + 0.75 * a * a * ddd;                     // 3z2-r2,3z2-r2
}

// d_t_dxz_dx2_y2 exactly as the reference evaluates it (TB_Koster_Slater.f90:839-853),
// i.e. with l2m2 = l*l - n*n and "+ temp1*Vdd_delta" in the second term.
__host__ __device__ __forceinline__ double ref_d_dxz_dx2y2(double l, double dl, double m, double dm, double n, double dn,
                                                          double Vs, double dVs, double Vp, double dVp, double Vd, double dVd) {
    const double ln = l * n, q = l * l - n * n;
    const double dq = 2.0 * (dl * l - dm * m);
    const double t0 = 1.0 - 2.0 * q, t1 = 1.0 - 0.5 * q;
    double E = (1.5 * (dq * Vs + q * dVs) + (-2.0 * dq * Vp + t0 * dVp) - (t1 * dVd - 0.5 * dq * Vd)) * ln;
    E += (1.5 * q * Vs + t0 * Vp + t1 * Vd) * (dl * n + l * dn);
    return E;
}

// sign of the lower-triangle entry (b,a) relative to the upper (a,b): (-1)^(l_a+l_b)
__host__ __device__ __forceinline__ double lower_sign(int a, int b) {
    const int la = a == 0 ? 0 : (a <= 3 ? 1 : 2);
    const int lb = b == 0 ? 0 : (b <= 3 ? 1 : 2);
    return ((la + lb) & 1) ? -1.0 : 1.0;
}

// ---- sp3 (Pettifor) 4x4 block: KS_sp3 convention is fixed in pettifor.cuh ----

}  // namespace sk
