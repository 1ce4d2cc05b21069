// NRL (Mehl-Papaconstantopoulos) radial functions, device side.
// Reference: Source_files/TB_NRL.f90:2345-2405 (On_site_NRL, Overlap_function_NRL, exp_F_NRL, F_NRL, d_F_NRL),
//            :3201-3229 (d_Overlap_function_NRL, d_exp_F_NRL), channel map :2164-2201.
#pragma once
#include "xtb_ctx.h"

namespace nrl {

// channel c (ss,sp,sd,pps,ppp,pds,pdp,dds,ddp,ddd) <- parameter index (file order ss,sp,pps,ppp,sd,pd..)
__device__ __forceinline__ int ch2par(int c) { return c == 2 ? 4 : (c == 3 ? 2 : (c == 4 ? 3 : c)); }
// delta(l,l') of the overlap form Eq.(23): 1 for ss,pps,ppp,dds,ddp,ddd
__device__ __forceinline__ double ch_ll(int c) { return (c == 1 || c == 2 || c == 5 || c == 6) ? 0.0 : 1.0; }

struct Cut {   // F(r) and F'(r), r in bohr
    double F, dF;
};
__device__ __forceinline__ Cut cutoff(double r, double rc, double dl) {
    Cut o;
    if (r >= rc) { o.F = 0.0; o.dF = 0.0; return o; }
    const double e = exp((r - rc) / dl + 5.0);
    const double den = 1.0 + e;
    o.F = 1.0 / den;
    o.dF = -e / (dl * den * den);
    return o;
}

// value and r-derivative of one radial function: poly(r) * exp(-lam^2 r) * F(r)
// ind = 1: poly = ll + p r + q r^2 + s r^3 (Eq. 23); else poly = p + q r + s r^2 (Eq. 19/20)
__device__ __forceinline__ void radial(double r, double p, double q, double s, double lam, double rc, const Cut& ct,
                                       int ind, double ll, double& V, double& dV) {
    if (r >= rc) { V = 0.0; dV = 0.0; return; }
    const double l2 = lam * lam;
    const double ex = exp(-l2 * r);
    const double X = ex * ct.F;                    // exp_F_NRL
    const double dX = ex * (ct.dF - l2 * ct.F);    // d_exp_F_NRL
    double poly, dpoly;
    if (ind == 1) {
        const double r2 = r * r;
        poly = ll + p * r + q * r2 + s * r * r2;
        dpoly = p + 2.0 * q * r + 3.0 * s * r2;
    } else {
        poly = p + q * r + s * r * r;
        dpoly = q + 2.0 * s * r;
    }
    V = poly * X;
    dV = dpoly * X + poly * dX;
}

// all 10 channels of H (mat = 0) or S (mat = 1) for a pair with TB(KOA1,KOA2) = P at r [bohr]
__device__ __forceinline__ void radial10(const xtb_nrl_pair& P, int mat, double r, double* V, double* dV) {
    const Cut ct = cutoff(r, P.Rc, P.lden);
#pragma unroll
    for (int c = 0; c < 10; ++c) {
        const int p = ch2par(c);
        if (mat == 0) radial(r, P.ellm[p], P.fllm[p], P.gllm[p], P.hllm[p], P.Rc, ct, 0, 0.0, V[c], dV[c]);
        else          radial(r, P.pllm[p], P.qllm[p], P.rllm[p], P.sllm[p], P.Rc, ct, P.ind_overlap, ch_ll(c), V[c], dV[c]);
    }
}

// embedded-density envelope X(r; lambd) and its derivative
__device__ __forceinline__ void envelope(const xtb_nrl_pair& P, double r, double& X, double& dX) {
    if (r >= P.Rc) { X = 0.0; dX = 0.0; return; }
    const Cut ct = cutoff(r, P.Rc, P.lden);
    const double l2 = P.lambd * P.lambd;
    const double ex = exp(-l2 * r);
    X = ex * ct.F;
    dX = ex * (ct.dF - l2 * ct.F);
}

__device__ __forceinline__ int onsite_set(int o, int ind_split) {
    return o == 0 ? 0 : (o <= 3 ? 1 : (o <= 6 ? 2 : (ind_split == 1 ? 3 : 2)));
}
__device__ __forceinline__ double onsite(double rho, double a, double b, double c, double d) {
    const double r23 = pow(rho, 2.0 / 3.0);
    return a + b * r23 + c * r23 * r23 + d * rho * rho;
}
// dh_l/drho (d_On_site_NRL, TB_NRL.f90:3109-3111); 0 when rho <= 0
__device__ __forceinline__ double d_onsite(double rho, double b, double c, double d) {
    if (!(rho > 0.0)) return 0.0;
    const double r13 = pow(rho, 1.0 / 3.0);
    return (2.0 / 3.0) * b / r13 + (2.0 * (2.0 / 3.0)) * c * r13 + 2.0 * d * rho;
}

}  // namespace nrl
