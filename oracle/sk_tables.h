/*
 * ORACLE (test infrastructure only -- never linked into or called from the product path).
 *
 * Slater-Koster angular functions for the sp3d5 basis and their derivatives, restated in C
 * from the reference Fortran, term by term and in the same evaluation order:
 *   values      t_*   : Source_files/TB_Koster_Slater.f90:371-625
 *   derivatives d_t_* : Source_files/TB_Koster_Slater.f90:633-967
 *   block builders KS_sp3d5 :5234-5348, d_KS_sp3d5 :5352-5466, d_KS_sp3d5_TEST :5589-5703
 * including the reference's known deviations from the exact derivative
 * (d_t_dxz_dx2_y2, :846 and :851 -- see SURVEY.md section 9 item 2).
 * Orbital order: s, px, py, pz, dxy, dxz, dyz, dx2-y2, d3z2-r2.
 * Channel order V[0..9]: ss_sigma, sp_sigma, sd_sigma, pp_sigma, pp_pi, pd_sigma, pd_pi,
 *                        dd_sigma, dd_pi, dd_delta.
 * Blocks are 9x9 column-major: ts[a + 9*b] == ts(a+1,b+1) of the reference.
 */
#ifndef ORC_SK_TABLES_H
#define ORC_SK_TABLES_H
#include <math.h>

#define ORC_SQRT3      1.7320508075688772   /* sqrt(3.0d0), TB_Koster_Slater.f90:41 */
#define ORC_SQRT3_HALF (ORC_SQRT3 * 0.5)    /* :42 */

/* ---- values (TB_Koster_Slater.f90:371-625) ---- */
static inline double t_s_s(double Vss) { return Vss; }
static inline double t_s_px(double da, double Vsp) { return da * Vsp; }
static inline double t_pa_pa(double da, double Vpps, double Vppp) {
    double da2 = da * da;
    return (Vpps - Vppp) * da2 + Vppp;
}
static inline double t_pa_pb(double da, double db, double Vpps, double Vppp) {
    double dadb = da * db;
    return (Vpps - Vppp) * dadb;
}
static inline double t_s_dab(double da, double db, double Vsd) { return ORC_SQRT3 * da * db * Vsd; }
static inline double t_s_dx2_y2(double da, double db, double Vsd) {
    return ORC_SQRT3_HALF * (da * da - db * db) * Vsd;
}
static inline double t_s_dz2_r2(double l, double m, double n, double Vsd) {
    return (n * n - 0.5 * (l * l + m * m)) * Vsd;
}
static inline double t_px_dxy(double l, double m, double Vpds, double Vpdp) {
    double l2 = l * l;
    return (ORC_SQRT3 * l2 * Vpds + (1.0 - 2.0 * l2) * Vpdp) * m;
}
static inline double t_px_dyz(double l, double m, double n, double Vpds, double Vpdp) {
    return (ORC_SQRT3 * Vpds - 2.0 * Vpdp) * l * m * n;
}
static inline double t_px_dx2_y2(double l, double m, double Vpds, double Vpdp) {
    double l2_m2 = l * l - m * m;
    return ((ORC_SQRT3_HALF * Vpds - Vpdp) * l2_m2 + Vpdp) * l;
}
static inline double t_py_dx2_y2(double l, double m, double Vpds, double Vpdp) {
    double l2_m2 = l * l - m * m;
    return ((ORC_SQRT3_HALF * Vpds - Vpdp) * l2_m2 - Vpdp) * m;
}
static inline double t_pz_dx2_y2(double l, double m, double n, double Vpds, double Vpdp) {
    double l2_m2 = l * l - m * m;
    return (ORC_SQRT3_HALF * Vpds - Vpdp) * l2_m2 * n;
}
static inline double t_pa_d3z2_r2(double da, double l, double m, double n, double Vpds, double Vpdp) {
    double n2 = n * n;
    return ((n2 - 0.5 * (l * l + m * m)) * Vpds - ORC_SQRT3 * n2 * Vpdp) * da;
}
static inline double t_pz_d3z2_r2(double l, double m, double n, double Vpds, double Vpdp) {
    double l2_m2 = l * l + m * m;
    double n2 = n * n;
    return ((n2 - 0.5 * l2_m2) * Vpds + ORC_SQRT3 * l2_m2 * Vpdp) * n;
}
static inline double t_dab_dab(double l, double m, double n, double Vs, double Vp, double Vd) {
    double l2 = l * l, m2 = m * m;
    double l2m2 = l2 * m2;
    return 3.0 * l2m2 * Vs + (l2 + m2 - 4.0 * l2m2) * Vp + (n * n + l2m2) * Vd;
}
static inline double t_dab_dbg(double l, double m, double n, double Vs, double Vp, double Vd) {
    double ln = l * n, m2 = m * m;
    return (3.0 * m2 * Vs + (1.0 - 4.0 * m2) * Vp + (m2 - 1.0) * Vd) * ln;
}
static inline double t_dxy_dx2_y2(double l, double m, double Vs, double Vp, double Vd) {
    double lm = l * m, l2m2 = l * l - m * m;
    return (1.50 * Vs - 2.0 * Vp + 0.5 * Vd) * lm * l2m2;
}
static inline double t_dyz_dx2_y2(double l, double m, double n, double Vs, double Vp, double Vd) {
    double mn = m * n, l2m2 = l * l - m * m;
    return (1.50 * l2m2 * Vs - (1.0 + 2.0 * l2m2) * Vp + (1.0 + 0.5 * l2m2) * Vd) * mn;
}
static inline double t_dxz_dx2_y2(double l, double m, double n, double Vs, double Vp, double Vd) {
    double ln = l * n, l2m2 = l * l - m * m;
    return (1.50 * l2m2 * Vs + (1.0 - 2.0 * l2m2) * Vp - (1.0 - 0.5 * l2m2) * Vd) * ln;
}
static inline double t_dxy_d3z2_r2(double l, double m, double n, double Vs, double Vp, double Vd) {
    double lm = l * m, l2m2 = l * l + m * m, n2 = n * n;
    return (ORC_SQRT3 * ((n2 - 0.5 * l2m2) * Vs - 2.0 * n2 * Vp) + ORC_SQRT3_HALF * (1.0 + n2) * Vd) * lm;
}
static inline double t_dyz_d3z2_r2(double l, double m, double n, double Vs, double Vp, double Vd) {
    double mn = m * n, l2m2 = l * l + m * m, n2 = n * n;
    return ((n2 - 0.5 * l2m2) * Vs + (l2m2 - n2) * Vp - 0.5 * l2m2 * Vd) * ORC_SQRT3 * mn;
}
static inline double t_dxz_d3z2_r2(double l, double m, double n, double Vs, double Vp, double Vd) {
    double ln = l * n, l2m2 = l * l + m * m, n2 = n * n;
    return ((n2 - 0.5 * l2m2) * Vs + (l2m2 - n2) * Vp - 0.5 * l2m2 * Vd) * ORC_SQRT3 * ln;
}
static inline double t_dx2_y2_dx2_y2(double l, double m, double n, double Vs, double Vp, double Vd) {
    double l2m2 = l * l + m * m, l2_m2 = l * l - m * m;
    double l2_m2_2 = l2_m2 * l2_m2, n2 = n * n;
    return 0.75 * l2_m2_2 * Vs + (l2m2 - l2_m2_2) * Vp + (n2 + 0.25 * l2_m2_2) * Vd;
}
static inline double t_dx2_y2_d3z2_r2(double l, double m, double n, double Vs, double Vp, double Vd) {
    double l2m2 = l * l + m * m, l2_m2 = l * l - m * m, n2 = n * n;
    return (0.5 * (n2 - 0.5 * l2m2) * Vs - n2 * Vp + 0.25 * (1.0 + n2) * Vd) * ORC_SQRT3 * l2_m2;
}
static inline double t_d3z2_r2_d3z2_r2(double l, double m, double n, double Vs, double Vp, double Vd) {
    double l2m2 = l * l + m * m, n2 = n * n;
    double temp = n2 - 0.5 * l2m2;
    return temp * temp * Vs + 3.0 * n2 * l2m2 * Vp + 0.75 * l2m2 * l2m2 * Vd;
}

/* ---- derivatives (TB_Koster_Slater.f90:633-967); d? = derivative of that cosine,
 *      dV? = derivative of that radial function, both w.r.t. the same variable ---- */
static inline double d_t_s_s(double dVss) { return dVss; }
static inline double d_t_s_px(double da, double dda, double Vsp, double dVsp) {
    return dda * Vsp + da * dVsp;
}
static inline double d_t_pa_pa(double da, double dda, double Vs, double dVs, double Vp, double dVp) {
    double da2 = da * da;
    return (dVs - dVp) * da2 + (Vs - Vp) * 2.0 * da * dda + dVp;
}
static inline double d_t_pa_pb(double da, double dda, double db, double ddb,
                               double Vs, double dVs, double Vp, double dVp) {
    return (dVs - dVp) * (da * db) + (Vs - Vp) * (dda * db + da * ddb);
}
static inline double d_t_s_dab(double da, double dda, double db, double ddb, double Vsd, double dVsd) {
    return ORC_SQRT3 * ((dda * db + da * ddb) * Vsd + da * db * dVsd);
}
static inline double d_t_s_dx2_y2(double da, double dda, double db, double ddb, double Vsd, double dVsd) {
    double E = 2.0 * (dda * da - ddb * db) * Vsd + (da * da - db * db) * dVsd;
    return ORC_SQRT3_HALF * E;
}
static inline double d_t_s_dz2_r2(double l, double dl, double m, double dm, double n, double dn,
                                  double Vsd, double dVsd) {
    return (2.0 * dn * n - (dl * l + dm * m)) * Vsd + (n * n - 0.5 * (l * l + m * m)) * dVsd;
}
static inline double d_t_px_dxy(double l, double dl, double m, double dm,
                                double Vs, double dVs, double Vp, double dVp) {
    double l2 = l * l, dll = 2.0 * dl * l;
    double temp = (1.0 - 2.0 * l2);
    double E = (ORC_SQRT3 * l2 * Vs + temp * Vp) * dm;
    E = E + (ORC_SQRT3 * (dll * Vs + l2 * dVs) - 2.0 * dll * Vp + temp * dVp) * m;
    return E;
}
static inline double d_t_px_dyz(double l, double dl, double m, double dm, double n, double dn,
                                double Vs, double dVs, double Vp, double dVp) {
    double mn = m * n;
    double E = (ORC_SQRT3 * dVs - 2.0 * dVp) * l * mn;
    E = E + (ORC_SQRT3 * Vs - 2.0 * Vp) * (dl * mn + l * dm * n + l * m * dn);
    return E;
}
static inline double d_t_px_dx2_y2(double l, double dl, double m, double dm,
                                   double Vs, double dVs, double Vp, double dVp) {
    double l2_m2 = l * l - m * m;
    double dl2_m2 = 2.0 * (dl * l - dm * m);
    double temp = (ORC_SQRT3_HALF * Vs - Vp);
    double E = (temp * l2_m2 + Vp) * dl;
    E = E + ((ORC_SQRT3_HALF * dVs - dVp) * l2_m2 + temp * dl2_m2 + dVp) * l;
    return E;
}
static inline double d_t_py_dx2_y2(double l, double dl, double m, double dm,
                                   double Vs, double dVs, double Vp, double dVp) {
    double l2_m2 = l * l - m * m;
    double dl2_m2 = 2.0 * (dl * l - dm * m);
    double temp = (ORC_SQRT3_HALF * Vs - Vp);
    double E = (temp * l2_m2 - Vp) * dm;
    E = E + ((ORC_SQRT3_HALF * dVs - dVp) * l2_m2 + temp * dl2_m2 - dVp) * m;
    return E;
}
static inline double d_t_pz_dx2_y2(double l, double dl, double m, double dm, double n, double dn,
                                   double Vs, double dVs, double Vp, double dVp) {
    double l2_m2 = l * l - m * m;
    double dl2_m2 = 2.0 * (dl * l - dm * m);
    double E = (ORC_SQRT3_HALF * dVs - dVp) * l2_m2 * n;
    E = E + (ORC_SQRT3_HALF * Vs - Vp) * (dl2_m2 * n + l2_m2 * dn);
    return E;
}
static inline double d_t_pa_d3z2_r2(double da, double dda, double l, double dl, double m, double dm,
                                    double n, double dn, double Vs, double dVs, double Vp, double dVp) {
    double n2 = n * n, dn2 = 2.0 * dn * n;
    double temp = (n2 - 0.5 * (l * l + m * m));
    double E = (temp * Vs - ORC_SQRT3 * n2 * Vp) * dda;
    E = E + ((dn2 - (dl * l + dm * m)) * Vs + temp * dVs - ORC_SQRT3 * (dn2 * Vp + n2 * dVp)) * da;
    return E;
}
static inline double d_t_pz_d3z2_r2(double l, double dl, double m, double dm, double n, double dn,
                                    double Vs, double dVs, double Vp, double dVp) {
    double l2_m2 = l * l + m * m;
    double dl2_m2 = 2.0 * (dl * l + dm * m);
    double n2 = n * n;
    double temp = (n2 - 0.5 * l2_m2);
    double E = (temp * Vs + ORC_SQRT3 * l2_m2 * Vp) * dn;
    E = E + ((2.0 * dn * n - 0.5 * dl2_m2) * Vs + temp * dVs + ORC_SQRT3 * (dl2_m2 * Vp + l2_m2 * dVp)) * n;
    return E;
}
static inline double d_t_dab_dab(double l, double dl, double m, double dm, double n, double dn,
                                 double Vs, double dVs, double Vp, double dVp, double Vd, double dVd) {
    double l2 = l * l, m2 = m * m;
    double l2m2 = l2 * m2;
    double dl2m2 = 2.0 * (dl * l * m2 + l2 * m * dm);
    double E = 3.0 * (dl2m2 * Vs + l2m2 * dVs);
    E = E + 2.0 * (l * dl + m * dm - 2.0 * dl2m2) * Vp + (l2 + m2 - 4.0 * l2m2) * dVp;
    E = E + (2.0 * n * dn + dl2m2) * Vd + (n * n + l2m2) * dVd;
    return E;
}
static inline double d_t_dab_dbg(double l, double dl, double m, double dm, double n, double dn,
                                 double Vs, double dVs, double Vp, double dVp, double Vd, double dVd) {
    double ln = l * n, m2 = m * m, dm2 = 2.0 * m * dm;
    double temp = (1.0 - 4.0 * m2), temp1 = (m2 - 1.0);
    double E = 3.0 * (dm2 * Vs + m2 * dVs);
    E = E + temp * dVp - 4.0 * dm2 * Vp;
    E = E + temp1 * dVd + dm2 * Vd;
    E = E * ln;
    E = E + (3.0 * m2 * Vs + temp * Vp + temp1 * Vd) * (dl * n + l * dn);
    return E;
}
static inline double d_t_dxy_dx2_y2(double l, double dl, double m, double dm,
                                    double Vs, double dVs, double Vp, double dVp, double Vd, double dVd) {
    double lm = l * m, l2m2 = l * l - m * m;
    double E = (1.50 * dVs - 2.0 * dVp + 0.5 * dVd) * lm * l2m2;
    E = E + (1.50 * Vs - 2.0 * Vp + 0.5 * Vd) * ((dl * m + l * dm) * l2m2 + 2.0 * lm * (dl * l - dm * m));
    return E;
}
static inline double d_t_dyz_dx2_y2(double l, double dl, double m, double dm, double n, double dn,
                                    double Vs, double dVs, double Vp, double dVp, double Vd, double dVd) {
    double mn = m * n, l2m2 = l * l - m * m;
    double dl2m2 = 2.0 * (dl * l - dm * m);
    double temp = (1.0 + 2.0 * l2m2), temp1 = (1.0 + 0.5 * l2m2);
    double E = (1.50 * (dl2m2 * Vs + l2m2 * dVs) - (2.0 * dl2m2 * Vp + temp * dVp) + temp1 * dVd + 0.5 * dl2m2 * Vd) * mn;
    E = E + (1.50 * l2m2 * Vs - temp * Vp + temp1 * Vd) * (dm * n + m * dn);
    return E;
}
/* NOTE: verbatim reference behaviour (TB_Koster_Slater.f90:846,851): l2m2 uses n, and the
 * second term carries +temp1*Vd although the value function has a minus sign.
 * orc_strict_reference = 0 switches to the exact derivative of t_dxz_dx2_y2 -- used ONLY by the
 * finite-difference self-test of the oracle, never for parity. */
static int orc_strict_reference = 1;
static inline double d_t_dxz_dx2_y2(double l, double dl, double m, double dm, double n, double dn,
                                    double Vs, double dVs, double Vp, double dVp, double Vd, double dVd) {
    double ln = l * n, l2m2 = orc_strict_reference ? (l * l - n * n) : (l * l - m * m);
    double dl2m2 = 2.0 * (dl * l - dm * m);
    double temp = (1.0 - 2.0 * l2m2), temp1 = (1.0 - 0.5 * l2m2);
    double E = (1.50 * (dl2m2 * Vs + l2m2 * dVs) + (-2.0 * dl2m2 * Vp + temp * dVp) - (temp1 * dVd - 0.5 * dl2m2 * Vd)) * ln;
    if (orc_strict_reference)
        E = E + (1.50 * l2m2 * Vs + temp * Vp + temp1 * Vd) * (dl * n + l * dn);
    else
        E = E + (1.50 * l2m2 * Vs + temp * Vp - temp1 * Vd) * (dl * n + l * dn);
    return E;
}
static inline double d_t_dxy_d3z2_r2(double l, double dl, double m, double dm, double n, double dn,
                                     double Vs, double dVs, double Vp, double dVp, double Vd, double dVd) {
    double lm = l * m, l2m2 = l * l + m * m;
    double dl2m2 = 2.0 * (dl * l + dm * m);
    double n2 = n * n, dn2 = 2.0 * n * dn;
    double temp = (n2 - 0.5 * l2m2), temp1 = (1.0 + n2);
    double E = (ORC_SQRT3 * (((dn2 - 0.5 * dl2m2) * Vs + temp * dVs) - 2.0 * (dn2 * Vp + n2 * dVp)) + ORC_SQRT3_HALF * (dn2 * Vd + temp1 * dVd)) * lm;
    E = E + (ORC_SQRT3 * (temp * Vs - 2.0 * n2 * Vp) + ORC_SQRT3_HALF * temp1 * Vd) * (dl * m + l * dm);
    return E;
}
static inline double d_t_dyz_d3z2_r2(double l, double dl, double m, double dm, double n, double dn,
                                     double Vs, double dVs, double Vp, double dVp, double Vd, double dVd) {
    double mn = m * n, l2m2 = l * l + m * m;
    double dl2m2 = 2.0 * (l * dl + m * dm);
    double n2 = n * n, dn2 = 2.0 * n * dn;
    double temp = (n2 - 0.5 * l2m2), temp1 = (l2m2 - n2);
    double E = (temp * Vs + temp1 * Vp - 0.5 * l2m2 * Vd) * (dm * n + m * dn);
    E = E + ((dn2 - 0.5 * dl2m2) * Vs + temp * dVs + (dl2m2 - dn2) * Vp + temp1 * dVp - 0.5 * (dl2m2 * Vd + l2m2 * dVd)) * mn;
    return E * ORC_SQRT3;
}
static inline double d_t_dxz_d3z2_r2(double l, double dl, double m, double dm, double n, double dn,
                                     double Vs, double dVs, double Vp, double dVp, double Vd, double dVd) {
    double ln = l * n, l2m2 = l * l + m * m;
    double dl2m2 = 2.0 * (l * dl + m * dm);
    double n2 = n * n, dn2 = 2.0 * n * dn;
    double temp = (n2 - 0.5 * l2m2), temp1 = (l2m2 - n2);
    double E = (temp * Vs + temp1 * Vp - 0.5 * l2m2 * Vd) * (dl * n + l * dn);
    E = E + ((dn2 - 0.5 * dl2m2) * Vs + temp * dVs + (dl2m2 - dn2) * Vp + temp1 * dVp - 0.5 * (dl2m2 * Vd + l2m2 * dVd)) * ln;
    return E * ORC_SQRT3;
}
static inline double d_t_dx2_y2_dx2_y2(double l, double dl, double m, double dm, double n, double dn,
                                       double Vs, double dVs, double Vp, double dVp, double Vd, double dVd) {
    double l2 = l * l, m2 = m * m;
    double l2m2 = l2 + m2;
    double ldl = l * dl, mdm = m * dm;
    double dl2m2 = 2.0 * (ldl + mdm);
    double l2_m2 = l2 - m2;
    double l2_m2_2 = l2_m2 * l2_m2;
    double dl2_m2_2 = 4.0 * l2_m2 * (ldl - mdm);
    double n2 = n * n, dn2 = 2.0 * n * dn;
    double E = 0.75 * (dl2_m2_2 * Vs + l2_m2_2 * dVs);
    E = E + (dl2m2 - dl2_m2_2) * Vp + (l2m2 - l2_m2_2) * dVp;
    E = E + (dn2 + 0.25 * dl2_m2_2) * Vd + (n2 + 0.25 * l2_m2_2) * dVd;
    return E;
}
static inline double d_t_dx2_y2_d3z2_r2(double l, double dl, double m, double dm, double n, double dn,
                                        double Vs, double dVs, double Vp, double dVp, double Vd, double dVd) {
    double l2 = l * l, dl2 = 2.0 * l * dl;
    double m2 = m * m, dm2 = 2.0 * m * dm;
    double l2m2 = l2 + m2, dl2m2 = dl2 + dm2;
    double l2_m2 = l2 - m2, dl2_m2 = dl2 - dm2;
    double n2 = n * n, dn2 = 2.0 * n * dn;
    double temp = (n2 - 0.5 * l2m2), temp1 = (1.0 + n2);
    double E = (0.5 * temp * Vs - n2 * Vp + 0.25 * temp1 * Vd) * dl2_m2;
    E = E + (0.5 * ((dn2 - 0.5 * dl2m2) * Vs + temp * dVs) - (dn2 * Vp + n2 * dVp) + 0.25 * (dn2 * Vd + temp1 * dVd)) * l2_m2;
    return E * ORC_SQRT3;
}
static inline double d_t_d3z2_r2_d3z2_r2(double l, double dl, double m, double dm, double n, double dn,
                                         double Vs, double dVs, double Vp, double dVp, double Vd, double dVd) {
    double l2m2 = l * l + m * m;
    double dl2m2 = 2.0 * (dl * l + dm * m);
    double n2 = n * n, dn2 = 2.0 * n * dn;
    double temp = n2 - 0.5 * l2m2;
    double dtemp = dn2 - 0.5 * dl2m2;
    double E = (2.0 * dtemp * Vs + temp * dVs) * temp;
    E = E + 3.0 * ((dn2 * l2m2 + n2 * dl2m2) * Vp + n2 * l2m2 * dVp);
    E = E + 0.75 * (2.0 * dl2m2 * Vd + l2m2 * dVd) * l2m2;
    return E;
}

#define TS(a, b) ts[((a) - 1) + 9 * ((b) - 1)]

/* fill rows 5..9 / columns from the upper triangle, shared by all three builders
 * (TB_Koster_Slater.f90:5294-5347 pattern: s-d and d-d symmetric, p-d antisymmetric) */
static inline void orc_ks_fill_lower(double* ts) {
    TS(3, 2) = TS(2, 3);
    TS(4, 2) = TS(2, 4); TS(4, 3) = TS(3, 4);
    for (int d = 5; d <= 9; ++d) {
        TS(d, 1) = TS(1, d);
        TS(d, 2) = -TS(2, d); TS(d, 3) = -TS(3, d); TS(d, 4) = -TS(4, d);
        for (int e = 5; e < d; ++e) TS(d, e) = TS(e, d);
    }
}

/* KS_sp3d5, TB_Koster_Slater.f90:5234-5348 (s-p and p-d entries carry the minus sign) */
static inline void orc_KS_sp3d5(const double* V, double l, double m, double n, double* ts) {
    TS(1, 1) = t_s_s(V[0]);
    TS(1, 2) = -t_s_px(l, V[1]);
    TS(1, 3) = -t_s_px(m, V[1]);
    TS(1, 4) = -t_s_px(n, V[1]);
    TS(1, 5) = t_s_dab(l, m, V[2]);
    TS(1, 6) = t_s_dab(l, n, V[2]);
    TS(1, 7) = t_s_dab(m, n, V[2]);
    TS(1, 8) = t_s_dx2_y2(l, m, V[2]);
    TS(1, 9) = t_s_dz2_r2(l, m, n, V[2]);
    TS(2, 1) = -TS(1, 2);
    TS(2, 2) = t_pa_pa(l, V[3], V[4]);
    TS(2, 3) = t_pa_pb(l, m, V[3], V[4]);
    TS(2, 4) = t_pa_pb(l, n, V[3], V[4]);
    TS(2, 5) = -t_px_dxy(l, m, V[5], V[6]);
    TS(2, 6) = -t_px_dxy(l, n, V[5], V[6]);
    TS(2, 7) = -t_px_dyz(l, m, n, V[5], V[6]);
    TS(2, 8) = -t_px_dx2_y2(l, m, V[5], V[6]);
    TS(2, 9) = -t_pa_d3z2_r2(l, l, m, n, V[5], V[6]);
    TS(3, 1) = -TS(1, 3);
    TS(3, 3) = t_pa_pa(m, V[3], V[4]);
    TS(3, 4) = t_pa_pb(m, n, V[3], V[4]);
    TS(3, 5) = -t_px_dxy(m, l, V[5], V[6]);
    TS(3, 6) = -t_px_dyz(m, l, n, V[5], V[6]);
    TS(3, 7) = -t_px_dxy(m, n, V[5], V[6]);
    TS(3, 8) = -t_py_dx2_y2(l, m, V[5], V[6]);
    TS(3, 9) = -t_pa_d3z2_r2(m, l, m, n, V[5], V[6]);
    TS(4, 1) = -TS(1, 4);
    TS(4, 4) = t_pa_pa(n, V[3], V[4]);
    TS(4, 5) = -t_px_dyz(n, l, m, V[5], V[6]);
    TS(4, 6) = -t_px_dxy(n, l, V[5], V[6]);
    TS(4, 7) = -t_px_dxy(n, m, V[5], V[6]);
    TS(4, 8) = -t_pz_dx2_y2(l, m, n, V[5], V[6]);
    TS(4, 9) = -t_pz_d3z2_r2(l, m, n, V[5], V[6]);
    TS(5, 5) = t_dab_dab(l, m, n, V[7], V[8], V[9]);
    TS(5, 6) = t_dab_dbg(m, l, n, V[7], V[8], V[9]);
    TS(5, 7) = t_dab_dbg(l, m, n, V[7], V[8], V[9]);
    TS(5, 8) = t_dxy_dx2_y2(l, m, V[7], V[8], V[9]);
    TS(5, 9) = t_dxy_d3z2_r2(l, m, n, V[7], V[8], V[9]);
    TS(6, 6) = t_dab_dab(l, n, m, V[7], V[8], V[9]);
    TS(6, 7) = t_dab_dbg(l, n, m, V[7], V[8], V[9]);
    TS(6, 8) = t_dxz_dx2_y2(l, m, n, V[7], V[8], V[9]);
    TS(6, 9) = t_dxz_d3z2_r2(l, m, n, V[7], V[8], V[9]);
    TS(7, 7) = t_dab_dab(m, n, l, V[7], V[8], V[9]);
    TS(7, 8) = t_dyz_dx2_y2(l, m, n, V[7], V[8], V[9]);
    TS(7, 9) = t_dyz_d3z2_r2(l, m, n, V[7], V[8], V[9]);
    TS(8, 8) = t_dx2_y2_dx2_y2(l, m, n, V[7], V[8], V[9]);
    TS(8, 9) = t_dx2_y2_d3z2_r2(l, m, n, V[7], V[8], V[9]);
    TS(9, 9) = t_d3z2_r2_d3z2_r2(l, m, n, V[7], V[8], V[9]);
    orc_ks_fill_lower(ts);
}

/* sgn = -1: d_KS_sp3d5 (:5352-5466);  sgn = +1: d_KS_sp3d5_TEST (:5589-5703).
 * The two differ only in the sign carried by the s-p and p-d upper-triangle entries. */
static inline void orc_d_KS_sp3d5_sgn(const double* V, const double* dV, double l, double m, double n,
                                      double dl, double dm, double dn, double sgn, double* ts) {
    TS(1, 1) = d_t_s_s(dV[0]);
    TS(1, 2) = sgn * d_t_s_px(l, dl, V[1], dV[1]);
    TS(1, 3) = sgn * d_t_s_px(m, dm, V[1], dV[1]);
    TS(1, 4) = sgn * d_t_s_px(n, dn, V[1], dV[1]);
    TS(1, 5) = d_t_s_dab(l, dl, m, dm, V[2], dV[2]);
    TS(1, 6) = d_t_s_dab(l, dl, n, dn, V[2], dV[2]);
    TS(1, 7) = d_t_s_dab(m, dm, n, dn, V[2], dV[2]);
    TS(1, 8) = d_t_s_dx2_y2(l, dl, m, dm, V[2], dV[2]);
    TS(1, 9) = d_t_s_dz2_r2(l, dl, m, dm, n, dn, V[2], dV[2]);
    TS(2, 1) = -TS(1, 2);
    TS(2, 2) = d_t_pa_pa(l, dl, V[3], dV[3], V[4], dV[4]);
    TS(2, 3) = d_t_pa_pb(l, dl, m, dm, V[3], dV[3], V[4], dV[4]);
    TS(2, 4) = d_t_pa_pb(l, dl, n, dn, V[3], dV[3], V[4], dV[4]);
    TS(2, 5) = sgn * d_t_px_dxy(l, dl, m, dm, V[5], dV[5], V[6], dV[6]);
    TS(2, 6) = sgn * d_t_px_dxy(l, dl, n, dn, V[5], dV[5], V[6], dV[6]);
    TS(2, 7) = sgn * d_t_px_dyz(l, dl, m, dm, n, dn, V[5], dV[5], V[6], dV[6]);
    TS(2, 8) = sgn * d_t_px_dx2_y2(l, dl, m, dm, V[5], dV[5], V[6], dV[6]);
    TS(2, 9) = sgn * d_t_pa_d3z2_r2(l, dl, l, dl, m, dm, n, dn, V[5], dV[5], V[6], dV[6]);
    TS(3, 1) = -TS(1, 3);
    TS(3, 3) = d_t_pa_pa(m, dm, V[3], dV[3], V[4], dV[4]);
    TS(3, 4) = d_t_pa_pb(m, dm, n, dn, V[3], dV[3], V[4], dV[4]);
    TS(3, 5) = sgn * d_t_px_dxy(m, dm, l, dl, V[5], dV[5], V[6], dV[6]);
    TS(3, 6) = sgn * d_t_px_dyz(m, dm, l, dl, n, dn, V[5], dV[5], V[6], dV[6]);
    TS(3, 7) = sgn * d_t_px_dxy(m, dm, n, dn, V[5], dV[5], V[6], dV[6]);
    TS(3, 8) = sgn * d_t_py_dx2_y2(l, dl, m, dm, V[5], dV[5], V[6], dV[6]);
    TS(3, 9) = sgn * d_t_pa_d3z2_r2(m, dm, l, dl, m, dm, n, dn, V[5], dV[5], V[6], dV[6]);
    TS(4, 1) = -TS(1, 4);
    TS(4, 4) = d_t_pa_pa(n, dn, V[3], dV[3], V[4], dV[4]);
    TS(4, 5) = sgn * d_t_px_dyz(n, dn, l, dl, m, dm, V[5], dV[5], V[6], dV[6]);
    TS(4, 6) = sgn * d_t_px_dxy(n, dn, l, dl, V[5], dV[5], V[6], dV[6]);
    TS(4, 7) = sgn * d_t_px_dxy(n, dn, m, dm, V[5], dV[5], V[6], dV[6]);
    TS(4, 8) = sgn * d_t_pz_dx2_y2(l, dl, m, dm, n, dn, V[5], dV[5], V[6], dV[6]);
    TS(4, 9) = sgn * d_t_pz_d3z2_r2(l, dl, m, dm, n, dn, V[5], dV[5], V[6], dV[6]);
    TS(5, 5) = d_t_dab_dab(l, dl, m, dm, n, dn, V[7], dV[7], V[8], dV[8], V[9], dV[9]);
    TS(5, 6) = d_t_dab_dbg(m, dm, l, dl, n, dn, V[7], dV[7], V[8], dV[8], V[9], dV[9]);
    TS(5, 7) = d_t_dab_dbg(l, dl, m, dm, n, dn, V[7], dV[7], V[8], dV[8], V[9], dV[9]);
    TS(5, 8) = d_t_dxy_dx2_y2(l, dl, m, dm, V[7], dV[7], V[8], dV[8], V[9], dV[9]);
    TS(5, 9) = d_t_dxy_d3z2_r2(l, dl, m, dm, n, dn, V[7], dV[7], V[8], dV[8], V[9], dV[9]);
    TS(6, 6) = d_t_dab_dab(l, dl, n, dn, m, dm, V[7], dV[7], V[8], dV[8], V[9], dV[9]);
    TS(6, 7) = d_t_dab_dbg(l, dl, n, dn, m, dm, V[7], dV[7], V[8], dV[8], V[9], dV[9]);
    TS(6, 8) = d_t_dxz_dx2_y2(l, dl, m, dm, n, dn, V[7], dV[7], V[8], dV[8], V[9], dV[9]);
    TS(6, 9) = d_t_dxz_d3z2_r2(l, dl, m, dm, n, dn, V[7], dV[7], V[8], dV[8], V[9], dV[9]);
    TS(7, 7) = d_t_dab_dab(m, dm, n, dn, l, dl, V[7], dV[7], V[8], dV[8], V[9], dV[9]);
    TS(7, 8) = d_t_dyz_dx2_y2(l, dl, m, dm, n, dn, V[7], dV[7], V[8], dV[8], V[9], dV[9]);
    TS(7, 9) = d_t_dyz_d3z2_r2(l, dl, m, dm, n, dn, V[7], dV[7], V[8], dV[8], V[9], dV[9]);
    TS(8, 8) = d_t_dx2_y2_dx2_y2(l, dl, m, dm, n, dn, V[7], dV[7], V[8], dV[8], V[9], dV[9]);
    TS(8, 9) = d_t_dx2_y2_d3z2_r2(l, dl, m, dm, n, dn, V[7], dV[7], V[8], dV[8], V[9], dV[9]);
    TS(9, 9) = d_t_d3z2_r2_d3z2_r2(l, dl, m, dm, n, dn, V[7], dV[7], V[8], dV[8], V[9], dV[9]);
    orc_ks_fill_lower(ts);
}
#undef TS
#endif
