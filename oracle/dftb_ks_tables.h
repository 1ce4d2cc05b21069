/* ORACLE (test infrastructure): hetero-nuclear Slater-Koster blocks used by the DFTB branch.
 * Entry-by-entry restatement of
 *   KS_sp3_hetero       Source_files/TB_Koster_Slater.f90:4125-4155
 *   d_KS_sp3_hetero     :4159-4189
 *   KS_sp3d5_hetero     :4406-4520
 *   d_KS_sp3d5_hetero   :4524-4639
 *   KS_s / d_KS_s       :5051-5068
 * on top of the t_* / d_t_* primitives of sk_tables.h (TB_Koster_Slater.f90:371-967).
 * A = Vsr12, B = Vsr21 (0-based channels), dA = dVsr12, dB = dVsr21; ts is (n,n) column-major.
 * The device code does NOT use these tables: it composes the block from the homo-nuclear upper
 * triangle and a parity rule (xtant-3_b200/csrc/dftb.cu), so the two are independent. */
#ifndef ORC_DFTB_KS_TABLES_H
#define ORC_DFTB_KS_TABLES_H
#include "sk_tables.h"

#define TS(a, b) ts[((a) - 1) + 4 * ((b) - 1)]
static inline void orc_KS_sp3_hetero(const double* A, const double* B, double l, double m, double n, double* ts) {
    TS(1, 1) = t_s_s(A[0]);
    TS(1, 2) = -t_s_px(l, A[1]);
    TS(1, 3) = -t_s_px(m, A[1]);
    TS(1, 4) = -t_s_px(n, A[1]);
    TS(2, 1) = t_s_px(l, B[1]);
    TS(2, 2) = t_pa_pa(l, A[2], A[3]);
    TS(2, 3) = t_pa_pb(l, m, A[2], A[3]);
    TS(2, 4) = t_pa_pb(l, n, A[2], A[3]);
    TS(3, 1) = t_s_px(m, B[1]);
    TS(3, 2) = t_pa_pb(l, m, B[2], B[3]);
    TS(3, 3) = t_pa_pa(m, A[2], A[3]);
    TS(3, 4) = t_pa_pb(m, n, A[2], A[3]);
    TS(4, 1) = t_s_px(n, B[1]);
    TS(4, 2) = t_pa_pb(l, n, B[2], B[3]);
    TS(4, 3) = t_pa_pb(m, n, B[2], B[3]);
    TS(4, 4) = t_pa_pa(n, A[2], A[3]);
}
static inline void orc_d_KS_sp3_hetero(const double* A, const double* B, const double* dA, const double* dB,
                                       double l, double m, double n, double dl, double dm, double dn, double* ts) {
    TS(1, 1) = d_t_s_s(dA[0]);
    TS(1, 2) = d_t_s_px(l, dl, A[1], dA[1]);
    TS(1, 3) = d_t_s_px(m, dm, A[1], dA[1]);
    TS(1, 4) = d_t_s_px(n, dn, A[1], dA[1]);
    TS(2, 1) = -d_t_s_px(l, dl, B[1], dB[1]);
    TS(2, 2) = d_t_pa_pa(l, dl, A[2], dA[2], A[3], dA[3]);
    TS(2, 3) = d_t_pa_pb(l, dl, m, dm, A[2], dA[2], A[3], dA[3]);
    TS(2, 4) = d_t_pa_pb(l, dl, n, dn, A[2], dA[2], A[3], dA[3]);
    TS(3, 1) = -d_t_s_px(m, dm, B[1], dB[1]);
    TS(3, 2) = d_t_pa_pb(l, dl, m, dm, B[2], dB[2], B[3], dB[3]);
    TS(3, 3) = d_t_pa_pa(m, dm, A[2], dA[2], A[3], dA[3]);
    TS(3, 4) = d_t_pa_pb(m, dm, n, dn, A[2], dA[2], A[3], dA[3]);
    TS(4, 1) = -d_t_s_px(n, dn, B[1], dB[1]);
    TS(4, 2) = d_t_pa_pb(l, dl, n, dn, B[2], dB[2], B[3], dB[3]);
    TS(4, 3) = d_t_pa_pb(m, dm, n, dn, B[2], dB[2], B[3], dB[3]);
    TS(4, 4) = d_t_pa_pa(n, dn, A[2], dA[2], A[3], dA[3]);
}
#undef TS
#define TS(a, b) ts[((a) - 1) + 9 * ((b) - 1)]
static inline void orc_KS_sp3d5_hetero(const double* A, const double* B, double l, double m, double n, double* ts) {
    TS(1, 1) = t_s_s(A[0]);
    TS(1, 2) = -t_s_px(l, A[1]);
    TS(1, 3) = -t_s_px(m, A[1]);
    TS(1, 4) = -t_s_px(n, A[1]);
    TS(1, 5) = t_s_dab(l, m, A[2]);
    TS(1, 6) = t_s_dab(l, n, A[2]);
    TS(1, 7) = t_s_dab(m, n, A[2]);
    TS(1, 8) = t_s_dx2_y2(l, m, A[2]);
    TS(1, 9) = t_s_dz2_r2(l, m, n, A[2]);
    TS(2, 1) = t_s_px(l, B[1]);
    TS(2, 2) = t_pa_pa(l, A[3], A[4]);
    TS(2, 3) = t_pa_pb(l, m, A[3], A[4]);
    TS(2, 4) = t_pa_pb(l, n, A[3], A[4]);
    TS(2, 5) = -t_px_dxy(l, m, A[5], A[6]);
    TS(2, 6) = -t_px_dxy(l, n, A[5], A[6]);
    TS(2, 7) = -t_px_dyz(l, m, n, A[5], A[6]);
    TS(2, 8) = -t_px_dx2_y2(l, m, A[5], A[6]);
    TS(2, 9) = -t_pa_d3z2_r2(l, l, m, n, A[5], A[6]);
    TS(3, 1) = t_s_px(m, B[1]);
    TS(3, 2) = t_pa_pb(l, m, B[3], B[4]);
    TS(3, 3) = t_pa_pa(m, A[3], A[4]);
    TS(3, 4) = t_pa_pb(m, n, A[3], A[4]);
    TS(3, 5) = -t_px_dxy(m, l, A[5], A[6]);
    TS(3, 6) = -t_px_dyz(m, l, n, A[5], A[6]);
    TS(3, 7) = -t_px_dxy(m, n,  A[5], A[6]);
    TS(3, 8) = -t_py_dx2_y2(l, m, A[5], A[6]);
    TS(3, 9) = -t_pa_d3z2_r2(m, l, m, n, A[5], A[6]);
    TS(4, 1) = t_s_px(n, B[1]);
    TS(4, 2) = t_pa_pb(l, n, B[3], B[4]);
    TS(4, 3) = t_pa_pb(m, n, B[3], B[4]);
    TS(4, 4) = t_pa_pa(n, A[3], A[4]);
    TS(4, 5) = -t_px_dyz(n, l, m, A[5], A[6]);
    TS(4, 6) = -t_px_dxy(n, l, A[5], A[6]);
    TS(4, 7) = -t_px_dxy(n, m, A[5], A[6]);
    TS(4, 8) = -t_pz_dx2_y2(l, m, n, A[5], A[6]);
    TS(4, 9) = -t_pz_d3z2_r2(l, m, n, A[5], A[6]);
    TS(5, 1) = t_s_dab(l, m, B[2]);
    TS(5, 2) = t_px_dxy(l, m, B[5], B[6]);
    TS(5, 3) = t_px_dxy(m, l, B[5], B[6]);
    TS(5, 4) = t_px_dyz(n, l, m, B[5], B[6]);
    TS(5, 5) = t_dab_dab(l, m, n, A[7], A[8], A[9]);
    TS(5, 6) = t_dab_dbg(m, l, n, A[7], A[8], A[9]);
    TS(5, 7) = t_dab_dbg(l, m, n, A[7], A[8], A[9]);
    TS(5, 8) = t_dxy_dx2_y2(l, m, A[7], A[8], A[9]);
    TS(5, 9) = t_dxy_d3z2_r2(l, m, n, A[7], A[8], A[9]);
    TS(6, 1) = t_s_dab(l, n, B[2]);
    TS(6, 2) = t_px_dxy(l, n, B[5], B[6]);
    TS(6, 3) = t_px_dyz(m, l, n, B[5], B[6]);
    TS(6, 4) = t_px_dxy(n, l, B[5], B[6]);
    TS(6, 5) = t_dab_dbg(m, l, n, B[7], B[8], B[9]);
    TS(6, 6) = t_dab_dab(l, n, m, A[7], A[8], A[9]);
    TS(6, 7) = t_dab_dbg(l, n, m, A[7], A[8], A[9]);
    TS(6, 8) = t_dxz_dx2_y2(l, m, n, A[7], A[8], A[9]);
    TS(6, 9) = t_dxz_d3z2_r2(l, m, n, A[7], A[8], A[9]);
    TS(7, 1) = t_s_dab(m, n, B[2]);
    TS(7, 2) = t_px_dyz(l, m, n, B[5], B[6]);
    TS(7, 3) = t_px_dxy(m, n,  B[5], B[6]);
    TS(7, 4) = t_px_dxy(n, m, B[5], B[6]);
    TS(7, 5) = t_dab_dbg(l, m, n, B[7], B[8], B[9]);
    TS(7, 6) = t_dab_dbg(l, n, m, B[7], B[8], B[9]);
    TS(7, 7) = t_dab_dab(m, n, l, A[7], A[8], A[9]);
    TS(7, 8) = t_dyz_dx2_y2(l, m, n, A[7], A[8], A[9]);
    TS(7, 9) = t_dyz_d3z2_r2(l, m, n, A[7], A[8], A[9]);
    TS(8, 1) = t_s_dx2_y2(l, m, B[2]);
    TS(8, 2) = t_px_dx2_y2(l, m, B[5], B[6]);
    TS(8, 3) = t_py_dx2_y2(l, m, B[5], B[6]);
    TS(8, 4) = t_pz_dx2_y2(l, m, n, B[5], B[6]);
    TS(8, 5) = t_dxy_dx2_y2(l, m, B[7], B[8], B[9]);
    TS(8, 6) = t_dxz_dx2_y2(l, m, n, B[7], B[8], B[9]);
    TS(8, 7) = t_dyz_dx2_y2(l, m, n, B[7], B[8], B[9]);
    TS(8, 8) = t_dx2_y2_dx2_y2(l, m, n, A[7], A[8], A[9]);
    TS(8, 9) = t_dx2_y2_d3z2_r2(l, m, n, A[7], A[8], A[9]);
    TS(9, 1) = t_s_dz2_r2(l, m, n, B[2]);
    TS(9, 2) = t_pa_d3z2_r2(l, l, m, n, B[5], B[6]);
    TS(9, 3) = t_pa_d3z2_r2(m, l, m, n, B[5], B[6]);
    TS(9, 4) = t_pz_d3z2_r2(l, m, n, B[5], B[6]);
    TS(9, 5) = t_dxy_d3z2_r2(l, m, n, B[7], B[8], B[9]);
    TS(9, 6) = t_dxz_d3z2_r2(l, m, n, B[7], B[8], B[9]);
    TS(9, 7) = t_dyz_d3z2_r2(l, m, n, B[7], B[8], B[9]);
    TS(9, 8) = t_dx2_y2_d3z2_r2(l, m, n, B[7], B[8], B[9]);
    TS(9, 9) = t_d3z2_r2_d3z2_r2(l, m, n, A[7], A[8], A[9]);
}
static inline void orc_d_KS_sp3d5_hetero(const double* A, const double* B, const double* dA, const double* dB,
                                         double l, double m, double n, double dl, double dm, double dn, double* ts) {
    TS(1, 1) = d_t_s_s(dA[0]);
    TS(1, 2) = d_t_s_px(l, dl, A[1], dA[1]);
    TS(1, 3) = d_t_s_px(m, dm, A[1], dA[1]);
    TS(1, 4) = d_t_s_px(n, dn, A[1], dA[1]);
    TS(1, 5) = d_t_s_dab(l, dl, m, dm, A[2], dA[2]);
    TS(1, 6) = d_t_s_dab(l, dl, n, dn, A[2], dA[2]);
    TS(1, 7) = d_t_s_dab(m, dm, n, dn, A[2], dA[2]);
    TS(1, 8) = d_t_s_dx2_y2(l, dl, m, dm, A[2], dA[2]);
    TS(1, 9) = d_t_s_dz2_r2(l, dl, m, dm, n, dn, A[2], dA[2]);
    TS(2, 1) = -d_t_s_px(l, dl, B[1], dB[1]);
    TS(2, 2) = d_t_pa_pa(l, dl, A[3], dA[3], A[4], dA[4]);
    TS(2, 3) = d_t_pa_pb(l, dl, m, dm, A[3], dA[3], A[4], dA[4]);
    TS(2, 4) = d_t_pa_pb(l, dl, n, dn, A[3], dA[3], A[4], dA[4]);
    TS(2, 5) = d_t_px_dxy(l, dl, m, dm, A[5], dA[5], A[6], dA[6]);
    TS(2, 6) = d_t_px_dxy(l, dl, n, dn, A[5], dA[5], A[6], dA[6]);
    TS(2, 7) = d_t_px_dyz(l, dl, m, dm, n, dn, A[5], dA[5], A[6], dA[6]);
    TS(2, 8) = d_t_px_dx2_y2(l, dl, m, dm, A[5], dA[5], A[6], dA[6]);
    TS(2, 9) = d_t_pa_d3z2_r2(l, dl, l, dl, m, dm, n, dn, A[5], dA[5], A[6], dA[6]);
    TS(3, 1) = -d_t_s_px(m, dm, B[1], dB[1]);
    TS(3, 2) = d_t_pa_pb(l, dl, m, dm, B[3], dB[3], B[4], dB[4]);
    TS(3, 3) = d_t_pa_pa(m, dm, A[3], dA[3], A[4], dA[4]);
    TS(3, 4) = d_t_pa_pb(m, dm, n, dn, A[3], dA[3], A[4], dA[4]);
    TS(3, 5) = d_t_px_dxy(m, dm, l, dl, A[5], dA[5], A[6], dA[6]);
    TS(3, 6) = d_t_px_dyz(m, dm, l, dl, n, dn, A[5], dA[5], A[6], dA[6]);
    TS(3, 7) = d_t_px_dxy(m, dm, n, dn, A[5], dA[5], A[6], dA[6]);
    TS(3, 8) = d_t_py_dx2_y2(l, dl, m, dm, A[5], dA[5], A[6], dA[6]);
    TS(3, 9) = d_t_pa_d3z2_r2(m, dm, l, dl, m, dm, n, dn, A[5], dA[5], A[6], dA[6]);
    TS(4, 1) = -d_t_s_px(n, dn, B[1], dB[1]);
    TS(4, 2) = d_t_pa_pb(l, dl, n, dn, B[3], dB[3], B[4], dB[4]);
    TS(4, 3) = d_t_pa_pb(m, dm, n, dn, B[3], dB[3], B[4], dB[4]);
    TS(4, 4) = d_t_pa_pa(n, dn, A[3], dA[3], A[4], dA[4]);
    TS(4, 5) = d_t_px_dyz(n, dn, l, dl, m, dm, A[5], dA[5], A[6], dA[6]);
    TS(4, 6) = d_t_px_dxy(n, dn, l, dl, A[5], dA[5], A[6], dA[6]);
    TS(4, 7) = d_t_px_dxy(n, dn, m, dm, A[5], dA[5], A[6], dA[6]);
    TS(4, 8) = d_t_pz_dx2_y2(l, dl, m, dm, n, dn, A[5], dA[5], A[6], dA[6]);
    TS(4, 9) = d_t_pz_d3z2_r2(l, dl, m, dm, n, dn, A[5], dA[5], A[6], dA[6]);
    TS(5, 1) = d_t_s_dab(l, dl, m, dm, B[2], dB[2]);
    TS(5, 2) = -d_t_px_dxy(l, dl, m, dm, B[5], dB[5], B[6], dB[6]);
    TS(5, 3) = -d_t_px_dxy(m, dm, l, dl, B[5], dB[5], B[6], dB[6]);
    TS(5, 4) = -d_t_px_dyz(n, dn, l, dl, m, dm, B[5], dB[5], B[6], dB[6]);
    TS(5, 5) = d_t_dab_dab(l, dl, m, dm, n, dn, A[7], dA[7], A[8], dA[8], A[9], dA[9]);
    TS(5, 6) = d_t_dab_dbg(m, dm, l, dl, n, dn, A[7], dA[7], A[8], dA[8], A[9], dA[9]);
    TS(5, 7) = d_t_dab_dbg(l, dl, m, dm, n, dn, A[7], dA[7], A[8], dA[8], A[9], dA[9]);
    TS(5, 8) = d_t_dxy_dx2_y2(l, dl, m, dm, A[7], dA[7], A[8], dA[8], A[9], dA[9]);
    TS(5, 9) = d_t_dxy_d3z2_r2(l, dl, m, dm, n, dn, A[7], dA[7], A[8], dA[8], A[9], dA[9]);
    TS(6, 1) = d_t_s_dab(l, dl, n, dn, B[2], dB[2]);
    TS(6, 2) = -d_t_px_dxy(l, dl, n, dn, B[5], dB[5], B[6], dB[6]);
    TS(6, 3) = -d_t_px_dyz(m, dm, l, dl, n, dn, B[5], dB[5], B[6], dB[6]);
    TS(6, 4) = -d_t_px_dxy(n, dn, l, dl, B[5], dB[5], B[6], dB[6]);
    TS(6, 5) = d_t_dab_dbg(m, dm, l, dl, n, dn, B[7], dB[7], B[8], dB[8], B[9], dB[9]);
    TS(6, 6) = d_t_dab_dab(l, dl, n, dn, m, dm, A[7], dA[7], A[8], dA[8], A[9], dA[9]);
    TS(6, 7) = d_t_dab_dbg(l, dl, n, dn, m, dm, A[7], dA[7], A[8], dA[8], A[9], dA[9]);
    TS(6, 8) = d_t_dxz_dx2_y2(l, dl, m, dm, n, dn, A[7], dA[7], A[8], dA[8], A[9], dA[9]);
    TS(6, 9) = d_t_dxz_d3z2_r2(l, dl, m, dm, n, dn, A[7], dA[7], A[8], dA[8], A[9], dA[9]);
    TS(7, 1) = d_t_s_dab(m, dm, n, dn, B[2], dB[2]);
    TS(7, 2) = -d_t_px_dyz(l, dl, m, dm, n, dn, B[5], dB[5], B[6], dB[6]);
    TS(7, 3) = -d_t_px_dxy(m, dm, n, dn, B[5], dB[5], B[6], dB[6]);
    TS(7, 4) = -d_t_px_dxy(n, dn, m, dm, B[5], dB[5], B[6], dB[6]);
    TS(7, 5) = d_t_dab_dbg(l, dl, m, dm, n, dn, B[7], dB[7], B[8], dB[8], B[9], dB[9]);
    TS(7, 6) = d_t_dab_dbg(l, dl, n, dn, m, dm, B[7], dB[7], B[8], dB[8], B[9], dB[9]);
    TS(7, 7) = d_t_dab_dab(m, dm, n, dn, l, dl, A[7], dA[7], A[8], dA[8], A[9], dA[9]);
    TS(7, 8) = d_t_dyz_dx2_y2(l, dl, m, dm, n, dn, A[7], dA[7], A[8], dA[8], A[9], dA[9]);
    TS(7, 9) = d_t_dyz_d3z2_r2(l, dl, m, dm, n, dn, A[7], dA[7], A[8], dA[8], A[9], dA[9]);
    TS(8, 1) = d_t_s_dx2_y2(l, dl, m, dm, B[2], dB[2]);
    TS(8, 2) = -d_t_px_dx2_y2(l, dl, m, dm, B[5], dB[5], B[6], dB[6]);
    TS(8, 3) = -d_t_py_dx2_y2(l, dl, m, dm, B[5], dB[5], B[6], dB[6]);
    TS(8, 4) = -d_t_pz_dx2_y2(l, dl, m, dm, n, dn, B[5], dB[5], B[6], dB[6]);
    TS(8, 5) = d_t_dxy_dx2_y2(l, dl, m, dm, B[7], dB[7], B[8], dB[8], B[9], dB[9]);
    TS(8, 6) = d_t_dxz_dx2_y2(l, dl, m, dm, n, dn, B[7], dB[7], B[8], dB[8], B[9], dB[9]);
    TS(8, 7) = d_t_dyz_dx2_y2(l, dl, m, dm, n, dn, B[7], dB[7], B[8], dB[8], B[9], dB[9]);
    TS(8, 8) = d_t_dx2_y2_dx2_y2(l, dl, m, dm, n, dn, A[7], dA[7], A[8], dA[8], A[9], dA[9]);
    TS(8, 9) = d_t_dx2_y2_d3z2_r2(l, dl, m, dm, n, dn, A[7], dA[7], A[8], dA[8], A[9], dA[9]);
    TS(9, 1) = d_t_s_dz2_r2(l, dl, m, dm, n, dn, B[2], dB[2]);
    TS(9, 2) = -d_t_pa_d3z2_r2(l, dl, l, dl, m, dm, n, dn, B[5], dB[5], B[6], dB[6]);
    TS(9, 3) = -d_t_pa_d3z2_r2(m, dm, l, dl, m, dm, n, dn, B[5], dB[5], B[6], dB[6]);
    TS(9, 4) = -d_t_pz_d3z2_r2(l, dl, m, dm, n, dn, B[5], dB[5], B[6], dB[6]);
    TS(9, 5) = d_t_dxy_d3z2_r2(l, dl, m, dm, n, dn, B[7], dB[7], B[8], dB[8], B[9], dB[9]);
    TS(9, 6) = d_t_dxz_d3z2_r2(l, dl, m, dm, n, dn, B[7], dB[7], B[8], dB[8], B[9], dB[9]);
    TS(9, 7) = d_t_dyz_d3z2_r2(l, dl, m, dm, n, dn, B[7], dB[7], B[8], dB[8], B[9], dB[9]);
    TS(9, 8) = d_t_dx2_y2_d3z2_r2(l, dl, m, dm, n, dn, B[7], dB[7], B[8], dB[8], B[9], dB[9]);
    TS(9, 9) = d_t_d3z2_r2_d3z2_r2(l, dl, m, dm, n, dn, A[7], dA[7], A[8], dA[8], A[9], dA[9]);
}
#undef TS
#endif
