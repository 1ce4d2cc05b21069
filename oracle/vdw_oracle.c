/* ORACLE (test infrastructure; included by xtant_oracle.c) -- van der Waals pair term (class TB_vdW) and Coulomb term of a
 * charged supercell (class TB_Coulomb_cut), restated statement by statement from the reference:
 *   Source_files/Van_der_Waals.f90   get_mirror_cell_num :132-157, get_vdW_s :158-250 (OpenMP arm :223-246), Construct_B :251-284,
 *                                    cell_cycle_B :287-421 (OpenMP arm :384-420), vdW_energy :578-604, dvdW :608-627,
 *                                    double_truncated_LJ :650-660, d_double_truncated_LJ :663-675, ILJ :718-727, d_ILJ :730-745,
 *                                    General_LJ_potential :839-850, d_General_LJ_potential :853-864, Improved_LJ_potential :882-893,
 *                                    d_Improved_LJ_potential :896-907, large/short_range_cutoff :924-960,
 *                                    vdW_Girifalco :933-991, d_vdW_Girifalco :995-1057, LJ_potential :1116-1123, d_LJ_potential :1126-1134
 *   Source_files/TB.f90              vdW_forces :3599-3633, d_Forces_s :3638-3781 (OpenMP arm :3727-3778), d_Forces_Pressure :3785-3848
 *   Source_files/Atomic_tools.f90    get_near_neighbours (vdW arm) :636-641, get_number_of_image_cells :654-702,
 *                                    distance_to_given_cell_0 :4135-4166
 *   Source_files/Coulomb.f90         get_mirror_cell_num_C :435-452, get_Coulomb_s :454-550 (OpenMP arm :526-547: a = sum_a, NOT halved),
 *                                    Coulomb_pot :552-587, Coulomb_potential :589-594, d_Coulomb_potential :596-601, f_cut_L_C :610-614,
 *                                    d_f_cut_L_C :617-629 (both restated in wall_oracle.c), dCoulomb :641-672, Construct_B_C :715-880
 *                                    (zero arrays for Q <= 1e-12 :762; XijSupce contracted with the TRANSPOSE of supce :864-866);
 *                                    TB.f90 Coulomb_forces :3560-3596 -> the same d_Forces_s / d_Forces_Pressure
 *   Source_files/Little_subroutines.f90   Fermi_function :349-357, d_Fermi_function :360-381, fast_pow :436-516, count_3d
 * The reference has no test, fixture or golden vector for this term and there is no Fortran compiler here: PARITY UNPINNED, like
 * the rest of oracle/ (see the header of xtant_oracle.c).  What pins it instead (tests/test_vdw_cpu.py): the shipped Girifalco
 * files join their polynomial arms to the Lennard-Jones arm continuously (value and slope) only if the members are read in the
 * right order into the right formula; forces and supercell force are the exact derivatives of the energy (finite differences). */

static double vdw_Fermi_function(double rcut, double d, double r) {
    if (r >= rcut + 10.0 * d) return 0.0;
    return 1.0 / (1.0 + exp((r - rcut) / d));
}
static double vdw_d_Fermi_function(double rcut, double d, double r) {
    if (r >= rcut + 10.0 * d) return 0.0;
    const double exprd = exp((r - rcut) / d), exprd1 = 1.0 + exprd;
    return -exprd / (d * exprd1 * exprd1);
}
static double vdw_fast_pow6(double x) { const double x2 = x * x, x3 = x2 * x; return x3 * x3; }
static double vdw_LJ_potential(double C12, double C6, double r) {
    const double a_r6 = vdw_fast_pow6(r), a_r12 = a_r6 * a_r6;
    return C12 / a_r12 - C6 / a_r6;
}
static double vdw_d_LJ_potential(double C12, double C6, double r) {
    const double a_r6 = vdw_fast_pow6(r), a_r7 = a_r6 * r, a_r13 = a_r6 * a_r7;
    return -12.0 * C12 / a_r13 + 6.0 * C6 / a_r7;
}
static double vdw_Girifalco(const xtb_vdw_pair* W, double a_r) {
    if (a_r > W->d_cut || a_r < W->ds_cut) return 0.0;
    if (a_r > W->dm) { const double a_r2 = a_r * a_r; return W->a * a_r2 * a_r + W->b * a_r2 + W->c * a_r + W->d; }
    if (a_r < W->dsm) {
        const double a_r2 = a_r * a_r, a_r4 = a_r2 * a_r2;
        return W->as * a_r4 * a_r + W->bs * a_r4 + W->cs * a_r2 * a_r + W->ds * a_r2 + W->es * a_r + W->fs;
    }
    return vdw_LJ_potential(W->C12, W->C6, a_r);
}
static double vdw_d_Girifalco(const xtb_vdw_pair* W, double a_r) {
    if (a_r > W->d_cut || a_r < W->ds_cut) return 0.0;
    if (a_r > W->dm) return 3.0 * W->a * a_r * a_r + 2.0 * W->b * a_r + W->c;
    if (a_r < W->dsm) {
        const double a_r2 = a_r * a_r, a_r4 = a_r2 * a_r2;
        return 5.0 * W->as * a_r4 + 4.0 * W->bs * a_r2 * a_r + 3.0 * W->cs * a_r2 + 2.0 * W->ds * a_r + W->es;
    }
    return vdw_d_LJ_potential(W->C12, W->C6, a_r);
}
static double vdw_General_LJ(double eps, double r0, double n, double r) {
    if (!(r > 0.0)) return 0.0;
    const double r0r = r0 / r, r0rn = pow(r0r, n);
    return eps * (r0rn - 2.0) * r0rn;
}
static double vdw_d_General_LJ(double eps, double r0, double n, double r) {
    if (!(r > 0.0)) return 0.0;
    const double r0r = r0 / r, r0rn = pow(r0r, n);
    return -eps * 2.0 * n / r * r0rn * (r0rn - 1.0);
}
static double vdw_Improved_LJ(double eps, double r0, double n, double m, double r) {
    if (!(r > 0.0)) return 0.0;
    const double r0r = r0 / r, n_m = n - m;
    return eps * (m / n_m * pow(r0r, n) - n / n_m * pow(r0r, m));
}
static double vdw_d_Improved_LJ(double eps, double r0, double n, double m, double r) {
    if (!(r > 0.0)) return 0.0;
    const double r0r = r0 / r, n_m = n - m;
    return -eps * (pow(r0r, n) - pow(r0r, m)) * n * m / (n_m * r);
}
static double vdw_truncate(const xtb_vdw_pair* W, double V_LJ, double a_r) {
    const double f_long = vdw_Fermi_function(W->d0_cut, W->dd_cut, a_r);
    const double f_short = 1.0 - vdw_Fermi_function(W->d0_short, W->dd_short, a_r);
    return V_LJ * f_short * f_long;
}
static double vdw_d_truncate(const xtb_vdw_pair* W, double V_LJ, double d_V_LJ, double a_r) {
    const double f_long = vdw_Fermi_function(W->d0_cut, W->dd_cut, a_r);
    const double d_f_long = vdw_d_Fermi_function(W->d0_cut, W->dd_cut, a_r);
    const double f_short = 1.0 - vdw_Fermi_function(W->d0_short, W->dd_short, a_r);
    const double d_f_short = -vdw_d_Fermi_function(W->d0_short, W->dd_short, a_r);
    return V_LJ * (d_f_short * f_long + f_short * d_f_long) + d_V_LJ * (f_short * f_long);
}
/* class TB_Coulomb_cut rides on the same record: eps = k*Q*Q (Coulomb_potential's k*Q1*Q2), d0_cut = dm, dd_cut = dd */
#define ORC_KIND_COULOMB 4
static double coul_Coulomb_pot(const xtb_vdw_pair* W, double a_r) {
    if (a_r > W->d0_cut + W->dd_cut * 10.0) return 0.0;
    const double f_cut_large = wall_f_cut_L_C(a_r, W->d0_cut, W->dd_cut);
    return W->eps / a_r * f_cut_large;
}
static double coul_dCoulomb(const xtb_vdw_pair* W, double a_r) {
    if (a_r > W->d0_cut + W->dd_cut * 10.0) return 0.0;
    const double f_cut_large = wall_f_cut_L_C(a_r, W->d0_cut, W->dd_cut);
    const double E_C = W->eps / a_r;
    const double d_E_C = -W->eps / (a_r * a_r);
    const double d_f_large = wall_d_f_cut_L_C(a_r, W->d0_cut, W->dd_cut);
    return d_E_C * f_cut_large + E_C * d_f_large;
}

/* vdW_energy: `select type`; the caller passes TB_Waals(KOAi, KOAj) */
static double vdw_energy_one(int kind, const xtb_vdw_pair* W, double a_r) {
    switch (kind) {
        case ORC_KIND_COULOMB:  return coul_Coulomb_pot(W, a_r);
        case XTB_VDW_GIRIFALCO: return vdw_Girifalco(W, a_r);
        case XTB_VDW_LJ_CUT:    return vdw_truncate(W, vdw_General_LJ(W->eps, W->r0, W->n, a_r), a_r);
        case XTB_VDW_ILJ_CUT:   return vdw_truncate(W, vdw_Improved_LJ(W->eps, W->r0, W->n, W->m, a_r), a_r);
    }
    return 0.0;
}
/* dvdW: the caller passes TB_Waals(KOA1, KOA2) */
static double vdw_dvdW(int kind, const xtb_vdw_pair* W, double a_r) {
    switch (kind) {
        case ORC_KIND_COULOMB:  return coul_dCoulomb(W, a_r);
        case XTB_VDW_GIRIFALCO: return vdw_d_Girifalco(W, a_r);
        case XTB_VDW_LJ_CUT:    return vdw_d_truncate(W, vdw_General_LJ(W->eps, W->r0, W->n, a_r), vdw_d_General_LJ(W->eps, W->r0, W->n, a_r), a_r);
        case XTB_VDW_ILJ_CUT:   return vdw_d_truncate(W, vdw_Improved_LJ(W->eps, W->r0, W->n, W->m, a_r),
                                                      vdw_d_Improved_LJ(W->eps, W->r0, W->n, W->m, a_r), a_r);
    }
    return 0.0;
}

/* distance_to_given_cell_0; supce(a,b) = h[a + 3 b], S(ik) of atom i = S[3 i + ik] */
static double vdw_distance_to_cell(const double* S, const double* h, const double* zb, int i1, int j1, double* x, double* s) {
    double x0 = 0.0, y0 = 0.0, z0 = 0.0;
    for (int ik = 0; ik < 3; ++ik) {
        x0 = x0 + (S[3 * i1 + ik] - S[3 * j1 + ik] + zb[ik]) * h[ik];
        y0 = y0 + (S[3 * i1 + ik] - S[3 * j1 + ik] + zb[ik]) * h[ik + 3];
        z0 = z0 + (S[3 * i1 + ik] - S[3 * j1 + ik] + zb[ik]) * h[ik + 6];
    }
    if (x) { x[0] = x0; x[1] = y0; x[2] = z0; }
    if (s) for (int ik = 0; ik < 3; ++ik) s[ik] = S[3 * i1 + ik] - S[3 * j1 + ik] + zb[ik];
    return sqrt(x0 * x0 + y0 * y0 + z0 * z0);
}

/* get_mirror_cell_num; rm_tb = the cut-off get_near_neighbours derives from the TB part */
void orc_vdw_cells(int n_kinds, const xtb_vdw_pair* WP, double rm_tb, const double* S, const double* h, int* N, double* R_cut_out) {
    double r = WP[0].d0_cut, d = WP[0].dd_cut, rm = rm_tb;
    for (int q = 1; q < n_kinds * n_kinds; ++q) {
        if (WP[q].d0_cut > r) r = WP[q].d0_cut;
        if (WP[q].dd_cut > d) d = WP[q].dd_cut;
    }
    const double X = r + 10.0 * d;
    if (rm < X) rm = X;
    if (R_cut_out) *R_cut_out = rm;
    for (int a = 0; a < 3; ++a) {
        double zb[3] = {0.0, 0.0, 0.0}, R;
        int n = 0;
        zb[a] = 1.0;
        R = vdw_distance_to_cell(S, h, zb, 0, 0, NULL, NULL);
        while (R < 2.0 * rm && n < 1000) {
            n = n + 1;
            zb[a] = (double)n;
            R = vdw_distance_to_cell(S, h, zb, 0, 0, NULL, NULL);
        }
        N[a] = n < 1 ? 1 : n;
    }
}

/* get_mirror_cell_num_C: R_cut = TB_Coul(1,1)%dm */
void orc_coulomb_cells(double dm11, const double* S, const double* h, int* N) {
    for (int a = 0; a < 3; ++a) {
        double zb[3] = {0.0, 0.0, 0.0}, R;
        int n = 0;
        zb[a] = 1.0;
        R = vdw_distance_to_cell(S, h, zb, 0, 0, NULL, NULL);
        while (R < 2.0 * dm11 && n < 1000) {
            n = n + 1;
            zb[a] = (double)n;
            R = vdw_distance_to_cell(S, h, zb, 0, 0, NULL, NULL);
        }
        N[a] = n < 1 ? 1 : n;
    }
}

/* get_vdW_s (OpenMP arm): returns a = sum_a * 0.5; Epot[i] = the amount added to MDAtoms(i)%Epot.
 * kind = ORC_KIND_COULOMB: get_Coulomb_s, the same loops without either half (a = sum_a, Epot(i) += E_rep) */
double orc_vdw_energy(int na, int kind, int n_kinds, const int* koa, const double* S, const double* h, const xtb_vdw_pair* WP,
                      const int* N, double* Epot) {
    double sum_a = 0.0;
    const double half = kind == ORC_KIND_COULOMB ? 1.0 : 0.5;
    if (Epot) for (int i = 0; i < na; ++i) Epot[i] = 0.0;
    for (int x_cell = -N[0]; x_cell <= N[0]; ++x_cell)
        for (int y_cell = -N[1]; y_cell <= N[1]; ++y_cell)
            for (int z_cell = -N[2]; z_cell <= N[2]; ++z_cell) {
                const double zb[3] = {(double)x_cell, (double)y_cell, (double)z_cell};
                const int origin_cell = x_cell == 0 && y_cell == 0 && z_cell == 0;
                for (int i1 = 0; i1 < na; ++i1) {
                    double E_pot = 0.0;
                    for (int j1 = 0; j1 < na; ++j1)
                        if (j1 != i1 || !origin_cell) {
                            const double a_r = vdw_distance_to_cell(S, h, zb, i1, j1, NULL, NULL);
                            const double E_pot_one = vdw_energy_one(kind, WP + koa[i1] + n_kinds * koa[j1], a_r);
                            sum_a = sum_a + E_pot_one;
                            E_pot = E_pot + E_pot_one;
                        }
                    if (Epot) Epot[i1] = Epot[i1] + E_pot * half;
                }
            }
    return sum_a * half;
}

/* one (cell, i1, j1) entry of the arrays cell_cycle_B fills (:394-408) */
typedef struct { double Bij, A_rij, X[3], XS[3], SX[3]; } vdw_entry;
static void vdw_B_entry(vdw_entry* e, int kind, int n_kinds, const int* koa, const double* S, const double* h, const xtb_vdw_pair* WP,
                        const double* zb, int i1, int j1) {
    e->A_rij = vdw_distance_to_cell(S, h, zb, i1, j1, e->X, e->SX);
    if (kind == ORC_KIND_COULOMB)     /* Construct_B_C :864-866: x*supce(1,k) + y*supce(2,k) + z*supce(3,k) */
        for (int k = 0; k < 3; ++k) e->XS[k] = e->X[0] * h[3 * k] + e->X[1] * h[3 * k + 1] + e->X[2] * h[3 * k + 2];
    else
        for (int k = 0; k < 3; ++k) e->XS[k] = e->X[0] * h[k] + e->X[1] * h[k + 3] + e->X[2] * h[k + 6];  /* supce(k,1..3) */
    e->Bij = vdw_dvdW(kind, WP + koa[j1] + n_kinds * koa[i1], e->A_rij);                                 /* TB_Waals(KOA(j1), KOA(i1)) */
}

/* Construct_B: the (cells, n, n) arrays the reference allocates; index (cell, i1, j1) -> cell + ncell * (i1 + n * j1) */
typedef struct { int ncell, n, N[3]; double *Bij, *A_rij, *XS[3]; } vdw_B;
static int vdw_count_3d(const int* N, int x, int y, int z) { return ((x + N[0]) * (2 * N[1] + 1) + (y + N[1])) * (2 * N[2] + 1) + (z + N[2]); }
static int vdw_Construct_B(vdw_B* B, int na, int kind, int n_kinds, const int* koa, const double* S, const double* h,
                           const xtb_vdw_pair* WP, const int* N) {
    B->n = na; B->ncell = (2 * N[0] + 1) * (2 * N[1] + 1) * (2 * N[2] + 1);
    memcpy(B->N, N, sizeof B->N);
    const size_t cnt = (size_t)B->ncell * na * na;
    double* all = (double*)calloc(cnt * 5, sizeof(double));      /* the arrays d_Forces_s reads (of the eleven) */
    if (!all) return 1;
    B->Bij = all; B->A_rij = all + cnt;
    for (int k = 0; k < 3; ++k) B->XS[k] = all + cnt * (2 + k);
    for (size_t q = 0; q < cnt; ++q) B->A_rij[q] = 1.0e30;
    for (int x_cell = -N[0]; x_cell <= N[0]; ++x_cell)
        for (int y_cell = -N[1]; y_cell <= N[1]; ++y_cell)
            for (int z_cell = -N[2]; z_cell <= N[2]; ++z_cell) {
                const int coun_cell = vdw_count_3d(N, x_cell, y_cell, z_cell);
                const double zb[3] = {(double)x_cell, (double)y_cell, (double)z_cell};
                const int origin_cell = x_cell == 0 && y_cell == 0 && z_cell == 0;
                for (int i1 = 0; i1 < na; ++i1)
                    for (int j1 = 0; j1 < na; ++j1)
                        if (j1 != i1 || !origin_cell) {
                            const size_t e = coun_cell + (size_t)B->ncell * (i1 + (size_t)na * j1);
                            vdw_entry t;
                            vdw_B_entry(&t, kind, n_kinds, koa, S, h, WP, zb, i1, j1);
                            B->A_rij[e] = t.A_rij; B->Bij[e] = t.Bij;
                            for (int k = 0; k < 3; ++k) B->XS[k][e] = t.XS[k];
                        }
            }
    return 0;
}

/* vdW_forces -> d_Forces_s with (XijSupce, YijSupce, ZijSupce): F(3,na) = the term added to forces%rep.
 * literal != 0: the reference's arrays and loop nest as written, O(Na^3 cells) time and O(Na^2 cells) memory.
 * literal == 0: the same sums in the same order with the (i1, j1) pairs that multiply (dik - djk) = 0 left out -- they add
 * +-0.0 to a partial sum, which cannot change it -- and the array entries computed where they are used: O(Na^2 cells), no arrays. */
int orc_vdw_forces(int na, int kind, int n_kinds, const int* koa, const double* S, const double* h, const xtb_vdw_pair* WP,
                   const int* N, int literal, double* F) {
    vdw_B B;
    B.Bij = NULL;
    if (literal && vdw_Construct_B(&B, na, kind, n_kinds, koa, S, h, WP, N)) return 1;
#pragma omp parallel for schedule(dynamic, 4) if (!literal)
    for (int ian = 0; ian < na; ++ian) {
        double Erx_s[3] = {0.0, 0.0, 0.0};
        for (int i1 = 0; i1 < na; ++i1) {
            const int dik = ian == i1;
            double dpsi[3] = {0.0, 0.0, 0.0};
            for (int j1 = 0; j1 < na; ++j1) {
                const int djk = ian == j1;
                if (!literal && dik == djk) continue;
                for (int x_cell = -N[0]; x_cell <= N[0]; ++x_cell)
                    for (int y_cell = -N[1]; y_cell <= N[1]; ++y_cell)
                        for (int z_cell = -N[2]; z_cell <= N[2]; ++z_cell) {
                            const int origin_cell = x_cell == 0 && y_cell == 0 && z_cell == 0;
                            if (j1 != i1 || !origin_cell) {
                                double a_r, b, x1[3];
                                if (literal) {
                                    const size_t e = vdw_count_3d(N, x_cell, y_cell, z_cell) + (size_t)B.ncell * (i1 + (size_t)na * j1);
                                    a_r = B.A_rij[e]; b = B.Bij[e];
                                    for (int k = 0; k < 3; ++k) x1[k] = B.XS[k][e];
                                } else {
                                    const double zb[3] = {(double)x_cell, (double)y_cell, (double)z_cell};
                                    vdw_entry t;
                                    vdw_B_entry(&t, kind, n_kinds, koa, S, h, WP, zb, i1, j1);
                                    a_r = t.A_rij; b = t.Bij;
                                    for (int k = 0; k < 3; ++k) x1[k] = t.XS[k];
                                }
                                const double ddlta = (double)(dik - djk) / a_r;
                                const double b_delta = b * ddlta;
                                for (int k = 0; k < 3; ++k) dpsi[k] = dpsi[k] + b_delta * x1[k];
                            }
                        }
            }
            for (int k = 0; k < 3; ++k) Erx_s[k] = Erx_s[k] + dpsi[k];
        }
        for (int k = 0; k < 3; ++k) F[3 * ian + k] = Erx_s[k] * 0.5;
    }
    if (literal) free(B.Bij);
    return 0;
}

/* vdW_forces -> d_Forces_Pressure: P(l + 3 k) = the term added to SCforce%rep(l,k); array entries computed where they are used */
int orc_vdw_pressure(int na, int kind, int n_kinds, const int* koa, const double* S, const double* h, const xtb_vdw_pair* WP,
                     const int* N, double* P) {
    double PForce[9] = {0};
    for (int x_cell = -N[0]; x_cell <= N[0]; ++x_cell)
        for (int y_cell = -N[1]; y_cell <= N[1]; ++y_cell)
            for (int z_cell = -N[2]; z_cell <= N[2]; ++z_cell) {
                const int origin_cell = x_cell == 0 && y_cell == 0 && z_cell == 0;
                const double zb[3] = {(double)x_cell, (double)y_cell, (double)z_cell};
                for (int i = 0; i < na; ++i) {
                    double Rep_Pr[9] = {0};
                    for (int j = 0; j < na; ++j)
                        if (j != i || !origin_cell) {
                            vdw_entry t;
                            vdw_B_entry(&t, kind, n_kinds, koa, S, h, WP, zb, i, j);
                            const double r = t.A_rij, dpsy = t.Bij;
                            for (int k = 0; k < 3; ++k)
                                for (int l = 0; l < 3; ++l) Rep_Pr[l + 3 * k] = Rep_Pr[l + 3 * k] + dpsy * t.X[k] * t.SX[l] / r;
                        }
                    for (int k = 0; k < 3; ++k)
                        for (int l = 0; l < 3; ++l) PForce[l + 3 * k] = PForce[l + 3 * k] + Rep_Pr[l + 3 * k] * 0.5;
                }
            }
    memcpy(P, PForce, sizeof PForce);
    return 0;
}

/* single-pair probes for the continuity pins */
double orc_vdw_V(int kind, const xtb_vdw_pair* W, double r) { return vdw_energy_one(kind, W, r); }
double orc_vdw_dV(int kind, const xtb_vdw_pair* W, double r) { return vdw_dvdW(kind, W, r); }
