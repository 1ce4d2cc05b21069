/* ORACLE (test infrastructure; included by xtant_oracle.c) -- short-range "exponential wall", Fermi occupations at fixed
 * electron temperature and the velocity-Verlet halves, restated statement by statement from the reference:
 *   Source_files/Exponential_wall.f90  get_Exp_wall_s :990-1067, Exp_wall_pot :1070-1092, Wall_potential :1095-1105,
 *                                      d_Wall_potential :1108-1119, d_Exp_wall_pot :1140-1166, d_Exp_wall_pot_s :1200-1329
 *                                      (OpenMP arm :1270-1322), d_Exp_wall_Pressure_s :1334-1385
 *   Source_files/Coulomb.f90           f_cut_L_C :610-614, d_f_cut_L_C :617-629
 *   Source_files/Electron_tools.f90    Electron_Fixed_Te :2660-2706, get_N_tot :2758-2801, Fermi_distribution :1896-1909,
 *                                      set_Fermi :1913-1969
 *   Source_files/Atomic_tools.f90      make_time_step_atoms_SC :1897-1974, get_GGdot :1843-1861, save_last_timestep :1177-1192,
 *                                      check_boundaries :2639-2674 / update_atom_crossing_boundary (periodic arm) :2690-2693
 * PARITY UNPINNED, like the rest of oracle/ (see the header of xtant_oracle.c). */

static double wall_f_cut_L_C(double a_r, double r_L, double d_L) { return 1.0 / (1.0 + exp((a_r - r_L) / d_L)); }
static double wall_d_f_cut_L_C(double a_r, double r_L, double d_L) {
    const double arg = (a_r - r_L) / d_L;
    if (arg >= log(DBL_MAX)) return 0.0;
    const double exp_r = exp(arg), exp_r2 = 1.0 + exp_r;
    return -exp_r / (d_L * exp_r2 * exp_r2);
}
static double wall_Wall_potential(double C, double r0, double r) {
    const double arg = r - r0;
    if (fabs(arg) <= log(DBL_MIN)) return C >= 0 ? 1.0e25 : -1.0e25;      /* as written: log(TINY) < 0, never taken */
    return C * exp(1.0 / arg);
}
static double wall_d_Wall_potential(double C, double r0, double r) {
    const double arg = r - r0;
    if (fabs(arg) <= log(DBL_MIN)) return 0.0;
    const double rr0 = 1.0 / arg;
    return -C * exp(rr0) * rr0 * rr0;
}
static double wall_Exp_wall_pot(const xtb_wall_pair* W, double a_r) {
    if (a_r > W->d0 + W->dd * 10.0) return 0.0;
    const double f_cut_large = wall_f_cut_L_C(a_r, W->d0, W->dd);
    return wall_Wall_potential(W->C, W->r0, a_r) * f_cut_large;
}
static double wall_d_Exp_wall_pot(const xtb_wall_pair* W, double a_r) {
    if (a_r > W->d0 + W->dd * 10.0) return 0.0;
    const double E_C = wall_Wall_potential(W->C, W->r0, a_r);
    const double d_E_C = wall_d_Wall_potential(W->C, W->r0, a_r);
    const double f_cut_large = wall_f_cut_L_C(a_r, W->d0, W->dd);
    const double d_f_large = wall_d_f_cut_L_C(a_r, W->d0, W->dd);
    return d_E_C * f_cut_large + E_C * d_f_large;
}

/* get_Exp_wall_s: returns a = sum_a * 0.5; Epot[i] = E_pot * 0.5 (the amount added to MDAtoms(i)%Epot) */
double orc_wall_energy(int na, int cap, int n_kinds, const int* koa, const int* nn_size, const int* nn_list,
                       const double* nn_dist, const xtb_wall_pair* WP, double* Epot) {
    double sum_a = 0.0;
    for (int i = 0; i < na; ++i) {
        double E_pot = 0.0;
        for (int a2 = 0; a2 < nn_size[i]; ++a2) {
            const size_t e = (size_t)i * cap + a2;
            const int j = nn_list[e];
            const double Expwall_pot = wall_Exp_wall_pot(WP + koa[i] + n_kinds * koa[j], nn_dist[4 * e + 3]);
            sum_a = sum_a + Expwall_pot;
            E_pot = E_pot + Expwall_pot;
        }
        if (Epot) Epot[i] = E_pot * 0.5;
    }
    return sum_a * 0.5;
}

/* d_Exp_wall_pot_s, verbatim loop nest (every ian, every i1, whole list of i1).  F[3*ian + al] = Erx_s(al,ian) * 0.5 */
void orc_wall_forces(int na, int cap, int n_kinds, const int* koa, const double* h, const int* nn_size, const int* nn_list,
                     const double* nn_dist, const xtb_wall_pair* WP, double* F) {
#pragma omp parallel for schedule(dynamic, 4)
    for (int ian = 0; ian < na; ++ian) {
        double Erx[3] = {0.0, 0.0, 0.0};
        for (int i1 = 0; i1 < na; ++i1) {
            const int dik = ian == i1;
            double dpsi[3] = {0.0, 0.0, 0.0};
            for (int a2 = 0; a2 < nn_size[i1]; ++a2) {
                const size_t e = (size_t)i1 * cap + a2;
                const int j1 = nn_list[e];
                const int djk = ian == j1;
                if ((dik - djk) != 0) {
                    const double x = nn_dist[4 * e], y = nn_dist[4 * e + 1], z = nn_dist[4 * e + 2];
                    double x1[3];
                    x1[0] = x * H_(1, 1) + y * H_(1, 2) + z * H_(1, 3);
                    x1[1] = x * H_(2, 1) + y * H_(2, 2) + z * H_(2, 3);
                    x1[2] = x * H_(3, 1) + y * H_(3, 2) + z * H_(3, 3);
                    const double a_r = nn_dist[4 * e + 3];
                    const double b = wall_d_Exp_wall_pot(WP + koa[i1] + n_kinds * koa[j1], a_r);
                    const double ddlta = (double)(dik - djk) / a_r;
                    const double b_delta = b * ddlta;
                    dpsi[0] = dpsi[0] + b_delta * x1[0];
                    dpsi[1] = dpsi[1] + b_delta * x1[1];
                    dpsi[2] = dpsi[2] + b_delta * x1[2];
                }
            }
            Erx[0] = Erx[0] + dpsi[0]; Erx[1] = Erx[1] + dpsi[1]; Erx[2] = Erx[2] + dpsi[2];
        }
        F[3 * ian] = Erx[0] * 0.5; F[3 * ian + 1] = Erx[1] * 0.5; F[3 * ian + 2] = Erx[2] * 0.5;
    }
}

/* d_Exp_wall_Pressure_s.  P[l + 3*k] = PForce(l+1, k+1) */
void orc_wall_pressure(int na, int cap, int n_kinds, const int* koa, const int* nn_size, const int* nn_list,
                       const double* nn_dist, const double* nn_dist_s, const xtb_wall_pair* WP, double* P) {
    for (int e9 = 0; e9 < 9; ++e9) P[e9] = 0.0;
    for (int i = 0; i < na; ++i) {
        double Rep_Pr[9] = {0};
        for (int a2 = 0; a2 < nn_size[i]; ++a2) {
            const size_t e = (size_t)i * cap + a2;
            const int j = nn_list[e];
            const double r = nn_dist[4 * e + 3];
            const double dpsy = wall_d_Exp_wall_pot(WP + koa[i] + n_kinds * koa[j], r);
            for (int k = 0; k < 3; ++k)
                for (int l = 0; l < 3; ++l)
                    Rep_Pr[l + 3 * k] = Rep_Pr[l + 3 * k] + dpsy * nn_dist[4 * e + k] * nn_dist_s[3 * e + l] / r;
        }
        for (int e9 = 0; e9 < 9; ++e9) P[e9] = P[e9] + Rep_Pr[e9] * 0.5;
    }
}

/* ---- Fermi occupations ---- */
static double fermi_get_N_tot(int n, const double* wrD, double mu, double Te) {
    double N = 0.0;
    if (Te > 1.0e-12) {
        for (int i = 0; i < n; ++i)
            if ((wrD[i] - mu) / Te < log(DBL_MAX)) N = N + 2.0 / (1.0 + exp((wrD[i] - mu) / Te));
    } else {
        int cnt = 0;
        for (int i = 0; i < n; ++i) cnt += wrD[i] <= mu;
        N = 2.0 * cnt;
    }
    return N;
}
/* Electron_Fixed_Te + set_Fermi; returns mu */
double orc_fermi(int n, const double* wrD, double Netot, double Te, double* fe) {
    double a = wrD[0] - 50.0, b = wrD[n - 1] + 50.0, Ncur = 0.0, mu = 0.0;
    while (fabs(Ncur - Netot) > Netot * 1e-12) {
        mu = (a + b) / 2.0;
        Ncur = fermi_get_N_tot(n, wrD, mu, Te);
        if (Ncur > Netot) b = mu; else a = mu;
        if (fabs(a - b) < 1e-12) break;
    }
    for (int i = 0; i < n; ++i) fe[i] = 0.0;
    if (Te > 0.0) {
        for (int i = 0; i < n; ++i) {
            const double arg = (wrD[i] - mu) / Te;
            if (arg <= log(DBL_MAX)) fe[i] = 2.0 / (1.0 + exp(arg));
        }
    } else {
        for (int i = 0; i < n; ++i) fe[i] = wrD[i] <= mu ? 2.0 : 0.0;
    }
    return mu;
}

/* ---- velocity-Verlet half: save_last_timestep, then make_time_step_atoms_SC(ind) with the periodic boundary arm ---- */
static void orc_m3_mul(const double* A, const double* B, double* Cm) {      /* col-major C = A B (Two_Matr_mult) */
    for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) {
        double s = 0.0;
        for (int k = 0; k < 3; ++k) s += A[r + 3 * k] * B[k + 3 * c];
        Cm[r + 3 * c] = s;
    }
}
static void orc_m3_tr(const double* A, double* T) { for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) T[r + 3 * c] = A[c + 3 * r]; }
static void orc_m3_inv(const double* m, double* o) {
    const double a = m[0], b = m[3], c = m[6], d = m[1], e = m[4], f = m[7], g = m[2], h = m[5], i = m[8];
    const double det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
    o[0] = (e * i - f * h) / det; o[3] = (c * h - b * i) / det; o[6] = (b * f - c * e) / det;
    o[1] = (f * g - d * i) / det; o[4] = (a * i - c * g) / det; o[7] = (c * d - a * f) / det;
    o[2] = (d * h - e * g) / det; o[5] = (b * g - a * h) / det; o[8] = (a * e - b * d) / det;
}
void orc_md_half_step(int na, int ind, const int* koa, const double* mass_kg, const double* supce, const double* Vsupce,
                      double dt, const double* Ftot, double* S, double* S0, double* SV, double* SV0) {
    double dsupce[9], tempM[9], gg2[9], GinvGdot[9] = {0};
    orc_m3_tr(supce, dsupce); orc_m3_mul(dsupce, supce, tempM); orc_m3_inv(tempM, gg2);
    if (Vsupce) {
        double TV[9], Gdot1[9], Gdot2[9], Gdot[9];
        orc_m3_tr(Vsupce, TV); orc_m3_mul(TV, supce, Gdot1); orc_m3_mul(dsupce, Vsupce, Gdot2);
        for (int x = 0; x < 9; ++x) Gdot[x] = Gdot2[x] + Gdot1[x];
        orc_m3_mul(gg2, Gdot, GinvGdot);
    }
    const double dtsqare = dt * dt / 2.0, halfdt = dt / 2.0;
    for (int k = 0; k < na; ++k) {
        for (int x = 0; x < 3; ++x) { S0[3 * k + x] = S[3 * k + x]; SV0[3 * k + x] = SV[3 * k + x]; }
        double x0[3], tempV[3], Fors_s[3];
        for (int x = 0; x < 3; ++x) x0[x] = gg2[x] * Ftot[3 * k] + gg2[x + 3] * Ftot[3 * k + 1] + gg2[x + 6] * Ftot[3 * k + 2];
        const double fctr = -1.0 / mass_kg[koa[k]] * 1.602176634e-19 * 1e-10;
        for (int x = 0; x < 3; ++x) tempV[x] = GinvGdot[x] * SV0[3 * k] + GinvGdot[x + 3] * SV0[3 * k + 1] + GinvGdot[x + 6] * SV0[3 * k + 2];
        for (int x = 0; x < 3; ++x) Fors_s[x] = fctr * x0[x] - tempV[x];
        for (int x = 0; x < 3; ++x) {
            if (ind == 2) S[3 * k + x] = S0[3 * k + x] + dt * SV0[3 * k + x] + Fors_s[x] * dtsqare;
            SV[3 * k + x] = SV0[3 * k + x] + Fors_s[x] * halfdt;
        }
    }
    if (ind == 2)
        for (int k = 0; k < na; ++k)
            for (int x = 0; x < 3; ++x)
                if (S[3 * k + x] > 1.0 || S[3 * k + x] < -0.0) {
                    S[3 * k + x] = S[3 * k + x] - floor(S[3 * k + x]);
                    S0[3 * k + x] = S0[3 * k + x] - floor(S[3 * k + x]);      /* as written: FLOOR of the already wrapped S */
                }
}
