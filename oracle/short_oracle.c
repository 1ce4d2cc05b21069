/* ORACLE (test infrastructure; included by xtant_oracle.c) -- the 'General' short-range class (type TB_Short_Rep, Objects.f90:437-454),
 * restated statement by statement from the reference:
 *   Source_files/Exponential_wall.f90   get_short_range_rep_s :48-127 (OpenMP arm :100-122), Shortrange_pot :130-157,
 *                                       d_Short_range_pot :160-203, d_Short_range_pot_s :519-659 (OpenMP arm :600-650),
 *                                       d_Short_range_Pressure_s :664-719, tabulated_potential :726-763, d_tabulated_potential :766-806,
 *                                       power_function :850-862, d_power_function :864-876, inv_exp_function :894-905,
 *                                       d_inv_exp_function :907-918, exp_function :935-950, d_exp_function :952-967
 *                                       (Wall_potential / d_Wall_potential, f_cut_L_C / d_f_cut_L_C: see wall_oracle.c)
 *   Source_files/ZBL_potential.f90      constants :43-52, NLH_pot :150-161, d_NLH_pot :164-178, ZBL_pot :366-373, d_ZBL_pot :376-386,
 *                                       d_ZBL_phi :406-417, ZBL_phi :434-445, ZBL_a :448-452
 *   Source_files/Little_subroutines.f90 Find_monotonous_LE :1184-1228, linear_interpolation :1031-1125 (default arm)
 *   Source_files/Algebra_tools.f90      make_cubic_splines :115-190, get_first_derivative :193-219, get_second_derivative :222-254,
 *                                       cubic_function :256-263, d_cubic_function :266-273
 * One departure, stated: at r == R(Nsiz) exactly the reference's finite-difference arm indexes one element past the grid
 * (i_closest + 1 = Nsiz + 1); here, as in the product, the index is held at Nsiz.
 * No reference test or vector exists for this class: PARITY UNPINNED (see xtant_oracle.c).  Pins: tests/test_short_cpu.py. */

static const double sr_m_a_u = 0.8854, sr_m_phi1 = 0.1818, sr_m_phi2 = 0.5099, sr_m_phi3 = 0.2802, sr_m_phi4 = 0.02817;
static const double sr_m_exp1 = -3.2, sr_m_exp2 = -0.9423, sr_m_exp3 = -0.4028, sr_m_exp4 = -0.2016;
static const double sr_g_Pi = 3.1415926535897932384626433832795, sr_g_e0 = 8.854187817620e-12, sr_g_e = 1.602176634e-19, sr_g_a0 = 0.5291772085936;
#define SR_M_K (1.0 / (4.0 * sr_g_Pi * sr_g_e0))

static double sr_ZBL_a(double Z1, double Z2) { return sr_m_a_u * sr_g_a0 / (pow(Z1, 0.23) + pow(Z2, 0.23)); }
static double sr_ZBL_phi(double Z1, double Z2, double r) {
    const double a = sr_ZBL_a(Z1, Z2), x = r / a;
    return sr_m_phi1 * exp(sr_m_exp1 * x) + sr_m_phi2 * exp(sr_m_exp2 * x) + sr_m_phi3 * exp(sr_m_exp3 * x) + sr_m_phi4 * exp(sr_m_exp4 * x);
}
static double sr_d_ZBL_phi(double Z1, double Z2, double r) {
    const double a = sr_ZBL_a(Z1, Z2), x = r / a;
    return (sr_m_phi1 * sr_m_exp1 * exp(sr_m_exp1 * x) + sr_m_phi2 * sr_m_exp2 * exp(sr_m_exp2 * x) + sr_m_phi3 * sr_m_exp3 * exp(sr_m_exp3 * x) +
            sr_m_phi4 * sr_m_exp4 * exp(sr_m_exp4 * x)) / a;
}
static double sr_ZBL_pot(double Z1, double Z2, double r) {
    const double phi = sr_ZBL_phi(Z1, Z2, r);
    return SR_M_K * Z1 * Z2 * sr_g_e / (1.0e-10 * r) * phi;
}
static double sr_d_ZBL_pot(double Z1, double Z2, double r) {
    const double V_ZBL_part = SR_M_K * Z1 * Z2 * sr_g_e / (1.0e-10 * r);
    const double phi = sr_ZBL_phi(Z1, Z2, r), d_phi = sr_d_ZBL_phi(Z1, Z2, r);
    return V_ZBL_part * (d_phi - phi / r);
}
static double sr_NLH_pot(double Z1, double Z2, double r, const double* a, const double* b) {
    double phi = 0.0;
    for (int k = 0; k < 3; ++k) phi += a[k] * exp(-b[k] * r);
    return SR_M_K * Z1 * Z2 * sr_g_e / (1.0e-10 * r) * phi;
}
static double sr_d_NLH_pot(double Z1, double Z2, double r, const double* a, const double* b) {
    const double V_NLH_part = SR_M_K * Z1 * Z2 * sr_g_e / (1.0e-10 * r);
    double phi = 0.0, d_phi = 0.0;
    for (int k = 0; k < 3; ++k) { phi += a[k] * exp(-b[k] * r); d_phi += -b[k] * a[k] * exp(-b[k] * r); }
    return V_NLH_part * (d_phi - phi / r);
}
static double sr_exp_function(double r, int use_it, double Phi, double r0, double a) {
    double Pot = 0.0;
    if (use_it) { const double eps = 1.0e-5; if (fabs(a) > eps) Pot = Phi * exp(-(r - r0) / a); }
    return Pot;
}
static double sr_d_exp_function(double r, int use_it, double Phi, double r0, double a) {
    double Pot = 0.0;
    if (use_it) { const double eps = 1.0e-5; if (fabs(a) > eps) Pot = -Phi * exp(-(r - r0) / a) / a; }
    return Pot;
}
static double sr_power_function(double r, const xtb_short_pair* W) {
    double Pot = 0.0;
    for (int i = 0; i < W->n_pow; ++i) Pot = Pot + W->pow_Phi[i] * pow(r / W->pow_r0[i], W->pow_m[i]);
    return Pot;
}
static double sr_d_power_function(double r, const xtb_short_pair* W) {
    double Pot = 0.0;
    for (int i = 0; i < W->n_pow; ++i) Pot = Pot + W->pow_Phi[i] * pow(r / W->pow_r0[i], W->pow_m[i]) * W->pow_m[i] / r;
    return Pot;
}
/* Find_monotonous_LE, 1-based result as in the reference */
static int sr_Find_monotonous_LE(const double* Array, int N, double Value0) {
    int i_cur;
    if (Value0 <= Array[0]) i_cur = 1;
    else if (Value0 >= Array[N - 1]) i_cur = N;
    else {
        int i_1 = 1, i_2 = N;
        i_cur = (int)floor((i_1 + i_2) / 2.0);
        double temp_val = Array[i_cur - 1];
        for (;;) {
            if (Value0 >= Array[i_cur - 1] && Value0 < Array[i_cur]) break;
            if (temp_val < Value0) i_1 = i_cur; else i_2 = i_cur;
            i_cur = (int)floor((i_1 + i_2) / 2.0);
            temp_val = Array[i_cur - 1];
        }
    }
    return i_cur;
}
/* linear_interpolation without x0, y0, replac; i is 1-based (>= 2 on this path) */
static double sr_linear_interpolation(const double* xarray, const double* yarray, double x, int i) {
#define XA(k) xarray[(k) - 1]
#define YA(k) yarray[(k) - 1]
    if (i > 1) {
        if (fabs(YA(i) - YA(i - 1)) < 1.0e-6 * fabs(YA(i))) return YA(i);
        if (fabs(XA(i) - XA(i - 1)) < 1.0e-6 * fabs(XA(i))) return (YA(i - 1) + YA(i)) * 0.5;
    }
    if (i > 1) {
        if (fabs(x - XA(i - 1)) >= 1.0e-6 * fabs(XA(i - 1))) return YA(i - 1) + (YA(i) - YA(i - 1)) / (XA(i) - XA(i - 1)) * (x - XA(i - 1));
        if (fabs(XA(i)) > 1.0e-8) return YA(i - 1) + (YA(i) - YA(i - 1)) / (XA(i) - XA(i - 1)) * (x - XA(i - 1));
        return YA(i);
    }
    if (fabs(XA(i + 1) - XA(i)) > 1.0e-8) return YA(i) - (YA(i + 1) - YA(i)) / (XA(i + 1) - XA(i)) * (XA(i) - x);
    return YA(i);
#undef XA
#undef YA
}
static double sr_tabulated_potential(double r, const xtb_short_pair* W) {
    double E;
    if (W->n_tab > 0) {
        const int Nsiz = W->n_tab;
        if (r > W->tab_R[Nsiz - 1]) E = 0.0;
        else {
            int i_closest = sr_Find_monotonous_LE(W->tab_R, Nsiz, r);
            if (W->tab_spline) {
                if (i_closest == Nsiz) i_closest = Nsiz - 1;
                const double x = r - W->tab_R[i_closest - 1], x2 = x * x;
                E = W->tab_a[i_closest - 1] + W->tab_b[i_closest - 1] * x + W->tab_c[i_closest - 1] * x2 + W->tab_d[i_closest - 1] * x * x2;
            } else {
                i_closest = i_closest + 1;
                if (i_closest > Nsiz) i_closest = Nsiz;          /* stated departure: the reference reads R(Nsiz+1) here */
                E = sr_linear_interpolation(W->tab_R, W->tab_E, r, i_closest);
            }
        }
    } else E = 0.0;
    return E;
}
static double sr_d_tabulated_potential(double r, const xtb_short_pair* W) {
    double E;
    if (W->n_tab > 0) {
        const int Nsiz = W->n_tab;
        if (r > W->tab_R[Nsiz - 1]) E = 0.0;
        else {
            int i_closest = sr_Find_monotonous_LE(W->tab_R, Nsiz, r);
            if (i_closest == 1) E = (W->tab_E[1] - W->tab_E[0]) / (W->tab_R[1] - W->tab_R[0]);
            else if (W->tab_spline) {
                if (i_closest == Nsiz) i_closest = Nsiz - 1;
                const double x = r - W->tab_R[i_closest - 1], x2 = x * x;
                E = W->tab_b[i_closest - 1] + 2.0 * W->tab_c[i_closest - 1] * x + 3.0 * W->tab_d[i_closest - 1] * x2;
            } else {
                i_closest = i_closest + 1;
                if (i_closest > Nsiz) i_closest = Nsiz;
                E = (W->tab_E[i_closest - 1] - W->tab_E[i_closest - 2]) / (W->tab_R[i_closest - 1] - W->tab_R[i_closest - 2]);
            }
        }
    } else E = 0.0;
    return E;
}
static double sr_Shortrange_pot(const xtb_short_pair* W, double a_r, double Z1, double Z2) {
    double Pot = 0.0;
    if (a_r < W->d0 + W->dd * 10.0) {
        const double f_cut_large = wall_f_cut_L_C(a_r, W->d0, W->dd);
        const double f_exp = sr_exp_function(a_r, W->use_exp, W->exp_Phi, W->exp_r0, W->exp_a);
        const double f_invexp = W->use_inv_exp ? wall_Wall_potential(W->inv_C, W->inv_r0, a_r) : 0.0;
        const double f_pow = sr_power_function(a_r, W);
        const double f_ZBL = W->use_zbl ? sr_ZBL_pot(Z1, Z2, a_r) : 0.0;
        const double f_NLH = W->use_nlh ? sr_NLH_pot(Z1, Z2, a_r, W->nlh_a, W->nlh_b) : 0.0;
        const double f_tab = sr_tabulated_potential(a_r, W);
        Pot = f_invexp + f_exp + f_pow + f_ZBL + f_NLH + f_tab;
        Pot = Pot * f_cut_large;
    }
    return Pot;
}
static double sr_d_Short_range_pot(const xtb_short_pair* W, double a_r, double Z1, double Z2) {
    double dPot = 0.0;
    if (a_r < W->d0 + W->dd * 10.0) {
        const double f_cut_large = wall_f_cut_L_C(a_r, W->d0, W->dd), d_f_large = wall_d_f_cut_L_C(a_r, W->d0, W->dd);
        const double f_exp = sr_exp_function(a_r, W->use_exp, W->exp_Phi, W->exp_r0, W->exp_a);
        const double d_f_exp = sr_d_exp_function(a_r, W->use_exp, W->exp_Phi, W->exp_r0, W->exp_a);
        const double f_invexp = W->use_inv_exp ? wall_Wall_potential(W->inv_C, W->inv_r0, a_r) : 0.0;
        const double d_f_invexp = W->use_inv_exp ? wall_d_Wall_potential(W->inv_C, W->inv_r0, a_r) : 0.0;
        const double f_pow = sr_power_function(a_r, W), d_f_pow = sr_d_power_function(a_r, W);
        const double f_ZBL = W->use_zbl ? sr_ZBL_pot(Z1, Z2, a_r) : 0.0, d_f_ZBL = W->use_zbl ? sr_d_ZBL_pot(Z1, Z2, a_r) : 0.0;
        const double f_NLH = W->use_nlh ? sr_NLH_pot(Z1, Z2, a_r, W->nlh_a, W->nlh_b) : 0.0;
        const double d_f_NLH = W->use_nlh ? sr_d_NLH_pot(Z1, Z2, a_r, W->nlh_a, W->nlh_b) : 0.0;
        const double f_tab = sr_tabulated_potential(a_r, W), d_f_tab = sr_d_tabulated_potential(a_r, W);
        const double Pot = f_invexp + f_exp + f_pow + f_ZBL + f_NLH + f_tab;
        const double d_Pot = d_f_invexp + d_f_exp + d_f_pow + d_f_ZBL + d_f_NLH + d_f_tab;
        dPot = Pot * d_f_large + d_Pot * f_cut_large;
    }
    return dPot;
}

/* make_cubic_splines: a, b, c, d [n] from x, y [n] (n > 1) */
void orc_make_cubic_splines(int Nsiz, const double* x_array, const double* y_array, double* a, double* b, double* c, double* d) {
    double* h = (double*)malloc(sizeof(double) * 4 * (size_t)Nsiz);
    double *h_inv = h + Nsiz, *f_prime = h + 2 * Nsiz, *f_prime2 = h + 3 * Nsiz;
    for (int i = 1; i <= Nsiz - 1; ++i) h[i - 1] = x_array[i] - x_array[i - 1];
    h[Nsiz - 1] = h[Nsiz - 2];
    for (int i = 0; i < Nsiz; ++i) h_inv[i] = 1.0 / h[i];
    for (int i = 1; i <= Nsiz; ++i) {
        int j = i <= 1 ? 1 : (i >= Nsiz ? Nsiz - 1 : i);                                   /* get_first_derivative */
        f_prime[i - 1] = (y_array[j] - y_array[j - 1]) / (x_array[j] - x_array[j - 1]);
        j = i <= 2 ? 2 : (i >= Nsiz - 1 ? Nsiz - 1 : i);                                   /* get_second_derivative */
        const double xnn1 = x_array[j] - x_array[j - 1], xn1n2 = x_array[j - 1] - x_array[j - 2], xnn2 = x_array[j] - x_array[j - 2];
        f_prime2[i - 1] = 2.0 * (y_array[j] * xn1n2 - y_array[j - 1] * xnn2 + y_array[j - 2] * xnn1) / (xnn2 * xn1n2 * xnn1);
    }
    for (int i = 0; i < Nsiz; ++i) { a[i] = y_array[i]; b[i] = 0.0; c[i] = 0.0; d[i] = 0.0; }
    b[Nsiz - 1] = f_prime[Nsiz - 1];
    b[Nsiz - 2] = f_prime[Nsiz - 2];
    for (int i = Nsiz - 2; i >= 1; --i) b[i - 1] = f_prime[i] + (3.0 / 4.0) * f_prime2[i] * h[i - 1];
    for (int i = Nsiz - 2; i >= 1; --i) c[i - 1] = h_inv[i - 1] * (f_prime[i] - b[i - 1] - 0.5 * h[i - 1] * f_prime2[i]);
    for (int i = Nsiz - 2; i >= 1; --i) d[i - 1] = (1.0 / 6.0) * h_inv[i - 1] * (f_prime2[i] - 2.0 * c[i - 1]);
    free(h);
}

/* get_short_range_rep_s: returns a = sum_a * 0.5; Epot[i] = E_pot * 0.5 */
double orc_short_energy(int na, int cap, int n_kinds, const int* koa, const double* Z, const int* nn_size, const int* nn_list,
                        const double* nn_dist, const xtb_short_pair* WP, double* Epot) {
    double sum_a = 0.0;
    for (int i = 0; i < na; ++i) {
        double E_pot = 0.0;
        const double Z1 = Z[koa[i]];
        for (int a2 = 0; a2 < nn_size[i]; ++a2) {
            const size_t e = (size_t)i * cap + a2;
            const int j = nn_list[e];
            const double Z2 = Z[koa[j]];
            const double Short_range_pot = sr_Shortrange_pot(WP + koa[i] + n_kinds * koa[j], nn_dist[4 * e + 3], Z1, Z2);
            sum_a = sum_a + Short_range_pot;
            E_pot = E_pot + Short_range_pot;
        }
        if (Epot) Epot[i] = E_pot * 0.5;
    }
    return sum_a * 0.5;
}

/* d_Short_range_pot_s, verbatim loop nest */
void orc_short_forces(int na, int cap, int n_kinds, const int* koa, const double* Z, const double* h, const int* nn_size, const int* nn_list,
                      const double* nn_dist, const xtb_short_pair* WP, double* F) {
#pragma omp parallel for schedule(dynamic, 4)
    for (int ian = 0; ian < na; ++ian) {
        double Erx[3] = {0.0, 0.0, 0.0};
        for (int i1 = 0; i1 < na; ++i1) {
            const int dik = ian == i1;
            double dpsi[3] = {0.0, 0.0, 0.0};
            const double Z1 = Z[koa[i1]];
            for (int a2 = 0; a2 < nn_size[i1]; ++a2) {
                const size_t e = (size_t)i1 * cap + a2;
                const int j1 = nn_list[e];
                const int djk = ian == j1;
                if ((dik - djk) != 0) {
                    const double Z2 = Z[koa[j1]];
                    const double x = nn_dist[4 * e], y = nn_dist[4 * e + 1], z = nn_dist[4 * e + 2];
                    double x1[3];
                    x1[0] = x * H_(1, 1) + y * H_(1, 2) + z * H_(1, 3);
                    x1[1] = x * H_(2, 1) + y * H_(2, 2) + z * H_(2, 3);
                    x1[2] = x * H_(3, 1) + y * H_(3, 2) + z * H_(3, 3);
                    const double a_r = nn_dist[4 * e + 3];
                    const double b = sr_d_Short_range_pot(WP + koa[i1] + n_kinds * koa[j1], a_r, Z1, Z2);
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

/* d_Short_range_Pressure_s */
void orc_short_pressure(int na, int cap, int n_kinds, const int* koa, const double* Z, const int* nn_size, const int* nn_list,
                        const double* nn_dist, const double* nn_dist_s, const xtb_short_pair* WP, double* P) {
    for (int e9 = 0; e9 < 9; ++e9) P[e9] = 0.0;
    for (int i = 0; i < na; ++i) {
        double Rep_Pr[9] = {0};
        const double Z1 = Z[koa[i]];
        for (int a2 = 0; a2 < nn_size[i]; ++a2) {
            const size_t e = (size_t)i * cap + a2;
            const int j = nn_list[e];
            const double r = nn_dist[4 * e + 3];
            const double dpsy = sr_d_Short_range_pot(WP + koa[i] + n_kinds * koa[j], r, Z1, Z[koa[j]]);
            for (int k = 0; k < 3; ++k)
                for (int l = 0; l < 3; ++l)
                    Rep_Pr[l + 3 * k] = Rep_Pr[l + 3 * k] + dpsy * nn_dist[4 * e + k] * nn_dist_s[3 * e + l] / r;
        }
        for (int e9 = 0; e9 < 9; ++e9) P[e9] = P[e9] + Rep_Pr[e9] * 0.5;
    }
}

/* single-pair probes */
double orc_short_V(const xtb_short_pair* W, double r, double Z1, double Z2) { return sr_Shortrange_pot(W, r, Z1, Z2); }
double orc_short_dV(const xtb_short_pair* W, double r, double Z1, double Z2) { return sr_d_Short_range_pot(W, r, Z1, Z2); }
