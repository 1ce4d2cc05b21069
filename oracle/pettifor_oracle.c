/*
 * ORACLE (test infrastructure) -- Pettifor-type orthogonal sp3 branch (configs 1 and 4).
 * Included at the end of xtant_oracle.c (shares shortest_distance and the H_ macro); same rules:
 * only tests/, smoke() and bench.py's CPU legs may load it.  PARITY UNPINNED (see xtant_oracle.c).
 *
 * Restates, citing Source_files/TB_Pettifor.f90:
 *   Vs :1935-1976, dVs :1979-2022, Hopping :1872-1930, Hamilton_one :1813-1868, Hamil_tot :1658-1810 (non-MPI arm),
 *   Construct_M_Vs :227-350, dHopping_s :909-1130, dHamilton_one_s :571-618, dHamil_tot_s :463-521,
 *   Attract_TB_H3_near :354-399, dHij_s :141-223,
 *   dHopping_Press_h :1283-1423, dHamilton_one_Press :1253-1280, dHamil_tot_Press :1201-1250,
 *   Attract_TB_Forces_Press :1134-1198.
 * Parameter struct = xtb_pettifor_pair (type(TB_H_Pettifor), Objects.f90:170-174); TB[(k1) + nk*(k2)] = TB_Hamil(k1+1,k2+1).
 */

static double pet_Vs(const xtb_pettifor_pair* TB, int i, double r) {            /* :1935 (i is 0-based here) */
    if (r > TB->rm) return 0.0;
    if (r >= TB->r1) {                                                          /* `>=`  :1967 */
        double r_r1 = r - TB->r1, r_r12 = r_r1 * r_r1;
        return TB->V0[i] * (TB->c0[i] + TB->c1[i] * r_r1 + TB->c2[i] * r_r12 + TB->c3[i] * r_r12 * r_r1);
    }
    return TB->V0[i] * pow(TB->r0 / r, TB->n) *
           exp(TB->n * (-pow(r / TB->rc[i], TB->nc[i]) + pow(TB->r0 / TB->rc[i], TB->nc[i])));
}

static double pet_dVs(const xtb_pettifor_pair* TB, int i, double r, const double* Vs_precalc) {   /* :1979 */
    if (r > TB->rm) return 0.0;
    if (r > TB->r1) {                                                           /* `>`  :2010 (not `>=`: reference quirk) */
        double r_r1 = r - TB->r1;
        return TB->V0[i] * (TB->c1[i] + TB->c2[i] * r_r1 * 2.0 + TB->c3[i] * r_r1 * r_r1 * 3.0);
    }
    double V = Vs_precalc ? *Vs_precalc : pet_Vs(TB, i, r);
    return -TB->n * V / r * (1.0 + TB->nc[i] * pow(r / TB->rc[i], TB->nc[i]));
}

/* Hopping :1872-1930 ; ts[a + 4*b] = ts(a+1,b+1) */
static void pet_Hopping(const xtb_pettifor_pair* TB, double* ts, double x, double y, double z, double r) {
    if (r <= TB->rm) {
        double xr = x / r, yr = y / r, zr = z / r, V[4];
        for (int c = 0; c < 4; ++c) V[c] = pet_Vs(TB, c, r);
#define TS(a, b) ts[(a - 1) + 4 * (b - 1)]
        TS(1, 1) = V[0];
        TS(1, 2) = -xr * V[1]; TS(1, 3) = -yr * V[1]; TS(1, 4) = -zr * V[1];
        TS(2, 1) = xr * V[1];
        TS(2, 2) = xr * xr * (V[2] - V[3]) + V[3];
        TS(2, 3) = xr * yr * (V[2] - V[3]);
        TS(2, 4) = xr * zr * (V[2] - V[3]);
        TS(3, 1) = yr * V[1];
        TS(3, 2) = xr * yr * (V[2] - V[3]);
        TS(3, 3) = yr * yr * (V[2] - V[3]) + V[3];
        TS(3, 4) = yr * zr * (V[2] - V[3]);
        TS(4, 1) = zr * V[1];
        TS(4, 2) = xr * zr * (V[2] - V[3]);
        TS(4, 3) = yr * zr * (V[2] - V[3]);
        TS(4, 4) = zr * zr * (V[2] - V[3]) + V[3];
#undef TS
    } else {
        for (int e = 0; e < 16; ++e) ts[e] = 0.0;
    }
}

/* Hamil_tot :1752-1796 with Hamilton_one :1813-1868.  Dense H (N,N) column-major, N = 4*na, [eV].
 * Outer atom j, neighbour i (or i = j): block placed at rows of i, columns of j; the pair block is
 * evaluated from shortest_distance(i, j) -- recomputed, not read from the list (:1842). */
void orc_pet_assemble_H(int na, int cap, int n_kinds, const int* koa, const double* S, const double* h,
                        const int* nn_size, const int* nn_list, const xtb_pettifor_pair* TB, double* H) {
    const size_t N = (size_t)4 * na;
    memset(H, 0, sizeof(double) * N * N);                                       /* :1690 */
#pragma omp parallel for schedule(dynamic, 8)
    for (int j = 0; j < na; ++j) {
        const int m = nn_size[j];
        for (int atom_2 = 0; atom_2 <= m; ++atom_2) {
            const int i = atom_2 == 0 ? j : nn_list[(size_t)j * cap + atom_2 - 1];
            const xtb_pettifor_pair* P = TB + koa[j] + n_kinds * koa[i];        /* TB_Hamil(KOA(j), KOA(i)) :1769 */
            double blk[16];
            if (i == j) {
                for (int e = 0; e < 16; ++e) blk[e] = 0.0;
                blk[0] = P->Es; blk[5] = blk[10] = blk[15] = P->Ep;             /* :1831-1835 */
            } else {
                double r, x, y, z, sx, sy, sz;
                shortest_distance(S, h, i, j, &r, &x, &y, &z, &sx, &sy, &sz);   /* :1842 */
                pet_Hopping(P, blk, x, y, z, r);
            }
            for (int j1 = 0; j1 < 4; ++j1)
                for (int i1 = 0; i1 < 4; ++i1)
                    H[((size_t)i * 4 + i1) + N * ((size_t)j * 4 + j1)] = blk[i1 + 4 * j1];   /* Hij(k,l) = Hij1(i1,j1) :1776 */
        }
    }
}

/* dHopping_s :909-1130 for one list entry of atom i (neighbour j) and the sign b*r = dik - djk.
 * dts[al][a + 4*b] = dts{x,y,z}(a+1,b+1).  (x,y,z,r) = entry of i; TBp = TB_Hamil(KOA(i), KOA(j)). */
static void pet_dHopping_s(const xtb_pettifor_pair* P, double x, double y, double z, double r, const double* h,
                           double delta, double dts[3][16]) {
    const double b = delta / r;
    double x1[3], drds[3], dlds[3], dmds[3], dnds[3];
    for (int al = 1; al <= 3; ++al) x1[al - 1] = x * H_(al, 1) + y * H_(al, 2) + z * H_(al, 3);   /* M_x1, TB.f90:3505-3507 */
    const double xrr = x / r / r, yrr = y / r / r, zrr = z / r / r;             /* M_xrr = lmn/r, TB.f90:3516-3518 */
    for (int al = 0; al < 3; ++al) drds[al] = b * x1[al];
    for (int al = 1; al <= 3; ++al) {
        dlds[al - 1] = b * H_(al, 1) - xrr * drds[al - 1];
        dmds[al - 1] = b * H_(al, 2) - yrr * drds[al - 1];
        dnds[al - 1] = b * H_(al, 3) - zrr * drds[al - 1];
    }
    const double xr = x / r, yr = y / r, zr = z / r;
    double V[4], dV[4];
    for (int c = 0; c < 4; ++c) { V[c] = pet_Vs(P, c, r); dV[c] = pet_dVs(P, c, r, &V[c]); }     /* Construct_M_Vs :316-317 */
    for (int al = 0; al < 3; ++al) {
        double* d = dts[al];
#define D(a, b) d[(a - 1) + 4 * (b - 1)]
        D(1, 1) = dV[0] * drds[al];
        double temp = xr * dV[1];
        D(2, 1) = -(dlds[al] * V[1] + temp * drds[al]); D(1, 2) = -D(2, 1);
        temp = yr * dV[1];
        D(3, 1) = -(dmds[al] * V[1] + temp * drds[al]); D(1, 3) = -D(3, 1);
        temp = zr * dV[1];
        D(4, 1) = -(dnds[al] * V[1] + temp * drds[al]); D(1, 4) = -D(4, 1);
        temp = 2.0 * xr * (V[2] - V[3]);
        double temp1 = (xr * xr * (dV[2] - dV[3]) + dV[3]);
        D(2, 2) = temp * dlds[al] + temp1 * drds[al];
        temp = V[2] - V[3];
        temp1 = xr * yr * (dV[2] - dV[3]);
        D(3, 2) = (xr * dmds[al] + yr * dlds[al]) * temp + temp1 * drds[al]; D(2, 3) = D(3, 2);
        temp1 = xr * zr * (dV[2] - dV[3]);
        D(4, 2) = (xr * dnds[al] + zr * dlds[al]) * temp + temp1 * drds[al]; D(2, 4) = D(4, 2);
        double temp2 = 2.0 * yr * (V[2] - V[3]);
        temp1 = yr * yr * (dV[2] - dV[3]) + dV[3];
        D(3, 3) = temp2 * dmds[al] + temp1 * drds[al];
        temp1 = yr * zr * (dV[2] - dV[3]);
        D(4, 3) = (yr * dnds[al] + zr * dmds[al]) * temp + temp1 * drds[al]; D(3, 4) = D(4, 3);
        temp1 = zr * zr * (dV[2] - dV[3]) + dV[3];
        const double z_temp = 2.0 * zr * temp;
        D(4, 4) = z_temp * dnds[al] + temp1 * drds[al];
#undef D
    }
}

/* dHij_s :141-223, VERBATIM form: for every atom k the three dense matrices dHij{x,y,z}_s_all (N,N) are zeroed and
 * rebuilt from all list entries (dHamil_tot_s :463-521; non-zero only when k is one of the two atoms), then
 * contracted row by row with Aij on the neighbour pattern (Attract_TB_H3_near :354-399).
 * A = dense Aij (N,N) column-major.  F[3*k + al] = MDatoms(k)%forces%att(al).  O(na * N^2): small cells only. */
void orc_pet_forces_verbatim(int na, int cap, int n_kinds, const int* koa, const double* h, const int* nn_size,
                             const int* nn_list, const double* nn_dist, const xtb_pettifor_pair* TB,
                             const double* A, double* F) {
    const size_t N = (size_t)4 * na;
#pragma omp parallel
    {
        double* dH = (double*)malloc(sizeof(double) * 3 * N * N);
#pragma omp for schedule(dynamic, 1)
        for (int k = 0; k < na; ++k) {
            memset(dH, 0, sizeof(double) * 3 * N * N);                         /* :489-491 */
            for (int i = 0; i < na; ++i)
                for (int a2 = 0; a2 < nn_size[i]; ++a2) {
                    const size_t e = (size_t)i * cap + a2;
                    const int j = nn_list[e];
                    double dts[3][16];
                    const int dik = (i == k), djk = (k == j);
                    if (dik - djk != 0)
                        pet_dHopping_s(TB + koa[i] + n_kinds * koa[j], nn_dist[4 * e], nn_dist[4 * e + 1], nn_dist[4 * e + 2],
                                       nn_dist[4 * e + 3], h, (double)(dik - djk), dts);
                    else
                        memset(dts, 0, sizeof dts);                            /* :1125-1128 */
                    for (int al = 0; al < 3; ++al)
                        for (int j1 = 0; j1 < 4; ++j1)
                            for (int i1 = 0; i1 < 4; ++i1)                     /* dHijx(:,ki) = dtsx(ki,:) :597 ; placed (i4+i1, j4+j1) :510 */
                                dH[al * N * N + ((size_t)i * 4 + i1) + N * ((size_t)j * 4 + j1)] = dts[al][j1 + 4 * i1];
                }
            double E[3] = {0.0, 0.0, 0.0};                                     /* Attract_TB_H3_near */
            for (size_t i = 0; i < N; ++i) {
                const int i2 = (int)(i / 4);
                for (int kk = 0; kk < nn_size[i2]; ++kk) {
                    const size_t j = (size_t)nn_list[(size_t)i2 * cap + kk] * 4;
                    for (int al = 0; al < 3; ++al) {
                        const double* d = dH + al * N * N;
                        E[al] = E[al] + (d[i + N * j] * A[i + N * j]) + (d[i + N * (j + 1)] * A[i + N * (j + 1)]) +
                                (d[i + N * (j + 2)] * A[i + N * (j + 2)]) + (d[i + N * (j + 3)] * A[i + N * (j + 3)]);
                    }
                }
            }
            F[3 * k] = E[0]; F[3 * k + 1] = E[1]; F[3 * k + 2] = E[2];
        }
        free(dH);
    }
}

/* Sparse-equivalent form: the same terms in the same order, without the dense scratch.  rho_blk is the block
 * store of orc_density_blocks / orc_gather_blocks with n_orb = 4: blk[(i*(cap+1) + q)*16 + mu + 4*nu] = Aij(i mu, j nu). */
void orc_pet_forces_sparse(int na, int cap, int n_kinds, const int* koa, const double* h, const int* nn_size,
                           const int* nn_list, const double* nn_dist, const xtb_pettifor_pair* TB,
                           const double* rho_blk, int atom_first, int atom_last, double* F) {
#pragma omp parallel for schedule(dynamic, 4)
    for (int k = atom_first; k < atom_last; ++k) {
        double E[3] = {0.0, 0.0, 0.0};
        for (int i = 0; i < na; ++i) {                       /* rows ascending, as Attract_TB_H3_near */
            int lo = 0, hi = nn_size[i];
            if (i != k) {
                lo = find_entry(nn_size, nn_list, cap, i, k);
                if (lo < 0) continue;
                hi = lo + 1;
            }
            for (int i1 = 0; i1 < 4; ++i1)
                for (int a2 = lo; a2 < hi; ++a2) {
                    const size_t e = (size_t)i * cap + a2;
                    const int j = nn_list[e];
                    double d[3][16];
                    pet_dHopping_s(TB + koa[i] + n_kinds * koa[j], nn_dist[4 * e], nn_dist[4 * e + 1], nn_dist[4 * e + 2],
                                   nn_dist[4 * e + 3], h, (double)((i == k) - (k == j)), d);
                    const double* Ab = rho_blk + ((size_t)i * (cap + 1) + a2 + 1) * 16;
                    for (int al = 0; al < 3; ++al)
                        E[al] = E[al] + (d[al][0 + 4 * i1] * Ab[i1]) + (d[al][1 + 4 * i1] * Ab[i1 + 4]) +
                                (d[al][2 + 4 * i1] * Ab[i1 + 8]) + (d[al][3 + 4 * i1] * Ab[i1 + 12]);
                }
        }
        F[3 * k] = E[0]; F[3 * k + 1] = E[1]; F[3 * k + 2] = E[2];
    }
}

/* dHopping_Press_h :1283-1423 for one list entry; dts[k1][a + 4*b] = dtsx(k1+1, a+1, b+1), k1 = (gamma-1)*3 + delta-1 */
static void pet_dHopping_Press(const xtb_pettifor_pair* P, const double* d4, const double* s3, double dts[9][16]) {
    const double x = d4[0], y = d4[1], z = d4[2], r = d4[3];
    const double sr[3] = {s3[0] / r, s3[1] / r, s3[2] / r};
    const double xyz[3] = {x, y, z};
    const double x_r = x / r, y_r = y / r, z_r = z / r;
    const double Vs2r = pet_Vs(P, 1, r), dVs2r = pet_dVs(P, 1, r, NULL), Vs3r = pet_Vs(P, 2, r), dVs3r = pet_dVs(P, 2, r, NULL);
    const double Vs4r = pet_Vs(P, 3, r), dVs4r = pet_dVs(P, 3, r, NULL), dVs1r = pet_dVs(P, 0, r, NULL);
    for (int k1 = 0; k1 < 9; ++k1) {
        const int ga = k1 / 3, de = k1 % 3;
        double dc[3] = {0.0, 0.0, 0.0};
        const double drijdhgd = xyz[de] * sr[ga];
        dc[de] = sr[ga];
        const double dl = dc[0] - x * drijdhgd / (r * r);
        const double dm = dc[1] - y * drijdhgd / (r * r);
        const double dn = dc[2] - z * drijdhgd / (r * r);
        double* d = dts[k1];
#define D(a, b) d[(a - 1) + 4 * (b - 1)]
        D(1, 1) = dVs1r * drijdhgd;
        D(2, 1) = -(dl * Vs2r + x_r * dVs2r * drijdhgd);
        D(3, 1) = -(dm * Vs2r + y_r * dVs2r * drijdhgd);
        D(4, 1) = -(dn * Vs2r + z_r * dVs2r * drijdhgd);
        D(1, 2) = -D(2, 1);
        D(2, 2) = 2.0 * x_r * (dl) * (Vs3r - Vs4r) + (x_r * x_r * (dVs3r - dVs4r) + dVs4r) * drijdhgd;
        D(3, 2) = (x_r * (dm) + y_r * (dl)) * (Vs3r - Vs4r) + x_r * y_r * (dVs3r - dVs4r) * drijdhgd;
        D(4, 2) = (x_r * (dn) + z_r * (dl)) * (Vs3r - Vs4r) + x_r * z_r * (dVs3r - dVs4r) * drijdhgd;
        D(1, 3) = -D(3, 1);
        D(2, 3) = D(3, 2);
        D(3, 3) = 2.0 * y_r * (dm) * (Vs3r - Vs4r) + (y_r * y_r * (dVs3r - dVs4r) + dVs4r) * drijdhgd;
        D(4, 3) = (y_r * (dn) + z_r * (dm)) * (Vs3r - Vs4r) + y_r * z_r * (dVs3r - dVs4r) * drijdhgd;
        D(1, 4) = -D(4, 1);
        D(2, 4) = D(4, 2);
        D(3, 4) = D(4, 3);
        D(4, 4) = 2.0 * z_r * (dn) * (Vs3r - Vs4r) + (z_r * z_r * (dVs3r - dVs4r) + dVs4r) * drijdhgd;
#undef D
    }
}

/* Attract_TB_Forces_Press :1134-1198 (sparse-equivalent: dHij is non-zero only on list blocks).
 * dwr_press(k1, row) = sum_cols dHij(k1,row,:)*Aij(row,:) ; SCforce%att(k, i) = SUM(dwr_press((i-1)*3+k, :)).
 * att[k + 3*i] column-major (3,3). */
void orc_pet_pressure(int na, int cap, int n_kinds, const int* koa, const int* nn_size, const int* nn_list,
                      const double* nn_dist, const double* nn_dist_s, const xtb_pettifor_pair* TB,
                      const double* rho_blk, double* att) {
    double* dwr = (double*)calloc((size_t)9 * 4 * na, sizeof(double));
#pragma omp parallel for schedule(dynamic, 4)
    for (int i = 0; i < na; ++i)
        for (int a2 = 0; a2 < nn_size[i]; ++a2) {
            const size_t e = (size_t)i * cap + a2;
            const int j = nn_list[e];
            double dts[9][16];
            pet_dHopping_Press(TB + koa[i] + n_kinds * koa[j], nn_dist + 4 * e, nn_dist_s + 3 * e, dts);
            const double* Ab = rho_blk + ((size_t)i * (cap + 1) + a2 + 1) * 16;
            for (int k1 = 0; k1 < 9; ++k1)
                for (int i1 = 0; i1 < 4; ++i1) {
                    double s = 0.0;                          /* dHij(:,i2,j2) = dHij1(:,i1,j1) = dts_press(:,j1,i1) :1275,1243 */
                    for (int j1 = 0; j1 < 4; ++j1) s += dts[k1][j1 + 4 * i1] * Ab[i1 + 4 * j1];
                    dwr[k1 + 9 * ((size_t)i * 4 + i1)] += s;
                }
        }
    for (int i = 0; i < 3; ++i)
        for (int k = 0; k < 3; ++k) {
            double s = 0.0;
            for (size_t row = 0; row < (size_t)4 * na; ++row) s += dwr[(i * 3 + k) + 9 * row];
            att[k + 3 * i] = s;
        }
    free(dwr);
}

/* ===== repulsive part: TB_Pettifor.f90 fx :2108-2123, phi :2126-2158, dphi :2695-2734, dfx :2776-2792,
 * get_Erep_s :2074-2105, dErdr_s :2162-2306 (OpenMP arm), dErdr_Pressure_s :2634-2690.
 * REP[(k1) + nk*(k2)] = TB_Repuls(k1+1, k2+1); struct xtb_pettifor_rep_pair = type(TB_Rep_Pettifor), Objects.f90:95-100. */
static double pet_fx(const xtb_pettifor_rep_pair* R, double x) {                 /* :2108 */
    const double x2 = x * x;
    return R->a0[0] + R->a0[1] * x + R->a0[2] * x2 + R->a0[3] * x2 * x + R->a0[4] * x2 * x2;
}

static double pet_dfx(const xtb_pettifor_rep_pair* R, double x) {                /* :2776 */
    const double x2 = x * x;
    return R->a0[1] + R->a0[2] * x * 2.0 + R->a0[3] * x2 * 3.0 + R->a0[4] * x * x2 * 4.0;
}

static double pet_phi(const xtb_pettifor_rep_pair* R, double r) {                /* :2126 */
    if (r > R->dm) return 0.0;
    if (r > R->d1) {
        const double x1 = r - R->d1, x2 = x1 * x1;
        return R->c0[0] + R->c0[1] * x1 + R->c0[2] * x2 + R->c0[3] * x1 * x2;
    }
    if (r <= 0.0) return 0.0;
    return R->phi0 * pow(R->d0 / r, R->m) * exp(R->m * (-pow(r / R->dc, R->mc) + pow(R->d0 / R->dc, R->mc)));
}

static double pet_dphi(const xtb_pettifor_rep_pair* R, double r) {               /* :2695 (Phi_precalc absent at both call sites) */
    const double x1 = r - R->d1;
    if (r > R->dm) return 0.0;
    if (r > R->d1) return R->c0[1] + R->c0[2] * x1 * 2.0 + R->c0[3] * x1 * x1 * 3.0;
    if (r <= 0.0) return 0.0;
    return -R->m * pet_phi(R, r) / r * (1.0 + R->mc * pow(r / R->dc, R->mc));
}

/* get_Erep_s :2074.  Epot[i] = MDAtoms(i)%Epot; returns the total. */
double orc_pet_erep(int na, int cap, int n_kinds, const int* koa, const int* nn_size, const int* nn_list,
                    const double* nn_dist, const xtb_pettifor_rep_pair* REP, double* Epot) {
    double a = 0.0;
    for (int i1 = 0; i1 < na; ++i1) {
        double b = 0.0;
        for (int a2 = 0; a2 < nn_size[i1]; ++a2) {
            const size_t e = (size_t)i1 * cap + a2;
            const int j1 = nn_list[e];
            if (j1 != i1) b = b + pet_phi(REP + koa[j1] + n_kinds * koa[i1], nn_dist[4 * e + 3]);
        }
        const xtb_pettifor_rep_pair* R = REP + koa[i1] + n_kinds * koa[i1];
        const double E_pot = pet_fx(R, b) + R->E0_TB;
        a = a + E_pot;
        if (Epot) Epot[i1] = E_pot;
    }
    return a;
}

/* dErdr_s :2232-2300 verbatim: for every ian, every i1, the whole list of i1 (O(na^2 nn)). F[3*ian + al] = forces%rep(al). */
void orc_pet_rep_forces_verbatim(int na, int cap, int n_kinds, const int* koa, const double* h, const int* nn_size,
                                 const int* nn_list, const double* nn_dist, const xtb_pettifor_rep_pair* REP, double* F) {
#pragma omp parallel for schedule(dynamic, 4)
    for (int ian = 0; ian < na; ++ian) {
        double Erx[3] = {0.0, 0.0, 0.0};
        for (int i1 = 0; i1 < na; ++i1) {
            double dpsi[3] = {0.0, 0.0, 0.0}, psi = 0.0;
            for (int a2 = 0; a2 < nn_size[i1]; ++a2) {
                const size_t e = (size_t)i1 * cap + a2;
                const int j1 = nn_list[e];
                const int dik = ian == i1, djk = ian == j1;
                if (j1 != i1 || (dik - djk) != 0) {
                    const double x = nn_dist[4 * e], y = nn_dist[4 * e + 1], z = nn_dist[4 * e + 2], a_r = nn_dist[4 * e + 3];
                    const xtb_pettifor_rep_pair* R = REP + koa[j1] + n_kinds * koa[i1];
                    psi = psi + pet_phi(R, a_r);
                    double x1[3];
                    x1[0] = x * H_(1, 1) + y * H_(1, 2) + z * H_(1, 3);
                    x1[1] = x * H_(2, 1) + y * H_(2, 2) + z * H_(2, 3);
                    x1[2] = x * H_(3, 1) + y * H_(3, 2) + z * H_(3, 3);
                    const double b = pet_dphi(R, a_r);
                    const double ddlta = (double)(dik - djk) / a_r;
                    const double b_delta = b * ddlta;
                    dpsi[0] = dpsi[0] + b_delta * x1[0];
                    dpsi[1] = dpsi[1] + b_delta * x1[1];
                    dpsi[2] = dpsi[2] + b_delta * x1[2];
                }
            }
            const double a = pet_dfx(REP + koa[i1] + n_kinds * koa[i1], psi);
            Erx[0] = Erx[0] + a * dpsi[0];
            Erx[1] = Erx[1] + a * dpsi[1];
            Erx[2] = Erx[2] + a * dpsi[2];
        }
        F[3 * ian] = Erx[0]; F[3 * ian + 1] = Erx[1]; F[3 * ian + 2] = Erx[2];
    }
}

/* sparse-equivalent: psi_i once per atom; for ian only i1 = ian (all its entries) and the i1 that list ian contribute. */
void orc_pet_rep_forces_sparse(int na, int cap, int n_kinds, const int* koa, const double* h, const int* nn_size,
                               const int* nn_list, const double* nn_dist, const xtb_pettifor_rep_pair* REP,
                               int atom_first, int atom_last, double* F) {
    double* dfpsi = (double*)malloc(sizeof(double) * (size_t)na);
    for (int i1 = 0; i1 < na; ++i1) {
        double psi = 0.0;
        for (int a2 = 0; a2 < nn_size[i1]; ++a2) {
            const size_t e = (size_t)i1 * cap + a2;
            psi = psi + pet_phi(REP + koa[nn_list[e]] + n_kinds * koa[i1], nn_dist[4 * e + 3]);
        }
        dfpsi[i1] = pet_dfx(REP + koa[i1] + n_kinds * koa[i1], psi);
    }
#pragma omp parallel for schedule(dynamic, 8)
    for (int ian = atom_first; ian < atom_last; ++ian) {
        double Erx[3] = {0.0, 0.0, 0.0};
        for (int i1 = 0; i1 < na; ++i1) {                   /* i1 ascending, as the reference accumulates */
            int lo = 0, hi = nn_size[i1];
            if (i1 != ian) {
                lo = find_entry(nn_size, nn_list, cap, i1, ian);
                if (lo < 0) continue;
                hi = lo + 1;
            }
            double dpsi[3] = {0.0, 0.0, 0.0};
            for (int a2 = lo; a2 < hi; ++a2) {
                const size_t e = (size_t)i1 * cap + a2;
                const int j1 = nn_list[e];
                const double x = nn_dist[4 * e], y = nn_dist[4 * e + 1], z = nn_dist[4 * e + 2], a_r = nn_dist[4 * e + 3];
                const double b = pet_dphi(REP + koa[j1] + n_kinds * koa[i1], a_r);
                const double b_delta = b * ((double)((ian == i1) - (ian == j1)) / a_r);
                dpsi[0] = dpsi[0] + b_delta * (x * H_(1, 1) + y * H_(1, 2) + z * H_(1, 3));
                dpsi[1] = dpsi[1] + b_delta * (x * H_(2, 1) + y * H_(2, 2) + z * H_(2, 3));
                dpsi[2] = dpsi[2] + b_delta * (x * H_(3, 1) + y * H_(3, 2) + z * H_(3, 3));
            }
            Erx[0] = Erx[0] + dfpsi[i1] * dpsi[0];
            Erx[1] = Erx[1] + dfpsi[i1] * dpsi[1];
            Erx[2] = Erx[2] + dfpsi[i1] * dpsi[2];
        }
        F[3 * ian] = Erx[0]; F[3 * ian + 1] = Erx[1]; F[3 * ian + 2] = Erx[2];
    }
    free(dfpsi);
}

/* dErdr_Pressure_s :2634-2690.  rep[l + 3*k] = SCforce%rep(l+1, k+1) (l: relative-coordinate index, k: Cartesian index). */
void orc_pet_rep_pressure(int na, int cap, int n_kinds, const int* koa, const int* nn_size, const int* nn_list,
                          const double* nn_dist, const double* nn_dist_s, const xtb_pettifor_rep_pair* REP, double* rep) {
    for (int e9 = 0; e9 < 9; ++e9) rep[e9] = 0.0;
    for (int i = 0; i < na; ++i) {
        double Rep_Pr[9] = {0}, psi = 0.0;
        for (int a2 = 0; a2 < nn_size[i]; ++a2) {
            const size_t e = (size_t)i * cap + a2;
            const int j = nn_list[e];
            if (i == j) continue;
            const double r = nn_dist[4 * e + 3];
            const xtb_pettifor_rep_pair* R = REP + koa[j] + n_kinds * koa[i];
            psi = psi + pet_phi(R, r);
            const double dpsy = pet_dphi(R, r);
            for (int k = 0; k < 3; ++k)
                for (int l = 0; l < 3; ++l)
                    Rep_Pr[l + 3 * k] = Rep_Pr[l + 3 * k] + dpsy * nn_dist[4 * e + k] * nn_dist_s[3 * e + l] / r;
        }
        const double df_psy = pet_dfx(REP + koa[i] + n_kinds * koa[i], psi);
        for (int e9 = 0; e9 < 9; ++e9) rep[e9] = rep[e9] + df_psy * Rep_Pr[e9];
    }
}
