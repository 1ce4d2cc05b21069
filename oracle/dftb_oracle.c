/* ORACLE (test infrastructure) -- DFTB branch.  PARITY UNPINNED: the reference
 * ships no golden vectors for this path and cannot be compiled here (Fortran); see DESIGN.md section 3.
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may use this file.
 *
 * CPU restatement, function by function, of Source_files/TB_DFTB.f90:
 *   DFTB_radial_function / d_DFTB_radial_function   :231-280   (linear_interpolation: Little_subroutines.f90:1031-1125)
 *   Construct_Vij_DFTB                               :56-228    (Fermi_function / d_Fermi_function: Little_subroutines.f90:349-381)
 *   Hamil_tot_DFTB (assembly part)                   :319-503
 *   Hamilton_one_DFTB, Hopping_DFTB                  :673-735
 *   Get_overlap_S_matrix_DFTB                        :739-775
 *   get_forces_DFTB, Attract_TB_forces_DFTB          :829-972
 *   d_Hamilton_one_DFTB                              :977-1127
 *   Attract_TB_Forces_Press_DFTB, dHamil_tot_Press_DFTB, dHamilton_one_Press_DFTB, dHopping_Press_DFTB   :1135-1420
 * with the hetero-nuclear Slater-Koster tables of dftb_ks_tables.h and the geometric derivatives
 * drij_dska / ddija_dskb_kd / drij_dhab / dda_dhgd (TB_Koster_Slater.f90:59-161) already restated in xtant_oracle.c.
 *
 * Compact tables: rad[(e*4 + t)*10 + c], e = j*cap + atom_2 (list entry of atom j for neighbour i),
 *   t = 0..3 -> M_Vij(j,i,:), M_dVij(j,i,:), M_SVij(j,i,:), M_dSVij(j,i,:), c < nv (1, 4 or 10).
 * Included at the end of xtant_oracle.c (shares its helpers). */
#include "dftb_ks_tables.h"

static int dftb_n_orb(int basis) { return basis == 0 ? 1 : (basis == 1 ? 4 : 9); }     /* identify_DFTB_orbitals_per_atom */
static int dftb_n_v(int basis) { return basis == 0 ? 1 : (basis == 1 ? 4 : 10); }       /* identify_DFTB_basis_size */

/* linear_interpolation(xarray, yarray, x, y, i) without the optional arguments, Little_subroutines.f90:1031-1125;
 * i is 1-based.  xa, ya are the grid and ONE column of the table. */
static double lin_interp(const double* xa, const double* ya, double x, int i) {
#define XA(k) xa[(k) - 1]
#define YA(k) ya[(k) - 1]
    if (i > 1) {
        if (fabs(YA(i) - YA(i - 1)) < 1.0e-6 * fabs(YA(i))) return YA(i);                       /* :1042-1046 */
        if (fabs(XA(i) - XA(i - 1)) < 1.0e-6 * fabs(XA(i))) return (YA(i - 1) + YA(i)) * 0.5;   /* :1049-1052 */
        if (fabs(x - XA(i - 1)) >= 1.0e-6 * fabs(XA(i - 1)))                                    /* :1057-1058 */
            return YA(i - 1) + (YA(i) - YA(i - 1)) / (XA(i) - XA(i - 1)) * (x - XA(i - 1));
        if (fabs(XA(i)) > 1.0e-8)                                                               /* :1069-1071 */
            return YA(i - 1) + (YA(i) - YA(i - 1)) / (XA(i) - XA(i - 1)) * (x - XA(i - 1));
        return YA(i);
    }
    if (fabs(XA(i + 1) - XA(i)) > 1.0e-8)                                                       /* :1088-1091 */
        return YA(i) - (YA(i + 1) - YA(i)) / (XA(i + 1) - XA(i)) * (XA(i) - x);
    return YA(i);
#undef XA
#undef YA
}

/* DFTB_radial_function, TB_DFTB.f90:231-257 */
static double DFTB_radial_function(double r, const double* Rr, const double* col, int n_r) {
    double dr = Rr[0];
    int i_array = (int)ceil(r / dr);
    if (i_array > n_r) return col[n_r - 1];
    return lin_interp(Rr, col, r, i_array);
}
/* d_DFTB_radial_function, TB_DFTB.f90:260-280 */
static double d_DFTB_radial_function(double r, const double* Rr, const double* col, int n_r) {
    double dr = Rr[0];
    int i_array = (int)ceil(r / dr);
    if (i_array > n_r) return 0.0;
    if (i_array == 1) return (col[1] - col[0]) / dr;
    return (col[i_array - 1] - col[i_array - 2]) / dr;
}
/* Fermi_function / d_Fermi_function, Little_subroutines.f90:349-381 */
static double Fermi_function(double rcut, double d, double r) {
    if (r >= rcut + 10.0 * d) return 0.0;
    return 1.0 / (1.0 + exp((r - rcut) / d));
}
static double d_Fermi_function(double rcut, double d, double r) {
    if (r >= rcut + 10.0 * d) return 0.0;
    double exprd = exp((r - rcut) / d);
    double exprd1 = 1.0 + exprd;
    return -exprd / (d * exprd1 * exprd1);
}

/* table channel (0-based column of Vr/Sr) of compact channel c: sp3 uses columns 1,2,4,5 (TB_DFTB.f90:175-182) */
static int dftb_col(int basis, int c) {
    static const int sp3[4] = {0, 1, 3, 4};
    return basis == 1 ? sp3[c] : c;
}

/* Construct_Vij_DFTB, TB_DFTB.f90:158-219 (OpenMP arm): entry (j, atom_2 -> i) uses TB(KOA(i), KOA(j)) */
void orc_dftb_radial(int na, int cap, const int* nn_size, const int* nn_list, const double* nn_dist, const int* koa,
                     int n_kinds, const xtb_dftb_pair* TB, int basis, double* rad) {
    const int nv = dftb_n_v(basis);
#pragma omp parallel for schedule(static)
    for (int j = 0; j < na; ++j)
        for (int a2 = 0; a2 < nn_size[j]; ++a2) {
            size_t e = (size_t)j * cap + a2;
            int i = nn_list[e];
            const xtb_dftb_pair* P = &TB[koa[i] + n_kinds * koa[j]];     /* KOA1 = kind(i), KOA2 = kind(j) */
            double r = nn_dist[e * 4 + 3];
            double* o = rad + e * 40;
            for (int c = 0; c < 40; ++c) o[c] = 0.0;
            for (int c = 0; c < nv; ++c) {
                int col = dftb_col(basis, c);
                o[c]      = DFTB_radial_function(r, P->Rr, P->Vr + (size_t)P->n_r * col, P->n_r);
                o[20 + c] = DFTB_radial_function(r, P->Rr, P->Sr + (size_t)P->n_r * col, P->n_r);
                o[10 + c] = d_DFTB_radial_function(r, P->Rr, P->Vr + (size_t)P->n_r * col, P->n_r);
                o[30 + c] = d_DFTB_radial_function(r, P->Rr, P->Sr + (size_t)P->n_r * col, P->n_r);
            }
            double Fermi = Fermi_function(P->rcut, P->d, r);
            double dFermi = d_Fermi_function(P->rcut, P->d, r);
            for (int c = 0; c < nv; ++c) {                                /* :204-207 */
                o[10 + c] = o[10 + c] * Fermi + o[c] * dFermi;
                o[30 + c] = o[30 + c] * Fermi + o[20 + c] * dFermi;
                o[c] = o[c] * Fermi;
                o[20 + c] = o[20 + c] * Fermi;
            }
        }
}

typedef struct {
    int na, cap, n_kinds, basis, n_orb, nv;
    const int* nn_size; const int* nn_list; const double* nn_dist; const double* nn_dist_s;
    const double* lmn; const double* rad; const int* koa; const xtb_dftb_pair* TB; const double* h;
} DSys;

static const double dftb_zero10[10] = {0};
/* M_X(i,j,:) of the dense reference tables = compact entry of j in the list of i; zeros if j is not listed */
static const double* dftb_M(const DSys* s, int i, int j, int t) {
    for (int a2 = 0; a2 < s->nn_size[i]; ++a2)
        if (s->nn_list[(size_t)i * s->cap + a2] == j) return s->rad + ((size_t)i * s->cap + a2) * 40 + 10 * t;
    return dftb_zero10;
}

/* Hopping_DFTB, TB_DFTB.f90:707-735; ts is (n_orb,n_orb) column-major */
static void Hopping_DFTB(int basis, const double* V12, const double* V21, const double* lmn, double* ts) {
    if (basis == 0) ts[0] = t_s_s(V12[0]);                                   /* KS_s */
    else if (basis == 1) orc_KS_sp3_hetero(V12, V21, lmn[0], lmn[1], lmn[2], ts);
    else orc_KS_sp3d5_hetero(V12, V21, lmn[0], lmn[1], lmn[2], ts);
}

/* Hamilton_one_DFTB (TB_DFTB.f90:673-704) + Get_overlap_S_matrix_DFTB (:739-775) for the call
 * `Hamilton_one_DFTB(basis, j, i, TB(KOA(j),KOA(i)), ...)` of Hamil_tot_DFTB (:448-451): inside, the first atom is j.
 * a2 = 0: on-site (i == j); else list entry a2-1 of atom j (i = that neighbour). */
static void dftb_block_HS(const DSys* s, int j, int i, int a2, double* Hij1, double* Sij1) {
    const int no = s->n_orb;
    for (int q = 0; q < no * no; ++q) { Hij1[q] = 0.0; Sij1[q] = 0.0; }
    if (a2 == 0) {
        const xtb_dftb_pair* P = &s->TB[s->koa[j] + s->n_kinds * s->koa[i]];
        Hij1[0] = P->Es;
        if (s->basis > 0) for (int o = 1; o < 4; ++o) Hij1[o + no * o] = P->Ep;
        if (s->basis > 1) for (int o = 4; o < 9; ++o) Hij1[o + no * o] = P->Ed;
        for (int o = 0; o < no; ++o) Sij1[o + no * o] = 1.0;
        return;
    }
    size_t e = (size_t)j * s->cap + (a2 - 1);
    const double* lmn = s->lmn + e * 3;                      /* M_lmn(:,j,i) */
    Hopping_DFTB(s->basis, s->rad + e * 40, dftb_M(s, i, j, 0), lmn, Hij1);            /* M_Vij(j,i,:), M_Vij(i,j,:) */
    Hopping_DFTB(s->basis, s->rad + e * 40 + 20, dftb_M(s, i, j, 2), lmn, Sij1);       /* M_SVij(j,i,:), M_SVij(i,j,:) */
}

static DSys dftb_sys(int na, int cap, const int* nn_size, const int* nn_list, const double* nn_dist, const double* nn_dist_s,
                     const double* lmn, const double* rad, const int* koa, int n_kinds, const xtb_dftb_pair* TB,
                     int basis, const double* h) {
    DSys s; s.na = na; s.cap = cap; s.n_kinds = n_kinds; s.basis = basis; s.n_orb = dftb_n_orb(basis); s.nv = dftb_n_v(basis);
    s.nn_size = nn_size; s.nn_list = nn_list; s.nn_dist = nn_dist; s.nn_dist_s = nn_dist_s; s.lmn = lmn; s.rad = rad;
    s.koa = koa; s.TB = TB; s.h = h;
    return s;
}

/* Hamil_tot_DFTB assembly (TB_DFTB.f90:434-497, OpenMP arm): dense Hij [eV], Sij (N,N) column-major, zero-initialised by the caller */
void orc_dftb_assemble_HS(int na, int cap, const int* nn_size, const int* nn_list, const double* nn_dist, const double* lmn,
                          const double* rad, const int* koa, int n_kinds, const xtb_dftb_pair* TB, int basis,
                          double* Hij, double* Sij) {
    DSys s = dftb_sys(na, cap, nn_size, nn_list, nn_dist, NULL, lmn, rad, koa, n_kinds, TB, basis, NULL);
    const int no = s.n_orb;
    const size_t N = (size_t)na * no;
    const double epsylon = 1e-12;
#pragma omp parallel for schedule(dynamic, 4)
    for (int j = 0; j < na; ++j)
        for (int a2 = 0; a2 <= nn_size[j]; ++a2) {
            int i = a2 == 0 ? j : nn_list[(size_t)j * cap + a2 - 1];
            if (i < j) continue;
            double Hij1[81], Sij1[81];
            dftb_block_HS(&s, j, i, a2, Hij1, Sij1);
            for (int j1 = 0; j1 < no; ++j1) {
                size_t l = (size_t)j * no + j1;
                for (int i1 = 0; i1 < no; ++i1) {
                    size_t k = (size_t)i * no + i1;
                    Hij[k + N * l] = Hij1[i1 + no * j1];
                    double sv = Sij1[i1 + no * j1];
                    if (fabs(sv) <= epsylon) sv = 0.0;
                    Sij[k + N * l] = sv;
                }
            }
        }
#pragma omp parallel for schedule(dynamic, 4)
    for (int j = 0; j < na; ++j)                                /* :478-497 lower triangle by symmetry */
        for (int a2 = 1; a2 <= nn_size[j]; ++a2) {
            int i = nn_list[(size_t)j * cap + a2 - 1];
            if (i >= j) continue;
            for (int j1 = 0; j1 < no; ++j1) {
                size_t l = (size_t)j * no + j1;
                for (int i1 = 0; i1 < no; ++i1) {
                    size_t k = (size_t)i * no + i1;
                    Hij[k + N * l] = Hij[l + N * k];
                    Sij[k + N * l] = Sij[l + N * k];
                }
            }
        }
}

/* d_Hamilton_one_DFTB, TB_DFTB.f90:977-1127.  a2 = 0: on-site (zeros); else list entry a2-1 of atom i (j = neighbour).
 * dH, dS: [alpha][mu + n_orb*nu] */
static void d_Hamilton_one_DFTB(const DSys* s, int k, int i, int j, int a2, double dH[3][81], double dS[3][81]) {
    const int no = s->n_orb, nv = s->nv;
    memset(dH, 0, sizeof(double) * 3 * 81);
    memset(dS, 0, sizeof(double) * 3 * 81);
    if (i == j) return;
    size_t e = (size_t)i * s->cap + (a2 - 1);
    double x1 = s->nn_dist[e * 4], y1 = s->nn_dist[e * 4 + 1], z1 = s->nn_dist[e * 4 + 2], r1 = s->nn_dist[e * 4 + 3];
    double drij_dsk[3], M_dlmn[9];
    for (int a = 1; a <= 3; ++a) drij_dsk[a - 1] = drij_dska(i, j, k, x1, y1, z1, r1, s->h, a);
    for (int c = 1; c <= 3; ++c)
        for (int b = 1; b <= 3; ++b)
            M_dlmn[(c - 1) * 3 + (b - 1)] = ddija_dskb_kd(i, j, k, x1, y1, z1, r1, s->h, c, b, drij_dsk[b - 1]);
    /* vec_M_Vij21 = M_Vij(i,j,:) ; vec_M_Vij12 = M_Vij(j,i,:)   (:1029-1032) */
    const double* V21 = s->rad + e * 40, *dV21r = V21 + 10, *S21 = V21 + 20, *dS21r = V21 + 30;
    const double* V12 = dftb_M(s, j, i, 0), *dV12r = dftb_M(s, j, i, 1), *S12 = dftb_M(s, j, i, 2), *dS12r = dftb_M(s, j, i, 3);
    double l = s->lmn[e * 3], m = s->lmn[e * 3 + 1], n = s->lmn[e * 3 + 2];
    (void)no;
    for (int a = 0; a < 3; ++a) {
        double dV12[10], dV21[10], dS12[10], dS21[10];
        for (int c = 0; c < nv; ++c) {
            dV21[c] = dV21r[c] * drij_dsk[a]; dV12[c] = dV12r[c] * drij_dsk[a];
            dS21[c] = dS21r[c] * drij_dsk[a]; dS12[c] = dS12r[c] * drij_dsk[a];
        }
        double dl = M_dlmn[a], dm = M_dlmn[3 + a], dn = M_dlmn[6 + a];
        if (s->basis == 0) {
            /* the reference passes the VALUE vector to d_KS_s here (TB_DFTB.f90:1047-1048), reproduced as written */
            dH[a][0] = d_t_s_s(V12[0]);
            dS[a][0] = d_t_s_s(S12[0]);
        } else if (s->basis == 1) {
            orc_d_KS_sp3_hetero(V12, V21, dV12, dV21, l, m, n, dl, dm, dn, dH[a]);
            orc_d_KS_sp3_hetero(S12, S21, dS12, dS21, l, m, n, dl, dm, dn, dS[a]);
        } else {
            orc_d_KS_sp3d5_hetero(V12, V21, dV12, dV21, l, m, n, dl, dm, dn, dH[a]);
            orc_d_KS_sp3d5_hetero(S12, S21, dS12, dS21, l, m, n, dl, dm, dn, dS[a]);
        }
    }
}

/* VERBATIM form: get_forces_DFTB (:829-912) + Attract_TB_forces_DFTB (:916-972): dense dH,dS(3,N,N) per atom k */
void orc_dftb_forces_verbatim(int na, int cap, const int* nn_size, const int* nn_list, const double* nn_dist, const double* lmn,
                              const double* rad, const int* koa, int n_kinds, const xtb_dftb_pair* TB, int basis,
                              const double* h, const double* Aij, const double* AijE, double* F /*(3,na)*/) {
    DSys s = dftb_sys(na, cap, nn_size, nn_list, nn_dist, NULL, lmn, rad, koa, n_kinds, TB, basis, h);
    const int n_orb = s.n_orb;
    const size_t N = (size_t)na * n_orb;
#pragma omp parallel
    {
        double* dH = (double*)malloc(sizeof(double) * 3 * N * N);
        double* dS = (double*)malloc(sizeof(double) * 3 * N * N);
#define DH(a, r, c) dH[(a) + 3 * ((r) + N * (c))]
#define DS(a, r, c) dS[(a) + 3 * ((r) + N * (c))]
#pragma omp for schedule(dynamic, 1)
        for (int k = 0; k < na; ++k) {
            memset(dH, 0, sizeof(double) * 3 * N * N);
            memset(dS, 0, sizeof(double) * 3 * N * N);
            for (int i = 0; i < na; ++i) {                       /* :858-883 */
                size_t i4 = (size_t)i * n_orb;
                for (int a2 = 0; a2 <= nn_size[i]; ++a2) {
                    int j = a2 == 0 ? i : nn_list[(size_t)i * cap + a2 - 1];
                    if (j < i) continue;
                    size_t j4 = (size_t)j * n_orb;
                    double dH1[3][81], dS1[3][81];
                    d_Hamilton_one_DFTB(&s, k, i, j, a2, dH1, dS1);
                    for (int j1 = 0; j1 < n_orb; ++j1)
                        for (int i1 = 0; i1 < n_orb; ++i1)
                            for (int a = 0; a < 3; ++a) {
                                DH(a, i4 + i1, j4 + j1) = dH1[a][i1 + n_orb * j1];
                                DS(a, i4 + i1, j4 + j1) = dS1[a][i1 + n_orb * j1];
                            }
                }
            }
            for (int i = 0; i < na; ++i) {                       /* :888-906 */
                size_t i4 = (size_t)i * n_orb;
                for (int a2 = 1; a2 <= nn_size[i]; ++a2) {
                    int j = nn_list[(size_t)i * cap + a2 - 1];
                    if (j >= i) continue;
                    size_t j4 = (size_t)j * n_orb;
                    for (int j1 = 0; j1 < n_orb; ++j1)
                        for (int i1 = 0; i1 < n_orb; ++i1)
                            for (int a = 0; a < 3; ++a) {
                                DH(a, i4 + i1, j4 + j1) = DH(a, j4 + j1, i4 + i1);
                                DS(a, i4 + i1, j4 + j1) = DS(a, j4 + j1, i4 + i1);
                            }
                }
            }
            double E[3] = {0.0, 0.0, 0.0};                       /* Attract_TB_forces_DFTB :931-962 */
            for (size_t i = 0; i < N; ++i) {
                int i2 = (int)(i / n_orb);
                for (int kk = 0; kk <= nn_size[i2]; ++kk) {
                    int j1 = kk == 0 ? i2 : nn_list[(size_t)i2 * cap + kk - 1];
                    size_t j = (size_t)j1 * n_orb;
                    for (int a = 0; a < 3; ++a) {
                        double sum = 0.0;
                        for (int c = 0; c < n_orb; ++c)
                            sum += DH(a, i, j + c) * Aij[i + N * (j + c)] - DS(a, i, j + c) * AijE[i + N * (j + c)];
                        E[a] = E[a] + sum;
                    }
                }
            }
            F[3 * k] = E[0]; F[3 * k + 1] = E[1]; F[3 * k + 2] = E[2];
        }
#undef DH
#undef DS
        free(dH); free(dS);
    }
}

/* SPARSE-EQUIVALENT forces: the same per-block formulas; for atom k only the pair blocks with k in {i,j} are non-zero
 * (on-site derivatives vanish, :999-1001).  rho_blk / rhoE_blk: layout of orc_density_blocks. */
void orc_dftb_forces_sparse(int na, int cap, const int* nn_size, const int* nn_list, const double* nn_dist, const double* lmn,
                            const double* rad, const int* koa, int n_kinds, const xtb_dftb_pair* TB, int basis,
                            const double* h, const double* rho_blk, const double* rhoE_blk, int atom_first, int atom_last,
                            double* F /*(3,na)*/) {
    DSys s = dftb_sys(na, cap, nn_size, nn_list, nn_dist, NULL, lmn, rad, koa, n_kinds, TB, basis, h);
    const int no = s.n_orb, nb2 = no * no;
#pragma omp parallel for schedule(dynamic, 1)
    for (int k = atom_first; k < atom_last; ++k) {
        double E[3] = {0.0, 0.0, 0.0};
        for (int q = 0; q < nn_size[k]; ++q) {
            int b = nn_list[(size_t)k * cap + q];
            /* the block is evaluated from the list of the smaller index atom i for the larger j (:868) */
            int i = k < b ? k : b, j = k < b ? b : k;
            int a2 = -1, a2r = -1;
            for (int t = 0; t < nn_size[i]; ++t) if (nn_list[(size_t)i * cap + t] == j) { a2 = t + 1; break; }
            for (int t = 0; t < nn_size[j]; ++t) if (nn_list[(size_t)j * cap + t] == i) { a2r = t + 1; break; }
            if (a2 < 0) continue;
            double dH1[3][81], dS1[3][81];
            d_Hamilton_one_DFTB(&s, k, i, j, a2, dH1, dS1);
            const double* Aij = rho_blk + ((size_t)i * (cap + 1) + a2) * nb2;     /* Aij(i-orbitals, j-orbitals) */
            const double* AijE = rhoE_blk + ((size_t)i * (cap + 1) + a2) * nb2;
            for (int a = 0; a < 3; ++a) {
                double sum = 0.0;
                for (int j1 = 0; j1 < no; ++j1)
                    for (int i1 = 0; i1 < no; ++i1)
                        sum += dH1[a][i1 + no * j1] * Aij[i1 + no * j1] - dS1[a][i1 + no * j1] * AijE[i1 + no * j1];
                E[a] += sum;
            }
            if (a2r > 0) {   /* mirrored block dH(j-orbitals, i-orbitals) = dH(i-orb, j-orb)  contracted with Aij(j-orb, i-orb) */
                const double* Aji = rho_blk + ((size_t)j * (cap + 1) + a2r) * nb2;
                const double* AjiE = rhoE_blk + ((size_t)j * (cap + 1) + a2r) * nb2;
                for (int a = 0; a < 3; ++a) {
                    double sum = 0.0;
                    for (int j1 = 0; j1 < no; ++j1)
                        for (int i1 = 0; i1 < no; ++i1)
                            sum += dH1[a][i1 + no * j1] * Aji[j1 + no * i1] - dS1[a][i1 + no * j1] * AjiE[j1 + no * i1];
                    E[a] += sum;
                }
            }
        }
        F[3 * k] = E[0]; F[3 * k + 1] = E[1]; F[3 * k + 2] = E[2];
    }
}

/* dHopping_Press_DFTB, TB_DFTB.f90:1339-1420: directed entry (i, atom_2 >= 1); dH, dS: [k1][mu + n_orb*nu] */
static void dHopping_Press_DFTB(const DSys* s, int i, int a2, double dH[9][81], double dS[9][81]) {
    const int nv = s->nv;
    memset(dH, 0, sizeof(double) * 9 * 81);
    memset(dS, 0, sizeof(double) * 9 * 81);
    size_t e = (size_t)i * s->cap + (a2 - 1);
    int j = s->nn_list[e];
    double rij[3] = {s->nn_dist[e * 4], s->nn_dist[e * 4 + 1], s->nn_dist[e * 4 + 2]};
    double r = s->nn_dist[e * 4 + 3];
    double sij[3] = {s->nn_dist_s[e * 3], s->nn_dist_s[e * 3 + 1], s->nn_dist_s[e * 3 + 2]};
    const double* V21 = s->rad + e * 40, *dV21r = V21 + 10, *S21 = V21 + 20, *dS21r = V21 + 30;      /* M_X(i,j,:)  :1372-1375 */
    const double* V12 = dftb_M(s, j, i, 0), *dV12r = dftb_M(s, j, i, 1), *S12 = dftb_M(s, j, i, 2), *dS12r = dftb_M(s, j, i, 3);
    double l = s->lmn[e * 3], m = s->lmn[e * 3 + 1], n = s->lmn[e * 3 + 2];
    for (int i1 = 1; i1 <= 3; ++i1)          /* gamma */
        for (int j1 = 1; j1 <= 3; ++j1) {    /* delta */
            int k1 = (i1 - 1) * 3 + j1 - 1;
            double M_dlmn[3];
            for (int c = 1; c <= 3; ++c) M_dlmn[c - 1] = dda_dhgd(c, j1, rij[c - 1], rij[j1 - 1], sij[i1 - 1], r);
            double drij_dh = drij_dhab(rij[j1 - 1], sij[i1 - 1], r);
            double dV12[10], dV21[10], dS12[10], dS21[10];
            for (int c = 0; c < nv; ++c) {
                dV21[c] = dV21r[c] * drij_dh; dV12[c] = dV12r[c] * drij_dh;
                dS21[c] = dS21r[c] * drij_dh; dS12[c] = dS12r[c] * drij_dh;
            }
            if (s->basis == 0) {             /* value vector passed to d_KS_s, as written (:1393-1394) */
                dH[k1][0] = d_t_s_s(V12[0]);
                dS[k1][0] = d_t_s_s(S12[0]);
            } else if (s->basis == 1) {
                orc_d_KS_sp3_hetero(V12, V21, dV12, dV21, l, m, n, M_dlmn[0], M_dlmn[1], M_dlmn[2], dH[k1]);
                orc_d_KS_sp3_hetero(S12, S21, dS12, dS21, l, m, n, M_dlmn[0], M_dlmn[1], M_dlmn[2], dS[k1]);
            } else {
                orc_d_KS_sp3d5_hetero(V12, V21, dV12, dV21, l, m, n, M_dlmn[0], M_dlmn[1], M_dlmn[2], dH[k1]);
                orc_d_KS_sp3d5_hetero(S12, S21, dS12, dS21, l, m, n, M_dlmn[0], M_dlmn[1], M_dlmn[2], dS[k1]);
            }
        }
}

/* Attract_TB_Forces_Press_DFTB (:1135-1199) + dHamil_tot_Press_DFTB (:1203-1290), evaluated block by block:
 * dHij(:, i-orbitals, j-orbitals) = dHij1 for every directed entry (on-site blocks are zero, :1318-1320), then
 * dwr_press(k1, row) = sum_cols dHij(k1,row,:)*Aij(row,:) and SCforce%att(k,i) = sum_rows (dwr - dS) with k1 = (i-1)*3+k.
 * att is (3,3) column-major: att[k + 3*i]. */
void orc_dftb_pressure(int na, int cap, const int* nn_size, const int* nn_list, const double* nn_dist, const double* nn_dist_s,
                       const double* lmn, const double* rad, const int* koa, int n_kinds, const xtb_dftb_pair* TB, int basis,
                       const double* rho_blk, const double* rhoE_blk, int atom_first, int atom_last, double* att) {
    DSys s = dftb_sys(na, cap, nn_size, nn_list, nn_dist, nn_dist_s, lmn, rad, koa, n_kinds, TB, basis, NULL);
    const int no = s.n_orb, nb2 = no * no;
    double tot[9] = {0};
    for (int i = atom_first; i < atom_last; ++i)
        for (int a2 = 1; a2 <= nn_size[i]; ++a2) {
            double dH[9][81], dS[9][81];
            dHopping_Press_DFTB(&s, i, a2, dH, dS);
            const double* A = rho_blk + ((size_t)i * (cap + 1) + a2) * nb2;
            const double* AE = rhoE_blk + ((size_t)i * (cap + 1) + a2) * nb2;
            for (int k1 = 0; k1 < 9; ++k1) {
                double sum = 0.0;
                for (int q = 0; q < nb2; ++q) sum += dH[k1][q] * A[q] - dS[k1][q] * AE[q];
                tot[k1] += sum;
            }
        }
    /* k1 = (i-1)*3 + k  ->  att(k,i): identical flat index k + 3*i */
    for (int k1 = 0; k1 < 9; ++k1) att[k1] = tot[k1];
}
