/* Plain-C caller of the C ABI (include/xtb_b200.h), making the call sequence the Fortran shim makes for one MD step of the
 * NRL branch (xtant-3_b200/fortran/xtant_b200_shim.f90; reference caller get_Hamilonian_and_E, Source_files/TB.f90:227-272):
 *     xtb_create -> xtb_set_params_nrl -> xtb_set_system -> xtb_neighbours -> xtb_assemble_HS (host H_non, Sij)
 *     -> [eigensolve outside: here xtb_eigensolve + xtb_fermi_occupations, or host arrays read from a file]
 *     -> xtb_density (once: rho and rho_E) -> xtb_forces -> xtb_pressure
 * and then the same step as ONE call (xtb_step), comparing the two bit for bit.  With -DXTB_DRIVER_MPI_LIKE=<n> the program
 * forks n processes, one per GPU, that exchange the NCCL id through a file exactly as MPI ranks would through MPI_Bcast, and
 * run the partitioned step (xtb_comm_init + xtb_step); rank 0 compares the gathered forces with its own single-GPU run.
 * No Fortran compiler exists on either machine, so this is the closest thing to compiling the shim's sequence.
 *
 * usage: abi_driver <params.bin> <system.bin> [nranks]
 *   params.bin = one xtb_nrl_pair;  system.bin = int32 na, double supce[9], double S[3*na], double n_electrons, double Te
 * prints "OK key=value ..." lines; exit code 0 only if every check passed. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <unistd.h>
#include <sys/wait.h>
#include "../include/xtb_b200.h"

#define CK(call) do { int rc_ = (call); if (rc_ != 0) { fprintf(stderr, "FAIL %s -> %d: %s\n", #call, rc_, ctx ? xtb_last_error(ctx) : "?"); return 10; } } while (0)

typedef struct { int na; double supce[9]; double* S; double ne, te; xtb_nrl_pair p; } problem;

static int load(const char* fp, const char* fs, problem* q) {
    FILE* f = fopen(fp, "rb");
    if (!f || fread(&q->p, sizeof q->p, 1, f) != 1) return 1;
    fclose(f);
    f = fopen(fs, "rb");
    if (!f || fread(&q->na, 4, 1, f) != 1 || fread(q->supce, 8, 9, f) != 9) return 2;
    q->S = (double*)malloc(sizeof(double) * 3 * q->na);
    if (fread(q->S, 8, 3 * (size_t)q->na, f) != 3 * (size_t)q->na) return 3;
    if (fread(&q->ne, 8, 1, f) != 1 || fread(&q->te, 8, 1, f) != 1) return 4;
    fclose(f);
    return 0;
}

/* an artificial Lennard-Jones pair table for the metal cell of the driver (type(TB_vdW_LJ_cut), Objects.f90:336-343) */
static xtb_vdw_pair vdw_table(void) {
    xtb_vdw_pair W;
    memset(&W, 0, sizeof W);
    W.eps = 0.05; W.r0 = 3.2; W.n = 6.0; W.d0_short = 2.0; W.dd_short = 0.03; W.d0_cut = 5.0; W.dd_cut = 0.1;
    return W;
}

static int single(const problem* q, double* F_out, double* P_out) {
    xtb_ctx* ctx = NULL;
    const int na = q->na, N = 9 * na;
    int* koa = (int*)malloc(sizeof(int) * na);
    for (int i = 0; i < na; ++i) koa[i] = 1;
    CK(xtb_create(&ctx, 0));
    CK(xtb_set_params_nrl(ctx, 1, &q->p));
    CK(xtb_set_system(ctx, na, koa, q->S, q->supce));
    /* the shim's sequence */
    int* nn_size = (int*)malloc(sizeof(int) * na);
    CK(xtb_neighbours(ctx, 0.0, nn_size, NULL, NULL, NULL));
    double* H = (double*)malloc(sizeof(double) * (size_t)N * N);
    double* S = (double*)malloc(sizeof(double) * (size_t)N * N);
    CK(xtb_assemble_HS(ctx, H, S));
    double asym = 0.0, diag = 0.0;
    for (int i = 0; i < N; ++i) {
        diag = fmax(diag, fabs(S[i + (size_t)N * i] - 1.0));
        for (int j = 0; j < i; ++j) asym = fmax(asym, fabs(H[i + (size_t)N * j] - H[j + (size_t)N * i]));
    }
    double* Ei = (double*)malloc(sizeof(double) * N);
    double* fe = (double*)malloc(sizeof(double) * N);
    double mu = 0.0, eb = 0.0;
    CK(xtb_eigensolve(ctx, Ei, NULL, NULL));
    CK(xtb_fermi_occupations(ctx, q->ne, q->te, fe, &mu, &eb));
    double nel = 0.0;
    for (int i = 0; i < N; ++i) nel += fe[i];
    CK(xtb_density(ctx, NULL, NULL));
    const long long l0 = xtb_counter(ctx, 6);
    double* F = (double*)malloc(sizeof(double) * 3 * na);
    double P[9];
    CK(xtb_forces(ctx, F));
    CK(xtb_pressure(ctx, P));
    const long long launches_forces = xtb_counter(ctx, 6) - l0;
    double fsum[3] = {0, 0, 0};
    for (int i = 0; i < na; ++i) for (int a = 0; a < 3; ++a) fsum[a] += F[3 * i + a];
    printf("OK staged na=%d N=%d max_nn=%lld blocks=%lld H_asym=%.3g S_diag_err=%.3g mu=%.10f E_band=%.10f Ne=%.12f launches_forces=%lld\n",
           na, N, xtb_counter(ctx, 3), xtb_counter(ctx, 4), asym, diag, mu, eb, nel, launches_forces);
    int bad = !(asym == 0.0) || !(diag == 0.0) || !(fabs(nel - q->ne) <= 1e-11 * q->ne);
    /* the same step as one call: occupations from the device eigensolve again */
    double* F2 = (double*)malloc(sizeof(double) * 3 * na);
    double P2[9], mu2 = 0.0, eb2 = 0.0;
    xtb_step_io io;
    memset(&io, 0, sizeof io);
    io.S = q->S; io.n_electrons = q->ne; io.Te_eV = q->te; io.F_att = F2; io.SCforce_att = P2; io.mu = &mu2; io.E_band = &eb2;
    for (int rep = 0; rep < 2; ++rep) {       /* sizing form, then steady-state form */
        CK(xtb_step(ctx, XTB_STEP_ALL | XTB_STEP_EIGEN | XTB_STEP_PRESSURE, &io));
        double dF = 0.0, dP = 0.0;
        for (int i = 0; i < 3 * na; ++i) dF = fmax(dF, fabs(F2[i] - F[i]));
        for (int i = 0; i < 9; ++i) dP = fmax(dP, fabs(P2[i] - P[i]));
        printf("OK step rep=%d dF=%.3g dP=%.3g dmu=%.3g step_ms: list=%.3f HS=%.3f eig=%.3f rho=%.3f F=%.3f\n", rep, dF, dP, fabs(mu2 - mu),
               xtb_stage_ms(ctx, 1), xtb_stage_ms(ctx, 2), xtb_stage_ms(ctx, 8), xtb_stage_ms(ctx, 3), xtb_stage_ms(ctx, 4));
        bad |= !(dF == 0.0) || !(dP == 0.0) || !(mu2 == mu);
    }
    if (F_out) memcpy(F_out, F, sizeof(double) * 3 * na);
    if (P_out) memcpy(P_out, P, sizeof P);
    /* a pair term that does not use the list: van der Waals (here a Lennard-Jones table on the same cell), staged calls vs
     * the fused step with XTB_STEP_PAIR_TERMS, bit for bit; F_out[3 na ..), P_out[9 ..] carry it to the partitioned run */
    {
        const xtb_vdw_pair W = vdw_table();
        double Ev = 0.0, Ev2 = 0.0, Pv[9], Pv2[9];
        double* Fv = (double*)malloc(sizeof(double) * 3 * na);
        double* Fv2 = (double*)malloc(sizeof(double) * 3 * na);
        int Nc[3]; double Rc = 0.0;
        CK(xtb_set_params_vdw(ctx, XTB_VDW_LJ_CUT, 1, &W));
        CK(xtb_vdw_image_cells(ctx, Nc, &Rc));
        CK(xtb_vdw_energy(ctx, &Ev, NULL));
        CK(xtb_vdw_forces(ctx, Fv));
        CK(xtb_vdw_pressure(ctx, Pv));
        io.E_vdw = &Ev2; io.F_vdw = Fv2; io.SCforce_vdw = Pv2;
        CK(xtb_step(ctx, XTB_STEP_ALL | XTB_STEP_EIGEN | XTB_STEP_PRESSURE | XTB_STEP_PAIR_TERMS, &io));
        double dF = 0.0, dP = 0.0, fs[3] = {0, 0, 0}, fmx = 0.0;
        for (int i = 0; i < 3 * na; ++i) { dF = fmax(dF, fabs(Fv2[i] - Fv[i])); fs[i % 3] += Fv[i]; fmx = fmax(fmx, fabs(Fv[i])); }
        for (int i = 0; i < 9; ++i) dP = fmax(dP, fabs(Pv2[i] - Pv[i]));
        printf("OK vdw cells=%d,%d,%d R_cut=%.3f E=%.12g Fmax=%.6g Fsum=%.3g dE=%.3g dF=%.3g dP=%.3g ms=%.3f\n", Nc[0], Nc[1], Nc[2], Rc, Ev, fmx,
               fmax(fabs(fs[0]), fmax(fabs(fs[1]), fabs(fs[2]))), fabs(Ev2 - Ev), dF, dP, xtb_stage_ms(ctx, 17));
        bad |= !(Ev2 == Ev) || !(dF == 0.0) || !(dP == 0.0) || !(fmx > 0.0) || !(fmax(fabs(fs[0]), fmax(fabs(fs[1]), fabs(fs[2]))) <= 1e-10 * fmx * na);
        if (F_out) memcpy(F_out + 3 * na, Fv, sizeof(double) * 3 * na);
        if (P_out) { memcpy(P_out + 9, Pv, sizeof Pv); P_out[18] = Ev; }
        free(Fv); free(Fv2);
    }
    /* error path: a stage before its inputs -> XTB_ESTATE with a message, as the shim forwards to Save_error_details */
    xtb_ctx* c2 = NULL;
    xtb_create(&c2, 0);
    bad |= xtb_density(c2, NULL, NULL) != XTB_ESTATE || strlen(xtb_last_error(c2)) == 0;
    xtb_destroy(c2);
    CK(xtb_destroy(ctx));
    free(koa); free(nn_size); free(H); free(S); free(Ei); free(fe); free(F); free(F2);
    return bad ? 20 : 0;
}

/* one "MPI rank": device = rank, id through a file written by rank 0 */
static int rank_main(const problem* q, int rank, int nranks, const char* idfile, const double* Fref, const double* Pref) {
    xtb_ctx* ctx = NULL;
    const int na = q->na;
    int* koa = (int*)malloc(sizeof(int) * na);
    for (int i = 0; i < na; ++i) koa[i] = 1;
    char id[XTB_UNIQUE_ID_BYTES];
    if (rank == 0) {
        if (xtb_comm_unique_id(id) != 0) { fprintf(stderr, "FAIL xtb_comm_unique_id\n"); return 11; }
        char tmp[512]; snprintf(tmp, sizeof tmp, "%s.tmp", idfile);
        FILE* f = fopen(tmp, "wb"); fwrite(id, 1, sizeof id, f); fclose(f); rename(tmp, idfile);
    } else {
        FILE* f = NULL;
        for (int t = 0; t < 600 && !(f = fopen(idfile, "rb")); ++t) usleep(100000);
        if (!f || fread(id, 1, sizeof id, f) != sizeof id) { fprintf(stderr, "FAIL rank %d: no id file\n", rank); return 12; }
        fclose(f);
    }
    CK(xtb_create(&ctx, rank));
    CK(xtb_set_params_nrl(ctx, 1, &q->p));
    CK(xtb_set_system(ctx, na, koa, q->S, q->supce));
    CK(xtb_comm_init(ctx, id, rank, nranks));
    const xtb_vdw_pair W = vdw_table();
    CK(xtb_set_params_vdw(ctx, XTB_VDW_LJ_CUT, 1, &W));
    double* F = (double*)malloc(sizeof(double) * 3 * na);
    double* Fv = (double*)malloc(sizeof(double) * 3 * na);
    double P[9], Pv[9], Ev = 0.0;
    xtb_step_io io;
    memset(&io, 0, sizeof io);
    io.n_electrons = q->ne; io.Te_eV = q->te; io.F_att = F; io.SCforce_att = P; io.E_vdw = &Ev; io.F_vdw = Fv; io.SCforce_vdw = Pv;
    int bad = 0;
    for (int rep = 0; rep < 2; ++rep) {
        CK(xtb_step(ctx, XTB_STEP_ALL | XTB_STEP_EIGEN | XTB_STEP_PRESSURE | XTB_STEP_PAIR_TERMS, &io));
        double dF = 0.0, dP = 0.0, dFv = 0.0, dPv = 0.0;
        for (int i = 0; i < 3 * na; ++i) { dF = fmax(dF, fabs(F[i] - Fref[i])); dFv = fmax(dFv, fabs(Fv[i] - Fref[3 * na + i])); }
        for (int i = 0; i < 9; ++i) {
            dP = fmax(dP, fabs(P[i] - Pref[i]) / (fabs(Pref[i]) + 1e-300));
            dPv = fmax(dPv, fabs(Pv[i] - Pref[9 + i]) / (fabs(Pref[9]) + 1e-300));
        }
        printf("OK rank %d/%d rep=%d dF=%.3g dP_rel=%.3g dF_vdw=%.3g dP_vdw_rel=%.3g dE_vdw=%.3g halo_send_blocks=%lld halo_recv_blocks=%lld xchg_HS_bytes=%lld xchg_rho_bytes=%lld\n",
               rank, nranks, rep, dF, dP, dFv, dPv, fabs(Ev - Pref[18]), xtb_counter(ctx, 14), xtb_counter(ctx, 15), xtb_counter(ctx, 12), xtb_counter(ctx, 13));
        /* every block / atom has one owner: forces bit-identical; the 9-vectors are summed in another order; E_vdw is replicated */
        bad |= !(dF == 0.0) || !(dP <= 1e-13) || !(dFv == 0.0) || !(dPv <= 1e-13) || !(Ev == Pref[18]);
    }
    free(Fv);
    CK(xtb_comm_destroy(ctx));
    CK(xtb_destroy(ctx));
    free(koa); free(F);
    return bad ? 21 : 0;
}

int main(int argc, char** argv) {
    if (argc < 3) { fprintf(stderr, "usage: %s params.bin system.bin [nranks]\n", argv[0]); return 2; }
    problem q;
    int rc = load(argv[1], argv[2], &q);
    if (rc) { fprintf(stderr, "FAIL reading inputs (%d)\n", rc); return 3; }
    const int nranks = argc > 3 ? atoi(argv[3]) : 1;
    double* Fref = (double*)malloc(sizeof(double) * 6 * q.na);      /* attractive | van der Waals */
    double Pref[19];                                                /* attractive | van der Waals | E_vdW */
    if (nranks <= 1) {
        rc = single(&q, Fref, Pref);
        if (rc == 0) printf("OK all\n");
        return rc;
    }
    /* reference result in a child process (CUDA must not be initialised before fork) */
    int pfd[2];
    if (pipe(pfd) != 0) return 4;
    fflush(stdout);
    pid_t pid = fork();
    if (pid == 0) {
        close(pfd[0]);
        int r = single(&q, Fref, Pref);
        if (r == 0) { if (write(pfd[1], Fref, sizeof(double) * 6 * q.na) < 0 || write(pfd[1], Pref, sizeof Pref) < 0) r = 5; }
        fflush(stdout);
        _exit(r);
    }
    close(pfd[1]);
    size_t got = 0, want = sizeof(double) * 6 * q.na;
    while (got < want) { ssize_t n = read(pfd[0], (char*)Fref + got, want - got); if (n <= 0) break; got += (size_t)n; }
    size_t gotp = 0;
    while (gotp < sizeof Pref) { ssize_t n = read(pfd[0], (char*)Pref + gotp, sizeof Pref - gotp); if (n <= 0) break; gotp += (size_t)n; }
    int st = 0;
    waitpid(pid, &st, 0);
    if (!WIFEXITED(st) || WEXITSTATUS(st) != 0 || got != want) { fprintf(stderr, "FAIL single-GPU reference run\n"); return 6; }
    char idfile[256];
    snprintf(idfile, sizeof idfile, "/tmp/xtb_nccl_id_%d", (int)getpid());
    unlink(idfile);
    fflush(stdout);
    pid_t kids[16];
    for (int r = 0; r < nranks && r < 16; ++r) {
        kids[r] = fork();
        if (kids[r] == 0) { int rr = rank_main(&q, r, nranks, idfile, Fref, Pref); fflush(stdout); _exit(rr); }
    }
    int bad = 0;
    for (int r = 0; r < nranks && r < 16; ++r) { waitpid(kids[r], &st, 0); bad |= !WIFEXITED(st) || WEXITSTATUS(st) != 0; }
    unlink(idfile);
    if (!bad) printf("OK all\n");
    return bad ? 7 : 0;
}
