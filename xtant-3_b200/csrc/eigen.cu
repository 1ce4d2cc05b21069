// Eigensolve glue on the device -- the step BETWEEN assembly and density that the graft does
// not own.  It exists so that an end-to-end MD step never moves the N x N matrices over PCIe: H_non and Sij stay where
// xtb_assemble_HS left them, the eigenvectors stay where xtb_density reads them.
//
// Same algebra as the reference, with cuSOLVER / cuBLAS library calls in place of LAPACK / MKL:
//   Loewdin_Orthogonalization   Source_files/TB_NRL.f90:985-1060   dsyev(S) -> w, U ; negative w -> |w| ;
//                                                                   X = U diag(1/sqrt(w), 0 where |w| <= 1e-8) ;
//                                                                   H' = X^T (H X)
//   sym_diagonalize(H')         Source_files/Algebra_tools.f90:882-966 (dsyevd)  -> Ei, V
//   Ha = X V                    Source_files/TB_NRL.f90:421 / TB_DFTB.f90:535
//   DFTB only: |H'| < 1e-10 -> 0 before the diagonalisation (TB_DFTB.f90:517-519)
//   orthogonal branch (Pettifor): sym_diagonalize(H) directly (TB_Pettifor.f90:1455)
// The libraries are bound at run time (dlopen), so libxtb_b200.so itself has no link-time dependency on them; when they
// are missing xtb_eigensolve reports XTB_ESTATE and the caller keeps using xtb_set_eigen with host LAPACK eigenvectors.
#include "xtb_ctx.h"
#include <cstdlib>
#include <dlfcn.h>

namespace {

typedef void* handle_t;
typedef int (*fn_create_t)(handle_t*);
typedef int (*fn_destroy_t)(handle_t);
typedef int (*fn_setstream_t)(handle_t, cudaStream_t);
typedef int (*fn_syevd_buf_t)(handle_t, int /*jobz*/, int /*uplo*/, int n, const double* A, int lda, const double* W, int* lwork);
typedef int (*fn_syevd_t)(handle_t, int, int, int n, double* A, int lda, double* W, double* work, int lwork, int* info);
typedef int (*fn_dgemm_t)(handle_t, int ta, int tb, int m, int n, int k, const double* alpha, const double* A, int lda,
                          const double* B, int ldb, const double* beta, double* Cm, int ldc);
// 64-bit interface (cusolverDnXsyevd): workspace sizes in bytes as size_t.  The legacy Dsyevd counts its workspace in an int,
// which the ~2 N^2 doubles of the eigenvector job overflow from N ~ 32 000 on -- below the north-star N = 36 864.
typedef int (*fn_params_create_t)(void**);
typedef int (*fn_params_destroy_t)(void*);
typedef int (*fn_xsyevd_buf_t)(handle_t, void* params, int jobz, int uplo, int64_t n, int typeA, const void* A, int64_t lda, int typeW,
                               const void* W, int compute, size_t* dev_bytes, size_t* host_bytes);
typedef int (*fn_xsyevd_t)(handle_t, void* params, int jobz, int uplo, int64_t n, int typeA, void* A, int64_t lda, int typeW, void* W,
                           int compute, void* dev_buf, size_t dev_bytes, void* host_buf, size_t host_bytes, int* info);

struct Libs {
    bool tried = false, ok = false;
    std::string why;
    void* so_solver = nullptr; void* so_blas = nullptr;
    fn_create_t sv_create = nullptr; fn_destroy_t sv_destroy = nullptr; fn_setstream_t sv_stream = nullptr;
    fn_syevd_buf_t sv_buf = nullptr; fn_syevd_t sv_syevd = nullptr;
    fn_create_t bl_create = nullptr; fn_destroy_t bl_destroy = nullptr; fn_setstream_t bl_stream = nullptr; fn_dgemm_t bl_dgemm = nullptr;
    fn_params_create_t x_create = nullptr; fn_params_destroy_t x_destroy = nullptr; fn_xsyevd_buf_t x_buf = nullptr; fn_xsyevd_t x_syevd = nullptr;
    bool x_ok = false;      // the 64-bit interface is present (optional: older libraries only carry the legacy one)
};
Libs g_libs;            // the two libraries and their entry points (process-wide); the HANDLES live in the context

void* open_any(const char* const* names) {
    for (int i = 0; names[i]; ++i) {
        void* h = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
        if (h) return h;
    }
    return nullptr;
}

bool bind_libs() {
    Libs& L = g_libs;
    if (L.tried) return L.ok;
    L.tried = true;
    static const char* const solver[] = {"libcusolver.so.11", "libcusolver.so", "/usr/local/cuda/lib64/libcusolver.so.11", nullptr};
    static const char* const blas[] = {"libcublas.so.12", "libcublas.so", "/usr/local/cuda/lib64/libcublas.so.12", nullptr};
    L.so_blas = open_any(blas);
    L.so_solver = open_any(solver);
    if (!L.so_solver || !L.so_blas) { L.why = "cuSOLVER / cuBLAS shared libraries not found"; return false; }
#define BIND(field, so, type, name) L.field = (type)dlsym(L.so, name); if (!L.field) { L.why = std::string("missing symbol ") + name; return false; }
    BIND(sv_create, so_solver, fn_create_t, "cusolverDnCreate")
    BIND(sv_destroy, so_solver, fn_destroy_t, "cusolverDnDestroy")
    BIND(sv_stream, so_solver, fn_setstream_t, "cusolverDnSetStream")
    BIND(sv_buf, so_solver, fn_syevd_buf_t, "cusolverDnDsyevd_bufferSize")
    BIND(sv_syevd, so_solver, fn_syevd_t, "cusolverDnDsyevd")
    BIND(bl_create, so_blas, fn_create_t, "cublasCreate_v2")
    BIND(bl_destroy, so_blas, fn_destroy_t, "cublasDestroy_v2")
    BIND(bl_stream, so_blas, fn_setstream_t, "cublasSetStream_v2")
    BIND(bl_dgemm, so_blas, fn_dgemm_t, "cublasDgemm_v2")
#undef BIND
    L.x_create = (fn_params_create_t)dlsym(L.so_solver, "cusolverDnCreateParams");
    L.x_destroy = (fn_params_destroy_t)dlsym(L.so_solver, "cusolverDnDestroyParams");
    L.x_buf = (fn_xsyevd_buf_t)dlsym(L.so_solver, "cusolverDnXsyevd_bufferSize");
    L.x_syevd = (fn_xsyevd_t)dlsym(L.so_solver, "cusolverDnXsyevd");
    L.x_ok = L.x_create && L.x_destroy && L.x_buf && L.x_syevd;
    L.ok = true;
    return true;
}

// cuSOLVER / cuBLAS handles of this context: created on the context's device at first use, destroyed with it
// (xtb_release_lib_handles, called by xtb_destroy), so two contexts on two GPUs never share a handle
int ctx_handles(xtb_ctx* c) {
    Libs& L = g_libs;
    cudaSetDevice(c->device);
    if (!c->h_cusolver && L.sv_create(&c->h_cusolver) != 0) XTB_FAIL(c, XTB_ECUDA, "cusolverDnCreate failed");
    if (!c->h_cublas && L.bl_create(&c->h_cublas) != 0) XTB_FAIL(c, XTB_ECUDA, "cublasCreate failed");
    return XTB_OK;
}

constexpr int JOB_VECTOR = 1;     // CUSOLVER_EIG_MODE_VECTOR
constexpr int FILL_UPPER = 1;     // CUBLAS_FILL_MODE_UPPER
constexpr int OP_N = 0, OP_T = 1; // CUBLAS_OP_N / CUBLAS_OP_T

// X(:, i) *= s_i with s_i = 1/sqrt(|w_i|) if |w_i| > eps else 0; w <- |w| when the spectrum has a negative / tiny value
__global__ void loewdin_scale_kernel(long long N, const double* __restrict__ w, const int* __restrict__ fix, double* __restrict__ X) {
    const long long col = blockIdx.y;
    double ev = w[col];
    if (*fix && ev < 0.0) ev = fabs(ev);                        // TB_NRL.f90:1022-1026
    const double s = fabs(ev) > 1e-8 ? 1.0 / sqrt(ev) : 0.0;     // :1031
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < N; r += (long long)gridDim.x * blockDim.x)
        X[r + N * col] *= s;
}
// fix = 1 iff any eigenvalue is negative or |w| < eps (N_neg > 0 or N_zero > 0, :1019-1026)
__global__ void loewdin_flag_kernel(long long N, const double* __restrict__ w, int* __restrict__ fix) {
    __shared__ int any;
    if (threadIdx.x == 0) any = 0;
    __syncthreads();
    for (long long i = threadIdx.x; i < N; i += blockDim.x)
        if (w[i] < 0.0 || fabs(w[i]) < 1e-8) any = 1;
    __syncthreads();
    if (threadIdx.x == 0) *fix = any;
}
// out(r, n) = in(n, r): 32 x 32 tiles through shared memory, both sides coalesced.  in: ld_in = rows of `in` (its fast index)
__global__ void __launch_bounds__(256) transpose_kernel(long long n_fast_in, long long n_slow_in, const double* __restrict__ in, double* __restrict__ out) {
    __shared__ double tile[32][33];
    const long long f0 = 32LL * blockIdx.x, s0 = 32LL * blockIdx.y;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int yy = ty; yy < 32; yy += 8) {
        const long long f = f0 + tx, sl = s0 + yy;
        tile[yy][tx] = (f < n_fast_in && sl < n_slow_in) ? in[f + n_fast_in * sl] : 0.0;
    }
    __syncthreads();
    for (int yy = ty; yy < 32; yy += 8) {
        const long long sl = s0 + tx, f = f0 + yy;          // out fast index = in slow index
        if (sl < n_slow_in && f < n_fast_in) out[sl + n_slow_in * f] = tile[tx][yy];
    }
}
__global__ void chop_kernel(long long n, double* __restrict__ A, double eps) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        if (fabs(A[i]) < eps) A[i] = 0.0;
}

// legacy interface up to LEGACY_MAX_N (its int workspace count still holds there), the 64-bit one beyond; XTB_EIGEN_X64=1 forces
// the 64-bit interface at any size (tests)
constexpr int LEGACY_MAX_N = 30000;
int syevd_x64(xtb_ctx* c, int n, double* A, double* W, const std::string& what) {
    Libs& L = g_libs;
    constexpr int R_64F = 1;      // CUDA_R_64F
    void* params = nullptr;
    if (L.x_create(&params) != 0 || !params) XTB_FAIL(c, XTB_ECUDA, "cusolverDnCreateParams failed (%s)", what.c_str());
    size_t dev_bytes = 0, host_bytes = 0;
    const int stb = L.x_buf(c->h_cusolver, params, JOB_VECTOR, FILL_UPPER, (int64_t)n, R_64F, A, (int64_t)n, R_64F, W, R_64F, &dev_bytes, &host_bytes);
    if (stb != 0) {
        L.x_destroy(params);
        XTB_FAIL(c, XTB_ECUDA, "cusolverDnXsyevd_bufferSize failed (%s): status %d at n = %d", what.c_str(), stb, n);
    }
    const size_t info_off = (dev_bytes + 15) & ~(size_t)15;
    int rc;
    if ((rc = xtb_reserve(c, c->d_eig_work, info_off + 16))) { L.x_destroy(params); return rc; }
    std::vector<char> host_buf(host_bytes ? host_bytes : 1);
    int* d_info = (int*)((char*)c->d_eig_work.p + info_off);
    const int st = L.x_syevd(c->h_cusolver, params, JOB_VECTOR, FILL_UPPER, (int64_t)n, R_64F, A, (int64_t)n, R_64F, W, R_64F,
                             c->d_eig_work.p, dev_bytes, host_buf.data(), host_bytes, d_info);
    int info = 0;
    cudaError_t e = cudaMemcpyAsync(&info, d_info, sizeof(int), cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);      // also keeps host_buf alive until the solver is done with it
    L.x_destroy(params);
    if (st != 0) XTB_FAIL(c, XTB_ECUDA, "cusolverDnXsyevd failed to launch (%s): status %d", what.c_str(), st);
    if (e != cudaSuccess) { cudaGetLastError(); XTB_FAIL(c, XTB_ECUDA, "cusolverDnXsyevd (%s): %s", what.c_str(), cudaGetErrorString(e)); }
    if (info != 0) XTB_FAIL(c, XTB_ECUDA, "Xsyevd(%s): info = %d", what.c_str(), info);
    return XTB_OK;
}

int syevd(xtb_ctx* c, int n, double* A, double* W, std::string what) {
    Libs& L = g_libs;
    const char* ev = getenv("XTB_EIGEN_X64");
    if (L.x_ok && (n > LEGACY_MAX_N || (ev && atoi(ev) != 0))) return syevd_x64(c, n, A, W, what);
    int lwork = 0;
    if (L.sv_buf(c->h_cusolver, JOB_VECTOR, FILL_UPPER, n, A, n, W, &lwork) != 0) XTB_FAIL(c, XTB_ECUDA, "cusolverDnDsyevd_bufferSize failed (%s)", what.c_str());
    int rc;
    if ((rc = xtb_reserve(c, c->d_eig_work, sizeof(double) * (size_t)lwork + 16))) return rc;
    int* d_info = (int*)((double*)c->d_eig_work.p + lwork);
    if (L.sv_syevd(c->h_cusolver, JOB_VECTOR, FILL_UPPER, n, A, n, W, (double*)c->d_eig_work.p, lwork, d_info) != 0)
        XTB_FAIL(c, XTB_ECUDA, "cusolverDnDsyevd failed to launch (%s)", what.c_str());
    int info = 0;
    XTB_CUDA(c, cudaMemcpyAsync(&info, d_info, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    if (info != 0) XTB_FAIL(c, XTB_ECUDA, "Dsyevd(%s): info = %d", what.c_str(), info);    // the reference records LAPACK failures the same way (TB_NRL.f90:410-416)
    return XTB_OK;
}

}  // namespace

// The Fortran-layout eigenvectors Ha(N,N) (column n = state n) on the device.  When the context only holds the state-minor
// copy HaT (xtb_eigensolve with 9 orbitals per atom, xtb_set_eigen_device_t) it is transposed into own_Ha here, once.
int xtb_ensure_Ha_std(xtb_ctx* c) {
    if (c->d_Ha) return XTB_OK;
    if (!c->d_HaT) XTB_FAIL(c, XTB_ESTATE, "no eigenvectors on the device");
    const long long N = (long long)c->na * c->n_orb, K = c->K > 0 ? c->K : N;
    int rc;
    if ((rc = xtb_reserve(c, c->own_Ha, sizeof(double) * (size_t)N * K))) return rc;
    // HaT: fast index = state (K of them), slow = row (N).  Ha: fast = row, slow = state.
    const dim3 grid((unsigned)((K + 31) / 32), (unsigned)((N + 31) / 32));
    transpose_kernel<<<grid, 256, 0, c->stream>>>(K, N, c->d_HaT, (double*)c->own_Ha.p);
    XTB_LAUNCH_CHECK(c);
    c->d_Ha = (const double*)c->own_Ha.p;
    return XTB_OK;
}

void xtb_release_lib_handles(xtb_ctx* c) {
    if (c->h_cusolver && g_libs.sv_destroy) g_libs.sv_destroy(c->h_cusolver);
    if (c->h_cublas && g_libs.bl_destroy) g_libs.bl_destroy(c->h_cublas);
    c->h_cusolver = nullptr; c->h_cublas = nullptr;
}

// C = op(A) op(B) through the run-time bound cuBLAS (plain library GEMM; also used by coupling.cu)
int xtb_blas_dgemm(xtb_ctx* c, bool ta, bool tb, int m, int n, int k, const double* A, int lda, const double* B, int ldb, double* Cm, int ldc) {
    if (!bind_libs()) XTB_FAIL(c, XTB_ESTATE, "cuBLAS not available: %s", g_libs.why.c_str());
    Libs& L = g_libs;
    { int rc = ctx_handles(c); if (rc) return rc; }
    L.bl_stream(c->h_cublas, c->stream);
    const double one = 1.0, zero = 0.0;
    if (L.bl_dgemm(c->h_cublas, ta ? OP_T : OP_N, tb ? OP_T : OP_N, m, n, k, &one, A, lda, B, ldb, &zero, Cm, ldc) != 0)
        XTB_FAIL(c, XTB_ECUDA, "cublasDgemm failed (m=%d n=%d k=%d)", m, n, k);
    return XTB_OK;
}

int xtb_impl_eigensolve(xtb_ctx* c) {
    if (!bind_libs()) XTB_FAIL(c, XTB_ESTATE, "xtb_eigensolve: %s", g_libs.why.c_str());
    Libs& L = g_libs;
    const long long N = (long long)c->na * c->n_orb;
    if (N > LEGACY_MAX_N && !L.x_ok)
        XTB_FAIL(c, XTB_EINVAL, "xtb_eigensolve: N = %lld needs the 64-bit cusolverDnXsyevd, which this cuSOLVER library does not export", N);
    if (N > 46000) XTB_FAIL(c, XTB_EINVAL, "xtb_eigensolve: N = %lld: N * N exceeds the 32-bit element counts of the GEMM glue", N);
    const int n = (int)N;
    const size_t nn = (size_t)N * N;
    const bool ortho = c->tb == XTB_TB_PETTIFOR;
    int rc;
    if ((rc = ctx_handles(c))) return rc;
    L.sv_stream(c->h_cusolver, c->stream); L.bl_stream(c->h_cublas, c->stream);
    if ((rc = xtb_reserve(c, c->own_Ha, sizeof(double) * nn))) return rc;
    if ((rc = xtb_reserve(c, c->own_Ei, sizeof(double) * N))) return rc;
    if ((rc = xtb_reserve(c, c->own_fe, sizeof(double) * N))) return rc;
    if ((rc = xtb_reserve(c, c->d_eigS, sizeof(double) * N + 16))) return rc;
    XTB_T0(c, 8);
    double* Ha = (double*)c->own_Ha.p;
    double* Ei = (double*)c->own_Ei.p;
    bool have_std = true;
    if (ortho) {      // TB_Pettifor.f90:1455: the assembled matrix is diagonalised in place
        XTB_CUDA(c, cudaMemcpyAsync(Ha, c->d_H.p, sizeof(double) * nn, cudaMemcpyDeviceToDevice, c->stream));
        rc = syevd(c, n, Ha, Ei, "H");
        if (rc) return rc;
    } else {
        if ((rc = xtb_reserve(c, c->d_eig_X, sizeof(double) * nn))) return rc;
        if ((rc = xtb_reserve(c, c->d_eig_T, sizeof(double) * nn))) return rc;
        double* X = (double*)c->d_eig_X.p; double* T = (double*)c->d_eig_T.p;
        double* wS = (double*)c->d_eigS.p; int* fix = (int*)(wS + N);
        const double one = 1.0, zero = 0.0;
        XTB_CUDA(c, cudaMemcpyAsync(X, c->d_S2.p, sizeof(double) * nn, cudaMemcpyDeviceToDevice, c->stream));
        rc = syevd(c, n, X, wS, "S");
        if (rc) return rc;                                      // :1012
        loewdin_flag_kernel<<<1, 1024, 0, c->stream>>>(N, wS, fix);
        XTB_LAUNCH_CHECK(c);
        loewdin_scale_kernel<<<dim3(8, n), 256, 0, c->stream>>>(N, wS, fix, X);             // X = U s^-1/2   :1036
        XTB_LAUNCH_CHECK(c);
        if (L.bl_dgemm(c->h_cublas, OP_N, OP_N, n, n, n, &one, (const double*)c->d_H.p, n, X, n, &zero, T, n) != 0)   // H X   (dsymm, :1049)
            XTB_FAIL(c, XTB_ECUDA, "cublasDgemm(H X) failed");
        if (L.bl_dgemm(c->h_cublas, OP_T, OP_N, n, n, n, &one, X, n, T, n, &zero, Ha, n) != 0)                       // X^T (H X)  :1054
            XTB_FAIL(c, XTB_ECUDA, "cublasDgemm(X^T H X) failed");
        if (c->tb == XTB_TB_DFTB) {
            chop_kernel<<<1184, 256, 0, c->stream>>>((long long)nn, Ha, 1e-10);
            XTB_LAUNCH_CHECK(c);
        }
        rc = syevd(c, n, Ha, Ei, "H'");
        if (rc) return rc;                                     // TB_NRL.f90:409
        // Ha = X V (:421).  With 9 orbitals per atom the product is written state-minor, HaT = V^T X^T (the same GEMM with
        // both operands transposed, same cost): the layout the density kernel streams through its tensor map, so no re-tiled
        // copy is ever made.  The Fortran-layout Ha is derived on demand (xtb_ensure_Ha_std).  XTB_EIGEN_STATE_MINOR=0: as before.
        bool state_minor = c->n_orb == 9;
        if (const char* ev = getenv("XTB_EIGEN_STATE_MINOR")) state_minor = state_minor && atoi(ev) != 0;
        if (state_minor) {
            if (L.bl_dgemm(c->h_cublas, OP_T, OP_T, n, n, n, &one, Ha, n, X, n, &zero, T, n) != 0)
                XTB_FAIL(c, XTB_ECUDA, "cublasDgemm(V^T X^T) failed");
            DevBuf tmp = c->own_HaT; c->own_HaT = c->d_eig_T; c->d_eig_T = tmp;      // T now holds HaT: it becomes own_HaT
            have_std = false;
        } else {
            if (L.bl_dgemm(c->h_cublas, OP_N, OP_N, n, n, n, &one, X, n, Ha, n, &zero, T, n) != 0)
                XTB_FAIL(c, XTB_ECUDA, "cublasDgemm(X V) failed");
            // the eigenvectors now live in T: swap the roles of the two buffers
            DevBuf tmp = c->own_Ha; c->own_Ha = c->d_eig_T; c->d_eig_T = tmp;
        }
    }
    XTB_T1(c, 8);
    if ((rc = xtb_stage_finish(c))) return rc;
    c->d_Ha = have_std ? (const double*)c->own_Ha.p : nullptr;
    c->d_HaT = have_std ? nullptr : (const double*)c->own_HaT.p;
    c->d_Ei = (const double*)c->own_Ei.p; c->d_fe = (const double*)c->own_fe.p;
    c->K = N;
    c->have_eigvec = true;
    c->have_eigen = false;          // occupations still to come (xtb_set_occupations)
    c->have_keep = false;
    c->have_density = false;
    return XTB_OK;
}
