// DFTB branch: Slater-Koster-table Hamiltonian / overlap assembly, Hellmann-Feynman forces and
// supercell force for the s / sp3 / sp3d5 bases, hetero-nuclear pairs included.
//
// Replaces Construct_Vij_DFTB (Source_files/TB_DFTB.f90:56-228), Hamil_tot_DFTB's assembly part (:319-503) with
// Hamilton_one_DFTB / Hopping_DFTB / Get_overlap_S_matrix_DFTB (:673-775), get_dHij_drij_DFTB -> get_forces_DFTB ->
// d_Hamilton_one_DFTB -> Attract_TB_forces_DFTB (:781-1127) and Attract_TB_Forces_Press_DFTB -> dHopping_Press_DFTB
// (:1135-1420).
//
// Same device data flow as the NRL branch (nrl_assemble.cu, nrl_forces.cu): no dense (Na,Na,nv) tables and no dense
// per-atom dH,dS(3,N,N); radial values are interpolated from the tables in registers where they are needed.
// Differences to NRL that shape the kernels:
//   * a pair needs TWO radial vectors, M_Vij(j,i,:) from TB(kind i, kind j) and M_Vij(i,j,:) from TB(kind j, kind i)
//     (hetero-nuclear Slater-Koster table: upper triangle from the first, lower from the second);
//   * on-site blocks are constants (Es, Ep, Ed), so forces have pair terms only;
//   * everything is already in eV and Angstrom.
// The s basis: H, S and the density matrices are provided; forces / supercell force are refused, because the
// reference passes the radial VALUE where the derivative belongs there (TB_DFTB.f90:1047-1048,1393-1394) -- see DESIGN.md.
#include "dftb_device.cuh"

namespace {

// entry e = (j, q) of the list; i = neighbour.  M_X(j,i,:) uses TB(kind i, kind j) at the distance of that entry.
template <int NO>
__device__ __forceinline__ void entry_radial(const DftbDev* __restrict__ TB, const double* __restrict__ tab, int n_kinds,
                                             const int* __restrict__ koa, int j, int i, double r, int mat, double* V, double* dV) {
    dftb::radial<NO>(TB[koa[i] + n_kinds * koa[j]], tab, mat, r, V, dV);
}

// one thread per (atom j, slot q, mat): block (rows = neighbour i or j itself, cols = j) of H or S
template <int NO>
__global__ void __launch_bounds__(128) dftb_blocks_kernel(int na, int a0, int a1, int cap, int n_kinds,
                                                         const int* __restrict__ koa, const int* __restrict__ nn_size,
                                                         const int* __restrict__ nn_list, const int* __restrict__ rev,
                                                         const double* __restrict__ nn_dist, const int* __restrict__ rowptr,
                                                         const DftbDev* __restrict__ TB, const double* __restrict__ tab,
                                                         double* __restrict__ Hblk, double* __restrict__ Sblk) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long per_atom = 2LL * (cap + 1);
    const int j = a0 + (int)(t / per_atom);
    if (j >= a1) return;
    const int rem = (int)(t % per_atom);
    const int q = rem >> 1, mat = rem & 1;
    if (q > nn_size[j]) return;
    double* out = (mat == 0 ? Hblk : Sblk) + ((size_t)rowptr[j] + q) * (NO * NO);
    for (int x = 0; x < NO * NO; ++x) out[x] = 0.0;
    if (q == 0) {   // Hamilton_one_DFTB on-site (:682-694): diag(Es, Ep x3, Ed x5) of TB(kind j, kind j); S = I (:769-772)
        const DftbDev& P = TB[koa[j] + n_kinds * koa[j]];
        for (int o = 0; o < NO; ++o) out[o * (NO + 1)] = mat == 1 ? 1.0 : (o == 0 ? P.Es : (o <= 3 ? P.Ep : P.Ed));
        return;
    }
    const size_t e = (size_t)j * cap + (q - 1);
    const int i = nn_list[e];
    const int rv = rev[e];
    if (rv < 0) return;                               // asymmetric list: the reference's dense tables hold zeros there
    const size_t ei = (size_t)i * cap + rv;           // entry of j in the list of i
    // the reference evaluates the pair from the list of the SMALLER index atom `o` for the larger `p` (:444-451):
    // Hopping_DFTB(M(o,p,:), M(p,o,:), M_lmn(:,o,p)) -> T(rows = orbitals of p, cols = orbitals of o)
    const bool jsmall = j < i;
    const size_t eo = jsmall ? e : ei, ep = jsmall ? ei : e;      // entries (o -> p) and (p -> o)
    const int o_ = jsmall ? j : i, p_ = jsmall ? i : j;
    double V12[10], V21[10], dump[10];
    entry_radial<NO>(TB, tab, n_kinds, koa, o_, p_, nn_dist[4 * eo + 3], mat, V12, dump);     // M(o,p,:)
    entry_radial<NO>(TB, tab, n_kinds, koa, p_, o_, nn_dist[4 * ep + 3], mat, V21, dump);     // M(p,o,:)
    const double4 d4 = *reinterpret_cast<const double4*>(nn_dist + 4 * eo);
    const double l = d4.x / d4.w, m = d4.y / d4.w, n = d4.z / d4.w;
    if (NO == 1) { double v = V12[0]; if (mat == 1 && fabs(v) <= 1e-12) v = 0.0; out[0] = v; return; }
    double up12[45], up21[45];
    sk::block_sp3d5_upper<double>(V12, l, m, n, up12);
    sk::block_sp3d5_upper<double>(V21, l, m, n, up21);
    // T(a,b): a = orbital of p (row), b = orbital of o (col).  Stored block = (rows i, cols j):
    //   j = o (i = p): stored(a,b) = T(a,b) ;  j = p (i = o): stored(a,b) = T(b,a)   (lower triangle = element transpose, :478-497)
#pragma unroll
    for (int a = 0; a < NO; ++a)
#pragma unroll
        for (int b = 0; b < NO; ++b) {
            double u = dftb::hetero(up12, up21, a, b);
            if (mat == 1 && fabs(u) <= 1e-12) u = 0.0;
            if (jsmall) out[a + NO * b] = u; else out[b + NO * a] = u;
        }
}

// compact per-entry tables for the drop-in outputs of xtb_radial: rad[40] = M_Vij, M_dVij, M_SVij, M_dSVij (10-channel
// positions compacted to nv), geo[9] = M_lmn, M_xrr, M_x1
template <int NO>
__global__ void dftb_tables_kernel(int na, int cap, int n_kinds, const int* __restrict__ koa, const int* __restrict__ nn_size,
                                   const int* __restrict__ nn_list, const double* __restrict__ nn_dist,
                                   const DftbDev* __restrict__ TB, const double* __restrict__ tab, const double* __restrict__ hmat,
                                   double* __restrict__ rad, double* __restrict__ geo) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (long long)na * cap) return;
    const int j = (int)(t / cap), q = (int)(t % cap);
    if (q >= nn_size[j]) return;
    const int i = nn_list[t];
    const double4 d4 = *reinterpret_cast<const double4*>(nn_dist + 4 * t);
    double* g = geo + 9 * t;
    const double l = d4.x / d4.w, m = d4.y / d4.w, n = d4.z / d4.w;
    g[0] = l; g[1] = m; g[2] = n;
    g[3] = l / d4.w; g[4] = m / d4.w; g[5] = n / d4.w;
    for (int a = 0; a < 3; ++a) g[6 + a] = d4.x * hmat[a] + d4.y * hmat[a + 3] + d4.z * hmat[a + 6];
    double* o = rad + 40 * t;
    for (int c = 0; c < 40; ++c) o[c] = 0.0;
    double V[10], dV[10];
    constexpr int NV = NO == 1 ? 1 : (NO == 4 ? 4 : 10);
    for (int mat = 0; mat < 2; ++mat) {
        entry_radial<NO>(TB, tab, n_kinds, koa, j, i, d4.w, mat, V, dV);
        for (int c = 0; c < NV; ++c) { o[20 * mat + c] = V[dftb::col(NO, c)]; o[20 * mat + 10 + c] = dV[dftb::col(NO, c)]; }
    }
}

constexpr int FT = 256;

// ---- forces: one CTA per owned atom k; thread -> (neighbour q, matrix, direction) ----
template <int NO>
__global__ void __launch_bounds__(FT) dftb_forces_kernel(int na, int a0, int cap, int mx, int n_kinds, const int* __restrict__ koa,
                                                        const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                        const int* __restrict__ rev, const double* __restrict__ nn_dist,
                                                        const int* __restrict__ rowptr, const DftbDev* __restrict__ TB,
                                                        const double* __restrict__ tab, const double* __restrict__ hmat,
                                                        const double* __restrict__ rho, const double* __restrict__ rhoE,
                                                        double* __restrict__ F) {
    extern __shared__ __align__(16) double smf[];
    double* part = smf;                          // [mx][6]
    __shared__ double h[9];
    const int k = a0 + blockIdx.x;
    const int m = nn_size[k];
    if (threadIdx.x < 9) h[threadIdx.x] = hmat[threadIdx.x];
    __syncthreads();
    const size_t b0 = (size_t)rowptr[k] * (NO * NO);
    for (int w = threadIdx.x; w < m * 6; w += FT) {
        const int q = w / 6, rem = w - q * 6;
        const int mat = rem / 3, al = rem - mat * 3;
        const size_t e = (size_t)k * cap + q;
        const int b = nn_list[e];
        const int rv = rev[e];
        double res = 0.0;
        if (rv >= 0) {
            // block (rows i = smaller index, cols j = larger) from the list entry of i (TB_DFTB.f90:868-872)
            const bool ksmall = k < b;
            const size_t eb = (size_t)b * cap + rv;
            const size_t es = ksmall ? e : eb, el = ksmall ? eb : e;       // entries (i -> j), (j -> i)
            const int ia = ksmall ? k : b, ja = ksmall ? b : k;
            const double4 d4 = *reinterpret_cast<const double4*>(nn_dist + 4 * es);
            const double x = d4.x, y = d4.y, z = d4.z, r = d4.w;
            const double delt = ksmall ? 1.0 : -1.0;                      // delta_ik - delta_jk
            double V12[10], dV12[10], V21[10], dV21[10];
            entry_radial<NO>(TB, tab, n_kinds, koa, ja, ia, nn_dist[4 * el + 3], mat, V12, dV12);    // M(j,i,:)  = vec_12 (:1029-1032)
            entry_radial<NO>(TB, tab, n_kinds, koa, ia, ja, r, mat, V21, dV21);                      // M(i,j,:)  = vec_21
            const double l = x / r, mm = y / r, n = z / r;
            // drij_dska (TB_Koster_Slater.f90:75) and ddija_dskb_kd (:102)
            const double drds = delt / r * (x * h[al] + y * h[al + 3] + z * h[al + 6]);
            const double dl = (delt * h[al + 0] - x / r * drds) / r;
            const double dm = (delt * h[al + 3] - y / r * drds) / r;
            const double dn = (delt * h[al + 6] - z / r * drds) / r;
#pragma unroll
            for (int c = 0; c < 10; ++c) { dV12[c] *= drds; dV21[c] *= drds; }
            double d12[45], d21[45];
            dftb::dupper(V12, dV12, l, mm, n, dl, dm, dn, d12);
            dftb::dupper(V21, dV21, l, mm, n, dl, dm, dn, d21);
            // rho(i-orbitals, j-orbitals): block (k,q) directly if k is the smaller index, else its transpose;
            // the mirrored block contributes the same amount (rho symmetric): factor 2
            const double* A = (mat == 0 ? rho : rhoE) + b0 + (size_t)(q + 1) * (NO * NO);
            res = 2.0 * dftb::contract_hetero<NO>(d12, d21, A, !ksmall);
        }
        part[q * 6 + rem] = res;
    }
    __syncthreads();
    if (threadIdx.x < 3) {   // fixed-order reduction over the list
        const int al = threadIdx.x;
        double s = 0.0;
        for (int q = 0; q < m; ++q) s += part[q * 6 + al] - part[q * 6 + 3 + al];
        F[3 * k + al] = s;
    }
}

// ---- supercell force: one CTA per owned atom i, all directed pairs (i, q); thread -> (q, mat, gamma), 3 deltas each ----
template <int NO>
__global__ void __launch_bounds__(FT) dftb_pressure_kernel(int na, int a0, int cap, int mx, int n_kinds, const int* __restrict__ koa,
                                                          const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                          const int* __restrict__ rev, const double* __restrict__ nn_dist,
                                                          const double* __restrict__ nn_dist_s, const int* __restrict__ rowptr,
                                                          const DftbDev* __restrict__ TB, const double* __restrict__ tab,
                                                          const double* __restrict__ rho, const double* __restrict__ rhoE,
                                                          double* __restrict__ partial /*[n_owned][9]*/) {
    extern __shared__ __align__(16) double smf[];
    double* part = smf;                        // [mx][18]
    const int i = a0 + blockIdx.x;
    const int m = nn_size[i];
    const size_t b0 = (size_t)rowptr[i] * (NO * NO);
    for (int w = threadIdx.x; w < m * 6; w += FT) {
        const int q = w / 6, rem = w - q * 6;
        const int mat = rem / 3, ga = rem - mat * 3;
        const size_t e = (size_t)i * cap + q;
        const int j = nn_list[e];
        const int rv = rev[e];
        const double4 d4 = *reinterpret_cast<const double4*>(nn_dist + 4 * e);
        const double rij[3] = {d4.x, d4.y, d4.z};
        const double r = d4.w;
        const double sg = nn_dist_s[3 * e + ga];
        double V12[10], dV12[10], V21[10], dV21[10];
        entry_radial<NO>(TB, tab, n_kinds, koa, i, j, r, mat, V21, dV21);                      // M(i,j,:) = vec_21 (:1372-1375)
        if (rv >= 0) entry_radial<NO>(TB, tab, n_kinds, koa, j, i, nn_dist[4 * ((size_t)j * cap + rv) + 3], mat, V12, dV12);
        else for (int c = 0; c < 10; ++c) { V12[c] = 0.0; dV12[c] = 0.0; }
        const double l = rij[0] / r, mm = rij[1] / r, n = rij[2] / r;
        const double* A = (mat == 0 ? rho : rhoE) + b0 + (size_t)(q + 1) * (NO * NO);
        const double r2 = r * r;
#pragma unroll
        for (int de = 0; de < 3; ++de) {
            // dda_dhgd (TB_Koster_Slater.f90:153-161) ; drij_dhab (:146-150)
            const double pre = sg / (r2 * r);
            const double dl = pre * (r2 * (de == 0 ? 1.0 : 0.0) - rij[0] * rij[de]);
            const double dm = pre * (r2 * (de == 1 ? 1.0 : 0.0) - rij[1] * rij[de]);
            const double dn = pre * (r2 * (de == 2 ? 1.0 : 0.0) - rij[2] * rij[de]);
            const double drdh = rij[de] * sg / r;
            double a12[10], a21[10];
#pragma unroll
            for (int c = 0; c < 10; ++c) { a12[c] = dV12[c] * drdh; a21[c] = dV21[c] * drdh; }
            double d12[45], d21[45];
            dftb::dupper(V12, a12, l, mm, n, dl, dm, dn, d12);
            dftb::dupper(V21, a21, l, mm, n, dl, dm, dn, d21);
            part[q * 18 + mat * 9 + ga * 3 + de] = dftb::contract_hetero<NO>(d12, d21, A, false);
        }
    }
    __syncthreads();
    if (threadIdx.x < 9) {
        const int k1 = threadIdx.x;
        double s = 0.0;
        for (int q = 0; q < m; ++q) s += part[q * 18 + k1] - part[q * 18 + 9 + k1];
        partial[(size_t)blockIdx.x * 9 + k1] = s;
    }
}

__global__ void dftb_sum9_kernel(int n, const double* __restrict__ partial, double* __restrict__ out) {
    __shared__ double buf[128];
    const int k1 = blockIdx.x;
    double s = 0.0;
    for (int a = threadIdx.x; a < n; a += 128) s += partial[(size_t)a * 9 + k1];
    buf[threadIdx.x] = s;
    __syncthreads();
    for (int off = 64; off > 0; off >>= 1) {
        if (threadIdx.x < off) buf[threadIdx.x] += buf[threadIdx.x + off];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[k1] = buf[0];
}

#define DFTB_ARGS_LIST (const int*)c->d_koa.p, (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p

template <int NO>
int assemble(xtb_ctx* c, bool dense) {
    const int na = c->na;
    int rc;
    if ((rc = xtb_reserve(c, c->d_Hblk, sizeof(double) * NO * NO * xtb_blocks_alloc(c)))) return rc;
    if ((rc = xtb_reserve(c, c->d_Sblk, sizeof(double) * NO * NO * xtb_blocks_alloc(c)))) return rc;
    if (dense) {
    }
    XTB_T0(c, 2);
    const int nown = c->a1 - c->a0;
    const long long nthreads = 2LL * (c->cap + 1) * nown;
    if (nthreads > 0) {
        dftb_blocks_kernel<NO><<<(unsigned)((nthreads + 127) / 128), 128, 0, c->stream>>>(
            na, c->a0, c->a1, c->cap, c->n_kinds, DFTB_ARGS_LIST, (const int*)c->d_rev.p, (const double*)c->d_nn_dist.p,
            (const int*)c->d_rowptr.p, (const DftbDev*)c->d_params.p, (const double*)c->d_dftb_tab.p, (double*)c->d_Hblk.p, (double*)c->d_Sblk.p);
        XTB_LAUNCH_CHECK(c);
    }
    XTB_T0(c, 6);      // dense panel kernel alone
    if (dense && (rc = xtb_dense_update(c, c->a0, c->a1))) return rc;
    XTB_T1(c, 6);
    XTB_T1(c, 2);
    if ((rc = xtb_stage_finish(c))) return rc;
    c->have_blocks = true;
    c->have_HS = dense;
    return XTB_OK;
}

// ---- s basis, AS WRITTEN in the reference -----------------------------------------------------------------------------------
// d_Hamilton_one_DFTB hands d_KS_s the VALUE vectors (vec_M_Vij12, vec_M_SVij12) where the other bases get the derivative
// vectors (TB_DFTB.f90:1047-1048, and again for y and z), so dH_alpha(i,j) = V_ss(r_ij), dS_alpha(i,j) = S_ss(r_ij) for EVERY
// atom k and direction alpha, whether k belongs to the pair or not.  get_forces_DFTB / Attract_TB_forces_DFTB (:829-972) then
// give every atom, in every direction, the same number  sum_i sum_{j in NN(i), listed both ways} [V rho(i,j) - S rho_E(i,j)].
// It is not a derivative; it is reproduced because the parity target is the reference's own output.  One warp per atom,
// fixed-order reductions; needs the blocks of all atoms (unpartitioned context).
__global__ void __launch_bounds__(128) dftb_s_aswritten_partial_kernel(int na, int cap, int n_kinds, const int* __restrict__ koa,
                                                                       const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                                       const int* __restrict__ rev, const double* __restrict__ nn_dist,
                                                                       const int* __restrict__ rowptr, const DftbDev* __restrict__ TB,
                                                                       const double* __restrict__ tab, const double* __restrict__ rho,
                                                                       const double* __restrict__ rhoE, double* __restrict__ part) {
    const int i2 = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i2 >= na) return;
    const int m = nn_size[i2];
    double acc = 0.0;
    for (int q = lane; q < m; q += 32) {
        const size_t e = (size_t)i2 * cap + q;
        const int j1 = nn_list[e], rv = rev[e];
        if (rv < 0) continue;                                   // the block of the ordered pair was never written (:868)
        const bool small = i2 < j1;
        const int ia = small ? i2 : j1, ja = small ? j1 : i2;   // block rows = smaller index i, cols = larger j
        const size_t e_ji = small ? (size_t)j1 * cap + rv : e;  // list entry of j pointing at i: M(j,i,:) = vec_12 (:1029-1032)
        double V[10], dV[10], S[10];
        entry_radial<1>(TB, tab, n_kinds, koa, ja, ia, nn_dist[4 * e_ji + 3], 0, V, dV);
        entry_radial<1>(TB, tab, n_kinds, koa, ja, ia, nn_dist[4 * e_ji + 3], 1, S, dV);
        const size_t b = (size_t)rowptr[i2] + 1 + q;
        acc += V[0] * rho[b] - S[0] * rhoE[b];
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
    if (lane == 0) part[i2] = acc;
}
__global__ void __launch_bounds__(256) dftb_s_aswritten_fill_kernel(int na, const double* __restrict__ part, double* __restrict__ F) {
    __shared__ double buf[256];
    double s = 0.0;
    for (int i = threadIdx.x; i < na; i += 256) s += part[i];
    buf[threadIdx.x] = s;
    __syncthreads();
    for (int off = 128; off > 0; off >>= 1) {
        if (threadIdx.x < off) buf[threadIdx.x] += buf[threadIdx.x + off];
        __syncthreads();
    }
    const double tot = buf[0];
    for (int i = threadIdx.x; i < 3 * na; i += 256) F[i] = tot;
}

static int forces_s_aswritten(xtb_ctx* c) {
    if (c->a0 != 0 || c->a1 != c->na) XTB_FAIL(c, XTB_EINVAL, "DFTB s basis: the reference's force expression sums over all atoms; unpartitioned context only");
    int rc;
    if ((rc = xtb_reserve(c, c->d_F, sizeof(double) * 3 * (size_t)(c->na + c->nranks)))) return rc;
    if ((rc = xtb_reserve(c, c->d_ppart, sizeof(double) * 9 * (size_t)c->na))) return rc;
    XTB_T0(c, 4);
    dftb_s_aswritten_partial_kernel<<<(c->na + 3) / 4, 128, 0, c->stream>>>(c->na, c->cap, c->n_kinds, DFTB_ARGS_LIST, (const int*)c->d_rev.p,
                                                                          (const double*)c->d_nn_dist.p, (const int*)c->d_rowptr.p,
                                                                          (const DftbDev*)c->d_params.p, (const double*)c->d_dftb_tab.p,
                                                                          (const double*)c->d_rho.p, (const double*)c->d_rhoE.p, (double*)c->d_ppart.p);
    XTB_LAUNCH_CHECK(c);
    dftb_s_aswritten_fill_kernel<<<1, 256, 0, c->stream>>>(c->na, (const double*)c->d_ppart.p, (double*)c->d_F.p);
    XTB_LAUNCH_CHECK(c);
    XTB_T1(c, 4);
    return xtb_stage_finish(c);
}

template <int NO>
int forces(xtb_ctx* c) {
    const int nown = c->a1 - c->a0;
    int rc;
    if ((rc = xtb_reserve(c, c->d_F, sizeof(double) * 3 * (size_t)(c->na + c->nranks)))) return rc;
    XTB_T0(c, 4);
    if (nown > 0) {
        const int mx = c->max_nn > 0 ? c->max_nn : 1;
        const size_t smem = sizeof(double) * (size_t)mx * 6;
        XTB_CUDA(c, cudaFuncSetAttribute(dftb_forces_kernel<NO>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        dftb_forces_kernel<NO><<<nown, FT, smem, c->stream>>>(c->na, c->a0, c->cap, mx, c->n_kinds, DFTB_ARGS_LIST, (const int*)c->d_rev.p,
                                                              (const double*)c->d_nn_dist.p, (const int*)c->d_rowptr.p,
                                                              (const DftbDev*)c->d_params.p, (const double*)c->d_dftb_tab.p,
                                                              (const double*)c->d_hdev.p, (const double*)c->d_rho.p,
                                                              (const double*)c->d_rhoE.p, (double*)c->d_F.p);
        XTB_LAUNCH_CHECK(c);
    }
    XTB_T1(c, 4);
    if ((rc = xtb_stage_finish(c))) return rc;
    return XTB_OK;
}

template <int NO>
int pressure(xtb_ctx* c) {
    const int nown = c->a1 - c->a0;
    int rc;
    if ((rc = xtb_reserve(c, c->d_SC, sizeof(double) * 9))) return rc;
    if ((rc = xtb_reserve(c, c->d_ppart, sizeof(double) * 9 * (size_t)(nown > 0 ? nown : 1)))) return rc;
    XTB_T0(c, 5);
    if (nown > 0) {
        const int mx = c->max_nn > 0 ? c->max_nn : 1;
        const size_t smem = sizeof(double) * (size_t)mx * 18;
        XTB_CUDA(c, cudaFuncSetAttribute(dftb_pressure_kernel<NO>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        dftb_pressure_kernel<NO><<<nown, FT, smem, c->stream>>>(c->na, c->a0, c->cap, mx, c->n_kinds, DFTB_ARGS_LIST, (const int*)c->d_rev.p,
                                                                (const double*)c->d_nn_dist.p, (const double*)c->d_nn_dist_s.p,
                                                                (const int*)c->d_rowptr.p, (const DftbDev*)c->d_params.p,
                                                                (const double*)c->d_dftb_tab.p, (const double*)c->d_rho.p,
                                                                (const double*)c->d_rhoE.p, (double*)c->d_ppart.p);
        XTB_LAUNCH_CHECK(c);
        dftb_sum9_kernel<<<9, 128, 0, c->stream>>>(nown, (const double*)c->d_ppart.p, (double*)c->d_SC.p);
        XTB_LAUNCH_CHECK(c);
    } else {
        XTB_CUDA(c, cudaMemsetAsync(c->d_SC.p, 0, sizeof(double) * 9, c->stream));
    }
    XTB_T1(c, 5);
    if ((rc = xtb_stage_finish(c))) return rc;
    return XTB_OK;
}

}  // namespace

int xtb_impl_dftb_assemble(xtb_ctx* c, bool dense) {
    if (c->n_orb == 9) return assemble<9>(c, dense);
    if (c->n_orb == 4) return assemble<4>(c, dense);
    return assemble<1>(c, dense);
}

int xtb_impl_dftb_forces(xtb_ctx* c) {
    if (c->n_orb == 9) return forces<9>(c);
    if (c->n_orb == 4) return forces<4>(c);
    return forces_s_aswritten(c);       // the reference's expression as written (not a derivative, TB_DFTB.f90:1047-1048)
}

int xtb_impl_dftb_pressure(xtb_ctx* c) {
    if (c->n_orb == 9) return pressure<9>(c);
    if (c->n_orb == 4) return pressure<4>(c);
    XTB_FAIL(c, XTB_EINVAL, "DFTB s basis: the reference's supercell-force formula is not a derivative (TB_DFTB.f90:1393-1394); not provided");
}

int xtb_impl_dftb_tables(xtb_ctx* c, double* d_rad, double* d_geo) {
    const long long tot = (long long)c->na * c->cap;
    const unsigned grid = (unsigned)((tot + 127) / 128);
#define DFTB_TAB(NO) dftb_tables_kernel<NO><<<grid, 128, 0, c->stream>>>(c->na, c->cap, c->n_kinds, DFTB_ARGS_LIST, (const double*)c->d_nn_dist.p, \
        (const DftbDev*)c->d_params.p, (const double*)c->d_dftb_tab.p, (const double*)c->d_hdev.p, d_rad, d_geo)
    if (c->n_orb == 9) DFTB_TAB(9); else if (c->n_orb == 4) DFTB_TAB(4); else DFTB_TAB(1);
#undef DFTB_TAB
    XTB_LAUNCH_CHECK(c);
    return XTB_OK;
}
