// S3 -- occupation-weighted density matrices rho = C f C^T and rho_E = C (f E) C^T on the neighbour pattern.
//
// Replaces Construct_Aij (Source_files/TB.f90:4328-4398) and Construct_Aij_x_En (TB.f90:4401-4469):
//     Atemp(i,:) = w(:)*Ha(i,:) ; Aij = Atemp * Ha^T   (dgemm 'N','T'),  w = fe  resp.  fe*Ei.
// The reference forms the full N x N products (2 x 2N^3 flop); every consumer reads only the blocks
// (atom i, {i} U NN(i)) (TB.f90:250-299,5544-5573; TB_NRL.f90:2594-2596,2794-2796), so only those are
// evaluated, and each unordered pair only once (rho is symmetric): the pair (i,j) is evaluated by its
// "owner" (balanced half-list rule below) and mirrored into the other block afterwards.
//
// B200 mapping (FP64 tensor cores; tcgen05 has no FP64 kind, the native op is DMMA.8x8x4):
//   prep   : state compaction (skip f_n <= f_cut), weights, and a tiled transposed copy of the eigenvectors
//            Ct[chunk of 16 states][atom][4 k-steps][n_orb rows][4 states]  -- one contiguous tile
//            (1152 B for n_orb = 9) per (chunk, atom), laid out so that DMMA fragment loads from shared
//            memory are bank-conflict free without padding.
//   main   : persistent CTAs, one atom i at a time.  A producer warp streams the tiles of i and of its
//            owned neighbours through a multi-stage shared-memory ring with cp.async.bulk (TMA bulk
//            copies, mbarrier complete_tx); a second helper warp weights the B operand (f_n C[i], f_n E_n C[i])
//            once per chunk; 8 consumer warps issue mma.sync.m8n8k4.f64:
//                D[row = orbital nu of neighbour j][col = orbital mu < 8 of i] += C[j nu, n] * (w_n C[i mu, n])
//            once with w = f (rho) and once with w = f E (rho_E) -- M = packed neighbour rows, K = states --
//            and the ninth orbital of i as two DFMA per row tile.  rho and rho_E share the A operand, so one
//            pass over the eigenvectors produces both.
//   mirror : non-owner blocks filled by transposition, so S4/S5 stream complete rows of blocks.
// Algorithmic flops (roofline numerator): 2 matrices * 2 * n_orb^2 * K_kept * (n_pairs/2 + na).
// mirror_blocks_kernel takes the rho stores twice (read-only view of the owner blocks, writable view of the others: disjoint)
#pragma GCC diagnostic ignored "-Wrestrict"
#include "xtb_ctx.h"
#include <cuda.h>            // CUtensorMap and its enums only; the encoder is fetched through the runtime (no -lcuda)
#include <cstdint>
#include <cstdlib>
#include <type_traits>

namespace {

#ifndef XTB_KC
#define XTB_KC 16
#endif
constexpr int KC = XTB_KC;             // states per chunk
// K-split of the first row tile of the on-site block over the four sub-partitions (see density_dmma_kernel); experiment, off by default
#ifdef XTB_KSPLIT
constexpr bool KSPLIT = true;
#else
constexpr bool KSPLIT = false;
#endif
constexpr int KS = KC / 4;             // DMMA k-steps per chunk
constexpr int CT_AT = 512 / KC;        // atoms per CTA of the re-tiling kernel (static shared memory budget)

// pair ownership (exactly one of the two atoms owns an unordered pair) comes from the neighbour stage: own[e],
// see nn_rev_kernel in neighbours.cu (positive half-space orientation: balanced work per atom).

// ---------------- prep ----------------
// order-preserving compaction of the kept states; single CTA (K <= a few 1e5)
__global__ void keep_kernel(long long K, const double* __restrict__ fe, const double* __restrict__ Ei, double f_cut,
                            int* __restrict__ keep, double* __restrict__ wf, double* __restrict__ wfe, int* __restrict__ nkeep) {
    __shared__ int buf[1024];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (long long base = 0; base < K; base += 1024) {
        const long long n = base + threadIdx.x;
        const int flag = (n < K && fe[n] > f_cut) ? 1 : 0;
        buf[threadIdx.x] = flag;
        __syncthreads();
        for (int off = 1; off < 1024; off <<= 1) {
            int t = threadIdx.x >= off ? buf[threadIdx.x - off] : 0;
            __syncthreads();
            buf[threadIdx.x] += t;
            __syncthreads();
        }
        if (flag) {
            const int pos = carry + buf[threadIdx.x] - 1;
            keep[pos] = (int)n;
            wf[pos] = fe[n];                 // Atemp = distre * Ha           TB.f90:4378
            wfe[pos] = fe[n] * Ei[n];        // Atemp = distre * En * Ha      TB.f90:4459
        }
        __syncthreads();
        if (threadIdx.x == 1023) carry += buf[1023];
        __syncthreads();
    }
    if (threadIdx.x == 0) *nkeep = carry;
}

__global__ void pad_weights_kernel(int nkeep, int kpad, double* __restrict__ wf, double* __restrict__ wfe) {
    const int t = nkeep + blockIdx.x * blockDim.x + threadIdx.x;
    if (t < kpad) { wf[t] = 0.0; wfe[t] = 0.0; }
}

// Ct[chunk][atom][s][r][t] = Ha[atom*NO + r, keep[chunk*16 + 4 s + t]]  (0 beyond nkeep)
// one CTA per (chunk, group of CT_AT atoms): reads KC columns x (CT_AT*NO) contiguous rows, writes contiguous tiles.
// HBM-bound (8*N*K bytes in, the same out): every thread keeps KC independent 128-bit loads in flight, the transpose
// goes through shared memory with a row pitch that makes both phases conflict-free, stores are 128-bit and coalesced.
constexpr int CT_THREADS = 160;
// partitioned run: groups of CT_AT atoms whose tiles some owned atom reads (itself or a neighbour).  Benign race: all write 1.
__global__ void need_groups_kernel(int a0, int a1, int cap, const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                   int* __restrict__ need) {
    const int k = a0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= a1) return;
    need[k / CT_AT] = 1;
    const int m = nn_size[k];
    for (int q = 0; q < m; ++q) need[nn_list[(size_t)k * cap + q] / CT_AT] = 1;
}
template <int NO>
__global__ void __launch_bounds__(CT_THREADS) ct_tile_kernel(int na, long long N, int nkeep, const int* __restrict__ keep,
                                                            const int* __restrict__ need /*per atom group, or null = all*/,
                                                            const double* __restrict__ Ha, double* __restrict__ Ct) {
    constexpr int AT = CT_AT;
    if (need && !need[blockIdx.y]) return;           // partitioned run: no owned atom reads the tiles of this group
    constexpr int P = AT * NO + 4;                   // pitch = 4 mod 8: the 64-bit banks of (t, t+2) x 8 rows are disjoint
    __shared__ __align__(16) double tile[KC][P];
    __shared__ long long colbase[KC];
    const int chunk = blockIdx.x;
    const int atom0 = blockIdx.y * AT;
    const int nat = min(AT, na - atom0);
    const int nrow = nat * NO;
    if (threadIdx.x < KC) {
        const int ks = chunk * KC + threadIdx.x;
        colbase[threadIdx.x] = ks < nkeep ? (long long)keep[ks] * N + (long long)atom0 * NO : -1;
    }
    __syncthreads();
    if ((N & 1) == 0 && (nrow & 1) == 0) {           // 16-byte aligned columns: AT is even, so atom0*NO is even
        for (int r2 = threadIdx.x; r2 < (nrow >> 1); r2 += CT_THREADS) {
            double2 v[KC];
#pragma unroll
            for (int kk = 0; kk < KC; ++kk) {
                const long long cb = colbase[kk];
                v[kk] = cb >= 0 ? __ldcs(reinterpret_cast<const double2*>(Ha + cb) + r2) : make_double2(0.0, 0.0);
            }
#pragma unroll
            for (int kk = 0; kk < KC; ++kk) *reinterpret_cast<double2*>(&tile[kk][2 * r2]) = v[kk];
        }
    } else {
        for (int r = threadIdx.x; r < nrow; r += CT_THREADS) {
            double v[KC];
#pragma unroll
            for (int kk = 0; kk < KC; ++kk) { const long long cb = colbase[kk]; v[kk] = cb >= 0 ? __ldcs(Ha + cb + r) : 0.0; }
#pragma unroll
            for (int kk = 0; kk < KC; ++kk) tile[kk][r] = v[kk];
        }
    }
    __syncthreads();
    double2* dst = reinterpret_cast<double2*>(Ct + ((size_t)chunk * na + atom0) * (NO * KC));     // 16-byte aligned: NO*KC is even
    const int tot2 = nat * NO * KC / 2;
    for (int e2 = threadIdx.x; e2 < tot2; e2 += CT_THREADS) {
        const int e = 2 * e2;
        const int a = e / (NO * KC);
        int rem = e - a * (NO * KC);
        const int s = rem / (NO * 4); rem -= s * (NO * 4);
        const int r = rem >> 2, t = rem & 3;          // t is 0 or 2
        dst[e2] = make_double2(tile[4 * s + t][a * NO + r], tile[4 * s + t + 1][a * NO + r]);
    }
}

// ---------------- PTX helpers ----------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
// non-blocking phase test
__device__ __forceinline__ uint32_t mbar_test(uint64_t* b, uint32_t parity) {
    uint32_t done;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(done) : "r"(smem_u32(b)), "r"(parity) : "memory");
    return done;
}
// polling wait (mbarrier.test_wait never suspends the warp): lowest wake-up latency, for the helper warps
__device__ __forceinline__ void mbar_spin(uint64_t* b, uint32_t parity, bool nap = false) {
    uint32_t done;
    do {
        if (nap) __nanosleep(40);
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n" : "=r"(done) : "r"(smem_u32(b)), "r"(parity) : "memory");
    } while (!done);
}
// TMA bulk copy global -> shared, completion signalled on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// TMA tensor copy (2-D tiled, swizzled) global -> shared through a tensor map that lives in global memory (SASS: UTMALDG)
__device__ __forceinline__ void tensor_g2s(void* dst, const void* tmap, int c0, int c1, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(smem_u32(dst)), "l"(tmap), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
        : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// ---------------- main kernel ----------------
// Shared memory per pipeline stage (doubles):
//   [slotcap][SLOT]   tiles, SLOT = 148: the 144-double (atom, chunk) tile [4 k-steps][NO rows][4 states] plus 4 pad
//                     doubles, so that row R = 9*slot + r of k-step s sits at bank16 = 4*((R + s) mod 4) + t and the
//                     8 rows x 4 states of a DMMA A fragment are conflict-free for every tile, also across slots;
//   [4][BROW]         B operand of atom i, already weighted: per k-step 32 doubles f*C[i,mu=0..7] in fragment order,
//                     32 doubles f*E*C[i,mu<8], then f*C[i,8,t] (4) and f*E*C[i,8,t] (4);  BROW = 72
//   [2][KC]           the weights f_n, f_n*E_n of the chunk
// Warp roles: warps 0..NCW-1 consume; warps NCW+1.. issue the TMA bulk copies (producers); warp NCW scales the B operand
// once per chunk (instead of 8 warps x 4 k-steps doing it redundantly on the FP64 pipe the DMMAs need).
// Barriers per stage: full (TMA bytes landed) -> bready (B operand scaled) -> consumers -> empty (NCW arrivals).
//
// Column packing: the 18 useful columns (rho | rho_E) x 9 orbitals of i are issued as two DMMA column tiles
// (mu = 0..7 for rho and for rho_E) plus two DFMA per row tile for mu = 8.  DMMA and DFMA share the FP64
// execution resource on B200 (profiles/r01_probe_fp64_peaks.json), so this issues exactly the useful flops,
// where a third, 2/8-full DMMA column tile would waste 25 % of the pipe.
constexpr int SLOT = 9 * KC + 4;       // 148 (KC = 16) / 292 (KC = 32): both = 4 mod 16
// ---- state-minor eigenvectors (TM = true): no re-tiled copy --------------------------------------------------------------
// When the eigenvectors are resident as HaT[row r][state n] (states contiguous: what xtb_eigensolve's last GEMM writes, or
// xtb_set_eigen_device_t), the (atom, chunk) tile -- 9 rows x 16 states = 9 x 128 bytes -- is ONE 2-D TMA box straight from
// HaT, with the 128-byte swizzle: the 16-byte piece c of smem row R lands at piece c ^ (R & 7).  Slots are 9 rows = 1152
// bytes apart from a 1024-byte aligned stage base, so R is simply the row index of the sub-task (9 * slot + r), the 8 rows
// of a DMMA row tile are 8 consecutive R, and the A fragment of lane (g, t) at k-step s sits at
//     R * 16 + 2 * ((2 s + (t >> 1)) ^ (R & 7)) + (t & 1)        [doubles]
// -- for the 32 lanes every 8-byte bank pair is hit exactly twice: the two wavefronts a 64-bit warp load needs anyway.
// Same bytes per tile, same number of TMA operations per chunk as the Ct path, and the 21.7 GB re-tiling pass is gone.
constexpr int SLOT_TM = 9 * KC;        // 144 doubles = 9 rows of 128 bytes (KC = 16)
__host__ __device__ constexpr int stage_doubles_tm(int slotcap) { return (slotcap * SLOT_TM + (KC / 4) * 72 + 2 * KC + 127) / 128 * 128; }
constexpr int BROW = 72;
#ifndef XTB_NPROD
#define XTB_NPROD 3
#endif
constexpr int NPROD = XTB_NPROD;                // producer warps (bulk-copy issue is serialised per lane within a warp); 8 + 1 + 3 = 12 warps = 3 per SM sub-partition -> 168 registers per thread

// ---- lock-step between CTAs ----
// The CTAs of one round work on spatially adjacent atoms, so at the same chunk index they read largely the same tiles of
// Ct: as long as they stay within a few chunks of each other the 126 MB L2 serves most of the ~40x reuse of every tile.
// Left alone they drift apart (tasks differ by a row tile or two) and every tile is fetched from HBM again (340 GB
// instead of ~50 GB per launch, measured).  The producers therefore publish their position (round, sub-task, chunk) and
// do not run more than `lead` chunks ahead of the slowest CTA.  The slowest CTA never waits; should a CTA not be
// resident (foreign work on an SM) the wait gives up after a bounded number of polls and throttling is switched off.
// "finished" marker of a CTA (and the identity of the min below).  Half of LLONG_MAX: `m + lead` must not wrap -- with
// ~LLONG_MAX a CTA that found every other CTA finished saw a NEGATIVE sum, polled until the 8192-try timeout (3.4 ms) and only
// then went on: 3.5 ms instead of 0.035 ms for the density stage of a 32-atom cell (profiles/r02l_probe_small_density_*.log).
#define XTB_PROG_DONE 0x3fffffffffffffffLL
constexpr int SYNC_EVERY = 16;      // position published / checked every this many chunks (one L2 round trip per check)
__device__ __forceinline__ void st_relaxed(long long* p, long long v) {
    asm volatile("st.relaxed.gpu.global.s64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ long long ld_relaxed(const long long* p) {
    long long v;
    asm volatile("ld.relaxed.gpu.global.s64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
// returns false if the wait was abandoned (a CTA did not move for ~10 ms, e.g. it is not resident): the caller then
// stops throttling -- lock-step is a bandwidth optimisation, never a correctness requirement
__device__ __forceinline__ bool throttle(long long* prog, long long pos, int lead, int lane) {
    if (lane == 0) st_relaxed(prog + blockIdx.x, pos);
    for (int tries = 0; tries < 8192; ++tries) {
        long long m = XTB_PROG_DONE;
        for (int b = lane; b < (int)gridDim.x; b += 32)
            if (b != (int)blockIdx.x) { const long long v = ld_relaxed(prog + b); m = v < m ? v : m; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { const long long v = __shfl_xor_sync(0xffffffffu, m, o); m = v < m ? v : m; }
        if (m + lead >= pos) return true;
    }
    return false;
}

// ring position shared by all roles: stage index and phase parity, advanced without integer division
struct Ring {
    uint32_t st, ph;
    __device__ __forceinline__ void next(uint32_t stages) { if (++st == stages) { st = 0; ph ^= 1u; } }
};

// operands of one k-step of one warp: MT A fragments + the B fragments (rho, rho_E, and the mu = 8 columns)
template <int MT>
struct Frag {
    double a[MT], b0, b1, b8f, b8e;
};
template <int NO, int MT, bool TM = false, int NCW = 8>
__device__ __forceinline__ void load_frag(Frag<MT>& f, const double* __restrict__ stg, const double* __restrict__ Bs,
                                          const int (&off)[MT], int s, int lane, int t, const int (&ks)[KS]) {
    f.b0 = Bs[s * BROW + lane];
    f.b1 = Bs[s * BROW + 32 + lane];
    f.b8f = Bs[s * BROW + 64 + t];
    f.b8e = Bs[s * BROW + 68 + t];
    // TM: the rows of this lane's tiles are 8 NCW rows = NCW KB apart, so one register per k-step (row of tile 0 + swizzled
    // position) and an immediate per tile address every fragment; rows past the sub-task (at most 7, in its last tile) read on
    // into the stage (finite or not, their results are never stored)
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) f.a[mt] = TM ? stg[ks[s] + mt * (NCW * 8 * 16)] : stg[off[mt] + NO * 4 * s];
}

// Consumer warp, one sub-task (atom i + up to slotcap-1 owned neighbours), MT row tiles.
// The chunk loop is software-pipelined across the stage boundary: the operands of k-step 0 of the NEXT chunk are
// fetched (after its barriers) before the math of the last k-step of the current chunk is issued, so the barrier
// round trip and the shared-memory latency hide behind 12 DMMAs instead of idling the FP64 pipe once per chunk.
template <int NO, int NCW, int MT, bool DBG, bool TM = false>
__device__ __forceinline__ void consume_subtask(const double* sbuf, int stage_doubles, int boff, uint32_t stages,
                                                uint64_t* full, uint64_t* bready, uint64_t* empty, Ring& rg, int nchunks,
                                                int nrows, int warp, int lane, const int* slot_q, size_t b0i,
                                                bool first_subtask, int dbg, double* __restrict__ rho, double* __restrict__ rhoE,
                                                bool ksp = false, double* spart = nullptr) {
    // (TM: the four rows of a half warp carry the swizzle keys 0..3, whose sixteen-byte pieces pair up on four positions of the
    // 128-byte line, so a fragment load takes four wavefronts instead of two -- ncu: 3.6e9 bank-conflict wavefronts per launch.
    // Giving the half warps the rows 0,2,4,6 / 1,3,5,7 removes them and was measured SLOWER, 69.7 against 69.1 ms: shared
    // memory is at a quarter of its wavefront rate either way, the allocation of the loop decides.)
    const int g = lane >> 2, t = lane & 3;
    double acc0[MT][2], acc1[MT][2], p8f[MT], p8e[MT];
    int off[MT];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
        acc0[mt][0] = acc0[mt][1] = acc1[mt][0] = acc1[mt][1] = 0.0;
        p8f[mt] = p8e[mt] = 0.0;
        const int R = 8 * (warp + (KSPLIT && ksp ? 1 : 0) + NCW * mt) + g;      // ksp: row tile 0 is the special tile, the regular ones follow
        const int sl = R / NO, r = R - sl * NO;
        off[mt] = (R < nrows) ? sl * SLOT + 4 * r + t : t;      // rows beyond the task read row 0 (results dropped)
        if (!TM) asm volatile("" : "+r"(off[mt]));             // keep it in a register (no per-chunk re-derivation)
    }
    int ks[KS];                                                // TM: row (8 warp + g) of tile 0 + swizzled position of (k-step s, t); R & 7 = g
#pragma unroll
    for (int s = 0; s < KS; ++s) {
        ks[s] = (8 * (warp + (KSPLIT && ksp ? 1 : 0)) + g) * 16 + 2 * ((2 * s + (t >> 1)) ^ g) + (t & 1);
        if (TM) asm volatile("" : "+r"(ks[s]));
    }
    // K-split special tile (rows 0..7 of the on-site block): warps 4..7 = one per sub-partition, each takes every fourth chunk
    const bool doS = KSPLIT && ksp && warp >= 4 && warp < 8;
    const int ks_sub = 128 * (warp + 1);               // TM: ks[s] - ks_sub = address of row g of tile 0
    double sA0[2] = {0.0, 0.0}, sA1[2] = {0.0, 0.0}, s8f = 0.0, s8e = 0.0;
    if (nchunks > 0) {
        const double* stg = sbuf + (size_t)rg.st * stage_doubles;
#ifdef XTB_EXP_STAGGER
        if (warp >= 4) __nanosleep(XTB_EXP_STAGGER);      // experiment: de-phase the two consumer warps of a sub-partition
#endif
        mbar_wait(&full[rg.st], rg.ph);
        mbar_wait(&bready[rg.st], rg.ph);
        Frag<MT> f0, f1;      // double buffer: even k-steps compute from f0 while f1 is loaded, odd k-steps the other way round
        load_frag<NO, MT, TM, NCW>(f0, stg, stg + boff, off, 0, lane, t, ks);
        static_assert(KS % 2 == 0, "the fragment double buffer needs an even number of k-steps per chunk");
        for (int ch = 0; ch < nchunks; ++ch) {
            Ring rn = rg;
            rn.next(stages);
            const double* stg_n = sbuf + (size_t)rn.st * stage_doubles;
            uint32_t ready = 0;
#pragma unroll
            for (int s = 0; s < KS; ++s) {
                Frag<MT>& cu = (s & 1) ? f1 : f0;
                Frag<MT>& nx = (s & 1) ? f0 : f1;
                if (s + 1 < KS) {
#ifdef XTB_EXP_HALF_LDS      // timing experiment (wrong results): odd k-steps reuse the A fragments of the even ones -- half the A loads
                    if ((s + 1) & 1) { const double k0 = nx.b0, k1 = nx.b1; nx = cu; nx.b0 = k0 + cu.b1; nx.b1 = k1 + cu.b0; } else
#endif
                    load_frag<NO, MT, TM, NCW>(nx, stg, stg + boff, off, s + 1, lane, t, ks);
                    // probe the barriers of the next stage one k-step early: the answer is in a register long before it is needed
                    if (s + 2 == KS && ch + 1 < nchunks) ready = mbar_test(&full[rn.st], rn.ph) & mbar_test(&bready[rn.st], rn.ph);
                } else if (ch + 1 < nchunks) {
                    if (!ready) { mbar_wait(&full[rn.st], rn.ph); mbar_wait(&bready[rn.st], rn.ph); }
                    load_frag<NO, MT, TM, NCW>(nx, stg_n, stg_n + boff, off, 0, lane, t, ks);
                }
                if (!DBG || !(dbg & 1)) {
#pragma unroll
                    for (int mt = 0; mt < MT; ++mt) {
                        dmma884(acc0[mt][0], acc0[mt][1], cu.a[mt], cu.b0);
                        dmma884(acc1[mt][0], acc1[mt][1], cu.a[mt], cu.b1);
                        if (!DBG || !(dbg & 8)) {
                            p8f[mt] = fma(cu.a[mt], cu.b8f, p8f[mt]);
                            p8e[mt] = fma(cu.a[mt], cu.b8e, p8e[mt]);
                        }
                    }
                }
            }
            if (KSPLIT && doS && ((ch & 3) == (warp & 3))) {
                const double* Bs = stg + boff;
#pragma unroll
                for (int s = 0; s < KS; ++s) {
                    const double a = TM ? stg[ks[s] - ks_sub] : stg[4 * g + t + NO * 4 * s];
                    dmma884(sA0[0], sA0[1], a, Bs[s * BROW + lane]);
                    dmma884(sA1[0], sA1[1], a, Bs[s * BROW + 32 + lane]);
                    s8f = fma(a, Bs[s * BROW + 64 + t], s8f);
                    s8e = fma(a, Bs[s * BROW + 68 + t], s8e);
                }
            }
            // every shared-memory read of this stage has been issued (in order, ahead of the arrive)
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[rg.st]);
            rg = rn;
            stg = stg_n;
        }
    }
    // ---- epilogue: D(row R -> (slot, nu), col -> mu) -> block (i, slot)[mu + NO*nu] ----
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
        // mu = 8: sum the four k-slices held by the lanes t = 0..3 of row g
        double f8 = p8f[mt], e8 = p8e[mt];
        f8 += __shfl_xor_sync(0xffffffffu, f8, 1); e8 += __shfl_xor_sync(0xffffffffu, e8, 1);
        f8 += __shfl_xor_sync(0xffffffffu, f8, 2); e8 += __shfl_xor_sync(0xffffffffu, e8, 2);
        const int R = 8 * (warp + (KSPLIT && ksp ? 1 : 0) + NCW * mt) + g;
        if (R < nrows) {
            const int sl = R / NO, nu = R - sl * NO;
            if (sl > 0 || first_subtask) {     // the on-site block belongs to the first sub-task only
                const size_t blk = (b0i + (sl == 0 ? 0 : 1 + slot_q[sl])) * (NO * NO) + NO * nu;
                rho[blk + 2 * t] = acc0[mt][0];  rho[blk + 2 * t + 1] = acc0[mt][1];
                rhoE[blk + 2 * t] = acc1[mt][0]; rhoE[blk + 2 * t + 1] = acc1[mt][1];
                if (t == 0) rho[blk + 8] = f8;
                if (t == 1) rhoE[blk + 8] = e8;
            }
        }
    }
    if (KSPLIT && doS) {      // partial sums of the special tile: spart[warp - 4][matrix][row g][col 0..8]; summed by warp 0 after the CTA barrier
        s8f += __shfl_xor_sync(0xffffffffu, s8f, 1); s8e += __shfl_xor_sync(0xffffffffu, s8e, 1);
        s8f += __shfl_xor_sync(0xffffffffu, s8f, 2); s8e += __shfl_xor_sync(0xffffffffu, s8e, 2);
        double* pr = spart + (size_t)(warp - 4) * 144 + 9 * g;
        pr[2 * t] = sA0[0]; pr[2 * t + 1] = sA0[1];
        pr[72 + 2 * t] = sA1[0]; pr[72 + 2 * t + 1] = sA1[1];
        if (t == 0) pr[8] = s8f;
        if (t == 1) pr[72 + 8] = s8e;
    }
}

// consumer warp with no row tile in this sub-task: it still takes part in the empty barriers
__device__ __forceinline__ void consume_idle(uint32_t stages, uint64_t* full, uint64_t* bready, uint64_t* empty, Ring& rg,
                                             int nchunks, int lane) {
    for (int ch = 0; ch < nchunks; ++ch) {
        mbar_wait(&full[rg.st], rg.ph);
        mbar_wait(&bready[rg.st], rg.ph);
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[rg.st]);
        rg.next(stages);
    }
}

template <int NO, int NCW, bool DBG, bool TM = false>      // TM: `Ct` carries the address of the tensor map of HaT (global memory)
__global__ void __launch_bounds__((NCW + 1 + NPROD) * 32, 1)
density_dmma_kernel(int na, int a0, int a1, int cap, int slotcap, int stages_, int nchunks, int dbg, int maxsub, int lead, int split, long long* __restrict__ prog,
                    const int* __restrict__ nn_size, const int* __restrict__ nn_list, const signed char* __restrict__ own,
                    const int* __restrict__ rowptr, const double* __restrict__ Ct, const double* __restrict__ wf,
                    const double* __restrict__ wfe, double* __restrict__ rho, double* __restrict__ rhoE) {
    static_assert(NO == 9, "column packing is written for the 9-orbital case (8 DMMA columns + 1 DFMA column per matrix)");
    constexpr int TILE = NO * KC;                       // doubles per (atom, chunk) tile in global memory
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const uint32_t stages = (uint32_t)stages_;
    const int stage_doubles = TM ? stage_doubles_tm(slotcap) : slotcap * SLOT + KS * BROW + 2 * KC;
    constexpr int SLOTW = TM ? SLOT_TM : SLOT;          // doubles between the slots of a stage
    double* sbuf = reinterpret_cast<double*>(smem_raw);
    if (TM) sbuf += ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u) >> 3;      // swizzle pattern = address bits: 1 KB aligned stages
    uint64_t* full = reinterpret_cast<uint64_t*>(sbuf + (size_t)stages * stage_doubles);
    uint64_t* bfull = full + stages;          // slot 0 (atom i) + weights landed: the B warp can start early
    uint64_t* bready = bfull + stages;
    uint64_t* empty = bready + stages;
    int* slot_q = reinterpret_cast<int*>(empty + stages);     // [slotcap] list ordinal of slot s (s >= 1)
    int* slot_atom = slot_q + slotcap;                        // [slotcap] atom of slot s (slot 0 = i)
    int* owned_q = slot_atom + slotcap;                       // [cap] ordinals of all owned neighbours of i
    // KSPLIT: partial sums of the special tile, [4 warps][2 matrices][8 rows][9 columns], 8-byte aligned behind the integer tables
    double* spart = reinterpret_cast<double*>((reinterpret_cast<uintptr_t>(owned_q + cap) + 7) & ~(uintptr_t)7);
    __shared__ int n_owned_sh;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // helper warps: with 8 consumer warps the row tiles of a sub-task fall 12 / 11 / 11 / 11 (45 tiles) or 12 / 12 / 12 / 11 (47) on
    // the four sub-partitions, warp w on sub-partition w & 3.  XTB_BWARP_LAST puts the B warp -- the only helper that uses the
    // FP64 pipe -- last (warp 11, sub-partition 3, the one that is never the fullest) and the producers first
#ifdef XTB_BWARP_LAST
    constexpr int P0 = NCW, BW = NCW + NPROD;
#else
    constexpr int P0 = NCW + 1, BW = NCW;
#endif

    if (threadIdx.x == 0) {
        for (uint32_t s = 0; s < stages; ++s) { mbar_init(&full[s], NPROD); mbar_init(&bfull[s], 1); mbar_init(&bready[s], 1); mbar_init(&empty[s], NCW); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    Ring rg; rg.st = 0; rg.ph = 0;     // running ring position (same sequence in every role)

    int round = 0;
    // work unit = (owned atom, part of its owned neighbours); split > 1 only when the owned range is too short to fill
    // the persistent grid evenly (few rounds): the unit then costs 1/split of an atom (+ the atom's own tile again)
    for (int u = blockIdx.x; u < (a1 - a0) * split; u += gridDim.x, ++round) {
        const int i = a0 + u / split, part = u - (u / split) * split;
        // ---- owned neighbours of i, in list order (warp 0, ordered compaction) ----
        __syncthreads();
        if (warp == 0) {
            const int m = nn_size[i];
            int cnt = 0;
            for (int base = 0; base < m; base += 32) {
                const int q = base + lane;
                const bool f = q < m && own[(size_t)i * cap + q] != 0;
                const unsigned mask = __ballot_sync(0xffffffffu, f);
                if (f) owned_q[cnt + __popc(mask & ((1u << lane) - 1u))] = q;
                cnt += __popc(mask);
            }
            if (lane == 0) n_owned_sh = cnt;
        }
        __syncthreads();
        const int n_all = n_owned_sh;
        const int per = (n_all + split - 1) / split;
        const int lo = part * per, n_owned = min(n_all, lo + per);      // this unit: owned neighbours lo .. n_owned-1
        if (part > 0 && lo >= n_owned) continue;                        // the on-site block belongs to part 0 (uniform per CTA)
        const size_t b0i = (size_t)rowptr[i];
        // KSPLIT rule of atom i (a function of the atom alone, so that the summation order of its blocks does not depend on the
        // work unit they are computed in): all owned neighbours fit one sub-task and the row tiles of the whole atom are
        // 4 m + 1 -- then the first tile (rows 0..7 of the on-site block) is summed in four interleaved quarters of the chunks,
        // one per sub-partition, and the other 4 m tiles fall m per sub-partition.
        const int t_full = (NO * (n_all + 1) + 7) >> 3;
        const bool ksp_atom = KSPLIT && n_all + 1 <= slotcap && (t_full & 3) == 1 && t_full >= 13;
        // ---- sub-tasks of at most slotcap-1 neighbours (slot 0 is always atom i itself) ----
        int sub = 0;
        for (int first = lo; first == lo || first < n_owned; first += slotcap - 1, ++sub) {
            const int nn = min(slotcap - 1, n_owned - first);
            const int nslots = nn + 1;
            __syncthreads();
            for (int s = threadIdx.x; s < nslots; s += blockDim.x) {
                if (s == 0) { slot_q[0] = -1; slot_atom[0] = i; }
                else { const int q = owned_q[first + s - 1]; slot_q[s] = q; slot_atom[s] = nn_list[(size_t)i * cap + q]; }
            }
            __syncthreads();
            const int nrows = nslots * NO;
            const int ntiles = (nrows + 7) >> 3;
            const bool ksp = KSPLIT && ksp_atom && first == 0;      // this sub-task holds the on-site block of a K-split atom

            if (warp >= P0 && warp < P0 + NPROD) {
                // ===== producer warps: stream the tiles of all slots for every chunk; slot s belongs to producer s % NPROD =====
                const int pw = warp - P0;
                const int ncopy = (DBG && (dbg & 2)) ? (nslots + 3) / 4 : nslots;     // debug: quarter of the data
                // slot 0 and the weights go to their own barrier (bfull) so the B warp overlaps with the other copies
                const int others = ncopy - 1;                                   // slots 1 .. ncopy-1
                const int mine = others > pw ? (others - pw + NPROD - 1) / NPROD : 0;
                const uint32_t bytes = (uint32_t)(mine * TILE) * 8u;
                const double* csrc = Ct;
                const double* w1 = wf; const double* w2 = wfe;
                // this lane's source / destination offsets do not change from chunk to chunk
                const int k0 = lane < mine ? 1 + lane * NPROD + pw : -1;
                const int k1 = lane + 32 < mine ? 1 + (lane + 32) * NPROD + pw : -1;
                const size_t src0 = k0 >= 0 ? (size_t)slot_atom[k0] * TILE : 0, src1 = k1 >= 0 ? (size_t)slot_atom[k1] * TILE : 0;
                const size_t srcB = (size_t)slot_atom[0] * TILE;
                const int row0 = k0 >= 0 ? slot_atom[k0] * NO : 0, row1 = k1 >= 0 ? slot_atom[k1] * NO : 0, rowB = slot_atom[0] * NO;   // TM: first row of the box
                const long long pos0 = ((long long)round * maxsub + sub) * nchunks;
                for (int ch = 0; ch < nchunks; ++ch) {
                    if (lead > 0 && pw == 0 && (ch % SYNC_EVERY) == 0 && !throttle(prog, pos0 + ch, lead, lane)) lead = 0;   // the other producers cannot pass warp 0 by more than the ring
                    mbar_spin(&empty[rg.st], rg.ph ^ 1u, DBG && (dbg & 4));
                    double* dstage = sbuf + (size_t)rg.st * stage_doubles;
                    if (lane == 0) {
                        if (pw == 0) {
                            mbar_expect_tx(&bfull[rg.st], (uint32_t)(TILE + 2 * KC) * 8u);
                            if (TM) tensor_g2s(dstage, Ct, ch * KC, rowB, &bfull[rg.st]);
                            else bulk_g2s(dstage, csrc + srcB, TILE * 8, &bfull[rg.st]);
                            bulk_g2s(dstage + slotcap * SLOTW + KS * BROW, w1, KC * 8, &bfull[rg.st]);
                            bulk_g2s(dstage + slotcap * SLOTW + KS * BROW + KC, w2, KC * 8, &bfull[rg.st]);
                        }
                        mbar_expect_tx(&full[rg.st], bytes);
                    }
                    __syncwarp();
                    if (TM) {
                        if (k0 >= 0) tensor_g2s(dstage + k0 * SLOT_TM, Ct, ch * KC, row0, &full[rg.st]);
                        if (k1 >= 0) tensor_g2s(dstage + k1 * SLOT_TM, Ct, ch * KC, row1, &full[rg.st]);
                        for (int k = lane + 64; k < mine; k += 32) {
                            const int s = 1 + k * NPROD + pw;
                            tensor_g2s(dstage + s * SLOT_TM, Ct, ch * KC, slot_atom[s] * NO, &full[rg.st]);
                        }
                    } else {
                    if (k0 >= 0) bulk_g2s(dstage + k0 * SLOT, csrc + src0, TILE * 8, &full[rg.st]);
                    if (k1 >= 0) bulk_g2s(dstage + k1 * SLOT, csrc + src1, TILE * 8, &full[rg.st]);
                    for (int k = lane + 64; k < mine; k += 32) {      // only for slot capacities beyond 64 * NPROD
                        const int s = 1 + k * NPROD + pw;
                        bulk_g2s(dstage + s * SLOT, csrc + (size_t)slot_atom[s] * TILE, TILE * 8, &full[rg.st]);
                    }
                    }
                    csrc += (size_t)na * TILE; w1 += KC; w2 += KC;
                    rg.next(stages);
                }
            } else if (warp == BW) {
                // ===== B warp: weighted B operand of atom i, once per chunk =====
                const int g = lane >> 2, t = lane & 3;
                for (int ch = 0; ch < nchunks; ++ch) {
                    mbar_spin(&bfull[rg.st], rg.ph, DBG && (dbg & 4));
                    double* stg = sbuf + (size_t)rg.st * stage_doubles;
                    double* Bs = stg + slotcap * SLOTW;
                    const double* wsm = Bs + KS * BROW;
#pragma unroll
                    for (int s = 0; s < KS; ++s) {
                        const double w_f = wsm[4 * s + t], w_e = wsm[KC + 4 * s + t];
                        const double ci = TM ? stg[g * 16 + 2 * ((2 * s + (t >> 1)) ^ g) + (t & 1)]      // row R = mu = g of slot 0, swizzled
                                             : stg[NO * 4 * s + 4 * g + t];       // C[i, mu = g][state 4 s + t]
                        Bs[s * BROW + lane] = ci * w_f;                           // Atemp = distre * Ha        TB.f90:4378
                        Bs[s * BROW + 32 + lane] = ci * w_e;                      // Atemp = distre * En * Ha   TB.f90:4459
                        if (g == 0) {
                            const double c8 = TM ? stg[8 * 16 + 4 * s + t]        // row 8: R & 7 = 0, no swizzle
                                                 : stg[NO * 4 * s + 32 + t];     // C[i, mu = 8][state]
                            Bs[s * BROW + 64 + t] = c8 * w_f;
                            Bs[s * BROW + 68 + t] = c8 * w_e;
                        }
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&bready[rg.st]);     // release: the stores above are visible to the waiters
                    rg.next(stages);
                }
            } else {
                // ===== consumer warps: row tiles warp, warp + NCW, ... of this sub-task =====
                const int nreg = ntiles - (ksp ? 1 : 0);      // regular row tiles (ksp: all but the first)
                int my_tiles = nreg > warp ? (nreg - warp + NCW - 1) / NCW : 0;
                if (ksp && warp >= 4 && warp < 8 && my_tiles == 0) my_tiles = 1;      // the special tile needs the loop (the regular tile is past the rows: dropped)
                const bool fs = first == 0;
                const int boff = slotcap * SLOTW;
                constexpr int MTMAX = NCW == 8 ? 7 : 5;      // row tiles per warp (register budget)
#define XTB_CASE(M) if constexpr (M <= MTMAX) { if (my_tiles == M) consume_subtask<NO, NCW, M, DBG, TM>(sbuf, stage_doubles, boff, stages, full, bready, empty, rg, nchunks, nrows, warp, lane, slot_q, b0i, fs, dbg, rho, rhoE, ksp, spart); }
                if (my_tiles == 0) consume_idle(stages, full, bready, empty, rg, nchunks, lane);
                XTB_CASE(1) XTB_CASE(2) XTB_CASE(3) XTB_CASE(4) XTB_CASE(5) XTB_CASE(6) XTB_CASE(7)
#undef XTB_CASE
            }
            if (KSPLIT && ksp) {      // (uniform per CTA) on-site rows 0..7: the four quarter sums in a fixed order
                __syncthreads();
                if (warp == 0) {
                    for (int x = lane; x < 144; x += 32) {
                        const double v = (spart[x] + spart[144 + x]) + (spart[288 + x] + spart[432 + x]);
                        const int mat = x / 72, y = x - mat * 72;          // y = 9 * row + col = NO * nu + mu
                        (mat == 0 ? rho : rhoE)[b0i * (NO * NO) + y] = v;
                    }
                }
            }
        }
    }
    // this CTA is done: it no longer holds anyone back
    if (warp == P0 && lane == 0) st_relaxed(prog + blockIdx.x, XTB_PROG_DONE);
}

#include "density_pair.cuh"

// ---------------- simple CUDA-core kernel (cross-check of the DMMA path; also the generic-n_orb path) ----------------
// one CTA per owned block; thread (mu,nu) accumulates over the kept states in their compacted order.
template <int NO>
__global__ void __launch_bounds__(128) density_simple_kernel(int na, int a0, int a1, int cap, long long N, int nkeep,
                                                            const int* __restrict__ keep, const int* __restrict__ nn_size,
                                                            const int* __restrict__ nn_list, const signed char* __restrict__ own,
                                                            const int* __restrict__ rowptr,
                                                            const double* __restrict__ Ha, const double* __restrict__ wf,
                                                            const double* __restrict__ wfe, double* __restrict__ rho,
                                                            double* __restrict__ rhoE) {
    const int i = a0 + blockIdx.y;
    const int q = blockIdx.x;        // 0 = on-site
    if (q > nn_size[i]) return;
    const int j = q == 0 ? i : nn_list[(size_t)i * cap + q - 1];
    if (q > 0 && !own[(size_t)i * cap + q - 1]) return;
    if (threadIdx.x >= NO * NO) return;
    const int mu = threadIdx.x % NO, nu = threadIdx.x / NO;
    const double* ci = Ha + (size_t)i * NO + mu;
    const double* cj = Ha + (size_t)j * NO + nu;
    double s = 0.0, e = 0.0;
    for (int kk = 0; kk < nkeep; ++kk) {
        const size_t o = (size_t)keep[kk] * N;
        const double a = ci[o], b = cj[o];
        s += (wf[kk] * a) * b;
        e += (wfe[kk] * a) * b;
    }
    const size_t blk = ((size_t)rowptr[i] + q) * (NO * NO);
    rho[blk + mu + NO * nu] = s;
    rhoE[blk + mu + NO * nu] = e;
}

// ---------------- small-basis kernels (sp3: Pettifor-type and DFTB sp sets, n_orb <= 4) ----------------
// prep: X[atom][kept state kk][NO rows] = Ha[atom*NO + r, keep[kk]] -- the rows of one atom for consecutive states are
// consecutive in memory.  (In Ha itself consecutive states of an atom are N*8 bytes apart, so a warp whose lanes stride
// the states touches 32 cache lines per load instruction: the first version of the kernel below ran at 71 % L1TEX
// throughput with every other unit idle, profiles/r01r_ncu_warp_full_summary.csv.)  CTA = (32 states, DR_AT atoms),
// transposed through shared memory, both phases coalesced; HBM/L2-bound, 2 * 8 * N * K bytes.
constexpr int DR_AT = 16;
template <int NO>
__global__ void __launch_bounds__(256) rows_by_state_kernel(int na, long long N, int nkeep, const int* __restrict__ keep,
                                                            const double* __restrict__ Ha, double* __restrict__ X) {
    constexpr int W = DR_AT * NO;
    __shared__ double tile[32][W + 1];
    __shared__ long long colbase[32];
    const int k0 = blockIdx.x * 32, atom0 = blockIdx.y * DR_AT;
    const int nat = min(DR_AT, na - atom0), ns = min(32, nkeep - k0);
    if (threadIdx.x < 32) colbase[threadIdx.x] = threadIdx.x < ns ? (long long)(keep ? keep[k0 + threadIdx.x] : k0 + threadIdx.x) * N + (long long)atom0 * NO : -1;
    __syncthreads();
    const int w = nat * NO;
    for (int idx = threadIdx.x; idx < 32 * w; idx += 256) {
        const int st = idx / w, e = idx - st * w;
        if (st < ns) tile[st][e] = Ha[colbase[st] + e];
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < nat * ns * NO; idx += 256) {
        const int a = idx / (ns * NO), rem = idx - a * (ns * NO);
        const int st = rem / NO, r = rem - st * NO;
        X[((size_t)(atom0 + a) * nkeep + k0 + st) * NO + r] = tile[st][a * NO + r];
    }
}

// One warp per owned block (i, q).  Lanes stride the kept states; lane kk reads the NO rows of atom i and of atom j of
// its state from X (32 bytes each for NO = 4; the 32 lanes of a load instruction cover 1 KB contiguous = 8 cache lines).
// Each lane keeps the 2 x NO x NO partial sums of its states; fixed-order butterfly at the end (same bits on every run).
// Bound: L1/L2 bandwidth -- 2 * 8 * NO bytes per (block, state) for 4 NO^2 flop.
constexpr int DW_WARPS = 4;
template <int NO>
__global__ void __launch_bounds__(32 * DW_WARPS) density_warp_kernel(int a0, int a1, int cap, int nkeep,
                                                                     const int* __restrict__ nn_size,
                                                                     const int* __restrict__ nn_list, const signed char* __restrict__ own,
                                                                     const int* __restrict__ rowptr, const double* __restrict__ X,
                                                                     const double* __restrict__ wf, const double* __restrict__ wfe,
                                                                     double* __restrict__ rho, double* __restrict__ rhoE) {
    const int i = a0 + blockIdx.y;
    const int q = blockIdx.x * DW_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;        // 0 = on-site
    if (i >= a1 || q > nn_size[i]) return;
    if (q > 0 && !own[(size_t)i * cap + q - 1]) return;
    const int j = q == 0 ? i : nn_list[(size_t)i * cap + q - 1];
    const double* ci = X + (size_t)i * nkeep * NO;
    const double* cj = X + (size_t)j * nkeep * NO;
    double s[NO * NO], e[NO * NO];
#pragma unroll
    for (int x = 0; x < NO * NO; ++x) { s[x] = 0.0; e[x] = 0.0; }
#pragma unroll 2
    for (int kk = lane; kk < nkeep; kk += 32) {
        const double w1 = wf[kk], w2 = wfe[kk];
        double a[NO], b[NO];
        if constexpr (NO % 2 == 0) {          // rows of (atom, state) start at a multiple of 16 bytes
#pragma unroll
            for (int m = 0; m < NO / 2; ++m) {
                const double2 va = __ldg(reinterpret_cast<const double2*>(ci + (size_t)kk * NO) + m);
                const double2 vb = __ldg(reinterpret_cast<const double2*>(cj + (size_t)kk * NO) + m);
                a[2 * m] = va.x; a[2 * m + 1] = va.y; b[2 * m] = vb.x; b[2 * m + 1] = vb.y;
            }
        } else {
#pragma unroll
            for (int m = 0; m < NO; ++m) { a[m] = __ldg(ci + (size_t)kk * NO + m); b[m] = __ldg(cj + (size_t)kk * NO + m); }
        }
#pragma unroll
        for (int mu = 0; mu < NO; ++mu) {
            const double af = w1 * a[mu], ae = w2 * a[mu];          // Atemp = distre * Ha ; distre * En * Ha   TB.f90:4378,4459
#pragma unroll
            for (int nu = 0; nu < NO; ++nu) {
                s[mu + NO * nu] = fma(af, b[nu], s[mu + NO * nu]);
                e[mu + NO * nu] = fma(ae, b[nu], e[mu + NO * nu]);
            }
        }
    }
#pragma unroll
    for (int x = 0; x < NO * NO; ++x) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            s[x] += __shfl_xor_sync(0xffffffffu, s[x], off);
            e[x] += __shfl_xor_sync(0xffffffffu, e[x], off);
        }
    }
    const size_t blk = ((size_t)rowptr[i] + q) * (NO * NO);
#pragma unroll
    for (int x = 0; x < NO * NO; ++x)
        if (lane == (x & 31)) { rho[blk + x] = s[x]; rhoE[blk + x] = e[x]; }
}

// non-owner blocks = transpose of the owner's block.  One CTA per atom i: warp 0 lists the entries i does not own (list
// order) with the block index of their source, then the threads stride the elements of those blocks with four independent
// load / store groups in flight (one small CTA per list entry, as before, spent 0.21 ms at 4096 atoms on CTA launches and
// one dependent load chain each).
template <int NO>
__global__ void __launch_bounds__(256) mirror_blocks_kernel(int i0, int cap, const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                                                           const int* __restrict__ rev, const signed char* __restrict__ own, const int* __restrict__ rowptr,
                                                           const double* __restrict__ rho_src, const double* __restrict__ rhoE_src,
                                                           double* __restrict__ rho, double* __restrict__ rhoE) {
    // (rho_src / rhoE_src are the same stores: the blocks read -- owned by the partner -- and the blocks written are disjoint,
    // and separate restrict pointers let the loads of the unrolled groups be issued ahead of the stores)
    extern __shared__ int msm[];                     // [2 * m]: list ordinal of the destination, block index of the source
    __shared__ int cnt_sh;
    const int i = i0 + blockIdx.x;
    const int m = nn_size[i];
    int* dq = msm; int* sb = msm + m;
    if (threadIdx.x < 32) {
        int cnt = 0;
        for (int base = 0; base < m; base += 32) {
            const int q = base + threadIdx.x;
            bool f = false; int src = 0;
            if (q < m) {
                const size_t e = (size_t)i * cap + q;
                const int rv = rev[e];
                f = own[e] == 0 && rv >= 0;
                if (f) src = rowptr[nn_list[e]] + 1 + rv;
            }
            const unsigned mask = __ballot_sync(0xffffffffu, f);
            if (f) { const int k = cnt + __popc(mask & ((1u << threadIdx.x) - 1u)); dq[k] = q; sb[k] = src; }
            cnt += __popc(mask);
        }
        if (threadIdx.x == 0) cnt_sh = cnt;
    }
    __syncthreads();
    const int tot = cnt_sh * (NO * NO);
    const size_t b0 = (size_t)rowptr[i] + 1;
#pragma unroll 4
    for (int idx = threadIdx.x; idx < tot; idx += 256) {
        const int k = idx / (NO * NO), e = idx - k * (NO * NO);
        const int nu = e / NO, mu = e - nu * NO;
        const size_t dst = (b0 + dq[k]) * (NO * NO) + e;
        const size_t src = (size_t)sb[k] * (NO * NO) + nu + NO * mu;
        rho[dst] = rho_src[src];
        rhoE[dst] = rhoE_src[src];
    }
}

}  // namespace

// Kept states (fe > f_cut) and their weights f_n, f_n E_n.  Runs when the occupations are SET (xtb_set_occupations,
// xtb_set_eigen*, xtb_set_density_cutoff), not inside the density stage: the count sizes the launches of that stage, and
// reading it back there would put a host round trip in the middle of the step.
int xtb_impl_prepare_occupations(xtb_ctx* c) {
    const long long K = c->K;
    int rc;
    if ((rc = xtb_reserve(c, c->d_keep, sizeof(int) * K))) return rc;
    if ((rc = xtb_reserve(c, c->d_wf, sizeof(double) * (K + KC)))) return rc;
    if ((rc = xtb_reserve(c, c->d_wfe, sizeof(double) * (K + KC)))) return rc;
    if ((rc = xtb_reserve(c, c->d_nkeep, 16))) return rc;
    keep_kernel<<<1, 1024, 0, c->stream>>>(K, c->d_fe, c->d_Ei, c->f_cut, (int*)c->d_keep.p, (double*)c->d_wf.p,
                                           (double*)c->d_wfe.p, (int*)c->d_nkeep.p);
    XTB_LAUNCH_CHECK(c);
    int nkeep = 0, last = -1;
    XTB_CUDA(c, cudaMemcpyAsync(&nkeep, c->d_nkeep.p, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    if (nkeep > 0) XTB_CUDA(c, cudaMemcpy(&last, (const int*)c->d_keep.p + (nkeep - 1), sizeof(int), cudaMemcpyDeviceToHost));
    c->k_kept = nkeep;
    c->keep_prefix = nkeep == 0 || last == nkeep - 1;      // kept states are ascending and distinct: a prefix iff the last one is nkeep-1
    const int kpad = (nkeep + KC - 1) / KC * KC;
    if (kpad > nkeep) {
        pad_weights_kernel<<<1, KC, 0, c->stream>>>(nkeep, kpad, (double*)c->d_wf.p, (double*)c->d_wfe.p);
        XTB_LAUNCH_CHECK(c);
    }
    c->have_keep = true;
    return XTB_OK;
}

namespace {

// tensor map of HaT (rows x states, states contiguous): boxes of 16 states x 9 rows, 128-byte swizzle; kept in global memory
int ensure_tensor_map(xtb_ctx* c, long long nrows, long long nstates) {
    if (c->tmap_for == (const void*)c->d_HaT && c->tmap_rows == nrows && c->tmap_states == nstates && c->d_tmap.p) return XTB_OK;
    typedef CUresult (*encode_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                 const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static encode_t enc = nullptr;
    if (!enc) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn) {
            cudaGetLastError();
            XTB_FAIL(c, XTB_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
        }
        enc = (encode_t)fn;
    }
    alignas(64) CUtensorMap tm;
    const cuuint64_t dims[2] = {(cuuint64_t)nstates, (cuuint64_t)nrows};
    const cuuint64_t strides[1] = {(cuuint64_t)nstates * 8};
    const cuuint32_t box[2] = {(cuuint32_t)KC, 9};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, (void*)c->d_HaT, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) XTB_FAIL(c, XTB_ECUDA, "cuTensorMapEncodeTiled failed (%d): HaT %p, %lld rows x %lld states", (int)r, (const void*)c->d_HaT, nrows, nstates);
    int rc;
    if ((rc = xtb_reserve(c, c->d_tmap, sizeof tm))) return rc;
    XTB_CUDA(c, cudaMemcpyAsync(c->d_tmap.p, &tm, sizeof tm, cudaMemcpyHostToDevice, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));      // `tm` is a stack object
    c->tmap_for = (const void*)c->d_HaT; c->tmap_rows = nrows; c->tmap_states = nstates;
    return XTB_OK;
}

template <int NO>
int run_density(xtb_ctx* c) {
    const int na = c->na;
    const long long N = (long long)na * NO;
    int rc;
    if (!c->have_keep && (rc = xtb_impl_prepare_occupations(c))) return rc;
    const int nkeep = (int)c->k_kept;
    const size_t blk_bytes = sizeof(double) * NO * NO * xtb_blocks_alloc(c);
    if ((rc = xtb_reserve(c, c->d_rho, blk_bytes))) return rc;
    if ((rc = xtb_reserve(c, c->d_rhoE, blk_bytes))) return rc;
    XTB_T0(c, 3);
    const int nchunks = (nkeep + KC - 1) / KC;
    const int kpad = nchunks * KC;
    const int nown = c->a1 - c->a0;
    // blocks that no kernel below writes (atoms outside [a0,a1)) must not hold garbage; on a full range every block is
    // written by its owner or by the mirror pass
    // (a communicating context reads only its own rows and what the halo exchange delivers: nothing to clear there)
    if ((nown < na && !(c->comm && c->nranks > 1)) || nkeep == 0) {
        XTB_CUDA(c, cudaMemsetAsync(c->d_rho.p, 0, blk_bytes, c->stream));
        XTB_CUDA(c, cudaMemsetAsync(c->d_rhoE.p, 0, blk_bytes, c->stream));
    }
    bool use_dmma = (c->density_impl >= 1) && NO == 9 && nown > 0;
    // state-minor eigenvectors on the device and the kept states a prefix: the tiles come straight from HaT (tensor-map TMA),
    // no re-tiled copy.  XTB_DENSITY_TM=0 forces the Ct path (needs the Fortran-layout Ha: transposed on demand).
    bool use_tm = use_dmma && c->density_impl == 1 && !c->density_dbg && c->density_ncw != 12 && c->d_HaT && c->keep_prefix && KC == 16 &&
                  (c->K % 2) == 0;       // the row pitch of the tensor map (K * 8 bytes) must be a multiple of 16
    if (const char* ev = getenv("XTB_DENSITY_TM")) { if (atoi(ev) == 0) use_tm = false; }
    if (use_dmma && !use_tm && (rc = xtb_ensure_Ha_std(c))) return rc;
    if (!use_dmma && nown > 0 && (rc = xtb_ensure_Ha_std(c))) return rc;
    if (use_tm) {
        if ((rc = ensure_tensor_map(c, N, c->K))) return rc;
        XTB_T0(c, 7);
    }
    if (use_dmma) {
        if (!use_tm && (rc = xtb_reserve(c, c->d_Ct, sizeof(double) * (size_t)kpad * N + 1024))) return rc;
        if (nchunks > 0 && !use_tm) {
            dim3 grid(nchunks, (na + CT_AT - 1) / CT_AT);
            const int* need = nullptr;
            if (nown < na) {      // re-tile only what the owned atoms read (own rows + halo)
                if ((rc = xtb_reserve(c, c->d_ctneed, sizeof(int) * grid.y))) return rc;
                XTB_CUDA(c, cudaMemsetAsync(c->d_ctneed.p, 0, sizeof(int) * grid.y, c->stream));
                need_groups_kernel<<<(nown + 127) / 128, 128, 0, c->stream>>>(c->a0, c->a1, c->cap, (const int*)c->d_nn_size.p,
                                                                             (const int*)c->d_nn_list.p, (int*)c->d_ctneed.p);
                XTB_LAUNCH_CHECK(c);
                need = (const int*)c->d_ctneed.p;
            }
            XTB_T0(c, 16);        // the re-tiling kernel alone (HBM-bound: reads the kept columns of Ha, writes Ct)
            ct_tile_kernel<NO><<<grid, CT_THREADS, 0, c->stream>>>(na, N, nkeep, (const int*)c->d_keep.p, need, c->d_Ha, (double*)c->d_Ct.p);
            XTB_LAUNCH_CHECK(c);
            XTB_T1(c, 16);
        }
        if (!use_tm) XTB_T0(c, 7);             // the contraction kernel alone (stage 7)
        if (c->density_impl == 2) {
            // ---- two column atoms per CTA (density_pair.cuh) ----
            const int gcap = 2 * c->max_nn + 4;
            const size_t budget = 227 * 1024 - 1024;
            // slot capacity: the union of the two owned sets of a pair of neighbours is ~2/3 of the sum (half-balls displaced by
            // one bond), + the two atoms; never more than 8 * MUMAX row tiles per column atom can be dealt, and 3 stages must fit
            int slotcap = (int)(0.7 * c->max_nn) + 6;
            if (slotcap > c->max_nn + 3) slotcap = c->max_nn + 3;
            if (slotcap < 4) slotcap = 4;
            while (slotcap > 4 && pairk::smem_bytes(slotcap, 3, gcap) > budget) --slotcap;
            int stages = 3;
            while (stages < 6 && pairk::smem_bytes(slotcap, stages + 1, gcap) <= budget) ++stages;
            const size_t smem = pairk::smem_bytes(slotcap, stages, gcap);
            if (smem > budget) XTB_FAIL(c, XTB_ECAPACITY, "density: %d neighbours per atom do not fit the unit tables in shared memory", c->max_nn);
            // units: matching of the owned atoms on the device (no host round trip; the kernel reads the unit count from memory)
            if ((rc = xtb_reserve(c, c->d_units, sizeof(int) * (2 * (size_t)nown + 4 * (size_t)na + 4)))) return rc;
            if ((rc = xtb_reserve(c, c->d_pairw, sizeof(unsigned) * (size_t)na * c->cap))) return rc;
            int* d_units = (int*)c->d_units.p; int* d_nunits = d_units + 2 * (size_t)nown;
            int* d_partner = d_nunits + 4; int* d_prop = d_partner + na; int* d_nown = d_prop + na;
            pairk::own_count_kernel<<<(na + 127) / 128, 128, 0, c->stream>>>(na, c->cap, (const int*)c->d_nn_size.p, (const signed char*)c->d_own.p, d_nown);
            XTB_LAUNCH_CHECK(c);
            pairk::pair_weights_kernel<<<(nown + 3) / 4, 128, 0, c->stream>>>(c->a0, c->a1, c->cap, slotcap, (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p,
                                                                              (const signed char*)c->d_own.p, d_nown, (unsigned*)c->d_pairw.p);
            XTB_LAUNCH_CHECK(c);
            pairk::pair_match_kernel<<<1, 1024, 0, c->stream>>>(c->a0, c->a1, c->cap, c->density_pairing ? 16 : 0, (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p,
                                                                (const unsigned*)c->d_pairw.p, d_partner, d_prop, d_units, d_nunits);
            XTB_LAUNCH_CHECK(c);
            const int wmin = slotcap - 2 < 40 ? slotcap - 2 : 40;       // every window but the last holds at least this many row atoms
            const int maxsub = (2 * c->max_nn + 2 + wmin - 1) / wmin + 1;
            int grid = c->sm_count < nown ? c->sm_count : nown;          // units <= owned atoms; surplus CTAs leave at once
            if ((rc = xtb_reserve(c, c->d_prog, sizeof(long long) * (size_t)c->sm_count))) return rc;
            XTB_CUDA(c, cudaMemsetAsync(c->d_prog.p, 0, sizeof(long long) * (size_t)c->sm_count, c->stream));
            pairk::KArgs A;
            A.na = na; A.nunits = d_nunits; A.cap = c->cap; A.slotcap = slotcap; A.stages = stages; A.nchunks = nchunks; A.maxsub = maxsub;
            A.lead = c->density_lead; A.dbg = c->density_dbg; A.prog = (long long*)c->d_prog.p; A.units = d_units;
            A.nn_size = (const int*)c->d_nn_size.p; A.nn_list = (const int*)c->d_nn_list.p; A.own = (const signed char*)c->d_own.p;
            A.rowptr = (const int*)c->d_rowptr.p; A.Ct = (const double*)c->d_Ct.p; A.wf = (const double*)c->d_wf.p; A.wfe = (const double*)c->d_wfe.p;
            A.rho = (double*)c->d_rho.p; A.rhoE = (double*)c->d_rhoE.p;
            XTB_T0(c, 7);         // the contraction kernel alone (stage 7)
            XTB_CUDA(c, cudaFuncSetAttribute(pairk::density_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            pairk::density_pair_kernel<<<grid, pairk::THREADS, smem, c->stream>>>(A, gcap);
            XTB_LAUNCH_CHECK(c);
            c->density_slotcap = slotcap; c->density_stages = stages;
        } else {
        // slot capacity / stages from the shared-memory budget (227 KB per CTA)
        const int NCW = (c->density_ncw == 12 && !c->density_dbg) ? 12 : 8;     // the debug instantiation exists for 8 consumer warps only
        const int MTMAX = NCW == 8 ? 7 : 5;
        int slotcap = (c->max_nn + 1) / 2 + 6;                 // balanced ownership: half the neighbours (+ slack), + self
        if (slotcap > MTMAX * NCW * 8 / NO) slotcap = MTMAX * NCW * 8 / NO;     // rows must fit MTMAX tiles per warp
        if (slotcap > c->max_nn + 1) slotcap = c->max_nn + 1;
        if (slotcap < 2) slotcap = 2;
        auto stage_bytes = [&](int sc) { return use_tm ? (size_t)stage_doubles_tm(sc) * 8 : (size_t)(sc * SLOT + KS * BROW + 2 * KC) * 8; };
        auto fixed_bytes = [&](int sc, int st) { return (size_t)4 * 8 * st + sizeof(int) * (2 * (size_t)sc + c->cap) + 64 + (use_tm ? 1024 : 0) + (KSPLIT ? 4608 + 8 : 0); };
        const size_t budget = 227 * 1024 - 1024;
        int stages = 4;
        while (stages > 2 && stages * stage_bytes(slotcap) + fixed_bytes(slotcap, stages) > budget) {
            // prefer 4 stages with a slightly smaller slot capacity over 3 stages, as long as the mean fits
            if (slotcap > (c->max_nn + 1) / 2 + 2) --slotcap; else --stages;
        }
        while (stages * stage_bytes(slotcap) + fixed_bytes(slotcap, stages) > budget && slotcap > 2) --slotcap;
        const size_t smem = stages * stage_bytes(slotcap) + fixed_bytes(slotcap, stages);
        // few rounds on the persistent grid (partitioned runs, small cells) AND atoms that need two sub-tasks anyway (more
        // owned neighbours than slots, e.g. al500: 88 of 176): give the two halves to different CTAs.  Atoms that fit one
        // sub-task are NOT split: half the row tiles per warp costs 36 % per unit of work (measured, au4096: 94.6 vs 69.3 ms).
        int split = (nown < 6 * c->sm_count && nown > c->sm_count / 2 && c->max_nn / 2 >= slotcap - 1) ? 2 : 1;
        if (const char* ev = getenv("XTB_DENSITY_SPLIT")) { const int v = atoi(ev); if (v >= 1 && v <= 4) split = v; }
        int grid = c->sm_count < nown * split ? c->sm_count : nown * split;
        if ((rc = xtb_reserve(c, c->d_prog, sizeof(long long) * (size_t)c->sm_count))) return rc;
        XTB_CUDA(c, cudaMemsetAsync(c->d_prog.p, 0, sizeof(long long) * (size_t)c->sm_count, c->stream));
        const int maxsub = slotcap > 1 ? (c->max_nn + slotcap - 2) / (slotcap - 1) + 1 : 1;
        int a_na = na, a_a0 = c->a0, a_a1 = c->a1, a_cap = c->cap, a_slot = slotcap, a_st = stages, a_nch = nchunks, a_dbg = c->density_dbg, a_ms = maxsub, a_lead = c->density_lead, a_split = split;
        long long* a_prog = (long long*)c->d_prog.p;
        const int* a_sz = (const int*)c->d_nn_size.p; const int* a_li = (const int*)c->d_nn_list.p;
        const signed char* a_own = (const signed char*)c->d_own.p; const int* a_rp = (const int*)c->d_rowptr.p;
        const double* a_ct = use_tm ? (const double*)c->d_tmap.p : (const double*)c->d_Ct.p;      // TM: the tensor map travels in the Ct slot
        const double* a_wf = (const double*)c->d_wf.p; const double* a_wfe = (const double*)c->d_wfe.p;
        double* a_rho = (double*)c->d_rho.p; double* a_rhoE = (double*)c->d_rhoE.p;
        void* args[] = {&a_na, &a_a0, &a_a1, &a_cap, &a_slot, &a_st, &a_nch, &a_dbg, &a_ms, &a_lead, &a_split, &a_prog, &a_sz, &a_li, &a_own, &a_rp,
                        &a_ct, &a_wf, &a_wfe, &a_rho, &a_rhoE};
#define XTB_LAUNCH(W, D) { auto kern = density_dmma_kernel<9, W, D>; \
        XTB_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        XTB_CUDA(c, cudaLaunchKernel((const void*)kern, dim3(grid), dim3((W + 1 + NPROD) * 32), args, smem, c->stream)); }
        // Last round of the persistent grid: if its atoms fill at most half the CTAs, they run as a second launch of the
        // same kernel with two units per atom (512 owned atoms on 148 SMs: 3 + 0.68 rounds instead of 4).  The CTAs of a
        // launch run in lock-step and finish together, so the second launch starts on an idle GPU; the kernel itself is
        // untouched (an in-kernel tail decode cost 1.3 % of the single-GPU run through register allocation, measured).
        int tail_from = nown;                      // first owned atom (0-based within the range) of the split tail
        if (split == 1 && nown > grid) {
            const int full = (nown / grid) * grid, rest = nown - full;
            if (rest > 0 && 2 * rest <= grid) tail_from = full;
        }
        if (const char* ev = getenv("XTB_DENSITY_TAIL")) {      // tests: force the position of the tail (negative = none)
            const int v = atoi(ev);
            tail_from = (split == 1 && v >= 0 && v < nown) ? v : nown;
        }
#define XTB_LAUNCH_TM { auto kern = density_dmma_kernel<9, 8, false, true>; \
        XTB_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        XTB_CUDA(c, cudaLaunchKernel((const void*)kern, dim3(grid), dim3((8 + 1 + NPROD) * 32), args, smem, c->stream)); }
#define XTB_LAUNCH_ALL if (use_tm) XTB_LAUNCH_TM else if (c->density_dbg) XTB_LAUNCH(8, true) else if (NCW == 12) XTB_LAUNCH(12, false) else XTB_LAUNCH(8, false)
        if (tail_from > 0) {
            a_a1 = c->a0 + tail_from;
            grid = c->sm_count < tail_from * split ? c->sm_count : tail_from * split;
            XTB_LAUNCH_ALL
            XTB_LAUNCH_CHECK(c);
        }
        if (tail_from < nown) {
            if (tail_from > 0) XTB_CUDA(c, cudaMemsetAsync(c->d_prog.p, 0, sizeof(long long) * (size_t)c->sm_count, c->stream));
            a_a0 = c->a0 + tail_from; a_a1 = c->a1; a_split = 2;
            grid = c->sm_count < (nown - tail_from) * 2 ? c->sm_count : (nown - tail_from) * 2;
            XTB_LAUNCH_ALL
            XTB_LAUNCH_CHECK(c);
        }
#undef XTB_LAUNCH_ALL
#undef XTB_LAUNCH_TM
#undef XTB_LAUNCH
        c->density_slotcap = slotcap; c->density_stages = stages; c->density_tm = use_tm ? 1 : 0;
        }
    } else if (nown > 0 && NO <= 4 && c->density_impl >= 1) {
        if constexpr (NO <= 4) {
            if ((rc = xtb_reserve(c, c->d_Ct, sizeof(double) * (size_t)(nkeep > 0 ? nkeep : 1) * N + 1024))) return rc;
            if (nkeep > 0) {
                dim3 g((nkeep + 31) / 32, (na + DR_AT - 1) / DR_AT);
                rows_by_state_kernel<NO><<<g, 256, 0, c->stream>>>(na, N, nkeep, (const int*)c->d_keep.p, c->d_Ha, (double*)c->d_Ct.p);
                XTB_LAUNCH_CHECK(c);
            }
            XTB_T0(c, 7);
            dim3 grid((c->max_nn + 1 + DW_WARPS - 1) / DW_WARPS, nown);
            density_warp_kernel<NO><<<grid, 32 * DW_WARPS, 0, c->stream>>>(c->a0, c->a1, c->cap, nkeep,
                                                                            (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p,
                                                                            (const signed char*)c->d_own.p, (const int*)c->d_rowptr.p, (const double*)c->d_Ct.p,
                                                                            (const double*)c->d_wf.p, (const double*)c->d_wfe.p,
                                                                            (double*)c->d_rho.p, (double*)c->d_rhoE.p);
        }
        XTB_LAUNCH_CHECK(c);
    } else if (nown > 0) {
        XTB_T0(c, 7);
        dim3 grid(c->max_nn + 1, nown);
        density_simple_kernel<NO><<<grid, 128, 0, c->stream>>>(na, c->a0, c->a1, c->cap, N, nkeep, (const int*)c->d_keep.p,
                                                               (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p,
                                                               (const signed char*)c->d_own.p, (const int*)c->d_rowptr.p, c->d_Ha, (const double*)c->d_wf.p,
                                                               (const double*)c->d_wfe.p, (double*)c->d_rho.p, (double*)c->d_rhoE.p);
        XTB_LAUNCH_CHECK(c);
    } else {
        XTB_T0(c, 7);
    }
    XTB_T1(c, 7);
    // mirror (complete only when every owner block is local: single-GPU, or after the caller's allgather)
    if (c->a0 == 0 && c->a1 == na && c->max_nn > 0) {
        mirror_blocks_kernel<NO><<<na, 256, sizeof(int) * 2 * (size_t)c->max_nn, c->stream>>>(
            0, c->cap, (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p, (const int*)c->d_rev.p, (const signed char*)c->d_own.p,
            (const int*)c->d_rowptr.p, (const double*)c->d_rho.p, (const double*)c->d_rhoE.p, (double*)c->d_rho.p, (double*)c->d_rhoE.p);
        XTB_LAUNCH_CHECK(c);
    }
    XTB_T1(c, 3);
    c->have_density = true;
    return xtb_stage_finish(c);
}

}  // namespace

// X[atom][state][n_orb rows] of ALL states of Ha (coupling.cu reads the eigenvector rows of an atom through it: a warp's load
// then covers 32 consecutive states of one atom = contiguous memory, where Ha itself has them a whole column apart)
int xtb_rows_by_state(xtb_ctx* c, const double* Ha, long long K, double* X) {
    const int na = c->na;
    const long long N = (long long)na * c->n_orb;
    dim3 g((unsigned)((K + 31) / 32), (na + DR_AT - 1) / DR_AT);
    if (c->n_orb == 9) rows_by_state_kernel<9><<<g, 256, 0, c->stream>>>(na, N, (int)K, nullptr, Ha, X);
    else if (c->n_orb == 4) rows_by_state_kernel<4><<<g, 256, 0, c->stream>>>(na, N, (int)K, nullptr, Ha, X);
    else rows_by_state_kernel<1><<<g, 256, 0, c->stream>>>(na, N, (int)K, nullptr, Ha, X);
    XTB_LAUNCH_CHECK(c);
    return XTB_OK;
}

int xtb_impl_density(xtb_ctx* c) {
    if (c->n_orb == 9) return run_density<9>(c);
    if (c->n_orb == 4) return run_density<4>(c);
    if (c->n_orb == 1) return run_density<1>(c);
    XTB_FAIL(c, XTB_EINVAL, "density: unsupported n_orb %d", c->n_orb);
}

int xtb_impl_mirror_density(xtb_ctx* c) {
    if (c->max_nn <= 0) return XTB_OK;
    // a communicating context holds the sources of its own rows only (halo exchange)
    const int i0 = c->comm ? c->a0 : 0, rows = c->comm ? c->a1 - c->a0 : c->na;
    if (rows <= 0) return XTB_OK;
    const size_t msm = sizeof(int) * 2 * (size_t)c->max_nn;
#define XTB_MIRROR(NO) mirror_blocks_kernel<NO><<<rows, 256, msm, c->stream>>>(i0, c->cap, (const int*)c->d_nn_size.p, (const int*)c->d_nn_list.p, \
        (const int*)c->d_rev.p, (const signed char*)c->d_own.p, (const int*)c->d_rowptr.p, (const double*)c->d_rho.p, (const double*)c->d_rhoE.p, (double*)c->d_rho.p, (double*)c->d_rhoE.p)
    if (c->n_orb == 1) XTB_MIRROR(1); else if (c->n_orb == 9) XTB_MIRROR(9); else XTB_MIRROR(4);
#undef XTB_MIRROR
    XTB_LAUNCH_CHECK(c);
    return XTB_OK;
}
