// density_pair.cuh -- rho / rho_E contraction, two column atoms per CTA (included by density.cu, inside its namespace).
//
// Same contraction as density_dmma_kernel (Construct_Aij / Construct_Aij_x_En, TB.f90:4328-4469, on the neighbour
// pattern), but the work unit is a PAIR of nearby atoms (i1, i2): the eigenvector tiles of the neighbours both atoms own
// are fetched once and their DMMA A fragments feed the columns of both atoms.  Measured motivation: the
// one-atom kernel is bound by its consumer loop (one LDS.64 per 2 DMMA + 2 DFMA, two barrier round trips per 16 states,
// row tiles dealt to warps in whole tiles: 12/11.25 imbalance between SM sub-partitions); per pair of adjacent atoms of
// the au4096 cell 27 of the 2 x 40 owned rows are common (52.5 instead of 79.8 tiles per chunk through L2 -> shared memory).
//
// Unit (i1, i2) [i2 = -1: single atom]:
//   row atoms  G = [class 1: needed by i1 only][class 3: needed by both][class 2: needed by i2 only]; an atom x is needed by
//              i_a iff x == i_a (on-site block) or i_a owns the pair (i_a, x).  Every entry carries its ordinal in the list
//              of i1 and of i2 (-1 = the atom itself), so the result goes to block rowptr[i_a] + 1 + ordinal.
//   sub-task   = window of G that fits the slot capacity and 48 row tiles per column atom; shared-memory slot = position
//              in the window, so consecutive rows stay consecutive in bank space (the conflict-free A fragment pattern of
//              the one-atom kernel); i1 / i2 are appended as extra slots when the window does not hold them (B operand).
//   halves     row tile tau x column atom a is one "half" = 2 DMMA + 2 DFMA per k-step.  Halves, not tiles, are dealt to
//              the 8 consumer warps: x-halves (columns of i1) round-robin from the last tile that has class-1/3 rows
//              downwards, y-halves from the same tile downwards through the class-3 tiles, then upwards through the
//              class-2 tiles, the remainder of the y deal going to the warps the x deal left short.  A warp holds up to
//              6 (x, y) slots; when both halves of a slot are the same tile the A fragment is loaded once.
#pragma once

namespace pairk {

constexpr int MUMAX = 6;                 // (x, y) slots per consumer warp: 12 accumulator doubles each (register budget 232)
constexpr int NCW = 8;                   // consumer warps
constexpr int BROW2 = 2 * BROW;          // B operand of both column atoms per k-step

// position of atom x in the (ascending) list of atom i, or -1
__device__ __forceinline__ int find_in_list(const int* __restrict__ L, int m, int x) {
    int lo = 0, hi = m - 1;
    while (lo <= hi) {
        const int mid = (lo + hi) >> 1, v = L[mid];
        if (v == x) return mid;
        if (v < x) lo = mid + 1; else hi = mid - 1;
    }
    return -1;
}

// ---- units: pairs of owned atoms that share as many row atoms as possible ----
// Two owned atoms i, j that are neighbours can form a unit if the union of their row sets fits the slot capacity (one
// window, one pass over the states).  weight(i, j) = number of common row atoms (ties broken by a symmetric hash), 0 if
// not allowed.  Matching: every free atom proposes to its heaviest free neighbour, mutual proposals become pairs (locally
// dominant edges: at least the heaviest remaining edge is matched in every round; on a lattice a dozen rounds leave a few
// per cent of the atoms single).  One CTA does the rounds and the ordered compaction: pairs first, then singles.
__device__ __forceinline__ unsigned pair_hash(int a, int b) {
    unsigned lo = a < b ? a : b, hi = a < b ? b : a;
    unsigned h = lo * 0x9E3779B1u ^ (hi + 0x7F4A7C15u) * 0x85EBCA6Bu;
    h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12;
    return h & 0xFFFFFu;
}

// one warp per owned atom i, lanes over its list entries: weight of the candidate pair (i, j)
__global__ void own_count_kernel(int na, int cap, const int* __restrict__ nn_size, const signed char* __restrict__ own, int* __restrict__ nown) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= na) return;
    int n = 0;
    for (int q = 0; q < nn_size[i]; ++q) n += own[(size_t)i * cap + q] != 0;
    nown[i] = n;
}
__global__ void __launch_bounds__(128) pair_weights_kernel(int a0, int a1, int cap, int slotcap, const int* __restrict__ nn_size,
                                                           const int* __restrict__ nn_list, const signed char* __restrict__ own,
                                                           const int* __restrict__ nown, unsigned* __restrict__ wgt) {
    const int i = a0 + blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= a1) return;
    const int mi = nn_size[i];
    const int* Li = nn_list + (size_t)i * cap; const signed char* Oi = own + (size_t)i * cap;
    const int ni = nown[i];
    for (int q = lane; q < mi; q += 32) {
        const int j = Li[q];
        unsigned w = 0;
        if (j >= a0 && j < a1 && j != i) {
            const int mj = nn_size[j];
            const int* Lj = nn_list + (size_t)j * cap; const signed char* Oj = own + (size_t)j * cap;
            // |own(i) & own(j)| by merging the two ascending lists; the atoms themselves add one common row atom
            // (exactly one of the two owns the pair (i, j)) -- or two if the pair is missing from one list (both own it)
            int x = 0, y = 0, common = 0;
            const int nj = nown[j];
            while (x < mi && y < mj) {
                const int vx = Li[x], vy = Lj[y];
                if (vx == vy) { common += (Oi[x] != 0) && (Oj[y] != 0); ++x; ++y; }
                else if (vx < vy) ++x; else ++y;
            }
            int selfs = Oi[q] != 0;                               // j is a row atom of i
            { const int p = find_in_list(Lj, mj, i); selfs += (p >= 0 && Oj[p] != 0); }
            common += selfs;
            const int uni = (ni + 1) + (nj + 1) - common;
            if (uni <= slotcap) w = ((unsigned)common << 20) | pair_hash(i, j);
        }
        wgt[(size_t)i * cap + q] = w;
    }
}

// single CTA: matching rounds + ordered compaction.  units[2u], units[2u+1] = (i1, i2 or -1); *nunits = number of units.
__global__ void __launch_bounds__(1024) pair_match_kernel(int a0, int a1, int cap, int rounds, const int* __restrict__ nn_size,
                                                          const int* __restrict__ nn_list, const unsigned* __restrict__ wgt,
                                                          int* __restrict__ partner, int* __restrict__ prop,
                                                          int* __restrict__ units, int* __restrict__ nunits) {
    __shared__ int scan[1024];
    __shared__ int carry;
    for (int i = a0 + threadIdx.x; i < a1; i += 1024) partner[i] = -1;
    __syncthreads();
    for (int r = 0; r < rounds; ++r) {
        for (int i = a0 + threadIdx.x; i < a1; i += 1024) {
            int best = -1; unsigned bw = 0;
            if (partner[i] < 0) {
                const int m = nn_size[i];
                for (int q = 0; q < m; ++q) {
                    const unsigned w = wgt[(size_t)i * cap + q];
                    if (w > bw) { const int j = nn_list[(size_t)i * cap + q]; if (partner[j] < 0) { bw = w; best = j; } }
                }
            }
            prop[i] = best;
        }
        __syncthreads();
        for (int i = a0 + threadIdx.x; i < a1; i += 1024) {
            const int j = prop[i];
            if (j >= 0 && prop[j] == i) partner[i] = j;
        }
        __syncthreads();
    }
    // compaction in atom order: pass 0 the pairs (at their smaller atom), pass 1 the singles
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int pass = 0; pass < 2; ++pass) {
        for (int base = a0; base < a1; base += 1024) {
            const int i = base + threadIdx.x;
            int flag = 0;
            if (i < a1) { const int pj = partner[i]; flag = pass == 0 ? (pj > i) : (pj < 0); }
            scan[threadIdx.x] = flag;
            __syncthreads();
            for (int off = 1; off < 1024; off <<= 1) {
                const int tv = threadIdx.x >= off ? scan[threadIdx.x - off] : 0;
                __syncthreads();
                scan[threadIdx.x] += tv;
                __syncthreads();
            }
            if (flag) {
                const int u = carry + scan[threadIdx.x] - 1;
                units[2 * u] = i; units[2 * u + 1] = pass == 0 ? partner[i] : -1;
            }
            __syncthreads();
            if (threadIdx.x == 1023) carry += scan[1023];
            __syncthreads();
        }
    }
    if (threadIdx.x == 0) *nunits = carry;
}

// Unit tables in shared memory (ints).  qq packs the two list ordinals: (q1 + 1) | (q2 + 1) << 16, 0 = "the atom itself".
struct UnitTab {
    int* g_atom;      // [gcap]  row atoms of the unit, class-ordered
    int* g_qq;        // [gcap]
    int* g_mask;      // [gcap]  1, 3 or 2
    int* cp_slot;     // [slotcap] window: slots that go to the `full` barrier (all but the two B slots)
    int* scal;        // [16] scalars, see S_*
};
enum { S_NG = 0, S_N1, S_NB, S_N2, S_W0, S_W1, S_POSB1, S_POSB2, S_NSLOTS, S_NCP, S_N1W, S_NBW, S_N2W };

// warp 0: build G for the unit (two passes over the same classification: count, then place)
__device__ __forceinline__ void build_unit(const UnitTab& T, int i1, int i2, int cap, const int* __restrict__ nn_size,
                                           const int* __restrict__ nn_list, const signed char* __restrict__ own, int lane) {
    const int m1 = nn_size[i1], m2 = i2 >= 0 ? nn_size[i2] : 0;
    const int* L1 = nn_list + (size_t)i1 * cap; const int* L2 = i2 >= 0 ? nn_list + (size_t)i2 * cap : L1;
    const signed char* O1 = own + (size_t)i1 * cap; const signed char* O2 = i2 >= 0 ? own + (size_t)i2 * cap : O1;
    int n1 = 0, nb = 0, n2 = 0;          // counts (pass 0) / running positions (pass 1)
    int base1 = 0, baseb = 0, base2 = 0;
    for (int pass = 0; pass < 2; ++pass) {
        if (pass == 1) { base1 = 0; baseb = n1; base2 = n1 + nb; n1 = nb = n2 = 0; }
        // (A) the owned neighbours of i1 (and i1 itself, as entry -1)
        for (int b0 = -1; b0 < m1; b0 += 32) {
            const int q = b0 + lane;                 // -1 = i1 itself
            int x = -1, cls = 0, qq = 0;
            if (q == -1) {
                x = i1;
                int p2 = i2 >= 0 ? find_in_list(L2, m2, i1) : -1;
                const bool in2 = p2 >= 0 && O2[p2] != 0;
                cls = in2 ? 3 : 1; qq = 0 | ((in2 ? p2 + 1 : 0) << 16);
            } else if (q < m1 && O1[q] != 0) {
                x = L1[q];
                int p2 = -1; bool in2 = false;
                if (i2 >= 0) { if (x == i2) { in2 = true; p2 = -1; } else { p2 = find_in_list(L2, m2, x); in2 = p2 >= 0 && O2[p2] != 0; } }
                cls = in2 ? 3 : 1; qq = (q + 1) | ((in2 ? p2 + 1 : 0) << 16);
            }
            const unsigned mk1 = __ballot_sync(0xffffffffu, cls == 1), mkb = __ballot_sync(0xffffffffu, cls == 3);
            const unsigned below = (1u << lane) - 1u;
            if (pass == 1 && cls) {
                const int pos = cls == 1 ? base1 + n1 + __popc(mk1 & below) : baseb + nb + __popc(mkb & below);
                T.g_atom[pos] = x; T.g_qq[pos] = qq; T.g_mask[pos] = cls;
            }
            n1 += __popc(mk1); nb += __popc(mkb);
        }
        // (B) the owned neighbours of i2 (and i2 itself) that i1 does not need
        if (i2 >= 0) {
            for (int b0 = -1; b0 < m2; b0 += 32) {
                const int q = b0 + lane;
                int x = -1, cls = 0, qq = 0;
                if (q == -1) {
                    const int p1 = find_in_list(L1, m1, i2);
                    if (!(p1 >= 0 && O1[p1] != 0)) { x = i2; cls = 2; qq = 0; }        // otherwise already placed by (A) as class 3
                } else if (q < m2 && O2[q] != 0) {
                    x = L2[q];
                    if (x != i1) {                                                     // i1 itself was placed by (A)
                        const int p1 = find_in_list(L1, m1, x);
                        if (!(p1 >= 0 && O1[p1] != 0)) { cls = 2; qq = (q + 1) << 16; }
                    }
                }
                const unsigned mk2 = __ballot_sync(0xffffffffu, cls == 2);
                if (pass == 1 && cls == 2) {
                    const int pos = base2 + n2 + __popc(mk2 & ((1u << lane) - 1u));
                    T.g_atom[pos] = x; T.g_qq[pos] = qq; T.g_mask[pos] = 2;
                }
                n2 += __popc(mk2);
            }
        }
    }
    if (lane == 0) { T.scal[S_NG] = n1 + nb + n2; T.scal[S_N1] = n1; T.scal[S_NB] = nb; T.scal[S_N2] = n2; }
}

// tiles of a window with n1 class-1, nb class-3 and n2 class-2 row atoms (in that order)
struct Tiling {
    int nrows, T, q, p, nX, nY;          // x-halves: tiles q-1 .. 0; y-halves: tiles q-1 .. p, then q .. T-1
    __device__ __forceinline__ Tiling(int n1, int nb, int n2) {
        nrows = 9 * (n1 + nb + n2);
        T = (nrows + 7) >> 3;
        q = (9 * (n1 + nb) + 7) >> 3;
        p = (nb + n2) > 0 ? (9 * n1) >> 3 : T;
        nX = q; nY = T - p;
    }
    __device__ __forceinline__ int xtile(int k) const { return q - 1 - k; }
    __device__ __forceinline__ int ytile(int k) const { const int nboth = q - p; return k < nboth ? q - 1 - k : q + (k - nboth); }
};

// lane 0 of warp 0: next window [w0, w1) of G starting at w0: as many row atoms as fit the slot capacity (the two column
// atoms need a slot each, inside or appended) and 8 * MUMAX tiles per column atom
__device__ __forceinline__ int next_window(const UnitTab& T, int w0, int slotcap, bool pair) {
    const int ng = T.scal[S_NG], N1 = T.scal[S_N1], NB = T.scal[S_NB];
    int w1 = w0, best = w0;
    int c1 = 0, cb = 0, c2 = 0, selfs = 0;
    while (w1 < ng) {
        const int cls = w1 < N1 ? 1 : (w1 < N1 + NB ? 3 : 2);
        c1 += cls == 1; cb += cls == 3; c2 += cls == 2;
        selfs += (T.g_qq[w1] & 0xffff) == 0 && (cls & 1) ? 1 : 0;               // i1 itself
        selfs += pair && (T.g_qq[w1] >> 16) == 0 && (cls & 2) ? 1 : 0;          // i2 itself
        ++w1;
        const int extra = (pair ? 2 : 1) - selfs;
        const Tiling tl(c1, cb, c2);
        if ((w1 - w0) + extra > slotcap || tl.nX > NCW * MUMAX || tl.nY > NCW * MUMAX) break;
        best = w1;
    }
    return best > w0 ? best : w0 + 1;     // never empty (slotcap >= 3 is guaranteed by the host)
}

}  // namespace pairk

namespace pairk {

// 32-bit shared-space addressing in the consumer loop (no 64-bit generic pointers live across it)
__device__ __forceinline__ double lds64(uint32_t addr) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void mbar_wait32(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP32:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE32;\n"
        "bra WAIT_LOOP32;\n"
        "WAIT_DONE32:\n"
        "}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ uint32_t mbar_test32(uint32_t bar, uint32_t parity) {
    uint32_t done;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    return done;
}
__device__ __forceinline__ void mbar_arrive32(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// compile-time loop over the (x, y) slots of a warp, so that slot numbers are immediates
template <int S, int N, class F>
__device__ __forceinline__ void static_for(F&& f) {
    if constexpr (S < N) { f(std::integral_constant<int, S>{}); static_for<S + 1, N>(f); }
}

struct WarpDeal {                        // what one consumer warp holds in this sub-task
    int xw, yw, fy, ry;                  // number of x / y items; full y cycles; remainder item index or -1
    __device__ __forceinline__ WarpDeal(const Tiling& tl, int w) {
        xw = tl.nX > w ? (tl.nX - w + NCW - 1) / NCW : 0;
        fy = tl.nY / NCW;
        const int rem = tl.nY - fy * NCW, ex = tl.nX % NCW;
        const int r = (w - ex + NCW) % NCW;
        ry = r < rem ? fy * NCW + r : -1;
        const int full_y = fy;           // items w + 8 s, s < fy
        yw = full_y + (ry >= 0 ? 1 : 0);
    }
    __device__ __forceinline__ int xitem(int s, int w) const { return s < xw ? w + NCW * s : -1; }
    __device__ __forceinline__ int yitem(int s, int w) const { return s < fy ? w + NCW * s : (s == fy ? ry : -1); }
};

// Consumer warp, one sub-task.  MU (x, y) slots; per k-step and slot: up to 4 DMMA + 4 DFMA from one or two A fragments.
// A fragments are single-buffered and re-loaded for the next k-step right behind the instructions that consumed them
// (a whole k-step of FP64 work ahead of their next use); the 32-lane B fragments are double-buffered, the ninth-orbital
// B values are loaded at the top of their k-step (their first use comes behind the first DMMAs).  The chunk loop is
// software-pipelined across the stage boundary exactly as in the one-atom kernel.  Everything the loop touches is addressed
// through 32-bit shared-space addresses; what only the epilogue needs is re-derived there (register budget 232).
struct ConsArgs {
    uint32_t sbuf32, bar32;              // shared-space addresses of the ring and of full[0]
    uint32_t stage_bytes, boff_bytes, stages;
    int nchunks;
};
template <int MU, bool PAIR>
__device__ __forceinline__ void consume_pair(const ConsArgs ca, Ring& rg, const UnitTab& T, int w0, int warp, int lane,
                                             size_t blk1, size_t blk2, double* __restrict__ rho, double* __restrict__ rhoE) {
    const int g = lane >> 2, t = lane & 3;
    double x0[MU][2], x1[MU][2], y0[MU][2], y1[MU][2], xf[MU], xe[MU], yf[MU], ye[MU];
    uint32_t off[MU];                    // byte offsets / 8 of this lane's row in the x tile (low half) and the y tile (high half)
    unsigned same = 0;
    {
        const Tiling tl(T.scal[S_N1W], T.scal[S_NBW], T.scal[S_N2W]);
        const WarpDeal wd(tl, warp);
#pragma unroll
        for (int s = 0; s < MU; ++s) {
            x0[s][0] = x0[s][1] = x1[s][0] = x1[s][1] = y0[s][0] = y0[s][1] = y1[s][0] = y1[s][1] = 0.0;
            xf[s] = xe[s] = yf[s] = ye[s] = 0.0;
            const int kx = wd.xitem(s, warp), ky = PAIR ? wd.yitem(s, warp) : -1;
            const int tx = kx >= 0 ? tl.xtile(kx) : -1, ty = ky >= 0 ? tl.ytile(ky) : -1;
            if (tx >= 0 && tx == ty) same |= 1u << s;
            int R = 8 * tx + g;
            const uint32_t ox = (tx >= 0 && R < tl.nrows) ? (R / 9) * SLOT + 4 * (R % 9) + t : t;      // absent halves / rows beyond the window read row 0 (results dropped)
            R = 8 * ty + g;
            const uint32_t oy = (ty >= 0 && R < tl.nrows) ? (R / 9) * SLOT + 4 * (R % 9) + t : t;
            off[s] = ox | (oy << 16);
            asm volatile("" : "+r"(off[s]));
        }
    }
    const uint32_t lane8 = ca.boff_bytes + 8u * lane, t8 = ca.boff_bytes + 8u * t;
    const uint32_t full0 = ca.bar32, bready0 = ca.bar32 + 16u * ca.stages, empty0 = ca.bar32 + 24u * ca.stages;
    if (ca.nchunks > 0) {
        uint32_t stg = ca.sbuf32 + rg.st * ca.stage_bytes;
#ifdef XTB_EXP_STAGGER
        if (warp >= 4) __nanosleep(XTB_EXP_STAGGER);      // experiment: de-phase the two consumer warps of a sub-partition
#endif
        mbar_wait32(full0 + 8u * rg.st, rg.ph);
        mbar_wait32(bready0 + 8u * rg.st, rg.ph);
        double ax[MU], ay[MU];
        double bA[4], bB[4], b8[4];      // 32-lane B fragments of the even / odd k-steps (i1 rho, i1 rho_E, i2 rho, i2 rho_E); ninth-orbital values
#define XTB_LOADB(dst, base, ks) { const uint32_t pb_ = (base) + lane8 + (ks) * (BROW2 * 8); \
        dst[0] = lds64(pb_); dst[1] = lds64(pb_ + 256); dst[2] = lds64(pb_ + BROW * 8); dst[3] = lds64(pb_ + BROW * 8 + 256); }
#define XTB_LOADB8(base, ks) { const uint32_t pt_ = (base) + t8 + (ks) * (BROW2 * 8); \
        b8[0] = lds64(pt_ + 512); b8[1] = lds64(pt_ + 544); b8[2] = lds64(pt_ + BROW * 8 + 512); b8[3] = lds64(pt_ + BROW * 8 + 544); }
        XTB_LOADB(bA, stg, 0)
#pragma unroll
        for (int s = 0; s < MU; ++s) {
            ax[s] = lds64(stg + ((off[s] & 0xffffu) << 3));
            if (PAIR) ay[s] = ((same >> s) & 1u) ? ax[s] : lds64(stg + ((off[s] >> 16) << 3));
        }
        static_assert(KS % 2 == 0, "the B fragment double buffer needs an even number of k-steps per chunk");
        for (int ch = 0; ch < ca.nchunks; ++ch) {
            Ring rn = rg;
            rn.next(ca.stages);
            const uint32_t stg_n = ca.sbuf32 + rn.st * ca.stage_bytes;
            uint32_t ready = 0;
#pragma unroll
            for (int ks = 0; ks < KS; ++ks) {
                double (&cu)[4] = (ks & 1) ? bB : bA;
                double (&nx)[4] = (ks & 1) ? bA : bB;
                uint32_t nbase;               // A operands of the next k-step
                bool have_next = true;
                XTB_LOADB8(stg, ks)
                if (ks + 1 < KS) {
                    nbase = stg + 288 * (ks + 1);
                    XTB_LOADB(nx, stg, ks + 1)
                    if (ks + 2 == KS && ch + 1 < ca.nchunks) ready = mbar_test32(full0 + 8u * rn.st, rn.ph) & mbar_test32(bready0 + 8u * rn.st, rn.ph);
                } else {
                    nbase = stg_n;
                    have_next = ch + 1 < ca.nchunks;
                    if (have_next) {
                        if (!ready) { mbar_wait32(full0 + 8u * rn.st, rn.ph); mbar_wait32(bready0 + 8u * rn.st, rn.ph); }
                        XTB_LOADB(nx, stg_n, 0)
                    }
                }
                // No predication on the DMMAs: a predicated mma.sync costs a WARPSYNC.ALL + NOP in front of it whether it
                // executes or not (measured: ~50 cycles per DMMA, 157 ms instead of 69 ms for single-atom units), so a slot
                // always does its 4 (pair) / 2 (single) DMMAs; halves a warp does not hold read row 0 and are dropped.
#pragma unroll
                for (int s = 0; s < MU; ++s) {
                    dmma884(x0[s][0], x0[s][1], ax[s], cu[0]); dmma884(x1[s][0], x1[s][1], ax[s], cu[1]);
                    xf[s] = fma(ax[s], b8[0], xf[s]); xe[s] = fma(ax[s], b8[1], xe[s]);
                    if (PAIR) {
                        dmma884(y0[s][0], y0[s][1], ay[s], cu[2]); dmma884(y1[s][0], y1[s][1], ay[s], cu[3]);
                        yf[s] = fma(ay[s], b8[2], yf[s]); ye[s] = fma(ay[s], b8[3], ye[s]);
                    }
                    if (have_next) {
                        ax[s] = lds64(nbase + ((off[s] & 0xffffu) << 3));
#ifdef XTB_EXP_LOADY
                        if (PAIR) ay[s] = lds64(nbase + ((off[s] >> 16) << 3));
#else
                        if (PAIR) ay[s] = ((same >> s) & 1u) ? ax[s] : lds64(nbase + ((off[s] >> 16) << 3));
#endif
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive32(empty0 + 8u * rg.st);
            rg = rn;
            stg = stg_n;
        }
#undef XTB_LOADB
#undef XTB_LOADB8
    }
    // ---- epilogue: row R of tile -> (row atom, nu); block of column atom a at rowptr[i_a] + 1 + ordinal (0 = on-site) ----
    const Tiling tl(T.scal[S_N1W], T.scal[S_NBW], T.scal[S_N2W]);
    const WarpDeal wd(tl, warp);
    const int* __restrict__ g_qq = T.g_qq; const int* __restrict__ g_mask = T.g_mask;
#pragma unroll
    for (int s = 0; s < MU; ++s) {
        double f8 = xf[s], e8 = xe[s], f9 = yf[s], e9 = ye[s];
        f8 += __shfl_xor_sync(0xffffffffu, f8, 1); e8 += __shfl_xor_sync(0xffffffffu, e8, 1);
        f8 += __shfl_xor_sync(0xffffffffu, f8, 2); e8 += __shfl_xor_sync(0xffffffffu, e8, 2);
        f9 += __shfl_xor_sync(0xffffffffu, f9, 1); e9 += __shfl_xor_sync(0xffffffffu, e9, 1);
        f9 += __shfl_xor_sync(0xffffffffu, f9, 2); e9 += __shfl_xor_sync(0xffffffffu, e9, 2);
        const int kx = wd.xitem(s, warp), ky = PAIR ? wd.yitem(s, warp) : -1;
        if (kx >= 0) {
            const int R = 8 * tl.xtile(kx) + g;
            if (R < tl.nrows) {
                const int sl = R / 9, nu = R - 9 * sl;
                if (g_mask[w0 + sl] & 1) {
                    const int q = (g_qq[w0 + sl] & 0xffff) - 1;
                    const size_t blk = (blk1 + (q < 0 ? 0 : 1 + q)) * 81 + 9 * nu;
                    rho[blk + 2 * t] = x0[s][0];  rho[blk + 2 * t + 1] = x0[s][1];
                    rhoE[blk + 2 * t] = x1[s][0]; rhoE[blk + 2 * t + 1] = x1[s][1];
                    if (t == 0) rho[blk + 8] = f8;
                    if (t == 1) rhoE[blk + 8] = e8;
                }
            }
        }
        if (ky >= 0) {
            const int R = 8 * tl.ytile(ky) + g;
            if (R < tl.nrows) {
                const int sl = R / 9, nu = R - 9 * sl;
                if (g_mask[w0 + sl] & 2) {
                    const int q = (g_qq[w0 + sl] >> 16) - 1;
                    const size_t blk = (blk2 + (q < 0 ? 0 : 1 + q)) * 81 + 9 * nu;
                    rho[blk + 2 * t] = y0[s][0];  rho[blk + 2 * t + 1] = y0[s][1];
                    rhoE[blk + 2 * t] = y1[s][0]; rhoE[blk + 2 * t + 1] = y1[s][1];
                    if (t == 0) rho[blk + 8] = f9;
                    if (t == 1) rhoE[blk + 8] = e9;
                }
            }
        }
    }
}

// consumer warp without a half in this sub-task: it still takes part in the barriers
__device__ __forceinline__ void consume_none(uint32_t stages, uint64_t* full, uint64_t* bready, uint64_t* empty, Ring& rg,
                                             int nchunks, int lane) {
    for (int ch = 0; ch < nchunks; ++ch) {
        mbar_wait(&full[rg.st], rg.ph);
        mbar_wait(&bready[rg.st], rg.ph);
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[rg.st]);
        rg.next(stages);
    }
}

template <int REGS> __device__ __forceinline__ void reg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;\n" ::"n"(REGS)); }
template <int REGS> __device__ __forceinline__ void reg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" ::"n"(REGS)); }

constexpr int THREADS = (NCW + 1 + NPROD) * 32;      // 384: warpgroups 0, 1 consume, warpgroup 2 = B warp + producers
constexpr int REG_CONS = 232, REG_HELP = 40;         // 256 * 232 + 128 * 40 = 64512 <= 65536

// all 384 threads; used from both role paths (after setmaxnreg the two paths never share code, so that ptxas can give
// each its own register budget), hence a named barrier instead of __syncthreads()
__device__ __forceinline__ void cta_sync() { asm volatile("bar.sync 1, %0;" ::"n"(THREADS) : "memory"); }

struct Smem {
    double* sbuf; uint64_t *full, *bfull, *bready, *empty; int* slot_atom; UnitTab T; int stage_doubles;
};
struct KArgs {
    int na, cap, slotcap, stages, nchunks, maxsub, lead, dbg;
    const int* nunits;                   // number of units (device memory: the matching runs on the device)
    long long* prog; const int* units; const int* nn_size; const int* nn_list; const signed char* own; const int* rowptr;
    const double* Ct; const double* wf; const double* wfe; double* rho; double* rhoE;
};

// ---- warpgroups 0, 1: unit tables (warp 0 / thread 0) and the contraction ----
__device__ __forceinline__ void consumer_main(const Smem& S, const KArgs& A, int warp, int lane) {
    const UnitTab& T = S.T;
    const uint32_t stages = (uint32_t)A.stages;
    Ring rg; rg.st = 0; rg.ph = 0;
    ConsArgs ca;
    ca.sbuf32 = smem_u32(S.sbuf); ca.bar32 = smem_u32(S.full);
    ca.stage_bytes = (uint32_t)S.stage_doubles * 8u; ca.boff_bytes = (uint32_t)(A.slotcap * SLOT) * 8u; ca.stages = stages; ca.nchunks = A.nchunks;
    const int nunits = *A.nunits;
    for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
        const int i1 = A.units[2 * u], i2 = A.units[2 * u + 1];
        cta_sync();                                            // everybody is done with the tables of the previous unit
        if (warp == 0) build_unit(T, i1, i2, A.cap, A.nn_size, A.nn_list, A.own, lane);
        cta_sync();
        const int ng = T.scal[S_NG], N1 = T.scal[S_N1], NB = T.scal[S_NB];
        const size_t blk1 = (size_t)A.rowptr[i1], blk2 = i2 >= 0 ? (size_t)A.rowptr[i2] : 0;
        for (int w0 = 0; w0 < ng;) {
            cta_sync();                                        // ... and with the window of the previous sub-task
            if (threadIdx.x == 0) {
                const int w1 = next_window(T, w0, A.slotcap, i2 >= 0);
                int p1 = -1, p2 = -1;
                for (int e = w0; e < w1; ++e) {
                    const int m = T.g_mask[e], qq = T.g_qq[e];
                    if ((m & 1) && (qq & 0xffff) == 0) p1 = e - w0;
                    if ((m & 2) && (qq >> 16) == 0 && i2 >= 0) p2 = e - w0;
                }
                int ns = w1 - w0;
                if (p1 < 0) p1 = ns++;
                if (i2 >= 0 && p2 < 0) p2 = ns++;
                T.scal[S_W1] = w1; T.scal[S_POSB1] = p1; T.scal[S_POSB2] = p2; T.scal[S_NSLOTS] = ns;
                const int n1w = max(0, min(w1, N1) - w0), nbw = max(0, min(w1, N1 + NB) - max(w0, N1));
                T.scal[S_N1W] = n1w; T.scal[S_NBW] = nbw; T.scal[S_N2W] = (w1 - w0) - n1w - nbw;
            }
            cta_sync();
            const int w1 = T.scal[S_W1], nw = w1 - w0, pB1 = T.scal[S_POSB1], pB2 = T.scal[S_POSB2], nslots = T.scal[S_NSLOTS];
            for (int s = threadIdx.x; s < nslots; s += NCW * 32) {
                S.slot_atom[s] = s < nw ? T.g_atom[w0 + s] : (s == pB1 ? i1 : i2);
                if (s < nw && s != pB1 && s != pB2) T.cp_slot[s - (pB1 < s ? 1 : 0) - ((pB2 >= 0 && pB2 < s) ? 1 : 0)] = s;
            }
            cta_sync();
            int mu;
            { const Tiling tl(T.scal[S_N1W], T.scal[S_NBW], T.scal[S_N2W]); const WarpDeal wd(tl, warp); mu = max(wd.xw, wd.yw); }
// Two instantiations only (pair / single-atom unit, 6 slots each; a warp with fewer halves runs dummy ones): with three
// or more, ptxas spills inside the consumer loops (checked with tools/sass_loops.py).
            if (mu == 0) consume_none(stages, S.full, S.bready, S.empty, rg, A.nchunks, lane);
            else if (i2 >= 0) consume_pair<MUMAX, true>(ca, rg, T, w0, warp, lane, blk1, blk2, A.rho, A.rhoE);
            else consume_pair<MUMAX, false>(ca, rg, T, w0, warp, lane, blk1, blk2, A.rho, A.rhoE);
#define XTB_PCASE(M)
#undef XTB_PCASE
            w0 = w1;
        }
    }
}

// ---- warpgroup 2: warp NCW weights the B operands, warps NCW+1.. issue the bulk copies ----
__device__ __forceinline__ void helper_main(const Smem& S, const KArgs& A, int warp, int lane) {
    constexpr int NO = 9;
    constexpr int TILE = NO * KC;
    const UnitTab& T = S.T;
    const uint32_t stages = (uint32_t)A.stages;
    const int slotcap = A.slotcap, nchunks = A.nchunks;
    int lead = A.lead;
    Ring rg; rg.st = 0; rg.ph = 0;
    int round = 0;
    const int nunits = *A.nunits;
    for (int u = blockIdx.x; u < nunits; u += gridDim.x, ++round) {
        const int i1 = A.units[2 * u], i2 = A.units[2 * u + 1];
        cta_sync();
        cta_sync();
        const int ng = T.scal[S_NG];
        int sub = 0;
        for (int w0 = 0; w0 < ng; ++sub) {
            cta_sync();
            cta_sync();
            const int w1 = T.scal[S_W1], nw = w1 - w0, pB1 = T.scal[S_POSB1], pB2 = T.scal[S_POSB2];
            const int ncp = nw - (pB1 < nw ? 1 : 0) - ((pB2 >= 0 && pB2 < nw) ? 1 : 0);
            cta_sync();
            if (warp > NCW) {
                // ===== producer warps: cp_slot[k] belongs to producer k % NPROD, lane k / NPROD =====
                const int pw = warp - NCW - 1;
                const int mine = ncp > pw ? (ncp - pw + NPROD - 1) / NPROD : 0;
                const uint32_t bytes = (uint32_t)(mine * TILE) * 8u;
                const double* csrc = A.Ct;
                const double* w1p = A.wf; const double* w2p = A.wfe;
                const int s0 = lane < mine ? T.cp_slot[lane * NPROD + pw] : -1;
                const int s1 = lane + 32 < mine ? T.cp_slot[(lane + 32) * NPROD + pw] : -1;
                const size_t src0 = s0 >= 0 ? (size_t)S.slot_atom[s0] * TILE : 0, src1 = s1 >= 0 ? (size_t)S.slot_atom[s1] * TILE : 0;
                const size_t srcB1 = (size_t)i1 * TILE, srcB2 = i2 >= 0 ? (size_t)i2 * TILE : 0;
                const long long pos0 = ((long long)round * A.maxsub + sub) * nchunks;
                for (int ch = 0; ch < nchunks; ++ch) {
                    if (lead > 0 && pw == 0 && (ch % SYNC_EVERY) == 0 && !throttle(A.prog, pos0 + ch, lead, lane)) lead = 0;
                    mbar_spin(&S.empty[rg.st], rg.ph ^ 1u);
                    double* dstage = S.sbuf + (size_t)rg.st * S.stage_doubles;
                    if (lane == 0) {
                        if (pw == 0) {
                            mbar_expect_tx(&S.bfull[rg.st], (uint32_t)((i2 >= 0 ? 2 : 1) * TILE + 2 * KC) * 8u);
                            bulk_g2s(dstage + pB1 * SLOT, csrc + srcB1, TILE * 8, &S.bfull[rg.st]);
                            if (i2 >= 0) bulk_g2s(dstage + pB2 * SLOT, csrc + srcB2, TILE * 8, &S.bfull[rg.st]);
                            bulk_g2s(dstage + slotcap * SLOT + KS * BROW2, w1p, KC * 8, &S.bfull[rg.st]);
                            bulk_g2s(dstage + slotcap * SLOT + KS * BROW2 + KC, w2p, KC * 8, &S.bfull[rg.st]);
                        }
                        mbar_expect_tx(&S.full[rg.st], bytes);
                    }
                    __syncwarp();
                    if (s0 >= 0) bulk_g2s(dstage + s0 * SLOT, csrc + src0, TILE * 8, &S.full[rg.st]);
                    if (s1 >= 0) bulk_g2s(dstage + s1 * SLOT, csrc + src1, TILE * 8, &S.full[rg.st]);
                    for (int k = lane + 64; k < mine; k += 32) {
                        const int s = T.cp_slot[k * NPROD + pw];
                        bulk_g2s(dstage + s * SLOT, csrc + (size_t)S.slot_atom[s] * TILE, TILE * 8, &S.full[rg.st]);
                    }
                    csrc += (size_t)A.na * TILE; w1p += KC; w2p += KC;
                    rg.next(stages);
                }
            } else {
                // ===== B warp: weighted B operands of i1 and i2, once per chunk =====
                const int g = lane >> 2, t = lane & 3;
                for (int ch = 0; ch < nchunks; ++ch) {
                    mbar_spin(&S.bfull[rg.st], rg.ph);
                    double* stg = S.sbuf + (size_t)rg.st * S.stage_doubles;
                    double* Bs = stg + slotcap * SLOT;
                    const double* wsm = Bs + KS * BROW2;
#pragma unroll
                    for (int a = 0; a < 2; ++a) {
                        const int pB = a == 0 ? pB1 : pB2;
                        if (pB < 0 || (A.dbg & 16)) continue;       // dbg 16 (timing experiment, wrong results): no B weighting
                        const double* ti = stg + pB * SLOT;
#pragma unroll
                        for (int s = 0; s < KS; ++s) {
                            const double w_f = wsm[4 * s + t], w_e = wsm[KC + 4 * s + t];
                            const double ci = ti[NO * 4 * s + 4 * g + t];             // C[i_a, mu = g][state 4 s + t]
                            double* o = Bs + s * BROW2 + a * BROW;
                            o[lane] = ci * w_f;                                       // Atemp = distre * Ha        TB.f90:4378
                            o[32 + lane] = ci * w_e;                                  // Atemp = distre * En * Ha   TB.f90:4459
                            if (g == 0) {
                                const double c8 = ti[NO * 4 * s + 32 + t];            // C[i_a, mu = 8][state]
                                o[64 + t] = c8 * w_f;
                                o[68 + t] = c8 * w_e;
                            }
                        }
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&S.bready[rg.st]);
                    rg.next(stages);
                }
            }
            w0 = w1;
        }
    }
    if (warp == NCW + 1 && lane == 0) st_relaxed(A.prog + blockIdx.x, XTB_PROG_DONE);
}

// Shared memory per stage (doubles): [slotcap][SLOT] tiles | [KS][2][BROW] weighted B operands of i1, i2 | [2][KC] weights.
__global__ void __launch_bounds__(THREADS, 1) density_pair_kernel(const KArgs A, int gcap) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem S;
    S.stage_doubles = A.slotcap * SLOT + KS * BROW2 + 2 * KC;
    S.sbuf = reinterpret_cast<double*>(smem_raw);
    S.full = reinterpret_cast<uint64_t*>(S.sbuf + (size_t)A.stages * S.stage_doubles);
    S.bfull = S.full + A.stages;
    S.bready = S.bfull + A.stages;
    S.empty = S.bready + A.stages;
    S.slot_atom = reinterpret_cast<int*>(S.empty + A.stages);      // [slotcap]
    S.T.cp_slot = S.slot_atom + A.slotcap;
    S.T.g_atom = S.T.cp_slot + A.slotcap;
    S.T.g_qq = S.T.g_atom + gcap;
    S.T.g_mask = S.T.g_qq + gcap;
    S.T.scal = S.T.g_mask + gcap;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < A.stages; ++s) { mbar_init(&S.full[s], NPROD); mbar_init(&S.bfull[s], 1); mbar_init(&S.bready[s], 1); mbar_init(&S.empty[s], NCW); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (warp < NCW) { reg_inc<REG_CONS>(); consumer_main(S, A, warp, lane); }
    else            { reg_dec<REG_HELP>(); helper_main(S, A, warp, lane); }
}

// shared memory of the kernel for a slot capacity / number of stages
inline size_t smem_bytes(int slotcap, int stages, int gcap) {
    return (size_t)stages * (slotcap * SLOT + KS * BROW2 + 2 * KC) * 8 + (size_t)4 * 8 * stages
           + sizeof(int) * (2 * (size_t)slotcap + 3 * (size_t)gcap + 16) + 64;
}

}  // namespace pairk
