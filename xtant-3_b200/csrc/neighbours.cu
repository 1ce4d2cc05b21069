// S1 -- neighbour list on the device.
//
// Replaces get_near_neighbours -> Find_nearest_neighbours (Source_files/Atomic_tools.f90:562,4230-4294)
// with the minimum-image search of shortest_distance_OLD (Atomic_tools.f90:4028-4131).
// Semantics kept bit-exactly:
//   * x = sum_k (S_i(k) - S_j(k) + zb(k)) * supce(k,c), summed k = 1..3 starting from 0 (:4098-4104),
//     no FMA contraction (explicit __dmul_rn / __dadd_rn);
//   * 27 images scanned zb1, zb2, zb3 ascending, replace if r1 <= a_r (:4107); start value = image 0 (:4064-4072);
//   * neighbour iff a_r < rm (:4255); appended in ascending j (ordered warp compaction).
// B200 mapping: one warp per atom i, lanes stride over j; the reference's serial O(27 Na^2)
// double loop becomes Na warps; sqrt is only taken when the squared distances cannot decide
// the `<=` (exact: sqrt is monotone, and a relative gap of 1e-14 in r^2 is far above 1 ulp).
#include "xtb_ctx.h"
#include <cmath>
#include <cstdlib>

namespace {

__device__ __forceinline__ double sum3(double a, double b, double c) {
    // ((0 + a) + b) + c  -- the reference accumulates into a zero-initialised variable
    return __dadd_rn(__dadd_rn(__dadd_rn(0.0, a), b), c);
}
__device__ __forceinline__ double norm2(double x, double y, double z) {
    return __dadd_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)), __dmul_rn(z, z));
}

struct MinImage {
    double x, y, z, r, sx, sy, sz;
};

// h[k + 3*c] = supce(k+1, c+1)
__device__ __forceinline__ MinImage min_image(const double Si[3], const double Sj[3], const double* __restrict__ h) {
    double ds[3] = {__dsub_rn(Si[0], Sj[0]), __dsub_rn(Si[1], Sj[1]), __dsub_rn(Si[2], Sj[2])};
    // products p[k][z][c] = (ds_k + zb) * supce(k,c), zb = -1,0,1 -> z = 0,1,2
    double dsz[3][3];
    double p[3][3][3];
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int z = 0; z < 3; ++z) {
            dsz[k][z] = __dadd_rn(ds[k], (double)(z - 1));
#pragma unroll
            for (int c = 0; c < 3; ++c) p[k][z][c] = __dmul_rn(dsz[k][z], h[k + 3 * c]);
        }
    // start: un-shifted image, from ds itself (:4064-4072)
    double bx = sum3(__dmul_rn(ds[0], h[0]), __dmul_rn(ds[1], h[1]), __dmul_rn(ds[2], h[2]));
    double by = sum3(__dmul_rn(ds[0], h[3]), __dmul_rn(ds[1], h[4]), __dmul_rn(ds[2], h[5]));
    double bz = sum3(__dmul_rn(ds[0], h[6]), __dmul_rn(ds[1], h[7]), __dmul_rn(ds[2], h[8]));
    double bq = norm2(bx, by, bz);
    double br = -1.0;                 // sqrt(bq) computed lazily
    int b1 = 1, b2 = 1, b3 = 1;       // image indices of the current best (z index); start = "no zb added"
    bool shifted = false;             // whether sx,sy,sz carry a zb (even zb = 0 adds "+0.0")
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                double x0 = sum3(p[0][i][0], p[1][j][0], p[2][k][0]);
                double y0 = sum3(p[0][i][1], p[1][j][1], p[2][k][1]);
                double z0 = sum3(p[0][i][2], p[1][j][2], p[2][k][2]);
                double q = norm2(x0, y0, z0);
                bool take;
                if (q <= bq) take = true;                        // sqrt monotone => r1 <= a_r
                else if (q > bq * (1.0 + 1e-14)) take = false;   // r1 > a_r for sure
                else {                                           // undecidable on squares: compare the rounded roots
                    if (br < 0.0) br = __dsqrt_rn(bq);
                    take = (__dsqrt_rn(q) <= br);
                }
                if (take) { bx = x0; by = y0; bz = z0; bq = q; br = -1.0; b1 = i; b2 = j; b3 = k; shifted = true; }
            }
    MinImage o;
    o.x = bx; o.y = by; o.z = bz; o.r = __dsqrt_rn(bq);
    o.sx = shifted ? dsz[0][b1] : ds[0];
    o.sy = shifted ? dsz[1][b2] : ds[1];
    o.sz = shifted ? dsz[2][b3] : ds[2];
    return o;
}

// Orthorhombic cell with every edge longer than 2 rm: the squared distance is separable per axis, so the image the
// 27-scan ends with is the per-axis minimum of |ds_k + zb| (ties to the later zb, as the scan's `<=`), and it is the
// only image that can pass `a_r < rm`.  Two images of one axis are never confused by rounding for a pair inside the
// cut-off: they would both have |component| ~ L/2 > rm.  The chosen image is evaluated with the same operations as in
// the scan, so every stored number has the same bits; pairs outside the cut-off are not stored at all.
__device__ __forceinline__ MinImage min_image_ortho(const double Si[3], const double Sj[3], const double* __restrict__ h) {
    double d[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double ds = __dsub_rn(Si[k], Sj[k]);
        double best = __dadd_rn(ds, -1.0);
#pragma unroll
        for (int z = 1; z < 3; ++z) {
            const double v = __dadd_rn(ds, (double)(z - 1));
            if (fabs(v) <= fabs(best)) best = v;
        }
        d[k] = best;
    }
    MinImage o;
    o.x = sum3(__dmul_rn(d[0], h[0]), __dmul_rn(d[1], h[1]), __dmul_rn(d[2], h[2]));
    o.y = sum3(__dmul_rn(d[0], h[3]), __dmul_rn(d[1], h[4]), __dmul_rn(d[2], h[5]));
    o.z = sum3(__dmul_rn(d[0], h[6]), __dmul_rn(d[1], h[7]), __dmul_rn(d[2], h[8]));
    o.r = __dsqrt_rn(norm2(o.x, o.y, o.z));
    o.sx = d[0]; o.sy = d[1]; o.sz = d[2];
    return o;
}

template <bool ORTHO>
__global__ void __launch_bounds__(256) nn_build_kernel(int na, const double* __restrict__ S, const double* __restrict__ hmat,
                                                      double rm, int cap, int* __restrict__ nn_size,
                                                      int* __restrict__ nn_list, double* __restrict__ nn_dist,
                                                      double* __restrict__ nn_dist_s, int* __restrict__ max_nn) {
    __shared__ double h[9];
    if (threadIdx.x < 9) h[threadIdx.x] = hmat[threadIdx.x];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= na) return;
    const double Si[3] = {S[3 * i], S[3 * i + 1], S[3 * i + 2]};
    int count = 0;
    for (int base = 0; base < na; base += 32) {
        const int j = base + lane;
        bool isn = false;
        MinImage mi;
        if (j < na && j != i) {
            const double Sj[3] = {S[3 * j], S[3 * j + 1], S[3 * j + 2]};
            mi = ORTHO ? min_image_ortho(Si, Sj, h) : min_image(Si, Sj, h);
            isn = mi.r < rm;
        }
        const unsigned mask = __ballot_sync(0xffffffffu, isn);
        if (isn) {
            const int pos = count + __popc(mask & ((1u << lane) - 1u));
            if (pos < cap) {
                const size_t e = (size_t)i * cap + pos;
                nn_list[e] = j;
                double4 d4 = make_double4(mi.x, mi.y, mi.z, mi.r);
                *reinterpret_cast<double4*>(nn_dist + 4 * e) = d4;
                nn_dist_s[3 * e] = mi.sx; nn_dist_s[3 * e + 1] = mi.sy; nn_dist_s[3 * e + 2] = mi.sz;
            }
        }
        count += __popc(mask);
    }
    if (lane == 0) {
        nn_size[i] = count < cap ? count : cap;
        atomicMax(max_nn, count);
    }
}

// rowptr[i] = first block of atom i (blocks: on-site, then list order); rowptr[na] = n_blocks.
// rev[e] = position of atom i in the list of its neighbour nn_list[e] (lists ascend in j) or -1.
// parts > 1 (xtb_set_block_layout): the blocks of the `parts` contiguous atom ranges of the row partition (ceil(na/parts)
// atoms each) start at multiples of the largest range's block count, so that the ranges are equal-length segments of every
// block store and one in-place all-gather exchanges them.  The gaps are never addressed (every access is rowptr[atom] + q).
// meta[1] = number of real blocks (na + number of list entries).
// Guard of the steady-state form (xtb_impl_neighbours_steady): if the largest neighbour count exceeds nn_bound or the block
// count exceeds blk_capacity -- the sizes the later stages of this step were launched with -- meta[2] is raised and the
// list is emptied, so that those stages touch only the na on-site blocks.
__global__ void nn_rowptr_kernel(int na, int* __restrict__ nn_size, int* __restrict__ rowptr, int parts, int* __restrict__ meta,
                                 int nn_bound, long long blk_capacity) {
    __shared__ int carry;
    __shared__ int buf[1024];
    __shared__ int bad;
    if (threadIdx.x == 0) {
        carry = 0;
        bad = meta[0] > nn_bound ? 1 : 0;
    }
    __syncthreads();
    if (!bad) {         // block count = na + sum of the list lengths
        long long part = 0;
        for (int i = threadIdx.x; i < na; i += 1024) part += nn_size[i] + 1;
        for (int off = 16; off > 0; off >>= 1) part += __shfl_xor_sync(0xffffffffu, part, off);
        __shared__ long long wsum[32];
        if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = part;
        __syncthreads();
        if (threadIdx.x == 0) {
            long long tot = 0;
            for (int w = 0; w < 32; ++w) tot += wsum[w];
            if (tot > blk_capacity) bad = 1;
        }
        __syncthreads();
    }
    if (bad) {
        for (int i = threadIdx.x; i < na; i += 1024) nn_size[i] = 0;
        if (threadIdx.x == 0) meta[2] = 1;
        __syncthreads();
    }
    for (int base = 0; base < na; base += 1024) {
        int i = base + threadIdx.x;
        int v = i < na ? nn_size[i] + 1 : 0;
        buf[threadIdx.x] = v;
        __syncthreads();
        for (int off = 1; off < 1024; off <<= 1) {
            int t = threadIdx.x >= off ? buf[threadIdx.x - off] : 0;
            __syncthreads();
            buf[threadIdx.x] += t;
            __syncthreads();
        }
        if (i < na) rowptr[i] = carry + buf[threadIdx.x] - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry += buf[1023];
        __syncthreads();
    }
    if (threadIdx.x == 0) { rowptr[na] = carry; meta[1] = carry; meta[3] = carry; }
    if (parts <= 1) return;
    __syncthreads();
    // segment bases and the largest segment (parts <= 1024)
    const int per = (na + parts - 1) / parts;
    __shared__ int mx;
    if (threadIdx.x == 0) mx = 0;
    __syncthreads();
    if ((int)threadIdx.x < parts) {
        const int b = min(na, (int)threadIdx.x * per), e = min(na, ((int)threadIdx.x + 1) * per);
        buf[threadIdx.x] = rowptr[b];
        atomicMax(&mx, rowptr[e] - rowptr[b]);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < na; i += 1024) {
        const int r = i / per;
        rowptr[i] = r * mx + (rowptr[i] - buf[r]);
    }
    if (threadIdx.x == 0) { rowptr[na] = parts * mx; meta[3] = parts * mx; }
}

// own[e] = 1 iff atom i "owns" the unordered pair (i, j = nn_list[e]) in the symmetric density pass: the pair is
// oriented by the sign of the minimum-image displacement stored in the SMALLER-index atom's entry (positive
// half-space of a generic direction), so exactly one side owns it and, on a lattice with inversion symmetry,
// every atom owns half of its neighbours (equal work per atom).  Pairs missing from the partner's list are owned.
__device__ __forceinline__ bool positive_halfspace(const double* d) {
    // orientation by the sign of d.v, v a generic direction: no low-index lattice plane is perpendicular to it, so
    // thermal displacements flip (and correlate) only the few neighbours with |d.v| below the noise; on a lattice
    // with inversion symmetry every atom then owns half of its neighbours.  Exact ties fall back to lexicographic order.
    const double p = d[0] + 0.754877666246693 * d[1] + 0.569840290998053 * d[2];
    if (p != 0.0) return p > 0.0;
    return d[0] > 0.0 || (d[0] == 0.0 && (d[1] > 0.0 || (d[1] == 0.0 && d[2] > 0.0)));
}

__global__ void nn_rev_kernel(int na, int cap, const int* __restrict__ nn_size, const int* __restrict__ nn_list,
                              const double* __restrict__ nn_dist, int* __restrict__ rev, signed char* __restrict__ own) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (long long)na * cap) return;
    const int i = (int)(t / cap), k = (int)(t % cap);
    if (k >= nn_size[i]) { rev[t] = -1; own[t] = 0; return; }
    const int j = nn_list[t];
    const int* L = nn_list + (size_t)j * cap;
    int lo = 0, hi = nn_size[j] - 1, pos = -1;
    while (lo <= hi) {
        int mid = (lo + hi) >> 1;
        int v = L[mid];
        if (v == i) { pos = mid; break; }
        if (v < i) lo = mid + 1; else hi = mid - 1;
    }
    rev[t] = pos;
    bool o = true;
    if (pos >= 0) o = (i < j) ? positive_halfspace(nn_dist + 4 * t) : !positive_halfspace(nn_dist + 4 * ((size_t)j * cap + pos));
    own[t] = o ? 1 : 0;
}

}  // namespace

// One launch sequence of the list: build -> rowptr (+ overflow guard) -> rev / own.  meta = {max count, real blocks,
// overflow flag, rowptr[na]}.
static int launch_list(xtb_ctx* c, double rm, int nn_bound, long long blk_capacity) {
    const int na = c->na, cap = c->cap;
    XTB_CUDA(c, cudaMemsetAsync(c->d_maxnn.p, 0, sizeof(int) * 4, c->stream));
    const int warps_per_block = 8;
    // single-image path: diagonal cell, every edge > 2 rm (see min_image_ortho); XTB_NN_SCAN27=1 forces the full scan
    const double* hh = c->h_supce;
    bool ortho = hh[1] == 0.0 && hh[2] == 0.0 && hh[3] == 0.0 && hh[5] == 0.0 && hh[6] == 0.0 && hh[7] == 0.0;
    for (int k = 0; k < 3 && ortho; ++k) ortho = fabs(hh[4 * k]) > 2.0 * rm * (1.0 + 1e-9);
    if (const char* ev = getenv("XTB_NN_SCAN27")) { if (atoi(ev) == 1) ortho = false; }
    const dim3 g((na + warps_per_block - 1) / warps_per_block), b(warps_per_block * 32);
    if (ortho)
        nn_build_kernel<true><<<g, b, 0, c->stream>>>(na, (const double*)c->d_S.p, (const double*)c->d_hdev.p, rm, cap, (int*)c->d_nn_size.p,
                                                      (int*)c->d_nn_list.p, (double*)c->d_nn_dist.p, (double*)c->d_nn_dist_s.p, (int*)c->d_maxnn.p);
    else
        nn_build_kernel<false><<<g, b, 0, c->stream>>>(na, (const double*)c->d_S.p, (const double*)c->d_hdev.p, rm, cap, (int*)c->d_nn_size.p,
                                                       (int*)c->d_nn_list.p, (double*)c->d_nn_dist.p, (double*)c->d_nn_dist_s.p, (int*)c->d_maxnn.p);
    XTB_LAUNCH_CHECK(c);
    nn_rowptr_kernel<<<1, 1024, 0, c->stream>>>(na, (int*)c->d_nn_size.p, (int*)c->d_rowptr.p, c->layout_parts, (int*)c->d_maxnn.p,
                                                nn_bound, blk_capacity);
    XTB_LAUNCH_CHECK(c);
    const long long tot = (long long)na * cap;
    nn_rev_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, c->stream>>>(na, cap, (const int*)c->d_nn_size.p,
                                                                       (const int*)c->d_nn_list.p, (const double*)c->d_nn_dist.p,
                                                                       (int*)c->d_rev.p, (signed char*)c->d_own.p);
    XTB_LAUNCH_CHECK(c);
    return XTB_OK;
}

static int reserve_list(xtb_ctx* c) {
    const int na = c->na;
    const size_t ne = (size_t)na * c->cap;
    int rc;
    if ((rc = xtb_reserve(c, c->d_nn_size, sizeof(int) * na))) return rc;
    if ((rc = xtb_reserve(c, c->d_nn_list, sizeof(int) * ne))) return rc;
    if ((rc = xtb_reserve(c, c->d_rev, sizeof(int) * ne))) return rc;
    if ((rc = xtb_reserve(c, c->d_own, ne))) return rc;
    if ((rc = xtb_reserve(c, c->d_nn_dist, sizeof(double) * 4 * ne))) return rc;
    if ((rc = xtb_reserve(c, c->d_nn_dist_s, sizeof(double) * 3 * ne))) return rc;
    if ((rc = xtb_reserve(c, c->d_rowptr, sizeof(int) * (na + 1)))) return rc;
    if ((rc = xtb_reserve(c, c->d_maxnn, sizeof(int) * 4))) return rc;
    return XTB_OK;
}

// Steady-state form used inside xtb_step (no host round trip): the list is rebuilt with the capacity, neighbour bound and
// block capacity of the previous step; nn_rowptr_kernel raises meta[2] and empties the list if one of them is exceeded
// (later stages then run on a harmless empty list), and xtb_step re-runs the step through the sizing path below when it
// finds the flag after its single synchronisation.  meta is copied to pinned host memory asynchronously.
int xtb_impl_neighbours_steady(xtb_ctx* c, double rm) {
    if (!c->have_shape || c->layout_parts != 1) XTB_FAIL(c, XTB_ESTATE, "neighbours (steady): no sized list yet");
    int rc;
    XTB_T0(c, 1);
    if ((rc = launch_list(c, rm, c->max_nn, c->blk_cap))) return rc;
    XTB_CUDA(c, cudaMemcpyAsync(c->h_meta, c->d_maxnn.p, sizeof(int) * 4, cudaMemcpyDeviceToHost, c->stream));
    XTB_T1(c, 1);
    c->rm = rm;
    c->have_neigh = true;
    c->have_HS = c->have_blocks = c->have_density = c->have_dens = false;
    return XTB_OK;
}

int xtb_impl_neighbours(xtb_ctx* c, double rm) {
    const int na = c->na;
    if (na <= 0) XTB_FAIL(c, XTB_ESTATE, "xtb_neighbours: no system set");
    // capacity guess from the mean density (grown and re-run if exceeded)
    if (c->cap <= 0) {
        const double* h = c->h_supce;
        double vol = h[0] * (h[4] * h[8] - h[5] * h[7]) - h[3] * (h[1] * h[8] - h[2] * h[7]) + h[6] * (h[1] * h[5] - h[2] * h[4]);
        vol = vol < 0 ? -vol : vol;
        double est = vol > 0 ? na * (4.18879020478639 * rm * rm * rm) / vol : na;
        int cap = (int)(est * 1.5) + 16;
        if (cap > na - 1) cap = na - 1;
        if (cap < 1) cap = 1;
        c->cap = cap;
    }
    int rc;
    XTB_T0(c, 1);
    int meta[4] = {0, 0, 0, 0};
    for (int attempt = 0; attempt < 4; ++attempt) {
        if ((rc = reserve_list(c))) return rc;
        if ((rc = launch_list(c, rm, 0x7fffffff, 0x7fffffffffffLL))) return rc;
        XTB_CUDA(c, cudaMemcpyAsync(meta, c->d_maxnn.p, sizeof(int) * 4, cudaMemcpyDeviceToHost, c->stream));
        XTB_T1(c, 1);
        XTB_CUDA(c, cudaStreamSynchronize(c->stream));
        if (meta[0] <= c->cap) break;
        if (attempt == 3) XTB_FAIL(c, XTB_ECAPACITY, "neighbour capacity %d exceeded (%d) after retries", c->cap, meta[0]);
        const int mx = meta[0];
        c->cap = mx + 8 > na - 1 ? na - 1 : mx + 8;
        if (c->cap < mx) c->cap = mx;
    }
    if ((rc = xtb_collect_timing(c))) return rc;
    c->max_nn = c->max_nn_seen = meta[0];
    c->n_blocks = meta[3];
    c->n_pairs = meta[1] - na;      // real list entries (n_blocks counts the layout gaps of xtb_set_block_layout as well)
    c->blk_cap = c->n_blocks;
    c->have_shape = false;          // xtb_step arms the steady-state form (with its slack) when it wants it
    c->rm = rm;
    c->have_neigh = true;
    c->have_HS = c->have_blocks = c->have_density = c->have_dens = c->have_halo = false;
    if (c->comm && c->nranks > 1 && (rc = xtb_impl_comm_lists(c))) return rc;      // halo transfer lists of this list
    return XTB_OK;
}
