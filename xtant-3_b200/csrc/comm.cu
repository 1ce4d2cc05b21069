// Multi-GPU exchanges of the tight-binding step INSIDE the library (one process / MPI rank per GPU).
//
// The reference's MPI build replicates every array, splits the atom loops round-robin and MPI_Allreduce(SUM)s whole dense
// matrices (TB_NRL.f90:267-327, TB.f90:4374, MPI_subroutines.f90:3965-3982).  Here atoms are row-partitioned into contiguous
// ranges, every block has one owner, and what travels over NVLink is
//     H, S blocks      in-place all-gather of the ranks' block segments (xtb_set_block_layout)  -> replicated dense H_non, Sij
//     rho, rho_E       HALO ONLY: a rank receives the owner blocks (j, q) whose row atom nn_list[j][q] it owns (they are
//                      transposed into its own rows by the mirror pass) and the on-site blocks of the neighbours of its atoms
//                      (the on-site force weight w_j, TB_NRL.f90:3102-3118) -- point-to-point ncclSend / ncclRecv of packed
//                      blocks between ranks whose ranges touch, instead of an all-gather that delivers every block to everyone
//     forces           in-place all-gather of 3 * ceil(na / G) doubles
//     supercell force  all-reduce of 9 doubles
// The transfer lists need no index exchange: positions and the neighbour list are replicated, so sender and receiver
// enumerate the same (j, q) in the same order.
// NCCL is bound at run time (dlopen, like cuSOLVER in eigen.cu): the library links only cudart, and a single-GPU user
// never loads NCCL.
#include "xtb_ctx.h"
#include <dlfcn.h>
#include <cstring>
#include <cstdlib>

namespace {

typedef void* nccl_comm_t;
struct nccl_uid { char internal[XTB_UNIQUE_ID_BYTES]; };
typedef int (*fn_uid_t)(nccl_uid*);
typedef int (*fn_init_t)(nccl_comm_t*, int, nccl_uid, int);
typedef int (*fn_destroy_t)(nccl_comm_t);
typedef int (*fn_allgather_t)(const void*, void*, size_t, int, nccl_comm_t, cudaStream_t);
typedef int (*fn_allreduce_t)(const void*, void*, size_t, int, int, nccl_comm_t, cudaStream_t);
typedef int (*fn_sendrecv_t)(void*, size_t, int, int, nccl_comm_t, cudaStream_t);
typedef int (*fn_group_t)(void);
typedef const char* (*fn_errstr_t)(int);
constexpr int NCCL_F64 = 8, NCCL_SUM = 0;          // ncclFloat64, ncclSum (nccl.h)

struct Nccl {
    bool tried = false, ok = false;
    std::string why;
    void* so = nullptr;
    fn_uid_t uid = nullptr; fn_init_t init = nullptr; fn_destroy_t destroy = nullptr;
    fn_allgather_t allgather = nullptr; fn_allreduce_t allreduce = nullptr;
    fn_sendrecv_t send = nullptr, recv = nullptr;
    fn_group_t gstart = nullptr, gend = nullptr;
    fn_errstr_t errstr = nullptr;
};
Nccl g_nccl;

bool bind_nccl() {
    Nccl& L = g_nccl;
    if (L.tried) return L.ok;
    L.tried = true;
    const char* names[] = {getenv("XTB_NCCL_LIB"), "libnccl.so.2", "libnccl.so", nullptr};
    for (int i = 0; i < 3 && !L.so; ++i)
        if (names[i] && names[i][0]) L.so = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
    if (!L.so) { L.why = "libnccl.so.2 not found (set XTB_NCCL_LIB)"; return false; }
#define BIND(field, type, name) L.field = (type)dlsym(L.so, name); if (!L.field) { L.why = std::string("missing symbol ") + name; return false; }
    BIND(uid, fn_uid_t, "ncclGetUniqueId")
    BIND(init, fn_init_t, "ncclCommInitRank")
    BIND(destroy, fn_destroy_t, "ncclCommDestroy")
    BIND(allgather, fn_allgather_t, "ncclAllGather")
    BIND(allreduce, fn_allreduce_t, "ncclAllReduce")
    BIND(send, fn_sendrecv_t, "ncclSend")
    BIND(recv, fn_sendrecv_t, "ncclRecv")
    BIND(gstart, fn_group_t, "ncclGroupStart")
    BIND(gend, fn_group_t, "ncclGroupEnd")
    BIND(errstr, fn_errstr_t, "ncclGetErrorString")
#undef BIND
    L.ok = true;
    return true;
}

#define XTB_NCCL(c, call)                                                                               \
    do {                                                                                                \
        int _r = (call);                                                                                \
        if (_r != 0) XTB_FAIL(c, XTB_ECOMM, "%s failed: %s (%s:%d)", #call, g_nccl.errstr(_r), __FILE__, __LINE__); \
    } while (0)

// ---- halo transfer lists ------------------------------------------------------------------------------------------------
// Transfer (s -> d): the slots rowptr[j] + q of the atoms j of range s, in ascending (j, q), that rank d needs:
//   q = 0  (on-site block of j)   iff some neighbour of j lies in range d  (lists are symmetric: j is then a neighbour of it)
//   q >= 1 (pair block)           iff atom j owns the pair (own flag) and the row atom nn_list[j][q-1] lies in range d.
// One warp per (atom j of range s, transfer t).  Pass 0 counts per atom, the single-CTA scan turns the counts into
// offsets, pass 1 writes the slots.
__global__ void halo_count_kernel(int ntr, const int* __restrict__ tr /*[ntr][4] = s0, s1, d0, d1*/, int cap, int maxatoms,
                                  const int* __restrict__ nn_size, const int* __restrict__ nn_list, const signed char* __restrict__ own,
                                  int* __restrict__ cnt /*[ntr][maxatoms]*/) {
    const int t = blockIdx.y;
    const int s0 = tr[4 * t], s1 = tr[4 * t + 1], d0 = tr[4 * t + 2], d1 = tr[4 * t + 3];
    const int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (w >= maxatoms) return;
    const int j = s0 + w;
    if (j >= s1) { if (lane == 0) cnt[(size_t)t * maxatoms + w] = 0; return; }
    const int n = nn_size[j];
    int own_in = 0, any_in = 0;
    for (int k = lane; k < n; k += 32) {
        const int i = nn_list[(size_t)j * cap + k];
        const bool in = i >= d0 && i < d1;
        any_in |= in ? 1 : 0;
        own_in += (in && own[(size_t)j * cap + k]) ? 1 : 0;
    }
    for (int off = 16; off > 0; off >>= 1) {
        own_in += __shfl_xor_sync(0xffffffffu, own_in, off);
        any_in |= __shfl_xor_sync(0xffffffffu, any_in, off);
    }
    if (lane == 0) cnt[(size_t)t * maxatoms + w] = own_in + (any_in ? 1 : 0);
}

// exclusive scan of cnt[t][0..maxatoms) per transfer (one CTA per transfer); total -> tot[t]
__global__ void halo_scan_kernel(int maxatoms, int* __restrict__ cnt, int* __restrict__ tot) {
    __shared__ int buf[1024];
    __shared__ int carry;
    int* c = cnt + (size_t)blockIdx.x * maxatoms;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < maxatoms; base += 1024) {
        const int i = base + threadIdx.x;
        const int v = i < maxatoms ? c[i] : 0;
        buf[threadIdx.x] = v;
        __syncthreads();
        for (int off = 1; off < 1024; off <<= 1) {
            const int x = threadIdx.x >= off ? buf[threadIdx.x - off] : 0;
            __syncthreads();
            buf[threadIdx.x] += x;
            __syncthreads();
        }
        if (i < maxatoms) c[i] = carry + buf[threadIdx.x] - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry += buf[1023];
        __syncthreads();
    }
    if (threadIdx.x == 0) tot[blockIdx.x] = carry;
}

__global__ void halo_fill_kernel(int ntr, const int* __restrict__ tr, int cap, int maxatoms, int listcap,
                                 const int* __restrict__ nn_size, const int* __restrict__ nn_list, const signed char* __restrict__ own,
                                 const int* __restrict__ rowptr, const int* __restrict__ off /*scanned cnt*/,
                                 int* __restrict__ slots /*[ntr][listcap]*/) {
    const int t = blockIdx.y;
    const int s0 = tr[4 * t], s1 = tr[4 * t + 1], d0 = tr[4 * t + 2], d1 = tr[4 * t + 3];
    const int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    const int j = s0 + w;
    if (w >= maxatoms || j >= s1) return;
    const int n = nn_size[j];
    int any_in = 0;
    for (int k = lane; k < n; k += 32) {
        const int i = nn_list[(size_t)j * cap + k];
        any_in |= (i >= d0 && i < d1) ? 1 : 0;
    }
    for (int o = 16; o > 0; o >>= 1) any_in |= __shfl_xor_sync(0xffffffffu, any_in, o);
    if (!any_in) return;
    int pos = off[(size_t)t * maxatoms + w];
    int* out = slots + (size_t)t * listcap;
    const int base = rowptr[j];
    if (lane == 0 && pos < listcap) out[pos] = base;        // on-site block
    ++pos;
    for (int k0 = 0; k0 < n; k0 += 32) {
        const int k = k0 + lane;
        bool need = false;
        if (k < n) {
            const int i = nn_list[(size_t)j * cap + k];
            need = own[(size_t)j * cap + k] && i >= d0 && i < d1;
        }
        const unsigned m = __ballot_sync(0xffffffffu, need);
        if (need) {
            const int p = pos + __popc(m & ((1u << lane) - 1u));
            if (p < listcap) out[p] = base + 1 + k;
        }
        pos += __popc(m);
    }
}

// pack: buf[(b * 2 + m) * nb2 + x] = store_m[slot[b] * nb2 + x]; unpack is the inverse
__global__ void halo_pack_kernel(int nblk, int nb2, const int* __restrict__ slots, const double* __restrict__ A, const double* __restrict__ B,
                                 double* __restrict__ buf) {
    const int b = blockIdx.x;
    if (b >= nblk) return;
    const size_t src = (size_t)slots[b] * nb2;
    for (int x = threadIdx.x; x < nb2; x += blockDim.x) {
        buf[((size_t)b * 2) * nb2 + x] = A[src + x];
        buf[((size_t)b * 2 + 1) * nb2 + x] = B[src + x];
    }
}
__global__ void halo_unpack_kernel(int nblk, int nb2, const int* __restrict__ slots, const double* __restrict__ buf,
                                   double* __restrict__ A, double* __restrict__ B) {
    const int b = blockIdx.x;
    if (b >= nblk) return;
    const size_t dst = (size_t)slots[b] * nb2;
    for (int x = threadIdx.x; x < nb2; x += blockDim.x) {
        A[dst + x] = buf[((size_t)b * 2) * nb2 + x];
        B[dst + x] = buf[((size_t)b * 2 + 1) * nb2 + x];
    }
}

inline void range_of(int na, int world, int r, int& b, int& e) {
    const int per = (na + world - 1) / world;
    b = r * per < na ? r * per : na;
    e = b + per < na ? b + per : na;
}

}  // namespace

// Transfer lists of this rank for the current neighbour list: sends (rank -> d) then receives (s -> rank), d, s != rank.
// Called at the end of the (sizing) neighbour stage of a communicating context; one small read-back of the counts.
int xtb_impl_comm_lists(xtb_ctx* c) {
    const int G = c->nranks, me = c->rank, na = c->na;
    if (G <= 1) return XTB_OK;
    const int ntr = 2 * (G - 1);
    const int per = (na + G - 1) / G;
    std::vector<int> tr(4 * ntr);
    int t = 0;
    for (int pass = 0; pass < 2; ++pass)
        for (int r = 0; r < G; ++r) {
            if (r == me) continue;
            int m0, m1, r0, r1;
            range_of(na, G, me, m0, m1); range_of(na, G, r, r0, r1);
            if (pass == 0) { tr[4 * t] = m0; tr[4 * t + 1] = m1; tr[4 * t + 2] = r0; tr[4 * t + 3] = r1; }     // send me -> r
            else           { tr[4 * t] = r0; tr[4 * t + 1] = r1; tr[4 * t + 2] = m0; tr[4 * t + 3] = m1; }     // recv r -> me
            ++t;
        }
    const int listcap = per * (c->max_nn + 1) > 0 ? per * (c->max_nn + 1) : 1;
    int rc;
    if ((rc = xtb_reserve(c, c->d_halo_tr, sizeof(int) * 4 * ntr))) return rc;
    if ((rc = xtb_reserve(c, c->d_halo_cnt, sizeof(int) * ((size_t)ntr * per + ntr)))) return rc;
    if ((rc = xtb_reserve(c, c->d_halo_slots, sizeof(int) * (size_t)ntr * listcap))) return rc;
    XTB_CUDA(c, cudaMemcpyAsync(c->d_halo_tr.p, tr.data(), sizeof(int) * 4 * ntr, cudaMemcpyHostToDevice, c->stream));
    int* cnt = (int*)c->d_halo_cnt.p; int* tot = cnt + (size_t)ntr * per;
    const dim3 grid((per + 7) / 8, ntr);
    halo_count_kernel<<<grid, 256, 0, c->stream>>>(ntr, (const int*)c->d_halo_tr.p, c->cap, per, (const int*)c->d_nn_size.p,
                                                   (const int*)c->d_nn_list.p, (const signed char*)c->d_own.p, cnt);
    XTB_LAUNCH_CHECK(c);
    halo_scan_kernel<<<ntr, 1024, 0, c->stream>>>(per, cnt, tot);
    XTB_LAUNCH_CHECK(c);
    halo_fill_kernel<<<grid, 256, 0, c->stream>>>(ntr, (const int*)c->d_halo_tr.p, c->cap, per, listcap, (const int*)c->d_nn_size.p,
                                                  (const int*)c->d_nn_list.p, (const signed char*)c->d_own.p, (const int*)c->d_rowptr.p, cnt,
                                                  (int*)c->d_halo_slots.p);
    XTB_LAUNCH_CHECK(c);
    c->halo_count.assign(ntr, 0);
    XTB_CUDA(c, cudaMemcpyAsync(c->halo_count.data(), tot, sizeof(int) * ntr, cudaMemcpyDeviceToHost, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));      // tr (host vector) is free, the counts are here
    c->halo_listcap = listcap;
    long long nsend = 0, nrecv = 0;
    for (int k = 0; k < G - 1; ++k) { nsend += c->halo_count[k]; nrecv += c->halo_count[G - 1 + k]; }
    c->halo_send_blocks = nsend; c->halo_recv_blocks = nrecv;
    c->have_halo = true;
    return XTB_OK;
}

// which = XTB_XCHG_HS: all-gather of the H and S block segments; XTB_XCHG_RHO: halo exchange of rho / rho_E (or the
// all-gather when XTB_RHO_EXCHANGE=allgather).  Everything is enqueued on the context stream; no host wait.
int xtb_impl_exchange_blocks(xtb_ctx* c, int which, bool side_stream) {
    const int G = c->nranks, me = c->rank;
    if (G <= 1) return XTB_OK;
    // side_stream (H/S all-gather only): the collective starts when the blocks are written (ev_comm0) and runs on
    // comm_stream; the caller keeps working on the step's stream and waits for ev_comm1 before it reads foreign blocks
    cudaStream_t st = c->stream;
    if (side_stream && which == XTB_XCHG_HS && c->comm_stream && !c->capturing) {
        XTB_CUDA(c, cudaEventRecord(c->ev_comm0, c->stream));
        XTB_CUDA(c, cudaStreamWaitEvent(c->comm_stream, c->ev_comm0, 0));
        st = c->comm_stream;
    } else side_stream = false;
    Nccl& L = g_nccl;
    const size_t nb2 = (size_t)c->n_orb * c->n_orb;
    if (c->layout_parts != G) XTB_FAIL(c, XTB_ESTATE, "exchange: block layout has %d segments for %d ranks", c->layout_parts, G);
    const size_t seg = (size_t)(c->n_blocks / G) * nb2;          // doubles per rank segment (rowptr[na] = G * largest segment)
    double* A = (double*)(which == XTB_XCHG_HS ? c->d_Hblk.p : c->d_rho.p);
    double* B = (double*)(which == XTB_XCHG_HS ? c->d_Sblk.p : c->d_rhoE.p);
    const bool two = !(which == XTB_XCHG_HS && c->tb == XTB_TB_PETTIFOR);      // the orthogonal branch has no S blocks
    const int stage = which == XTB_XCHG_HS ? 14 : 15;
    if (!side_stream) XTB_T0(c, stage);
    if (which == XTB_XCHG_HS || c->rho_allgather) {
        XTB_NCCL(c, L.gstart());
        XTB_NCCL(c, L.allgather(A + seg * me, A, seg, NCCL_F64, (nccl_comm_t)c->comm, st));
        if (two) XTB_NCCL(c, L.allgather(B + seg * me, B, seg, NCCL_F64, (nccl_comm_t)c->comm, st));
        XTB_NCCL(c, L.gend());
        if (side_stream) {
            XTB_CUDA(c, cudaEventRecord(c->ev_comm1, st));
            c->xchg_bytes[which] = (double)seg * 8 * (two ? 2 : 1) * (G - 1);
            return XTB_OK;          // timed by the caller (stage 14 brackets the overlapped region)
        }
        c->xchg_bytes[which] = (double)seg * 8 * (two ? 2 : 1) * (G - 1);      // received per rank
    } else {
        if (!c->have_halo) XTB_FAIL(c, XTB_ESTATE, "exchange: no halo transfer lists (run the neighbour stage after xtb_comm_init)");
        int rc;
        const size_t per_blk = 2 * nb2;
        if ((rc = xtb_reserve(c, c->d_halo_send, sizeof(double) * per_blk * (size_t)(c->halo_send_blocks > 0 ? c->halo_send_blocks : 1)))) return rc;
        if ((rc = xtb_reserve(c, c->d_halo_recv, sizeof(double) * per_blk * (size_t)(c->halo_recv_blocks > 0 ? c->halo_recv_blocks : 1)))) return rc;
        const int* slots = (const int*)c->d_halo_slots.p;
        double* sb = (double*)c->d_halo_send.p; double* rb = (double*)c->d_halo_recv.p;
        size_t so = 0;
        for (int k = 0; k < G - 1; ++k) {
            const int n = c->halo_count[k];
            if (n > 0) {
                halo_pack_kernel<<<n, 96, 0, c->stream>>>(n, (int)nb2, slots + (size_t)k * c->halo_listcap, A, B, sb + so * per_blk);
                XTB_LAUNCH_CHECK(c);
            }
            so += n;
        }
        XTB_NCCL(c, L.gstart());
        so = 0; size_t ro = 0;
        for (int k = 0, r = 0; r < G; ++r) {
            if (r == me) continue;
            const int ns = c->halo_count[k], nr = c->halo_count[G - 1 + k];
            if (ns > 0) XTB_NCCL(c, L.send(sb + so * per_blk, (size_t)ns * per_blk, NCCL_F64, r, (nccl_comm_t)c->comm, c->stream));
            if (nr > 0) XTB_NCCL(c, L.recv(rb + ro * per_blk, (size_t)nr * per_blk, NCCL_F64, r, (nccl_comm_t)c->comm, c->stream));
            so += ns; ro += nr; ++k;
        }
        XTB_NCCL(c, L.gend());
        ro = 0;
        for (int k = 0; k < G - 1; ++k) {
            const int n = c->halo_count[G - 1 + k];
            if (n > 0) {
                halo_unpack_kernel<<<n, 96, 0, c->stream>>>(n, (int)nb2, slots + (size_t)(G - 1 + k) * c->halo_listcap, rb + ro * per_blk, A, B);
                XTB_LAUNCH_CHECK(c);
            }
            ro += n;
        }
        c->xchg_bytes[which] = (double)c->halo_recv_blocks * per_blk * 8;
    }
    XTB_T1(c, stage);
    return XTB_OK;
}

// forces (3, na): in-place all-gather of the owners' ranges (equal counts of 3 * ceil(na / G); d_F is padded to G ranges)
int xtb_impl_exchange_forces(xtb_ctx* c, DevBuf& F) {
    const int G = c->nranks, me = c->rank;
    if (G <= 1) return XTB_OK;
    const size_t per = (size_t)((c->na + G - 1) / G) * 3;
    if (F.bytes < sizeof(double) * per * G) XTB_FAIL(c, XTB_ESTATE, "exchange_forces: force buffer not padded for %d ranks", G);
    double* p = (double*)F.p;
    XTB_NCCL(c, g_nccl.allgather(p + per * me, p, per, NCCL_F64, (nccl_comm_t)c->comm, c->stream));
    return XTB_OK;
}

int xtb_impl_allreduce9(xtb_ctx* c, double* d9) {
    if (c->nranks <= 1) return XTB_OK;
    XTB_NCCL(c, g_nccl.allreduce(d9, d9, 9, NCCL_F64, NCCL_SUM, (nccl_comm_t)c->comm, c->stream));
    return XTB_OK;
}

extern "C" {

int xtb_comm_unique_id(void* id_out) {
    if (!id_out) return XTB_EINVAL;
    if (!bind_nccl()) return XTB_ECOMM;
    nccl_uid id;
    if (g_nccl.uid(&id) != 0) return XTB_ECOMM;
    memcpy(id_out, &id, sizeof id);
    return XTB_OK;
}

int xtb_comm_init(xtb_ctx* c, const void* unique_id, int rank, int nranks) {
    if (!c) return XTB_EINVAL;
    if (!unique_id || nranks < 1 || rank < 0 || rank >= nranks) XTB_FAIL(c, XTB_EINVAL, "xtb_comm_init: rank %d of %d", rank, nranks);
    if (c->comm) XTB_FAIL(c, XTB_ESTATE, "xtb_comm_init: the context already has a communicator");
    if (!bind_nccl()) XTB_FAIL(c, XTB_ECOMM, "xtb_comm_init: %s", g_nccl.why.c_str());
    cudaSetDevice(c->device);
    nccl_uid id;
    memcpy(&id, unique_id, sizeof id);
    nccl_comm_t comm = nullptr;
    XTB_NCCL(c, g_nccl.init(&comm, nranks, id, rank));
    c->comm = comm; c->rank = rank; c->nranks = nranks;
    if (!c->comm_stream) {
        XTB_CUDA(c, cudaStreamCreateWithFlags(&c->comm_stream, cudaStreamNonBlocking));
        XTB_CUDA(c, cudaEventCreateWithFlags(&c->ev_comm0, cudaEventDisableTiming));
        XTB_CUDA(c, cudaEventCreateWithFlags(&c->ev_comm1, cudaEventDisableTiming));
    }
    if (const char* ev = getenv("XTB_RHO_EXCHANGE")) c->rho_allgather = strcmp(ev, "allgather") == 0;
    // block stores in one equal-capacity segment per rank; the owned range follows the rank
    c->layout_parts = nranks;
    c->have_neigh = c->have_HS = c->have_blocks = c->have_density = c->have_dens = c->have_halo = c->have_shape = false;
    if (c->na > 0) { range_of(c->na, nranks, rank, c->a0, c->a1); }
    return XTB_OK;
}

int xtb_comm_destroy(xtb_ctx* c) {
    if (!c) return XTB_EINVAL;
    if (c->comm) {
        cudaSetDevice(c->device);
        cudaStreamSynchronize(c->stream);
        if (c->comm_stream) cudaStreamSynchronize(c->comm_stream);
        g_nccl.destroy((nccl_comm_t)c->comm);
        c->comm = nullptr;
        if (c->comm_stream) { cudaStreamDestroy(c->comm_stream); cudaEventDestroy(c->ev_comm0); cudaEventDestroy(c->ev_comm1); c->comm_stream = nullptr; }
    }
    c->rank = 0; c->nranks = 1; c->have_halo = false;
    return XTB_OK;
}

int xtb_exchange_blocks(xtb_ctx* c, int which) {
    if (!c) return XTB_EINVAL;
    if (which != XTB_XCHG_HS && which != XTB_XCHG_RHO) XTB_FAIL(c, XTB_EINVAL, "xtb_exchange_blocks: which = %d", which);
    if (!c->comm) XTB_FAIL(c, XTB_ESTATE, "xtb_exchange_blocks: no communicator (xtb_comm_init)");
    if (which == XTB_XCHG_HS && !c->have_blocks) XTB_FAIL(c, XTB_ESTATE, "xtb_exchange_blocks: run xtb_assemble_blocks first");
    if (which == XTB_XCHG_RHO && !c->have_density) XTB_FAIL(c, XTB_ESTATE, "xtb_exchange_blocks: run xtb_density first");
    cudaSetDevice(c->device);
    int rc;
    if ((rc = xtb_impl_exchange_blocks(c, which))) return rc;
    return xtb_stage_finish(c);
}

int xtb_exchange_forces(xtb_ctx* c, double* F_att) {
    if (!c) return XTB_EINVAL;
    if (!c->comm) XTB_FAIL(c, XTB_ESTATE, "xtb_exchange_forces: no communicator (xtb_comm_init)");
    if (!c->d_F.p) XTB_FAIL(c, XTB_ESTATE, "xtb_exchange_forces: run xtb_forces first");
    cudaSetDevice(c->device);
    int rc;
    if ((rc = xtb_impl_exchange_forces(c, c->d_F))) return rc;
    if (F_att) XTB_CUDA(c, cudaMemcpyAsync(F_att, c->d_F.p, sizeof(double) * 3 * c->na, cudaMemcpyDeviceToHost, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    return XTB_OK;
}

int xtb_allreduce_scforce(xtb_ctx* c, double* SCforce_att) {
    if (!c) return XTB_EINVAL;
    if (!c->comm) XTB_FAIL(c, XTB_ESTATE, "xtb_allreduce_scforce: no communicator (xtb_comm_init)");
    if (!c->d_SC.p) XTB_FAIL(c, XTB_ESTATE, "xtb_allreduce_scforce: run xtb_pressure first");
    cudaSetDevice(c->device);
    int rc;
    if ((rc = xtb_impl_allreduce9(c, (double*)c->d_SC.p))) return rc;
    if (SCforce_att) XTB_CUDA(c, cudaMemcpyAsync(SCforce_att, c->d_SC.p, sizeof(double) * 9, cudaMemcpyDeviceToHost, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    return XTB_OK;
}

}  // extern "C"
