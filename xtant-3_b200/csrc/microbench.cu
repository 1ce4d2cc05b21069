// FP64 peak probes (not part of the ABI): DMMA.8x8x4 and DFMA issue-rate kernels used once to obtain the
// FP64 roofline denominators that MEASURED_PEAKS.json does not carry (it has HBM and bf16 only).
#include "xtb_ctx.h"

namespace {
__global__ void __launch_bounds__(256) dmma_peak_kernel(int iters, double* out) {
    double acc[8][2];
#pragma unroll
    for (int i = 0; i < 8; ++i) { acc[i][0] = 0.0; acc[i][1] = 0.0; }
    double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(acc[i][0]), "+d"(acc[i][1]) : "d"(a), "d"(b));
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += acc[i][0] + acc[i][1];
    if (s == 123.456) out[0] = s;
}
__global__ void __launch_bounds__(256) dfma_peak_kernel(int iters, double* out) {
    double acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x * 1e-3 + i;
    const double a = 1.0000001, b = 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = fma(acc[i], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i];
    if (s == 123.456) out[0] = s;
}
// mode 2: every warp interleaves 8 DMMA with 16 DFMA per iteration (are the two FP64 pipes independent?)
__global__ void __launch_bounds__(256) mixed_peak_kernel(int iters, double* out) {
    double acc[8][2], f[16];
#pragma unroll
    for (int i = 0; i < 8; ++i) { acc[i][0] = 0.0; acc[i][1] = 0.0; }
#pragma unroll
    for (int i = 0; i < 16; ++i) f[i] = threadIdx.x * 1e-3 + i;
    double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
    const double fa = 1.0000001, fb = 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(acc[i][0]), "+d"(acc[i][1]) : "d"(a), "d"(b));
            f[2 * i] = fma(f[2 * i], fa, fb);
            f[2 * i + 1] = fma(f[2 * i + 1], fa, fb);
        }
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += acc[i][0] + acc[i][1];
#pragma unroll
    for (int i = 0; i < 16; ++i) s += f[i];
    if (s == 123.456) out[0] = s;
}
// mode 3: warp-specialised: even warps DMMA, odd warps DFMA
__global__ void __launch_bounds__(256) split_peak_kernel(int iters, double* out) {
    double acc[8][2], f[16];
#pragma unroll
    for (int i = 0; i < 8; ++i) { acc[i][0] = 0.0; acc[i][1] = 0.0; }
#pragma unroll
    for (int i = 0; i < 16; ++i) f[i] = threadIdx.x * 1e-3 + i;
    double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
    const double fa = 1.0000001, fb = 1e-9;
    if (((threadIdx.x >> 5) & 4) == 0) {       // warps 0-3 (one per SMSP): DMMA; warps 4-7: DFMA
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                             : "+d"(acc[i][0]), "+d"(acc[i][1]) : "d"(a), "d"(b));
        }
    } else {
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 16; ++i) f[i] = fma(f[i], fa, fb);
        }
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += acc[i][0] + acc[i][1];
#pragma unroll
    for (int i = 0; i < 16; ++i) s += f[i];
    if (s == 123.456) out[0] = s;
}
// ---- consumer-loop probes: the DMMA stream of density.cu without barriers / TMA, to separate issue effects ----
// variant 0: MT row tiles x 2 column tiles per k-step, operands in distinct registers (no loads)
// variant 1: + A and B operands re-loaded from shared memory every k-step (same conflict-free pattern as the kernel)
// variant 2: + the two DFMA per row tile (compiler-scheduled)   variant 3: the same DFMAs grouped after the DMMAs of the k-step
template <int MT, int VAR>
__global__ void __launch_bounds__(512) dmma_loop_probe(int iters, double* out) {
    __shared__ double sm[4 * 512 + 4 * 72];
    for (int e = threadIdx.x; e < 4 * 512 + 4 * 72; e += blockDim.x) sm[e] = 1.0 + 1e-9 * e;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, t = lane & 3;
    double acc0[MT][2], acc1[MT][2], p8f[MT], p8e[MT], a[MT];
#pragma unroll
    for (int m = 0; m < MT; ++m) { acc0[m][0] = acc0[m][1] = acc1[m][0] = acc1[m][1] = 0.0; p8f[m] = p8e[m] = 0.0; a[m] = 1.0 + lane * 1e-9 + m; }
    double b0 = 1.0 - lane * 1e-9, b1 = 1.0 + lane * 2e-9, b8f = 0.5, b8e = 0.25;
    const double* Bs = sm + 4 * 512;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            if (VAR >= 1) {
                b0 = Bs[s * 72 + lane]; b1 = Bs[s * 72 + 32 + lane]; b8f = Bs[s * 72 + 64 + t]; b8e = Bs[s * 72 + 68 + t];
#pragma unroll
                for (int m = 0; m < MT; ++m) a[m] = sm[s * 512 + ((warp * 7 + m) & 15) * 32 + lane];
            }
#pragma unroll
            for (int m = 0; m < MT; ++m) {
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                             : "+d"(acc0[m][0]), "+d"(acc0[m][1]) : "d"(a[m]), "d"(b0));
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                             : "+d"(acc1[m][0]), "+d"(acc1[m][1]) : "d"(a[m]), "d"(b1));
                if (VAR == 2) { p8f[m] = fma(a[m], b8f, p8f[m]); p8e[m] = fma(a[m], b8e, p8e[m]); }
            }
            if (VAR == 3) {      // the DFMAs of the k-step grouped after its DMMAs (volatile: order kept)
#pragma unroll
                for (int m = 0; m < MT; ++m) {
                    asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(p8f[m]) : "d"(a[m]), "d"(b8f));
                    asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(p8e[m]) : "d"(a[m]), "d"(b8e));
                }
            }
        }
        if (VAR >= 1) asm volatile("" ::: "memory");
    }
    double r = 0.0;
#pragma unroll
    for (int m = 0; m < MT; ++m) r += acc0[m][0] + acc0[m][1] + acc1[m][0] + acc1[m][1] + p8f[m] + p8e[m];
    if (r == 123.456) out[0] = r;
}
}  // namespace

// returns TFLOP/s of the probe (DMMA flops + DFMA flops), best of reps; warps_per_smsp in 1..4, mt in {4,5,6}, variant 0..2
extern "C" double xtb_debug_dmma_loop(xtb_ctx* c, int variant, int warps_per_smsp, int mt, int reps) {
    if (!c) return -1.0;
    cudaSetDevice(c->device);
    if (xtb_reserve(c, c->d_scratch, 256)) return -1.0;
    const int threads = 128 * warps_per_smsp, blocks = c->sm_count, iters = 4000;
    double best = 0.0;
    for (int r = 0; r < reps + 1; ++r) {
        xtb_t0(c, 15);
#define PROBE(M, V) dmma_loop_probe<M, V><<<blocks, threads, 0, c->stream>>>(iters, (double*)c->d_scratch.p)
#define PROBE_M(M) { if (variant == 0) PROBE(M, 0); else if (variant == 1) PROBE(M, 1); else if (variant == 2) PROBE(M, 2); else PROBE(M, 3); }
        if (mt == 4) PROBE_M(4) else if (mt == 5) PROBE_M(5) else PROBE_M(6)
#undef PROBE_M
#undef PROBE
        xtb_t1(c, 15);
        cudaStreamSynchronize(c->stream);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, c->tev[15][0], c->tev[15][1]); c->t_pending &= ~(1u << 15);
        const double warps = (double)blocks * threads / 32.0;
        const int m = mt == 4 ? 4 : mt == 5 ? 5 : 6;
        const double flop = warps * iters * 4.0 * m * (2.0 * 512.0 + (variant >= 2 ? 2.0 * 64.0 : 0.0));
        const double tf = flop / (ms * 1e-3) * 1e-12;
        if (r > 0 && tf > best) best = tf;
    }
    c->launches += reps + 1;
    return best;
}

// mode 0: DMMA.8x8x4, mode 1: DFMA, mode 2: both interleaved in every warp, mode 3: half the warps each.  Returns TFLOP/s (2 flop per FMA), best of `reps`.
extern "C" double xtb_debug_fp64_peak(xtb_ctx* c, int mode, int warps_per_sm, int reps) {
    if (!c) return -1.0;
    cudaSetDevice(c->device);
    if (xtb_reserve(c, c->d_scratch, 256)) return -1.0;
    const int threads = 256;
    const int blocks = c->sm_count * ((warps_per_sm * 32 + threads - 1) / threads);
    const int iters = 20000;
    double best = 0.0;
    for (int r = 0; r < reps + 1; ++r) {
        xtb_t0(c, 15);
        if (mode == 0) dmma_peak_kernel<<<blocks, threads, 0, c->stream>>>(iters, (double*)c->d_scratch.p);
        else if (mode == 2) mixed_peak_kernel<<<blocks, threads, 0, c->stream>>>(iters, (double*)c->d_scratch.p);
        else if (mode == 3) split_peak_kernel<<<blocks, threads, 0, c->stream>>>(iters, (double*)c->d_scratch.p);
        else dfma_peak_kernel<<<blocks, threads, 0, c->stream>>>(iters, (double*)c->d_scratch.p);
        xtb_t1(c, 15);
        cudaStreamSynchronize(c->stream);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, c->tev[15][0], c->tev[15][1]); c->t_pending &= ~(1u << 15);
        const double warps = (double)blocks * threads / 32.0;
        const double f_dmma = warps * iters * 8.0 * (8 * 8 * 4 * 2.0), f_dfma = warps * 32.0 * iters * 16.0 * 2.0;
        const double flop = mode == 0 ? f_dmma : mode == 1 ? f_dfma : mode == 2 ? f_dmma + f_dfma : 0.5 * (f_dmma + f_dfma);
        const double tf = flop / (ms * 1e-3) * 1e-12;
        if (r > 0 && tf > best) best = tf;
    }
    c->launches += reps + 1;
    return best;
}
