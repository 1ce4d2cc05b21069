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
}  // namespace

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
        cudaEventRecord(c->ev0, c->stream);
        if (mode == 0) dmma_peak_kernel<<<blocks, threads, 0, c->stream>>>(iters, (double*)c->d_scratch.p);
        else if (mode == 2) mixed_peak_kernel<<<blocks, threads, 0, c->stream>>>(iters, (double*)c->d_scratch.p);
        else if (mode == 3) split_peak_kernel<<<blocks, threads, 0, c->stream>>>(iters, (double*)c->d_scratch.p);
        else dfma_peak_kernel<<<blocks, threads, 0, c->stream>>>(iters, (double*)c->d_scratch.p);
        cudaEventRecord(c->ev1, c->stream);
        cudaStreamSynchronize(c->stream);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, c->ev0, c->ev1);
        const double warps = (double)blocks * threads / 32.0;
        const double f_dmma = warps * iters * 8.0 * (8 * 8 * 4 * 2.0), f_dfma = warps * 32.0 * iters * 16.0 * 2.0;
        const double flop = mode == 0 ? f_dmma : mode == 1 ? f_dfma : mode == 2 ? f_dmma + f_dfma : 0.5 * (f_dmma + f_dfma);
        const double tf = flop / (ms * 1e-3) * 1e-12;
        if (r > 0 && tf > best) best = tf;
    }
    c->launches += reps + 1;
    return best;
}
