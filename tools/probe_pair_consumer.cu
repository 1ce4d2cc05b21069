// compile-only probe (nvcc -gencode arch=compute_100a,code=sm_100a -O3 -Xptxas -v -c): register footprint of a consumer loop that
// feeds TWO atoms i from one A fragment (4 DMMA + 4 DFMA per fragment) with the fragment double buffer of density.cu and
// setmaxnreg 232 for the two consumer warpgroups / 40 for the helper warpgroup.  Result: profiles/r01y_probe_pair_consumer_registers.json
#include <cstdint>
__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
template <int MT> struct Frag { double a[MT], b[4], b8[4]; };
template <int MT>
__device__ __forceinline__ void load(Frag<MT>& f, const double* stg, const double* Bs, const int (&off)[MT], int s, int lane, int t) {
#pragma unroll
    for (int c = 0; c < 4; ++c) { f.b[c] = Bs[s * 144 + 32 * c + lane]; f.b8[c] = Bs[s * 144 + 128 + 4 * c + t]; }
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) f.a[mt] = stg[off[mt] + 36 * s];
}
template <int MT>
__global__ void __launch_bounds__(384, 1) probe(const double* __restrict__ smem_like, int nchunks, const int* __restrict__ offs, double* __restrict__ out) {
    extern __shared__ double sm[];
    const int lane = threadIdx.x & 31, t = lane & 3;
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = smem_like[i];
    __syncthreads();
    if (threadIdx.x >= 256) { asm volatile("setmaxnreg.dec.sync.aligned.u32 40;"); return; }
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    double acc[MT][8], p8[MT][4];
    int off[MT];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) { off[mt] = offs[mt * 32 + lane];
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[mt][c] = 0.0;
#pragma unroll
        for (int c = 0; c < 4; ++c) p8[mt][c] = 0.0; }
    Frag<MT> f0, f1;
    load<MT>(f0, sm, sm + 2048, off, 0, lane, t);
    for (int ch = 0; ch < nchunks; ++ch) {
        const double* stg = sm + (ch & 3) * 64;
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            Frag<MT>& cu = (s & 1) ? f1 : f0;
            Frag<MT>& nx = (s & 1) ? f0 : f1;
            load<MT>(nx, stg, stg + 2048, off, (s + 1) & 3, lane, t);
#pragma unroll
            for (int mt = 0; mt < MT; ++mt) {
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    dmma884(acc[mt][2 * c], acc[mt][2 * c + 1], cu.a[mt], cu.b[c]);
                    p8[mt][c] = fma(cu.a[mt], cu.b8[c], p8[mt][c]);
                }
            }
        }
    }
    double r = 0.0;
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
#pragma unroll
        for (int c = 0; c < 8; ++c) r += acc[mt][c];
#pragma unroll
        for (int c = 0; c < 4; ++c) r += p8[mt][c]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template __global__ void probe<4>(const double*, int, const int*, double*);
template __global__ void probe<5>(const double*, int, const int*, double*);
template __global__ void probe<6>(const double*, int, const int*, double*);
template __global__ void probe<7>(const double*, int, const int*, double*);
template __global__ void probe<8>(const double*, int, const int*, double*);
template __global__ void probe<9>(const double*, int, const int*, double*);
