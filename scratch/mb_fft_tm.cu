// mb_fft_tm.cu -- chain latency and throughput of the tensor-memory transform pieces, G groups of 64 threads per CTA
// (one CTA per SM), clk per iteration per SM and per group:
//   mode 0: forward (single)        mode 1: forward2 (pair)        mode 2: inverse2 (pair)
//   mode 3: forward2 without the tensor-memory exchanges' data dependence (st/ld issued, results ignored) -- n/a
//   mode 4: one TMEM round trip alone (st 32x32b x16 doubles -> wait -> 2 ld 16x256b -> use)
//   mode 5: shared-memory exchange alone (8 STS.128, bar, 8 LDS.128)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I eoc_tfhe_b200/csrc -o scratch/mb_fft_tm scratch/mb_fft_tm.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "negacyclic_fft_tm.cuh"
using namespace b200;
constexpr int ITERS = 512;

__device__ __forceinline__ uint32_t tm_alloc_cta(uint32_t *slot) {
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t warp = threadIdx.x >> 5;
    return *slot + (((warp & 3u) * 32u) << 16) + (warp >> 2) * 128u;
}

template <int G, int MODE>
__global__ void __launch_bounds__(64 * G, 1) k_mb(double *sink, long long *cycles) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint32_t slot;
    const int g = threadIdx.x >> 6, t = threadIdx.x & 63;
    double2 *bufs = reinterpret_cast<double2 *>(smem) + g * 2 * kTmBufElems;
    for (int i = threadIdx.x; i < G * 2 * kTmBufElems; i += blockDim.x) reinterpret_cast<double2 *>(smem)[i] = make_double2(1e-3 * i, -1e-3 * i);
    const uint32_t tm = tm_alloc_cta(&slot);
    FftTm fft;
    fft.init(t, tm);
    const NamedBarrier64 bar{1 + g};
    double2 xa[8], xb[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) { xa[r] = make_double2(t + r, t - r); xb[r] = make_double2(t - r, t + r); }
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
        if (MODE == 0) { fft.forward(xa, bufs, bar); bar(); }
        else if (MODE == 1) { fft.forward2(xa, xb, bufs, bufs + kTmBufElems, bar); bar(); }
        else if (MODE == 2) {
            fft.template inverse2<true>(xa, xb, bufs, bufs + kTmBufElems, bar); bar();
#pragma unroll
            for (int r = 0; r < 8; ++r) { xa[r].x *= 1e-3; xa[r].y *= 1e-3; xb[r].x *= 1e-3; xb[r].y *= 1e-3; }
        } else if (MODE == 4) {
            Tm8 lo, hi;
            double rt[8];
            tm_exchange_fwd_store<1>(xa, tm);
            tm_wait_st();
            tm_exchange_fwd_load<0>(lo, hi, rt, tm, tm + (16u << 16), 0);
            tm_unpack(xa, lo, hi);
#pragma unroll
            for (int r = 0; r < 8; ++r) { xa[r].x += 1.0; xa[r].y += 1.0; }
        } else if (MODE == 5) {
#pragma unroll
            for (int r = 0; r < 8; ++r) bufs[tm_slot(t + 64 * r)] = xa[r];
            bar();
            const int base = tm_slot(fft.rd0);
#pragma unroll
            for (int r = 0; r < 8; ++r) xa[r] = bufs[base + 8 * r];
#pragma unroll
            for (int r = 0; r < 8; ++r) { xa[r].x += 1.0; xa[r].y += 1.0; }
            bar();
        } else if (MODE == 6) {
            fwd_radix8(xa, c_tw.p1); fwd_radix8(xa, c_tw.p1); fwd_radix8(xa, c_tw.p1);
        }
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int r = 0; r < 8; ++r) s += xa[r].x + xa[r].y + xb[r].x + xb[r].y;
    if (s == 123.456) sink[0] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(slot) : "memory");
}

template <int G, int MODE>
static void run(const char *name, int transforms, double *sink, long long *cyc) {
    const int smem = G * 2 * kTmBufElems * 16;
    cudaFuncSetAttribute(k_mb<G, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    k_mb<G, MODE><<<148, 64 * G, smem>>>(sink, cyc);
    cudaDeviceSynchronize();
    k_mb<G, MODE><<<148, 64 * G, smem>>>(sink, cyc);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0;
    for (int i = 0; i < 148; ++i) avg += h[i];
    const double clk = avg / 148 / ITERS;
    printf("%-34s G=%d: %8.1f clk/iter (chain)  %8.1f clk per transform-ct per SM  %s\n", name, G, clk, transforms ? clk / (transforms * G) : 0.0,
           e == cudaSuccess ? "" : cudaGetErrorString(e));
}
int main() {
    double *sink; long long *cyc;
    cudaMalloc(&sink, 8); cudaMalloc(&cyc, 148 * 8);
#define ALL(G) \
    run<G, 6>("3 radix-8 passes, registers only", 1, sink, cyc); run<G, 0>("forward (single)", 1, sink, cyc); \
    run<G, 1>("forward2 (pair)", 2, sink, cyc); run<G, 2>("inverse2 (pair)", 2, sink, cyc); \
    run<G, 4>("TMEM round trip alone", 0, sink, cyc); run<G, 5>("smem exchange + 2 barriers alone", 0, sink, cyc);
    ALL(1) ALL(2) ALL(4) ALL(6)
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
