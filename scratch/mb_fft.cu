// mb_fft.cu -- how fast do the pieces of the 64-thread negacyclic transform run in isolation on one SM?
//   mode 0: 3 radix-8 passes in registers, no exchange (216 DFMA / thread / iter)      -> FP64 issue rate
//   mode 1: negacyclic_forward (3 passes + 2 shared-memory exchanges, named barriers)
//   mode 2: negacyclic_inverse
//   mode 3: forward, then 64-DFMA spectral MAC against shared-memory rows (16 LDS.128)
// for G = 4 and 6 groups of 64 threads per CTA (8 / 12 warps per SM), one CTA per SM.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I eoc_tfhe_b200/csrc -o scratch/mb_fft scratch/mb_fft.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "negacyclic_fft.cuh"
using namespace b200;

constexpr int ITERS = 1024;

template <int G, int MODE>
__global__ void __launch_bounds__(64 * G, 1) k_mb(double *sink, long long *cycles) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int g = threadIdx.x >> 6, t = threadIdx.x & 63;
    double2 *bufs = reinterpret_cast<double2 *>(smem) + g * 2 * kBufElems;
    const double2 *row = reinterpret_cast<const double2 *>(smem) + 6 * 2 * kBufElems + t;   // 16 KiB of key after the buffers
    for (int i = threadIdx.x; i < 6 * 2 * kBufElems + 2 * kM; i += blockDim.x)
        reinterpret_cast<double2 *>(smem)[i] = make_double2(1e-3 * i, -1e-3 * i);
    const RootSet w2 = load_rootset(c_tw.p2[t >> 3]), w3 = load_rootset(c_tw.p3[t]);
    const NamedBarrier64 bar{1 + g};
    double2 x[8], acc0[8], acc1[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) { x[r] = make_double2(t + r, t - r); acc0[r] = acc1[r] = make_double2(0, 0); }
    __syncthreads();
    const long long t0 = clock64();
    int toggle = 0;
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
        if (MODE == 0) {
            fwd_radix8(x, c_tw.p1);
            fwd_radix8(x, w2);
            fwd_radix8(x, w3);
        } else if (MODE == 1 || MODE == 3) {
            negacyclic_forward(x, bufs + toggle * kBufElems, t, w2, w3, bar);
            toggle ^= 1;
            if (MODE == 3) {
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    const double2 b0 = row[r * 64], b1 = row[kM + r * 64];
                    acc0[r].x = fma(x[r].x, b0.x, acc0[r].x);
                    acc0[r].x = fma(-x[r].y, b0.y, acc0[r].x);
                    acc0[r].y = fma(x[r].x, b0.y, acc0[r].y);
                    acc0[r].y = fma(x[r].y, b0.x, acc0[r].y);
                    acc1[r].x = fma(x[r].x, b1.x, acc1[r].x);
                    acc1[r].x = fma(-x[r].y, b1.y, acc1[r].x);
                    acc1[r].y = fma(x[r].x, b1.y, acc1[r].y);
                    acc1[r].y = fma(x[r].y, b1.x, acc1[r].y);
                }
            }
        } else {
            negacyclic_inverse(x, bufs + toggle * kBufElems, t, w2, w3, bar);
            toggle ^= 1;
#pragma unroll
            for (int r = 0; r < 8; ++r) { x[r].x *= 1e-3; x[r].y *= 1e-3; }
        }
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int r = 0; r < 8; ++r) s += x[r].x + x[r].y + acc0[r].x + acc1[r].y;
    if (s == 123.456) sink[0] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = t1 - t0;
}

template <int G, int MODE>
static void run(const char *name, double *sink, long long *cyc, int dfma_per_thread) {
    const int smem = 6 * 2 * kBufElems * 16 + 2 * kM * 16;
    cudaFuncSetAttribute(k_mb<G, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    k_mb<G, MODE><<<148, 64 * G, smem>>>(sink, cyc);
    k_mb<G, MODE><<<148, 64 * G, smem>>>(sink, cyc);
    cudaError_t e = cudaDeviceSynchronize();
    long long c = 0;
    cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    const double per = (double)c / ITERS;
    const double floor_clk = dfma_per_thread * 2.0 * G / 2.0;   // warp-DFMA per SM / (2 per clk)
    printf("%-34s G=%d: %8.1f clk/iter per SM, %7.1f clk per transform-ct, FP64 floor %6.1f (%.0f%%)  %s\n", name, G, per, per / G,
           floor_clk, 100.0 * floor_clk / per, e == cudaSuccess ? "" : cudaGetErrorString(e));
}

int main() {
    static TwiddleTables h;
    double2 *p = reinterpret_cast<double2 *>(&h);
    for (size_t i = 0; i < sizeof(h) / sizeof(double2); ++i) p[i] = make_double2(0.7 + 1e-3 * i, 0.3 - 1e-3 * i);
    cudaMemcpyToSymbol(c_tw, &h, sizeof(h));
    double *sink;
    long long *cyc;
    cudaMalloc(&sink, 8);
    cudaMalloc(&cyc, 8);
    run<4, 0>("3 radix-8 passes, registers only", sink, cyc, 216);
    run<6, 0>("3 radix-8 passes, registers only", sink, cyc, 216);
    run<4, 1>("forward transform", sink, cyc, 216);
    run<6, 1>("forward transform", sink, cyc, 216);
    run<4, 2>("inverse transform (+16 DMUL)", sink, cyc, 288 + 16);
    run<6, 2>("inverse transform (+16 DMUL)", sink, cyc, 288 + 16);
    run<4, 3>("forward + MAC", sink, cyc, 216 + 64);
    run<6, 3>("forward + MAC", sink, cyc, 216 + 64);
    return 0;
}
