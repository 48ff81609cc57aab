// mb_issue.cu -- does a DFMA hold the warp scheduler's issue port for one cycle or two?
// 64 DFMA (8 chains) and/or NI integer ops (16 chains, LOP3/IADD3 mix the compiler cannot fold) per iteration.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scratch/mb_issue scratch/mb_issue.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
constexpr int ITERS = 2048;
template <int THREADS, int NI>
__global__ void __launch_bounds__(THREADS, 1) k(int mode, double *sink, long long *cycles, uint32_t seed) {
    const int t = threadIdx.x;
    double x[8];
    uint32_t v[16];
#pragma unroll
    for (int r = 0; r < 8; ++r) x[r] = t - r;
#pragma unroll
    for (int r = 0; r < 16; ++r) v[r] = t * 17 + r + seed;
    const double c = 1.0 - 1e-9, e = 1e-9;
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int k2 = 0; k2 < 8; ++k2) {
            if (mode & 1) {
#pragma unroll
                for (int r = 0; r < 8; ++r) x[r] = fma(x[r], c, e);
            }
            if (mode & 2) {
#pragma unroll
                for (int r = 0; r < NI / 8; ++r) v[r] = (v[r] ^ seed) + (v[r] >> 3);   // LOP3 + shift-add per element
            }
        }
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int r = 0; r < 8; ++r) s += x[r];
#pragma unroll
    for (int r = 0; r < 16; ++r) s += v[r];
    if (s == 123.456) sink[0] = s;
    if (t == 0) cycles[blockIdx.x] = t1 - t0;
}
// dependent DFMA chain of one warp per SM sub-partition: latency
__global__ void k_lat(double *sink, long long *cycles, int chains) {
    double x[8];
    for (int r = 0; r < 8; ++r) x[r] = threadIdx.x + r;
    const double c = 1.0 - 1e-9, e = 1e-9;
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
        if (chains == 1) { x[0] = fma(x[0], c, e); x[0] = fma(x[0], c, e); x[0] = fma(x[0], c, e); x[0] = fma(x[0], c, e); }
        else if (chains == 2) { x[0] = fma(x[0], c, e); x[1] = fma(x[1], c, e); x[0] = fma(x[0], c, e); x[1] = fma(x[1], c, e); }
        else if (chains == 4) { x[0] = fma(x[0], c, e); x[1] = fma(x[1], c, e); x[2] = fma(x[2], c, e); x[3] = fma(x[3], c, e); }
    }
    const long long t1 = clock64();
    double s = 0;
    for (int r = 0; r < 8; ++r) s += x[r];
    if (s == 123.456) sink[0] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}
template <int THREADS, int NI>
static void run(int mode, const char *name, double *sink, long long *cyc) {
    k<THREADS, NI><<<148, THREADS>>>(mode, sink, cyc, 12345u);
    cudaDeviceSynchronize();
    k<THREADS, NI><<<148, THREADS>>>(mode, sink, cyc, 12345u);
    cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0;
    for (int i = 0; i < 148; ++i) avg += h[i];
    printf("%-44s thr %3d: %8.1f clk/iter per SM\n", name, THREADS, avg / 148 / ITERS);
}
int main() {
    double *sink; long long *cyc;
    cudaMalloc(&sink, 8); cudaMalloc(&cyc, 148 * 8);
#define ALL(T) \
    run<T, 64>(1, "64 DFMA", sink, cyc); run<T, 64>(2, "64 int elements (about 128-192 instr)", sink, cyc); run<T, 64>(3, "64 DFMA + 64 int elements", sink, cyc); \
    run<T, 128>(2, "128 int elements", sink, cyc); run<T, 128>(3, "64 DFMA + 128 int elements", sink, cyc);
    ALL(128) ALL(256) ALL(384) ALL(512)
    for (int ch : {1, 2, 4}) {
        k_lat<<<148, 32>>>(sink, cyc, ch);
        cudaDeviceSynchronize();
        long long h[148];
        cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        printf("one warp, %d DFMA chain(s), 4 DFMA per iteration: %.1f clk per iteration\n", ch, (double)h[0] / ITERS);
    }
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
