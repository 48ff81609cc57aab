// mb_dfma.cu -- DFMA rate as a function of where its operands come from (register-file bandwidth?)
//   mode 0: x = fma(x, c, e)        c, e loop-invariant registers shared by all chains
//   mode 1: x = fma(a_r, b_r, x)    three distinct register pairs per DFMA
//   mode 2: x = fma(a_r, c, x)      two distinct + one shared
//   mode 3: a radix-2 (cos,tan) butterfly network on 8 complex registers (the kernel's own instruction pattern)
//   mode 4: x = fma(a_r, cbank, x)  one operand from the constant bank
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scratch/mb_dfma scratch/mb_dfma.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
constexpr int ITERS = 2048;
__constant__ double c_k[8];
__device__ __forceinline__ void bf(double2 &u, double2 &v, const double2 ct) {
    const double wr = fma(-ct.y, v.y, v.x), wi = fma(ct.y, v.x, v.y);
    const double ur = u.x, ui = u.y;
    u.x = fma(ct.x, wr, ur); u.y = fma(ct.x, wi, ui); v.x = fma(-ct.x, wr, ur); v.y = fma(-ct.x, wi, ui);
}
template <int THREADS>
__global__ void __launch_bounds__(THREADS, 1) k(int mode, double *sink, long long *cycles, double seed) {
    const int t = threadIdx.x;
    double x[8], a[8], b[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) { x[r] = t - r; a[r] = 1.0 - 1e-9 * (t + r) * seed; b[r] = 1e-9 * (r + 1) * seed; }
    double2 z[8];
    int iv[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) { z[r] = make_double2(t + r, t - r); iv[r] = t * 977 + r; }
    double2 w0 = make_double2(0.7 * seed, 0.3 * seed), w1 = make_double2(0.6 * seed, 0.2 * seed), w2 = make_double2(0.5 * seed, 0.1 * seed);
    const double c = 1.0 - 1e-9 * seed, e = 1e-9 * seed;
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
        if (mode == 0) {
#pragma unroll
            for (int k2 = 0; k2 < 9; ++k2)
#pragma unroll
                for (int r = 0; r < 8; ++r) x[r] = fma(x[r], c, e);
        } else if (mode == 1) {
#pragma unroll
            for (int k2 = 0; k2 < 9; ++k2)
#pragma unroll
                for (int r = 0; r < 8; ++r) x[r] = fma(a[r], b[(r + k2) & 7], x[r]);
        } else if (mode == 2) {
#pragma unroll
            for (int k2 = 0; k2 < 9; ++k2)
#pragma unroll
                for (int r = 0; r < 8; ++r) x[r] = fma(a[(r + k2) & 7], c, x[r]);
        } else if (mode == 3) {
            // 12 butterflies x 6 DFMA = 72 DFMA: one radix-8 pass
            bf(z[0], z[4], w0); bf(z[1], z[5], w0); bf(z[2], z[6], w0); bf(z[3], z[7], w0);
            bf(z[0], z[2], w1); bf(z[1], z[3], w1); bf(z[4], z[6], w1); bf(z[5], z[7], w1);
            bf(z[0], z[1], w2); bf(z[2], z[3], w2); bf(z[4], z[5], w2); bf(z[6], z[7], w2);
#pragma unroll
            for (int r = 0; r < 8; ++r) { z[r].x *= 0.25; z[r].y *= 0.25; }   // keep values bounded (16 DMUL extra)
        } else if (mode == 4) {
#pragma unroll
            for (int k2 = 0; k2 < 9; ++k2)
#pragma unroll
                for (int r = 0; r < 8; ++r) x[r] = fma(a[(r + k2) & 7], c_k[k2 & 7], x[r]);
        } else if (mode == 8) {          // 72 DADD, one shared operand
#pragma unroll
            for (int k2 = 0; k2 < 9; ++k2)
#pragma unroll
                for (int r = 0; r < 8; ++r) x[r] = x[r] + e;
        } else if (mode == 9) {          // 72 DADD, two distinct register operands
#pragma unroll
            for (int k2 = 0; k2 < 9; ++k2)
#pragma unroll
                for (int r = 0; r < 8; ++r) x[r] = x[r] + a[(r + k2) & 7];
        } else if (mode == 10) {         // 72 DMUL
#pragma unroll
            for (int k2 = 0; k2 < 9; ++k2)
#pragma unroll
                for (int r = 0; r < 8; ++r) x[r] = x[r] * c;
        } else if (mode == 11) {         // one inverse (GS) radix-2 butterfly network: 12 x (4 DADD + 2 DFMA + 2 DMUL)
#define GS(u, v, ct) { const double dr = u.x - v.x, di = u.y - v.y; u.x += v.x; u.y += v.y; const double wr = fma(ct.y, di, dr), wi = fma(-ct.y, dr, di); v.x = ct.x * wr; v.y = ct.x * wi; }
            GS(z[0], z[1], w2) GS(z[2], z[3], w2) GS(z[4], z[5], w2) GS(z[6], z[7], w2)
            GS(z[0], z[2], w1) GS(z[1], z[3], w1) GS(z[4], z[6], w1) GS(z[5], z[7], w1)
            GS(z[0], z[4], w0) GS(z[1], z[5], w0) GS(z[2], z[6], w0) GS(z[3], z[7], w0)
#pragma unroll
            for (int r = 0; r < 8; ++r) { z[r].x *= 0.125; z[r].y *= 0.125; }
        } else {
            // mode 5: 72 DFMA (as mode 0) + 16 I2F.F64.S32; mode 6: the 16 I2F alone; mode 7: 72 DFMA + 16 DADD conversions
            if (mode != 6) {
#pragma unroll
                for (int k2 = 0; k2 < 9; ++k2)
#pragma unroll
                    for (int r = 0; r < 8; ++r) x[r] = fma(x[r], c, e);
            }
            if (mode == 5 || mode == 6) {
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    iv[r] = iv[r] * 3 + it;
                    b[r] += (double)(iv[r] >> 20) + (double)((iv[r] << 10) >> 22);
                }
            } else {
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    iv[r] = iv[r] * 3 + it;
                    b[r] += (__hiloint2double(0x43300000, (iv[r] >> 20) & 1023) - 4503599627370496.0) +
                            (__hiloint2double(0x43300000, (iv[r] >> 10) & 1023) - 4503599627370496.0);
                }
            }
        }
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int r = 0; r < 8; ++r) s += x[r] + a[r] + b[r] + z[r].x + z[r].y + iv[r];
    if (s == 123.456) sink[0] = s;
    if (t == 0) cycles[blockIdx.x] = t1 - t0;
}
template <int THREADS>
static void run(int mode, const char *name, int nfp, double *sink, long long *cyc) {
    k<THREADS><<<148, THREADS>>>(mode, sink, cyc, 1.0);
    cudaDeviceSynchronize();
    k<THREADS><<<148, THREADS>>>(mode, sink, cyc, 1.0);
    cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0;
    for (int i = 0; i < 148; ++i) avg += h[i];
    const double clk = avg / 148 / ITERS;
    printf("%-52s thr %3d: %8.1f clk/iter per SM = %.2f clk per warp-level FP64 instr per scheduler\n", name, THREADS, clk,
           clk / (nfp * (THREADS / 32) / 4.0));
}
int main() {
    double *sink; long long *cyc;
    cudaMalloc(&sink, 8); cudaMalloc(&cyc, 148 * 8);
    double hk[8] = {0.999999999, 0.999999998, 0.999999997, 0.999999996, 0.999999995, 0.999999994, 0.999999993, 0.999999992};
    cudaMemcpyToSymbol(c_k, hk, sizeof(hk));
#define ALL(T) \
    run<T>(0, "72 DFMA x = fma(x, c, e)", 72, sink, cyc); run<T>(1, "72 DFMA x = fma(a_r, b_r, x), 3 distinct operands", 72, sink, cyc); \
    run<T>(2, "72 DFMA x = fma(a_r, c, x)", 72, sink, cyc); run<T>(3, "radix-8 pass: 72 DFMA + 16 DMUL", 88, sink, cyc); \
    run<T>(4, "72 DFMA x = fma(a_r, c[bank], x)", 72, sink, cyc); \
    run<T>(6, "16 I2F.F64.S32 + 16 DADD alone", 32, sink, cyc); run<T>(5, "72 DFMA + 16 I2F.F64.S32 + 16 DADD", 88, sink, cyc); \
    run<T>(7, "72 DFMA + 16 magic-number conversions + 24 DADD", 112, sink, cyc); \
    run<T>(8, "72 DADD x = x + e", 72, sink, cyc); run<T>(9, "72 DADD x = x + a_r", 72, sink, cyc); run<T>(10, "72 DMUL x = x * c", 72, sink, cyc); \
    run<T>(11, "inverse radix-8 pass: 48 DADD + 24 DFMA + 24+16 DMUL", 112, sink, cyc);
    ALL(128) ALL(256) ALL(384)
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
