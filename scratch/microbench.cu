// microbench.cu -- B200 pipe-rate probes that decide the blind-rotation kernel layout:
//   smem LDS/STS.128 wavefront rate, SHFL rate, TMEM (tcgen05.ld/st 32x32b) rate, DFMA rate, and how
//   these overlap when issued from the same warps; LDS.128 broadcast merging between half warps.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scratch/microbench scratch/microbench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

constexpr int THREADS = 256;
constexpr int ITERS = 2048;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// mode bits: 1 = smem (8 STS.128 + 8 LDS.128 per iter), 2 = DFMA (96 per iter), 4 = SHFL (32 per iter),
//            8 = TMEM (st x16 + ld x16 = 64 B each way per thread per iter), 16 = TMEM ld only x32 (128 B)
//            32 = LDS.128 broadcast pattern l / l+16, 64 = LDS.128 broadcast pattern l / l^1
__global__ void __launch_bounds__(THREADS, 1) k_probe(int mode, double *sink, long long *cycles) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ uint32_t tmem_base_s;
    double2 *buf = reinterpret_cast<double2 *>(smem);
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    if (mode & (8 | 16)) {
        if (warp == 0) {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base_s)) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
    const uint32_t taddr = (mode & (8 | 16)) ? tmem_base_s + (((uint32_t)(warp & 3) * 32u) << 16) + (uint32_t)(warp >> 2) * 256u : 0;
    double2 x[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) x[r] = make_double2(t + r, t - r);
    for (int i = t; i < 8 * THREADS + 64; i += THREADS) buf[i] = make_double2(i, -i);
    uint32_t v[32];
#pragma unroll
    for (int r = 0; r < 32; ++r) v[r] = t * 32 + r;
    if (mode & (8 | 16)) {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
                     ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
                       "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31]) : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    __syncthreads();
    const double c = 1.0 - 1e-9, d = 1e-9;
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
        if (mode & 1) {
#pragma unroll
            for (int r = 0; r < 8; ++r) buf[t + THREADS * r] = x[r];
            __syncwarp();
#pragma unroll
            for (int r = 0; r < 8; ++r) x[r] = buf[(t ^ 1) + THREADS * r];
        }
        if (mode & 32) {
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const double2 y = buf[warp * 16 + (lane & 15) + THREADS * r];
                x[r].x += y.x; x[r].y += y.y;
            }
        }
        if (mode & 64) {
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const double2 y = buf[warp * 16 + (lane >> 1) + THREADS * r];
                x[r].x += y.x; x[r].y += y.y;
            }
        }
        if (mode & 2) {
#pragma unroll
            for (int k = 0; k < 6; ++k)
#pragma unroll
                for (int r = 0; r < 8; ++r) { x[r].x = fma(x[r].x, c, d); x[r].y = fma(x[r].y, c, d); }
        }
        if (mode & 4) {
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                int lo = __double2loint(x[r].x), hi = __double2hiint(x[r].x);
                lo = __shfl_xor_sync(0xffffffffu, lo, 1); hi = __shfl_xor_sync(0xffffffffu, hi, 1);
                x[r].x = __hiloint2double(hi, lo);
                lo = __double2loint(x[r].y); hi = __double2hiint(x[r].y);
                lo = __shfl_xor_sync(0xffffffffu, lo, 2); hi = __shfl_xor_sync(0xffffffffu, hi, 2);
                x[r].y = __hiloint2double(hi, lo);
            }
        }
        if (mode & 8) {
            asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
                         ::"r"(taddr + 32 * (it & 1)), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                         : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                         : "r"(taddr + 16 * (it & 3)) : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int r = 0; r < 16; ++r) v[r] += 1;
        }
        if (mode & 16) {
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                         : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                           "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                         : "r"(taddr + 32 * (it & 3)) : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            x[0].x += (double)(v[0] ^ v[17] ^ v[31]);
        }
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int r = 0; r < 8; ++r) s += x[r].x + x[r].y;
#pragma unroll
    for (int r = 0; r < 32; ++r) s += v[r];
    if (s == 123.456) sink[0] = s;
    if (t == 0) cycles[blockIdx.x] = t1 - t0;
    if (mode & (8 | 16)) {
        __syncthreads();
        if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base_s) : "memory");
    }
}

int main() {
    double *sink;
    long long *cyc;
    CK(cudaMalloc(&sink, 8));
    CK(cudaMalloc(&cyc, 148 * 8));
    const int smem = (8 * THREADS + 64) * 16;
    CK(cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    struct { int mode; const char *name; } tests[] = {
        {1, "smem 8xSTS.128 + 8xLDS.128"}, {2, "96 DFMA"}, {4, "32 SHFL"}, {8, "TMEM st.x16 + ld.x16 (+waits)"},
        {16, "TMEM ld.x32"}, {32, "8xLDS.128 bcast l/l+16"}, {64, "8xLDS.128 bcast l/l^1"},
        {1 | 2, "smem + DFMA"}, {1 | 4, "smem + SHFL"}, {1 | 8, "smem + TMEM st/ld"}, {1 | 16, "smem + TMEM ld.x32"},
        {2 | 8, "DFMA + TMEM st/ld"}, {1 | 2 | 8, "smem + DFMA + TMEM st/ld"}, {2 | 4, "DFMA + SHFL"},
    };
    for (auto &tc : tests) {
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        k_probe<<<148, THREADS, smem>>>(tc.mode, sink, cyc);
        CK(cudaDeviceSynchronize());
        cudaEventRecord(e0);
        k_probe<<<148, THREADS, smem>>>(tc.mode, sink, cyc);
        cudaEventRecord(e1);
        CK(cudaDeviceSynchronize());
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        long long h[148];
        CK(cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost));
        double avg = 0;
        for (int i = 0; i < 148; ++i) avg += h[i];
        avg /= 148;
        printf("%-36s mode %3d: %8.1f clk/iter per SM (256 thr)   %.3f ms\n", tc.name, tc.mode, avg / ITERS, ms);
    }
    return 0;
}
