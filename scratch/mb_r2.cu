// mb_r2.cu -- round-2 pipe probes for the blind rotation:
//   (1) FP64 tensor cores: mma.sync.m8n8k4.f64 (DMMA) rate alone, next to DFMA from the same warp / from other
//       warps, next to shared-memory traffic; and the spectral MAC as a per-frequency real contraction
//       [8 ciphertexts x 8] . [8 x 4]  (tlwe-fft-operations.cpp:92-100 restated as a batched tiny GEMM).
//   (2) tensor memory as an exchange medium: tcgen05.st 32x32b -> tcgen05.ld 16x256b is a cross-lane permutation
//       inside a warp; its lane map, its rate, and how it overlaps LDS/STS and DFMA.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scratch/mb_r2 scratch/mb_r2.cu
// run:   scratch/mb_r2 <test>   (dmma | tmem | map)
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

constexpr int ITERS = 1024;
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void dmma(double &d0, double &d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// ------------------------------------------------------------------------------------------------
// DMMA probes.  mode bits: 1 = 16 DMMA per warp-iter (8 independent accumulator chains x 2),
//   2 = 64 DFMA per thread-iter (8 chains), 4 = smem 8 STS.128 + 8 LDS.128, 8 = warp-split: even warps do the
//   DMMA part, odd warps the DFMA part (instead of every warp doing both)
// ------------------------------------------------------------------------------------------------
template <int THREADS>
__global__ void __launch_bounds__(THREADS, 1) k_dmma(int mode, double *sink, long long *cycles) {
    extern __shared__ __align__(16) unsigned char smem[];
    double2 *buf = reinterpret_cast<double2 *>(smem);
    const int t = threadIdx.x, warp = t >> 5;
    double d[16], x[8];
#pragma unroll
    for (int r = 0; r < 16; ++r) d[r] = t + r;
#pragma unroll
    for (int r = 0; r < 8; ++r) x[r] = t - r;
    double2 y[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) y[r] = make_double2(t + r, t - r);
    for (int i = t; i < 8 * THREADS + 64; i += THREADS) buf[i] = make_double2(i, -i);
    double a = 1.0 + 1e-9 * t, b = 1.0 - 1e-9 * t;
    const double c = 1.0 - 1e-9, e = 1e-9;
    const bool split = mode & 8;
    const bool do_dmma = (mode & 1) && (!split || !(warp & 1));
    const bool do_dfma = (mode & 2) && (!split || (warp & 1));
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
        if (do_dmma) {
#pragma unroll
            for (int k = 0; k < 2; ++k)
#pragma unroll
                for (int r = 0; r < 8; ++r) dmma(d[2 * r], d[2 * r + 1], a, b);
        }
        if (do_dfma) {
#pragma unroll
            for (int k = 0; k < 8; ++k)
#pragma unroll
                for (int r = 0; r < 8; ++r) x[r] = fma(x[r], c, e);
        }
        if (mode & 4) {
#pragma unroll
            for (int r = 0; r < 8; ++r) buf[t + THREADS * r] = y[r];
            __syncwarp();
#pragma unroll
            for (int r = 0; r < 8; ++r) y[r] = buf[(t ^ 1) + THREADS * r];
        }
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int r = 0; r < 16; ++r) s += d[r];
#pragma unroll
    for (int r = 0; r < 8; ++r) s += x[r] + y[r].x + y[r].y;
    if (s == 123.456) sink[0] = s;
    if (t == 0) cycles[blockIdx.x] = t1 - t0;
}

// spectral MAC of one CMux step for 8 ciphertexts, per frequency a real [8 ct x 8] . [8 x 4] contraction:
//   mode 0: CUDA cores as the kernel does it: a thread owns 8 frequencies of ONE ciphertext, per (row pair) it
//           reads b0, b1 (LDS.128) and issues 8 DFMA per frequency;   64 threads per ciphertext
//   mode 1: DMMA: a warp owns 8 ciphertexts x (its frequencies): per frequency and row pair the A fragment is
//           [8 ct x (re,im,re',im')] (k = 4: one row pair's complex value as 2 reals x ... ), B = key 4 x 8
// Only the instruction mix matters here (operands come from registers / shared memory as they would).
template <int THREADS>
__global__ void __launch_bounds__(THREADS, 1) k_mac(int mode, double *sink, long long *cycles) {
    extern __shared__ __align__(16) unsigned char smem[];
    double2 *key = reinterpret_cast<double2 *>(smem);   // one row pair: 2 x 512 double2
    const int t = threadIdx.x, lane = t & 31;
    for (int i = t; i < 1024; i += THREADS) key[i] = make_double2(1.0 + 1e-9 * i, 1.0 - 1e-9 * i);
    __syncthreads();
    double2 x[8], acc0[8], acc1[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) { x[r] = make_double2(t + r, t - r); acc0[r] = acc1[r] = make_double2(0, 0); }
    double d[16];
#pragma unroll
    for (int r = 0; r < 16; ++r) d[r] = 0;
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
        if (mode == 0) {
            // one row pair for this thread's 8 frequencies: 16 LDS.128 + 64 DFMA
            const double2 *row = key + (t & 63);
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const double2 b0 = row[r * 64], b1 = row[512 + r * 64];
                acc0[r].x = fma(x[r].x, b0.x, acc0[r].x); acc0[r].x = fma(-x[r].y, b0.y, acc0[r].x);
                acc0[r].y = fma(x[r].x, b0.y, acc0[r].y); acc0[r].y = fma(x[r].y, b0.x, acc0[r].y);
                acc1[r].x = fma(x[r].x, b1.x, acc1[r].x); acc1[r].x = fma(-x[r].y, b1.y, acc1[r].x);
                acc1[r].y = fma(x[r].x, b1.y, acc1[r].y); acc1[r].y = fma(x[r].y, b1.x, acc1[r].y);
            }
        } else {
            // per frequency f (8 per "thread slot" -> a warp covers 8 ct x 32 frequencies = the same 256 (ct,f) pairs
            // 32 threads x 8 cover): one row pair contributes k = 2 reals (re, im) of x against a 2 x 4 real block
            // [[b0r, b0i, b1r, b1i], [-b0i, b0r, -b1i, b1r]]; two row pairs fill k = 4.  Here: one row pair per
            // iteration like mode 0, so two frequencies share a DMMA (k = 4 = 2 freq x 2 reals is NOT a valid
            // contraction; instead n = 8 = 2 frequencies x 4 outputs with a block-diagonal B, half the MACs wasted).
            // A fragment: lane holds x[ct = lane/4][k = lane%4]; B fragment: lane holds B[k = lane%4][n = lane/4].
            // Per warp and row pair: 32 frequencies -> 16 DMMA (2 frequencies each), B from shared memory (LDS.64).
            const double *kb = reinterpret_cast<const double *>(key) + lane;
#pragma unroll
            for (int r = 0; r < 16; ++r) {
                const double bb = kb[r * 64];
                dmma(d[2 * (r & 7)], d[2 * (r & 7) + 1], x[r & 7].x, bb);
            }
        }
        x[0].x += 1.0;
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int r = 0; r < 8; ++r) s += acc0[r].x + acc0[r].y + acc1[r].x + acc1[r].y;
#pragma unroll
    for (int r = 0; r < 16; ++r) s += d[r];
    if (s == 123.456) sink[0] = s;
    if (t == 0) cycles[blockIdx.x] = t1 - t0;
}

// ------------------------------------------------------------------------------------------------
// TMEM probes
// ------------------------------------------------------------------------------------------------
#define R32(v) "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
#define W32(v) "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
#define L32 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}"
#define S32 "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32}"

__device__ __forceinline__ void tm_st32(uint32_t taddr, const uint32_t (&v)[32]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], " S32 ";" ::"r"(taddr), R32(v) : "memory");
}
__device__ __forceinline__ void tm_ld32(uint32_t (&v)[32], uint32_t taddr) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 " L32 ", [%32];" : W32(v) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tm_st256(uint32_t taddr, const uint32_t (&v)[32]) {   // 16 lanes x 64 columns
    asm volatile("tcgen05.st.sync.aligned.16x256b.x8.b32 [%0], " S32 ";" ::"r"(taddr), R32(v) : "memory");
}
__device__ __forceinline__ void tm_ld256(uint32_t (&v)[32], uint32_t taddr) {
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x8.b32 " L32 ", [%32];" : W32(v) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tm_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tm_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// mode: 1 = ld 32x32b.x32 (x4 in flight), 2 = ld 16x256b.x8 (two halves: lanes 0-15, 16-31; x2 in flight each),
//       3 = st 32x32b.x32 x4, 4 = st 16x256b.x8 x4, 5 = exchange: st 32x32b.x32 -> wait -> 2 x ld 16x256b.x4... as .x8 on
//       two lane halves reading half the columns each (the 4x4 lane transpose of 8 double2 per thread)
//   bit 16 = + smem exchange (8 STS.128 + 8 LDS.128), bit 32 = + 96 DFMA
template <int THREADS>
__global__ void __launch_bounds__(THREADS, 1) k_tmem(int mode, double *sink, long long *cycles) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ uint32_t tmem_base_s;
    double2 *buf = reinterpret_cast<double2 *>(smem);
    const int t = threadIdx.x, warp = t >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base_s)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // each warp: its lane quarter, a private 128-column window (warps sharing a quarter use different columns)
    const uint32_t taddr = tmem_base_s + (((uint32_t)(warp & 3) * 32u) << 16) + (uint32_t)(warp >> 2) * 128u;
    uint32_t v[32];
#pragma unroll
    for (int r = 0; r < 32; ++r) v[r] = t * 32 + r;
    double2 y[8];
    double x[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) { y[r] = make_double2(t + r, t - r); x[r] = t - r; }
    for (int i = t; i < 8 * THREADS + 64; i += THREADS) buf[i] = make_double2(i, -i);
    tm_st32(taddr, v); tm_st32(taddr + 32, v); tm_st32(taddr + 64, v); tm_st32(taddr + 96, v);
    tm_wait_st();
    __syncthreads();
    const double c = 1.0 - 1e-9, e = 1e-9;
    const int m = mode & 15;
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
        if (m == 1) {
            uint32_t w[32], w2[32];
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                tm_ld32(w, taddr + 64 * k); tm_ld32(w2, taddr + 64 * k + 32); tm_wait_ld();
                v[k] ^= w[k] ^ w[31] ^ w2[k] ^ w2[31];
            }
        } else if (m == 2) {
            uint32_t w[32], w2[32];
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                tm_ld256(w, taddr + 64 * k); tm_ld256(w2, taddr + 64 * k + (16u << 16)); tm_wait_ld();
                v[k] ^= w[k] ^ w[31] ^ w2[k] ^ w2[31];
            }
        } else if (m == 3) {
#pragma unroll
            for (int k = 0; k < 4; ++k) tm_st32(taddr + 32 * k, v);
            tm_wait_st();
            v[0] += 1;
        } else if (m == 4) {
#pragma unroll
            for (int k = 0; k < 2; ++k) { tm_st256(taddr + 64 * k, v); tm_st256(taddr + 64 * k + (16u << 16), v); }
            tm_wait_st();
            v[0] += 1;
        } else if (m == 5) {
            // the exchange of one transform pass: 8 double2 per thread out, 8 double2 per thread back, permuted
            tm_st32(taddr + 32 * (it & 1), v);
            tm_wait_st();
            uint32_t w0[16], w1[16];
            asm volatile("tcgen05.ld.sync.aligned.16x256b.x4.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                         : "=r"(w0[0]), "=r"(w0[1]), "=r"(w0[2]), "=r"(w0[3]), "=r"(w0[4]), "=r"(w0[5]), "=r"(w0[6]), "=r"(w0[7]), "=r"(w0[8]), "=r"(w0[9]), "=r"(w0[10]), "=r"(w0[11]), "=r"(w0[12]), "=r"(w0[13]), "=r"(w0[14]), "=r"(w0[15])
                         : "r"(taddr + 32 * (it & 1)) : "memory");
            asm volatile("tcgen05.ld.sync.aligned.16x256b.x4.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                         : "=r"(w1[0]), "=r"(w1[1]), "=r"(w1[2]), "=r"(w1[3]), "=r"(w1[4]), "=r"(w1[5]), "=r"(w1[6]), "=r"(w1[7]), "=r"(w1[8]), "=r"(w1[9]), "=r"(w1[10]), "=r"(w1[11]), "=r"(w1[12]), "=r"(w1[13]), "=r"(w1[14]), "=r"(w1[15])
                         : "r"(taddr + 32 * (it & 1) + (16u << 16)) : "memory");
            tm_wait_ld();
#pragma unroll
            for (int r = 0; r < 16; ++r) { v[r] = w0[r] + 1; v[16 + r] = w1[r] + 1; }
        }
        else if (m == 6) {
            // the same exchange with the narrowest forms: 16 x st 32x32b.x2 (one double each), 8 x ld 16x256b.x1
#pragma unroll
            for (int d = 0; d < 16; ++d)
                asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1,%2};" ::"r"(taddr + 2 * d), "r"(v[2 * d]), "r"(v[2 * d + 1]) : "memory");
            tm_wait_st();
#pragma unroll
            for (int q = 0; q < 8; ++q)
                asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0,%1,%2,%3}, [%4];"
                             : "=r"(v[4 * q]), "=r"(v[4 * q + 1]), "=r"(v[4 * q + 2]), "=r"(v[4 * q + 3])
                             : "r"(taddr + 8 * (q & 3) + ((uint32_t)(16 * (q >> 2)) << 16)) : "memory");
            tm_wait_ld();
#pragma unroll
            for (int r = 0; r < 32; ++r) v[r] += 1;
        } else if (m == 7) {
            // st 32x32b.x4 x 8 (a double2 each), ld 16x256b.x2 x 4
#pragma unroll
            for (int d = 0; d < 8; ++d)
                asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%2,%3,%4};" ::"r"(taddr + 4 * d), "r"(v[4 * d]), "r"(v[4 * d + 1]), "r"(v[4 * d + 2]), "r"(v[4 * d + 3]) : "memory");
            tm_wait_st();
#pragma unroll
            for (int q = 0; q < 4; ++q)
                asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                             : "=r"(v[8 * q]), "=r"(v[8 * q + 1]), "=r"(v[8 * q + 2]), "=r"(v[8 * q + 3]), "=r"(v[8 * q + 4]), "=r"(v[8 * q + 5]), "=r"(v[8 * q + 6]), "=r"(v[8 * q + 7])
                             : "r"(taddr + 16 * (q & 1) + ((uint32_t)(16 * (q >> 1)) << 16)) : "memory");
            tm_wait_ld();
#pragma unroll
            for (int r = 0; r < 32; ++r) v[r] += 1;
        }
        if (mode & 16) {
#pragma unroll
            for (int r = 0; r < 8; ++r) buf[t + THREADS * r] = y[r];
            __syncwarp();
#pragma unroll
            for (int r = 0; r < 8; ++r) y[r] = buf[(t ^ 1) + THREADS * r];
        }
        if (mode & 32) {
#pragma unroll
            for (int k = 0; k < 12; ++k)
#pragma unroll
                for (int r = 0; r < 8; ++r) x[r] = fma(x[r], c, e);
        }
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int r = 0; r < 32; ++r) s += v[r];
#pragma unroll
    for (int r = 0; r < 8; ++r) s += x[r] + y[r].x + y[r].y;
    if (s == 123.456) sink[0] = s;
    if (t == 0) cycles[blockIdx.x] = t1 - t0;
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base_s) : "memory");
}

// lane map of st 32x32b -> ld 16x256b: every thread stores v[c] = lane * 100 + c in columns c = 0..31; one warp
// reads with 16x256b.x1 (4 registers) at lane offset 0 and 16 and column offsets 0 and 8 and prints what it got.
__global__ void __launch_bounds__(128, 1) k_map(uint32_t *out) {
    __shared__ uint32_t tmem_base_s;
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(&tmem_base_s)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t taddr = tmem_base_s + (((uint32_t)warp * 32u) << 16);
    uint32_t v[32];
#pragma unroll
    for (int c = 0; c < 32; ++c) v[c] = (uint32_t)(lane * 100 + c);
    tm_st32(taddr, v);
    tm_wait_st();
    __syncwarp();
    if (warp == 1) {   // a warp other than 0, to see that the lane quarter rule holds with sub-offsets
        for (int half = 0; half < 2; ++half)
            for (int cb = 0; cb < 2; ++cb) {
                uint32_t r0, r1, r2, r3;
                asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0,%1,%2,%3}, [%4];"
                             : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(taddr + ((uint32_t)(16 * half) << 16) + 8 * cb) : "memory");
                tm_wait_ld();
                uint32_t *o = out + ((half * 2 + cb) * 32 + lane) * 4;
                o[0] = r0; o[1] = r1; o[2] = r2; o[3] = r3;
            }
        // and the reverse: st 16x256b.x1 -> ld 32x32b.x8
        uint32_t a0 = lane * 100 + 0, a1 = lane * 100 + 1, a2 = lane * 100 + 2, a3 = lane * 100 + 3;
        asm volatile("tcgen05.st.sync.aligned.16x256b.x1.b32 [%0], {%1,%2,%3,%4};" ::"r"(taddr + 32), "r"(a0), "r"(a1), "r"(a2), "r"(a3) : "memory");
        tm_wait_st();
        uint32_t q[8];
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(q[0]), "=r"(q[1]), "=r"(q[2]), "=r"(q[3]), "=r"(q[4]), "=r"(q[5]), "=r"(q[6]), "=r"(q[7]) : "r"(taddr + 32) : "memory");
        tm_wait_ld();
        for (int c = 0; c < 8; ++c) out[512 + lane * 8 + c] = q[c];
    }
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tmem_base_s) : "memory");
}

template <class K>
static int run(K kern, int threads, int mode, int smem, const char *name, double *sink, long long *cyc) {
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    kern<<<148, threads, smem>>>(mode, sink, cyc);
    CK(cudaDeviceSynchronize());
    kern<<<148, threads, smem>>>(mode, sink, cyc);
    CK(cudaDeviceSynchronize());
    long long h[148];
    CK(cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost));
    double avg = 0;
    for (int i = 0; i < 148; ++i) avg += h[i];
    avg /= 148;
    printf("%-58s thr %3d mode %2d: %8.1f clk/iter per SM\n", name, threads, mode, avg / ITERS);
    fflush(stdout);
    return 0;
}

int main(int argc, char **argv) {
    const char *what = argc > 1 ? argv[1] : "dmma";
    double *sink;
    long long *cyc;
    CK(cudaMalloc(&sink, 8));
    CK(cudaMalloc(&cyc, 148 * 8));
    if (!strcmp(what, "dmma")) {
        for (int thr : {256, 384, 512}) {
            const int smem = (8 * thr + 64) * 16;
#define RUN(T) \
            run(k_dmma<T>, T, 1, smem, "16 DMMA m8n8k4 / warp", sink, cyc); \
            run(k_dmma<T>, T, 2, smem, "64 DFMA / thread", sink, cyc); \
            run(k_dmma<T>, T, 3, smem, "16 DMMA + 64 DFMA, same warps", sink, cyc); \
            run(k_dmma<T>, T, 11, smem, "16 DMMA (even warps) | 64 DFMA (odd warps)", sink, cyc); \
            run(k_dmma<T>, T, 4, smem, "smem 8 STS.128 + 8 LDS.128", sink, cyc); \
            run(k_dmma<T>, T, 5, smem, "16 DMMA + smem", sink, cyc); \
            run(k_dmma<T>, T, 6, smem, "64 DFMA + smem", sink, cyc); \
            run(k_dmma<T>, T, 7, smem, "16 DMMA + 64 DFMA + smem", sink, cyc);
            if (thr == 256) { RUN(256) } else if (thr == 384) { RUN(384) } else { RUN(512) }
#undef RUN
        }
        run(k_mac<384>, 384, 0, 16384, "MAC row pair, CUDA cores: 16 LDS.128 + 64 DFMA / thread", sink, cyc);
        run(k_mac<384>, 384, 1, 16384, "MAC row pair as contraction: 16 LDS.64 + 16 DMMA / warp", sink, cyc);
        run(k_mac<512>, 512, 0, 16384, "MAC row pair, CUDA cores: 16 LDS.128 + 64 DFMA / thread", sink, cyc);
        run(k_mac<512>, 512, 1, 16384, "MAC row pair as contraction: 16 LDS.64 + 16 DMMA / warp", sink, cyc);
    } else if (!strcmp(what, "tmem")) {
        for (int thr : {256, 384}) {
            const int smem = (8 * thr + 64) * 16;
#define RUN(T) \
            run(k_tmem<T>, T, 1, smem, "TMEM ld 32x32b.x32 x4, two in flight (512 B/thread)", sink, cyc); \
            run(k_tmem<T>, T, 2, smem, "TMEM ld 16x256b.x8 x4, two in flight (512 B/thread)", sink, cyc); \
            run(k_tmem<T>, T, 3, smem, "TMEM st 32x32b.x32 x4 (512 B/thread)", sink, cyc); \
            run(k_tmem<T>, T, 4, smem, "TMEM st 16x256b.x8 x4 (512 B/thread)", sink, cyc); \
            run(k_tmem<T>, T, 5, smem, "TMEM exchange st 32x32b.x32 -> 2 ld 16x256b.x4 (128 B each way)", sink, cyc); \
            run(k_tmem<T>, T, 6, smem, "TMEM exchange 16 st 32x32b.x2 -> 8 ld 16x256b.x1", sink, cyc); \
            run(k_tmem<T>, T, 7, smem, "TMEM exchange 8 st 32x32b.x4 -> 4 ld 16x256b.x2", sink, cyc); \
            run(k_tmem<T>, T, 6 + 32, smem, "TMEM exchange (narrow) + 96 DFMA", sink, cyc); \
            run(k_tmem<T>, T, 16, smem, "smem exchange 8 STS.128 + 8 LDS.128 (128 B each way)", sink, cyc); \
            run(k_tmem<T>, T, 32, smem, "96 DFMA", sink, cyc); \
            run(k_tmem<T>, T, 5 + 16, smem, "TMEM exchange + smem exchange", sink, cyc); \
            run(k_tmem<T>, T, 5 + 32, smem, "TMEM exchange + 96 DFMA", sink, cyc); \
            run(k_tmem<T>, T, 16 + 32, smem, "smem exchange + 96 DFMA", sink, cyc); \
            run(k_tmem<T>, T, 5 + 16 + 32, smem, "TMEM exchange + smem exchange + 96 DFMA", sink, cyc);
            if (thr == 256) { RUN(256) } else { RUN(384) }
#undef RUN
        }
    } else {
        uint32_t *out;
        CK(cudaMalloc(&out, 1024 * 4));
        CK(cudaMemset(out, 0xff, 1024 * 4));
        k_map<<<1, 128>>>(out);
        CK(cudaDeviceSynchronize());
        uint32_t h[1024];
        CK(cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost));
        printf("st 32x32b (value = lane*100 + column) -> ld 16x256b.x1: thread: reg0 reg1 reg2 reg3\n");
        for (int half = 0; half < 2; ++half)
            for (int cb = 0; cb < 2; ++cb) {
                printf(" lane offset %d, column offset %d\n", 16 * half, 8 * cb);
                for (int l = 0; l < 32; ++l) {
                    const uint32_t *o = h + ((half * 2 + cb) * 32 + l) * 4;
                    printf("  T%02d: %4u %4u %4u %4u%s", l, o[0], o[1], o[2], o[3], (l & 3) == 3 ? "\n" : "  |");
                }
            }
        printf("st 16x256b.x1 (value = thread*100 + reg) -> ld 32x32b.x8: lane: columns 0..7\n");
        for (int l = 0; l < 32; ++l) {
            printf("  L%02d:", l);
            for (int c = 0; c < 8; ++c) printf(" %10d", (int)h[512 + l * 8 + c]);
            printf("\n");
        }
    }
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
