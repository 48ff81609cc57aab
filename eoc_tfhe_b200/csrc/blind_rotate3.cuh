// blind_rotate3.cuh -- blind rotation "v3" (included by kernels.cu after blind_rotate16.cuh).
//
// Same transform and thread layout as k_blind_rotate (64 threads x 8 points per polynomial, two
// shared-memory exchanges per transform), re-cut around the two measured limits of that kernel
// (profiles/r1_v2_ncu.txt: shared-memory data pipe 72 %, FP64 51 %, 8 warps per SM at 248 registers):
//   * the two evaluation-domain accumulators of the external product live in TENSOR MEMORY
//     (tcgen05.st/ld, 64 columns per ciphertext and lane) instead of 64 registers per thread, so that
//   * C = 2 ciphertexts can share one 64-thread group: their transforms are interleaved instruction by
//     instruction (one ciphertext computes while the other's exchange is in flight) and every
//     bootstrapping-key element read from the shared-memory ring is used for both (key wavefronts per
//     ciphertext halve), or
//   * C = 1 runs G = 6 groups (12 warps) per SM inside a 168-register budget.
// Arithmetic is instruction-for-instruction that of cmux_step (same FMA order), so outputs are
// bit-identical to k_blind_rotate; tests compare the two kernels exactly.
//
// Reference semantics: lwe-bootstrapping-functions-fft.cpp:67-112 (CMux chain),
// tgsw-fft-operations.cpp:89-114 (external product), tgsw-functions.cpp:298-348 (decomposition).
#pragma once

namespace b200 {

constexpr int kTm3ColsPerCt = 64;   // acc0: 8 double2 = 32 columns, acc1: the next 32

// forward transforms of C polynomials held by one 64-thread group
template <int C, class Barrier>
__device__ __forceinline__ void forward3(double2 (&x)[C][8], double2 *bufs, int &toggle, const int t,
                                         const RootSet &w2, const RootSet &w3, const Barrier &bar) {
    if (C == 1) {
        negacyclic_forward(x[0], bufs + toggle * kBufElems, t, w2, w3, bar);
        toggle ^= 1;
        return;
    }
#pragma unroll
    for (int c = 0; c < C; ++c) {
        fwd_radix8(x[c], c_tw.p1);
        double2 *buf = bufs + c * kBufElems;
#pragma unroll
        for (int r = 0; r < 8; ++r) buf[pad_idx(t + 64 * r)] = x[c][r];
    }
    bar();
    const int base2 = pad_idx(64 * (t >> 3) + (t & 7));
#pragma unroll
    for (int c = 0; c < C; ++c)
#pragma unroll
        for (int r = 0; r < 8; ++r) x[c][r] = bufs[c * kBufElems + base2 + 9 * r];
#pragma unroll
    for (int c = 0; c < C; ++c) {
        fwd_radix8(x[c], w2);
#pragma unroll
        for (int r = 0; r < 8; ++r) bufs[c * kBufElems + base2 + 9 * r] = x[c][r];
    }
    __syncwarp();  // octet-local exchange (negacyclic_fft.cuh)
#pragma unroll
    for (int c = 0; c < C; ++c)
#pragma unroll
        for (int r = 0; r < 8; ++r) x[c][r] = bufs[c * kBufElems + 9 * t + r];
    bar();  // every thread has its last-pass inputs: the buffers may be overwritten by the next transforms
#pragma unroll
    for (int c = 0; c < C; ++c) fwd_radix8(x[c], w3);
}

template <int C, class Barrier>
__device__ __forceinline__ void inverse3(double2 (&x)[C][8], double2 *bufs, int &toggle, const int t,
                                         const RootSet &w2, const RootSet &w3, const Barrier &bar) {
    if (C == 1) {
        negacyclic_inverse(x[0], bufs + toggle * kBufElems, t, w2, w3, bar);
        toggle ^= 1;
        return;
    }
#pragma unroll
    for (int c = 0; c < C; ++c) {
        inv_radix8(x[c], w3);
#pragma unroll
        for (int r = 0; r < 8; ++r) bufs[c * kBufElems + 9 * t + r] = x[c][r];
    }
    __syncwarp();  // octet-local exchange
    const int base2 = pad_idx(64 * (t >> 3) + (t & 7));
#pragma unroll
    for (int c = 0; c < C; ++c)
#pragma unroll
        for (int r = 0; r < 8; ++r) x[c][r] = bufs[c * kBufElems + base2 + 9 * r];
#pragma unroll
    for (int c = 0; c < C; ++c) {
        inv_radix8(x[c], w2);
#pragma unroll
        for (int r = 0; r < 8; ++r) bufs[c * kBufElems + base2 + 9 * r] = x[c][r];
    }
    bar();
#pragma unroll
    for (int c = 0; c < C; ++c)
#pragma unroll
        for (int r = 0; r < 8; ++r) x[c][r] = bufs[c * kBufElems + pad_idx(t + 64 * r)];
    bar();
#pragma unroll
    for (int c = 0; c < C; ++c) inv_radix8(x[c], c_tw.p1);
}

// ------------------------------------------------------------------------------------------------
// one CMux step  ACC_c <- ACC_c + BK_i (x) ((X^{a_c} - 1) ACC_c)  for the C ciphertexts of a group.
// a_c == 0 needs no special case: the rotated difference is 0, every digit is exactly 0 and the
// accumulator is left unchanged (the reference skips the step, lwe-bootstrapping-functions-fft.cpp:101).
// ------------------------------------------------------------------------------------------------
template <int L, int BGBIT, int C, class Barrier, class Rows>
__device__ __forceinline__ void cmux_step3(int32_t *ACC, double2 *bufs, const int (&a)[C], const Rows &rows,
                                           const int t, const RootSet &w2, const RootSet &w3, const Barrier &bar,
                                           const uint32_t tm) {
    constexpr uint32_t BG = 1u << BGBIT, HALF = BG >> 1;
    constexpr int CH = (C == 1) ? 4 : 2;   // frequencies per MAC chunk (bounds the registers in flight)
    uint32_t offset = 0;  // tgsw.cpp:7-29
#pragma unroll
    for (int p = 0; p < L; ++p) offset += HALF << (32 - (p + 1) * BGBIT);

    int toggle = 0;
#pragma unroll 1
    for (int q = 0; q < 2; ++q) {
        uint32_t dv[C][16];  // (rotated difference) + decomposition offset, coefficient t + 64 rr
#pragma unroll
        for (int c = 0; c < C; ++c) {
            const int32_t *P = ACC + c * 2 * kN + q * kN;
#pragma unroll
            for (int rr = 0; rr < 16; ++rr) {
                const int j = t + 64 * rr;
                const int m = (j - a[c]) & (2 * kN - 1);
                int32_t v = P[m & (kN - 1)];
                if (m & kN) v = -v;
                dv[c][rr] = (uint32_t)(v - P[j]) + offset;
            }
        }
#pragma unroll 1
        for (int p = 0; p < L; ++p) {
            const int shift = 32 - (p + 1) * BGBIT;
            double2 x[C][8];
#pragma unroll
            for (int c = 0; c < C; ++c)
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    x[c][r] = make_double2(digit_to_double<BGBIT>(dv[c][r], shift), digit_to_double<BGBIT>(dv[c][r + 8], shift));
                }
            forward3<C>(x, bufs, toggle, t, w2, w3, bar);

            const double2 *row = rows.acquire(q * L + p) + t;
            const bool first = (q == 0 && p == 0);
            tmem_wait_st();  // the previous row's accumulator stores have landed
#pragma unroll
            for (int r0 = 0; r0 < 8; r0 += CH) {
                double2 a0[C][CH], a1[C][CH], b0[CH], b1[CH];
                if (!first) {
#pragma unroll
                    for (int c = 0; c < C; ++c)
#pragma unroll
                        for (int k = 0; k < CH; ++k) {
                            tmem_ld1(tm + c * kTm3ColsPerCt + 4 * (r0 + k), a0[c][k]);
                            tmem_ld1(tm + c * kTm3ColsPerCt + 32 + 4 * (r0 + k), a1[c][k]);
                        }
                } else {
#pragma unroll
                    for (int c = 0; c < C; ++c)
#pragma unroll
                        for (int k = 0; k < CH; ++k) a0[c][k] = a1[c][k] = make_double2(0.0, 0.0);
                }
#pragma unroll
                for (int k = 0; k < CH; ++k) {
                    b0[k] = rows.load(row + (r0 + k) * 64);
                    b1[k] = rows.load(row + kM + (r0 + k) * 64);
                }
                if (!first) tmem_wait_ld();
#pragma unroll
                for (int c = 0; c < C; ++c)
#pragma unroll
                    for (int k = 0; k < CH; ++k) {
                        const double2 xv = x[c][r0 + k];
                        a0[c][k].x = fma(xv.x, b0[k].x, a0[c][k].x);
                        a0[c][k].x = fma(-xv.y, b0[k].y, a0[c][k].x);
                        a0[c][k].y = fma(xv.x, b0[k].y, a0[c][k].y);
                        a0[c][k].y = fma(xv.y, b0[k].x, a0[c][k].y);
                        a1[c][k].x = fma(xv.x, b1[k].x, a1[c][k].x);
                        a1[c][k].x = fma(-xv.y, b1[k].y, a1[c][k].x);
                        a1[c][k].y = fma(xv.x, b1[k].y, a1[c][k].y);
                        a1[c][k].y = fma(xv.y, b1[k].x, a1[c][k].y);
                    }
#pragma unroll
                for (int c = 0; c < C; ++c)
#pragma unroll
                    for (int k = 0; k < CH; ++k) {
                        tmem_st1(tm + c * kTm3ColsPerCt + 4 * (r0 + k), a0[c][k]);
                        tmem_st1(tm + c * kTm3ColsPerCt + 32 + 4 * (r0 + k), a1[c][k]);
                    }
            }
            rows.release(q * L + p);
        }
    }

    constexpr double kScale = 1.0 / kM;
    tmem_wait_st();
#pragma unroll 1
    for (int col = 0; col < 2; ++col) {
        double2 x[C][8];
#pragma unroll
        for (int c = 0; c < C; ++c)
#pragma unroll
            for (int r = 0; r < 8; ++r) tmem_ld1(tm + c * kTm3ColsPerCt + 32 * col + 4 * r, x[c][r]);
        tmem_wait_ld();
        inverse3<C>(x, bufs, toggle, t, w2, w3, bar);
#pragma unroll
        for (int c = 0; c < C; ++c) {
            int32_t *P = ACC + c * 2 * kN + col * kN;
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const int j = t + 64 * r;
                P[j] += round_scaled_to_torus(x[c][r].x, kScale);
                P[j + kM] += round_scaled_to_torus(x[c][r].y, kScale);
            }
        }
    }
    bar();  // accumulators complete before the next step's rotated reads
}

// per group: C accumulators + 2 exchange buffers + C bara arrays
template <int C> __host__ __device__ constexpr int group3_smem_bytes() { return C * 2 * kN * 4 + 2 * kBufElems * 16 + C * 1280; }
template <int G, int C, int STAGES> __host__ __device__ constexpr int br3_smem_bytes() { return STAGES * kStageBytes + 128 + G * group3_smem_bytes<C>(); }
template <int G, int C> __host__ __device__ constexpr int br3_tmem_cols() {
    const int need = ((2 * G + 3) / 4) * kTm3ColsPerCt * C;
    return need <= 32 ? 32 : need <= 64 ? 64 : need <= 128 ? 128 : need <= 256 ? 256 : 512;
}

template <int L, int BGBIT, int G, int C, int STAGES>
__global__ void __launch_bounds__(64 * G, 1) k_blind_rotate3(const DeviceKey key, const BlindRotateArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int KPL = 2 * L;
    constexpr int TM_COLS = br3_tmem_cols<G, C>();
    unsigned char *ring = smem;
    RingState<STAGES> *st = reinterpret_cast<RingState<STAGES> *>(smem + STAGES * kStageBytes);
    uint32_t *tm_slot = reinterpret_cast<uint32_t *>(smem + STAGES * kStageBytes + 96);
    const int g = threadIdx.x >> 6, t = threadIdx.x & 63, warp = threadIdx.x >> 5;
    unsigned char *mine = smem + STAGES * kStageBytes + 128 + (size_t)g * group3_smem_bytes<C>();
    int32_t *ACC = reinterpret_cast<int32_t *>(mine);                                        // [C][2][1024]
    double2 *bufs = reinterpret_cast<double2 *>(mine + C * 2 * kN * 4);                      // [2][576]
    uint16_t *bara = reinterpret_cast<uint16_t *>(mine + C * 2 * kN * 4 + 2 * kBufElems * 16);  // [C][640]

    const int n = key.n, n1 = n + 1;
    const int64_t item0 = ((int64_t)blockIdx.x * G + g) * C;
    const int total = n * KPL;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(&st->full[s], 1); st->done[s] = 0; }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        for (int s = 0; s < STAGES && s < total; ++s)
            tma_load_row_pair(ring + (size_t)s * kStageBytes, key.bkfft + (size_t)s * 2 * kM, &st->full[s]);
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tm_slot)), "n"(TM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }

    // ---- gate prologue + modulus switch to 2N (numeric-functions.cpp:59-65)
#pragma unroll
    for (int c = 0; c < C; ++c) {
        const int64_t item = item0 + c;
        if (item < a.items) {
            const int32_t *r0, *r1;
            int s0, s1;
            int32_t cst;
            if (a.desc) {
                const RotDesc d = a.desc[item];
                r0 = a.pool + (size_t)d.w0 * n1;
                r1 = d.w1 >= 0 ? a.pool + (size_t)d.w1 * n1 : r0;
                s0 = (int16_t)(d.s0s1 & 0xffff);
                s1 = d.w1 >= 0 ? (int16_t)(d.s0s1 >> 16) : 0;
                cst = d.cst;
            } else {
                int second;
                const int64_t gate = (a.op == B200_MUX) ? (item >> 1) : item;
                gate_coeffs(a.op, (int)(item & 1), s0, s1, cst, second);
                r0 = a.in0 + (size_t)gate * n1;
                r1 = s1 ? ((second == 2 ? a.in2 : a.in1) + (size_t)gate * n1) : r0;
            }
            for (int i = t; i < n1; i += 64) {
                uint32_t x = (uint32_t)s0 * (uint32_t)r0[i] + (uint32_t)s1 * (uint32_t)r1[i];
                if (i == n) x += (uint32_t)cst;
                bara[c * 640 + i] = (uint16_t)((x + 0x00100000u) >> 21);
            }
        } else {
            for (int i = t; i < n1; i += 64) bara[c * 640 + i] = 0;
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();  // bara visible; barriers initialised; TMEM base published
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = *tm_slot + (((uint32_t)(warp & 3) * 32u) << 16) + (uint32_t)(warp >> 2) * (kTm3ColsPerCt * C);

    // ---- ACC = (0, X^{2N-barb} * mu*(1+X+...+X^{N-1}))   (lwe-bootstrapping-functions-fft.cpp:146-148)
#pragma unroll
    for (int c = 0; c < C; ++c) {
        const int barb = bara[c * 640 + n];
        const int32_t mu = a.mu;
#pragma unroll
        for (int rr = 0; rr < 16; ++rr) {
            const int j = t + 64 * rr;
            ACC[c * 2 * kN + j] = 0;
            ACC[c * 2 * kN + kN + j] = ((j + barb) & kN) ? -mu : mu;
        }
    }
    const RootSet w2 = load_rootset(c_tw.p2[t >> 3]), w3 = load_rootset(c_tw.p3[t]);
    const NamedBarrier64 bar{1 + g};
    bar();

    // ---- blind rotation (lwe-bootstrapping-functions-fft.cpp:88-112)
    RingRows<STAGES> rows{ring, st, key.bkfft, 0, total, 2 * G, (threadIdx.x & 31) == 0};
#pragma unroll 1
    for (int i = 0; i < n; ++i) {
        rows.seq0 = i * KPL;
        int ai[C];
        int any = 0;
#pragma unroll
        for (int c = 0; c < C; ++c) { ai[c] = bara[c * 640 + i]; any |= ai[c]; }
        if (any == 0) { rows.skip_step(KPL); continue; }
        cmux_step3<L, BGBIT, C>(ACC, bufs, ai, rows, t, w2, w3, bar, tm);
    }

    // ---- sample extract at index 0 (lwe.cpp:41-60)
#pragma unroll
    for (int c = 0; c < C; ++c) {
        const int64_t item = item0 + c;
        if (item < a.items) {
            const int32_t *P = ACC + c * 2 * kN;
            int32_t *u = a.u + (size_t)item * (kN + 1);
#pragma unroll
            for (int rr = 0; rr < 16; ++rr) {
                const int j = t + 64 * rr;
                u[j] = (j == 0) ? P[0] : -P[kN - j];
            }
            if (t == 0) u[kN] = P[kN];
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(*tm_slot), "n"(TM_COLS) : "memory");
}

template <int L, int BGBIT, int G, int C, int STAGES>
static cudaError_t br3_launch(const DeviceKey &key, const BlindRotateArgs &a, cudaStream_t s) {
    static bool done[16] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    constexpr int smem = br3_smem_bytes<G, C, STAGES>();
    static_assert(smem <= 227 * 1024, "shared memory budget");
    if (!(dev < 16 && done[dev])) {
        cudaError_t e = cudaFuncSetAttribute(k_blind_rotate3<L, BGBIT, G, C, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        cudaFuncSetAttribute(k_blind_rotate3<L, BGBIT, G, C, STAGES>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        if (dev < 16) done[dev] = true;
    }
    const unsigned grid = (unsigned)((a.items + G * C - 1) / (G * C));
    k_blind_rotate3<L, BGBIT, G, C, STAGES><<<grid, 64 * G, smem, s>>>(key, a);
    return cudaSuccess;
}

}  // namespace b200
