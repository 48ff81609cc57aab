// blind_rotate16.cuh -- the warp-per-ciphertext blind rotation (included by kernels.cu).
//
// One warp owns one ciphertext: its TLWE accumulator (int32, shared memory), one exchange buffer, and
// 256 columns of tensor memory (TMEM) that hold what does not fit in registers next to a 16-point
// transform: the two evaluation-domain accumulators of the external product and the level-0 outputs of
// the second/third gadget digit.  TMEM is private per lane, is read and written with tcgen05.ld/st
// (LDTM/STTM) and has its own data path, so this traffic does not compete with the shared-memory
// pipe that bounds the kernel (scratch/microbench.cu: smem+TMEM runs at the smem-only rate).
//
// Reference semantics: lwe-bootstrapping-functions-fft.cpp:67-112 (CMux chain),
// tgsw-fft-operations.cpp:89-114 (external product), tgsw-functions.cpp:298-348 (decomposition).
#pragma once

namespace b200 {

// ---- tensor memory ------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t *slot_in_smem, int ncols) {
    // one warp; the base address lands in shared memory
    if (ncols == 512)
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(slot_in_smem)) : "memory");
    else
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(slot_in_smem)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_free(uint32_t base, int ncols) {
    if (ncols == 512) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(base) : "memory");
    else              asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(base) : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// one complex double <-> 4 columns of the calling lane.  (The .x16 forms made ptxas stage every operand
// through a fixed 16-register block: 32 IMAD.MOV per call, 20 % of all issued instructions.)
__device__ __forceinline__ void tmem_st1(uint32_t addr, const double2 v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%2,%3,%4};"
                 ::"r"(addr), "r"(__double2loint(v.x)), "r"(__double2hiint(v.x)), "r"(__double2loint(v.y)), "r"(__double2hiint(v.y)) : "memory");
}
__device__ __forceinline__ void tmem_ld1(uint32_t addr, double2 &v) {
    int w0, w1, w2, w3;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(w0), "=r"(w1), "=r"(w2), "=r"(w3) : "r"(addr) : "memory");
    v.x = __hiloint2double(w1, w0);  // only consumed after tmem_wait_ld()
    v.y = __hiloint2double(w3, w2);
}
__device__ __forceinline__ void tmem_st4(uint32_t addr, const double2 (&v)[4]) {
#pragma unroll
    for (int k = 0; k < 4; ++k) tmem_st1(addr + 4 * k, v[k]);
}
__device__ __forceinline__ void tmem_ld4(uint32_t addr, double2 (&v)[4]) {
#pragma unroll
    for (int k = 0; k < 4; ++k) tmem_ld1(addr + 4 * k, v[k]);
}

// TMEM columns of one warp
constexpr int kTmAcc0 = 0, kTmAcc1 = 64, kTmStash = 128, kTm16ColsPerWarp = 256;

// Rows for the single-step test kernel: straight from global memory, layout [col][r][lane]
struct GlobalRows16 {
    const double2 *bk_i;
    __device__ __forceinline__ const double2 *acquire(int rp) const { return bk_i + (size_t)rp * 2 * kM; }
    __device__ __forceinline__ double2 load(const double2 *p) const { return __ldg(p); }
    __device__ __forceinline__ void release(int) const {}
};

// ------------------------------------------------------------------------------------------------
// one CMux step on a shared-memory accumulator, executed by one warp
//   MODE 0: ACC <- ACC + BK_i (x) ((X^a - 1) ACC)      (lwe-bootstrapping-functions-fft.cpp:67-76)
//   MODE 1: ACC <- BK_i (x) ACC                         (tgsw-fft-operations.cpp:89-114)
// ------------------------------------------------------------------------------------------------
template <int L, int BGBIT, int MODE, class Rows>
__device__ __forceinline__ void cmux_step16(int32_t *ACC, double2 *buf, const int a, const Rows &rows, const int lane,
                                            const RootSet16 &w1, const RootSet16 &w2, const uint32_t tm) {
    constexpr uint32_t BG = 1u << BGBIT, HALF = BG >> 1, MASK = BG - 1;
    uint32_t offset = 0;  // tgsw.cpp:7-29
#pragma unroll
    for (int p = 0; p < L; ++p) offset += HALF << (32 - (p + 1) * BGBIT);
    const int sub = lane >> 4, tt = lane & 15;
    const double sg = sub ? -kCos45 : kCos45;   // level-0 root sqrt(i) = cos45 (1 + i); the lower half warp takes a + s b

#pragma unroll 1
    for (int q = 0; q < 2; ++q) {
        const int32_t *P = ACC + q * kN;
        double2 x[16];
        // ---- operand load = rotation, gadget decomposition and level 0 of all L digit transforms
#pragma unroll
        for (int r4 = 0; r4 < 4; ++r4) {
            double2 y[L > 1 ? L - 1 : 1][4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int r = 4 * r4 + k;
                uint32_t dv[4];  // coefficients e, e+256, e+512, e+768 of (X^a - 1) P, plus the offset
#pragma unroll
                for (int h = 0; h < 4; ++h) {
                    const int j = tt + 16 * r + 256 * h;
                    if (MODE == 0) {
                        const int m = (j - a) & (2 * kN - 1);
                        int32_t v = P[m & (kN - 1)];
                        if (m & kN) v = -v;
                        dv[h] = (uint32_t)(v - P[j]) + offset;
                    } else {
                        dv[h] = (uint32_t)P[j] + offset;
                    }
                }
#pragma unroll
                for (int p = 0; p < L; ++p) {
                    const int shift = 32 - (p + 1) * BGBIT;
                    const int32_t ax = (int32_t)((dv[0] >> shift) & MASK) - (int32_t)HALF;   // z_e     = ax + i ay
                    const int32_t bx = (int32_t)((dv[1] >> shift) & MASK) - (int32_t)HALF;   // z_e+256 = bx + i by
                    const int32_t ay = (int32_t)((dv[2] >> shift) & MASK) - (int32_t)HALF;
                    const int32_t by = (int32_t)((dv[3] >> shift) & MASK) - (int32_t)HALF;
                    double2 v;
                    v.x = fma(sg, int_to_double(bx - by), int_to_double(ax));
                    v.y = fma(sg, int_to_double(bx + by), int_to_double(ay));
                    if (p == 0) x[r] = v;
                    else y[p - 1][k] = v;
                }
            }
#pragma unroll
            for (int p = 1; p < L; ++p) tmem_st4(tm + kTmStash + 64 * (p - 1) + 16 * r4, y[p - 1]);
        }
#pragma unroll 1
        for (int p = 0; p < L; ++p) {
            if (p > 0) {
                tmem_wait_st();
#pragma unroll
                for (int r4 = 0; r4 < 4; ++r4) tmem_ld4(tm + kTmStash + 64 * (p - 1) + 16 * r4, *reinterpret_cast<double2(*)[4]>(&x[4 * r4]));
                tmem_wait_ld();
            }
            forward16_tail(x, buf, lane, w1, w2);
            const double2 *row = rows.acquire(q * L + p) + lane;
            const bool first = (q == 0 && p == 0);
            tmem_wait_st();  // the previous row's accumulator stores have landed
#pragma unroll
            for (int r4 = 0; r4 < 4; ++r4) {
                double2 a0[4], a1[4], b0[4], b1[4];
                if (!first) {
                    tmem_ld4(tm + kTmAcc0 + 16 * r4, a0);
                    tmem_ld4(tm + kTmAcc1 + 16 * r4, a1);
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    b0[k] = rows.load(row + (4 * r4 + k) * 32);
                    b1[k] = rows.load(row + kM + (4 * r4 + k) * 32);
                }
                if (!first) tmem_wait_ld();
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const double2 xv = x[4 * r4 + k];
                    if (first) {
                        a0[k].x = xv.x * b0[k].x; a0[k].y = xv.x * b0[k].y;
                        a1[k].x = xv.x * b1[k].x; a1[k].y = xv.x * b1[k].y;
                    } else {
                        a0[k].x = fma(xv.x, b0[k].x, a0[k].x); a0[k].y = fma(xv.x, b0[k].y, a0[k].y);
                        a1[k].x = fma(xv.x, b1[k].x, a1[k].x); a1[k].y = fma(xv.x, b1[k].y, a1[k].y);
                    }
                    a0[k].x = fma(-xv.y, b0[k].y, a0[k].x); a0[k].y = fma(xv.y, b0[k].x, a0[k].y);
                    a1[k].x = fma(-xv.y, b1[k].y, a1[k].x); a1[k].y = fma(xv.y, b1[k].x, a1[k].y);
                }
                tmem_st4(tm + kTmAcc0 + 16 * r4, a0);
                tmem_st4(tm + kTmAcc1 + 16 * r4, a1);
            }
            rows.release(q * L + p);
        }
    }

    // ---- back to coefficients: levels 8..1 in registers, level 0 across the two half warps
    constexpr double kScale = 1.0 / kM;
    const double sgs = sg * kScale;
    tmem_wait_st();
#pragma unroll 1
    for (int c = 0; c < 2; ++c) {
        double2 x[16];
#pragma unroll
        for (int r4 = 0; r4 < 4; ++r4) tmem_ld4(tm + (c ? kTmAcc1 : kTmAcc0) + 16 * r4, *reinterpret_cast<double2(*)[4]>(&x[4 * r4]));
        tmem_wait_ld();
        inverse16_head(x, buf, lane, w1, w2);
        int32_t *P = ACC + c * kN;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            // the lower half warp finishes the even elements, the upper half the odd ones
            const double2 mine = sub ? x[2 * k + 1] : x[2 * k];
            const double2 send = sub ? x[2 * k] : x[2 * k + 1];
            double2 other;
            other.x = shfl_xor16(send.x);
            other.y = shfl_xor16(send.y);
            const double sx = mine.x + other.x, sy = mine.y + other.y;      // 512 z_e
            const double dx = mine.x - other.x, dy = mine.y - other.y;      // +-(U - V); z_{e+256} = conj(s0) (U - V)
            const int e = tt + 16 * (2 * k + sub);
            const int32_t c0 = round_scaled_to_torus(sx, kScale), c2 = round_scaled_to_torus(sy, kScale);
            const int32_t c1 = round_scaled_to_torus(dx + dy, sgs), c3 = round_scaled_to_torus(dy - dx, sgs);
            if (MODE == 0) { P[e] += c0; P[e + 256] += c1; P[e + 512] += c2; P[e + 768] += c3; }
            else           { P[e] = c0;  P[e + 256] = c1;  P[e + 512] = c2;  P[e + 768] = c3; }
        }
    }
    __syncwarp();  // accumulator complete before the next step's rotated reads
}

// per ciphertext (= per warp): ACC + exchange buffer + bara
constexpr int kWarpSmemBytes = 2 * kN * 4 + kBuf16Elems * 16 + 1280;
constexpr int kBr16Stages = 4;
constexpr int br16_smem_bytes(int G) { return kBr16Stages * kStageBytes + 128 + G * kWarpSmemBytes; }

// G ciphertexts per CTA (one warp each), walking i = 0..n-1 on one TMA-fed ring of key row pairs.
template <int L, int BGBIT, int G>
__global__ void __launch_bounds__(32 * G, 1) k_blind_rotate16(const DeviceKey key, const BlindRotateArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int KPL = 2 * L;
    constexpr int TM_COLS = G > 4 ? 512 : 256;
    unsigned char *ring = smem;
    RingState<kBr16Stages> *st = reinterpret_cast<RingState<kBr16Stages> *>(smem + kBr16Stages * kStageBytes);
    uint32_t *tm_slot = reinterpret_cast<uint32_t *>(smem + kBr16Stages * kStageBytes + 96);
    const int g = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned char *mine = smem + kBr16Stages * kStageBytes + 128 + (size_t)g * kWarpSmemBytes;
    int32_t *ACC = reinterpret_cast<int32_t *>(mine);
    double2 *buf = reinterpret_cast<double2 *>(mine + 2 * kN * 4);
    uint16_t *bara = reinterpret_cast<uint16_t *>(mine + 2 * kN * 4 + kBuf16Elems * 16);

    const int n = key.n, n1 = n + 1;
    const int64_t item = (int64_t)blockIdx.x * G + g;
    const bool live = item < a.items;
    const int total = n * KPL;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kBr16Stages; ++s) { mbar_init(&st->full[s], 1); st->done[s] = 0; }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        for (int s = 0; s < kBr16Stages && s < total; ++s)
            tma_load_row_pair(ring + (size_t)s * kStageBytes, key.bkfft16 + (size_t)s * 2 * kM, &st->full[s]);
    }
    if (g == 0) tmem_alloc(tm_slot, TM_COLS);

    // ---- gate prologue + modulus switch to 2N (numeric-functions.cpp:59-65)
    if (live) {
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
        for (int i = lane; i < n1; i += 32) {
            uint32_t x = (uint32_t)s0 * (uint32_t)r0[i] + (uint32_t)s1 * (uint32_t)r1[i];
            if (i == n) x += (uint32_t)cst;
            bara[i] = (uint16_t)((x + 0x00100000u) >> 21);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();  // bara visible; barriers initialised; TMEM base published
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = *tm_slot + (((uint32_t)(g & 3) * 32u) << 16) + (uint32_t)(g >> 2) * kTm16ColsPerWarp;

    // ---- ACC = (0, X^{2N-barb} * mu*(1+X+...+X^{N-1}))   (lwe-bootstrapping-functions-fft.cpp:146-148)
    if (live) {
        const int barb = bara[n];
        const int32_t mu = a.mu;
#pragma unroll
        for (int rr = 0; rr < 32; ++rr) {
            const int j = lane + 32 * rr;
            ACC[j] = 0;
            ACC[kN + j] = ((j + barb) & kN) ? -mu : mu;
        }
    }
    const RootSet16 w1 = load_rootset16(c_tw16.p1[lane >> 4]), w2 = load_rootset16(c_tw16.p2[lane]);
    __syncwarp();

    // ---- blind rotation (lwe-bootstrapping-functions-fft.cpp:88-112)
    RingRows<kBr16Stages> rows{ring, st, key.bkfft16, 0, total, G, (threadIdx.x & 31) == 0};
#pragma unroll 1
    for (int i = 0; i < n; ++i) {
        rows.seq0 = i * KPL;
        const int ai = live ? bara[i] : 0;
        if (ai == 0) {
            rows.skip_step(KPL);
            continue;
        }
        cmux_step16<L, BGBIT, 0>(ACC, buf, ai, rows, lane, w1, w2, tm);
    }

    // ---- sample extract at index 0 (lwe.cpp:41-60)
    if (live) {
        int32_t *u = a.u + (size_t)item * (kN + 1);
#pragma unroll
        for (int rr = 0; rr < 32; ++rr) {
            const int j = lane + 32 * rr;
            u[j] = (j == 0) ? ACC[0] : -ACC[kN - j];
        }
        if (lane == 0) u[kN] = ACC[kN];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (g == 0) tmem_free(*tm_slot, TM_COLS);
}

template <int L, int BGBIT>
__global__ void __launch_bounds__(32, 1) k_cmux_test16(const DeviceKey key, int32_t *acc, int bk_index, int barai, int mode) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint32_t tm_slot;
    int32_t *ACC = reinterpret_cast<int32_t *>(smem);
    double2 *buf = reinterpret_cast<double2 *>(smem + 2 * kN * 4);
    const int lane = threadIdx.x;
    int32_t *gacc = acc + (size_t)blockIdx.x * 2 * kN;
    for (int j = lane; j < 2 * kN; j += 32) ACC[j] = gacc[j];
    tmem_alloc(&tm_slot, 256);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = tm_slot;
    const RootSet16 w1 = load_rootset16(c_tw16.p1[lane >> 4]), w2 = load_rootset16(c_tw16.p2[lane]);
    const GlobalRows16 rows{key.bkfft16 + (size_t)bk_index * (2 * L * 2 * kM)};
    if (mode == 0) {
        if (barai != 0) cmux_step16<L, BGBIT, 0>(ACC, buf, barai, rows, lane, w1, w2, tm);
    } else {
        cmux_step16<L, BGBIT, 1>(ACC, buf, 0, rows, lane, w1, w2, tm);
    }
    __syncwarp();
    for (int j = lane; j < 2 * kN; j += 32) gacc[j] = ACC[j];
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    tmem_free(tm, 256);
}

// one-time key conversion into the [r][lane] evaluation layout of the 16-point transform
__global__ void __launch_bounds__(32) k_bk_convert16(const int32_t *__restrict__ coefs, double2 *__restrict__ out) {
    __shared__ double2 buf[kBuf16Elems];
    const int lane = threadIdx.x, sub = lane >> 4, tt = lane & 15;
    const int32_t *p = coefs + (size_t)blockIdx.x * kN;
    const double sg = sub ? -kCos45 : kCos45;
    double2 x[16];
#pragma unroll
    for (int r = 0; r < 16; ++r) {
        const int e = tt + 16 * r;
        const double ax = (double)p[e], bx = (double)p[e + 256], ay = (double)p[e + 512], by = (double)p[e + 768];
        x[r].x = fma(sg, bx - by, ax);
        x[r].y = fma(sg, bx + by, ay);
    }
    const RootSet16 w1 = load_rootset16(c_tw16.p1[sub]), w2 = load_rootset16(c_tw16.p2[lane]);
    forward16_tail(x, buf, lane, w1, w2);
    double2 *o = out + (size_t)blockIdx.x * kM;
#pragma unroll
    for (int r = 0; r < 16; ++r) o[r * 32 + lane] = x[r];
}

}  // namespace b200
