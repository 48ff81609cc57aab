// blind_rotate5.cuh -- blind rotation "v5": two warps per ciphertext, each a complete 16-point-per-thread
// transform engine (included by kernels.cu after blind_rotate16.cuh; l = 2 only).
//
// v2 (64 threads x 8 points) needs two shared-memory exchanges per transform, one of them across the two
// warps, and is bound by those wavefronts.  The warp transform of negacyclic_fft16.cuh
// needs ONE warp-local exchange, but a warp that owns a whole ciphertext cannot keep two spectral
// accumulators (128 registers) next to its 16 points and the roots.  Here the CMux step is split by
// POLYNOMIAL instead: warp h of a ciphertext
//   1. rotates and decomposes accumulator polynomial q = h and transforms its two digits (x0, x1 in
//      registers, level 0 folded into the operand load),
//   2. multiplies them with the two key rows of that polynomial: the products for output column h stay
//      in registers, the products for column 1-h go to the partner warp through shared memory,
//   3. adds the partner's contribution and runs the inverse transform of output column h.
// Per ciphertext and step: 6 x 128 exchange wavefronts (was 1536), 512 key, 256 partial sums, about 300
// accumulator; two 64-thread barriers (was 7), everything else is __syncwarp.  No accumulator ever
// lives across transforms, so nothing is kept in tensor memory.
// Summation order differs from cmux_step (row products are added pairwise), so results agree with v2
// to FFT rounding, not bit for bit; the parity suite (2 LSB per step, decrypted bits, noise variance)
// applies unchanged.
#pragma once

namespace b200 {

// per ciphertext: ACC (2 x 1024 int32) + one exchange buffer per warp (also the partner's inbox) + bara
constexpr int kV5GroupSmemBytes = 2 * kN * 4 + 2 * kBuf16Elems * 16 + 1280;
constexpr int kV5Stages = 7;
__host__ __device__ constexpr int br5_smem_bytes(int G) { return kV5Stages * kStageBytes + 128 + G * kV5GroupSmemBytes; }

template <int BGBIT, class Barrier>
__device__ __forceinline__ void cmux_step5(int32_t *ACC, double2 *mybuf, double2 *peerbuf, const int a, const int h,
                                           const RingRows<kV5Stages> &rows, const int lane,
                                           const RootSet16 &w1, const RootSet16 &w2, const Barrier &bar) {
    constexpr int L = 2;
    constexpr uint32_t BG = 1u << BGBIT, HALF = BG >> 1, MASK = BG - 1;
    constexpr double kBiasDigit = 4503599627370496.0 + (double)HALF;   // 2^52 + Bg/2: unsigned field -> centred digit
    constexpr double kBiasPair = 4503599627370496.0 + (double)BG;      // 2^52 + Bg  : bx' - by' + Bg  or  bx' + by'
    constexpr uint32_t offset = (HALF << (32 - BGBIT)) + (HALF << (32 - 2 * BGBIT));   // tgsw.cpp:7-29
    const int sub = lane >> 4, tt = lane & 15;
    const double sg = sub ? -kCos45 : kCos45;   // level-0 root sqrt(i) = cos45 (1 + i)

    // ---- 1. operand load of polynomial h: rotation, gadget decomposition, level 0 of both digit transforms
    const int32_t *P = ACC + h * kN;
    double2 x0[16], x1[16];
#pragma unroll
    for (int r = 0; r < 16; ++r) {
        uint32_t dv[4];  // coefficients e, e+256, e+512, e+768 of (X^a - 1) P, plus the offset
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int j = tt + 16 * r + 256 * c;
            const int m = (j - a) & (2 * kN - 1);
            int32_t v = P[m & (kN - 1)];
            if (m & kN) v = -v;
            dv[c] = (uint32_t)(v - P[j]) + offset;
        }
#pragma unroll
        for (int p = 0; p < L; ++p) {
            const int shift = 32 - (p + 1) * BGBIT;
            const uint32_t ax = (dv[0] >> shift) & MASK, bx = (dv[1] >> shift) & MASK;
            const uint32_t ay = (dv[2] >> shift) & MASK, by = (dv[3] >> shift) & MASK;
            const double dax = __hiloint2double(0x43300000, (int)ax) - kBiasDigit;
            const double day = __hiloint2double(0x43300000, (int)ay) - kBiasDigit;
            const double dmi = __hiloint2double(0x43300000, (int)(bx - by + BG)) - kBiasPair;   // bx - by
            const double dpl = __hiloint2double(0x43300000, (int)(bx + by)) - kBiasPair;        // bx + by
            double2 v;
            v.x = fma(sg, dmi, dax);
            v.y = fma(sg, dpl, day);
            if (p == 0) x0[r] = v; else x1[r] = v;
        }
    }
    forward16_tail(x0, mybuf, lane, w1, w2);
    forward16_tail(x1, mybuf, lane, w1, w2);
    bar();   // both warps are done with their exchange buffers: they now serve as inboxes

    // ---- 2. products with the two key rows of polynomial h
    const double2 *rowA = rows.acquire(h * L + 0) + lane;
    const double2 *rowB = rows.acquire(h * L + 1) + lane;
    const int mine_off = h * kM, other_off = (1 - h) * kM;
#pragma unroll
    for (int r = 0; r < 16; ++r) {
        const double2 am = rowA[mine_off + r * 32], bm = rowB[mine_off + r * 32];
        const double2 ao = rowA[other_off + r * 32], bo = rowB[other_off + r * 32];
        const double2 u = x0[r], v = x1[r];
        double2 m, o;
        m.x = u.x * am.x;              m.y = u.x * am.y;
        m.x = fma(-u.y, am.y, m.x);    m.y = fma(u.y, am.x, m.y);
        m.x = fma(v.x, bm.x, m.x);     m.y = fma(v.x, bm.y, m.y);
        m.x = fma(-v.y, bm.y, m.x);    m.y = fma(v.y, bm.x, m.y);
        o.x = u.x * ao.x;              o.y = u.x * ao.y;
        o.x = fma(-u.y, ao.y, o.x);    o.y = fma(u.y, ao.x, o.y);
        o.x = fma(v.x, bo.x, o.x);     o.y = fma(v.x, bo.y, o.y);
        o.x = fma(-v.y, bo.y, o.x);    o.y = fma(v.y, bo.x, o.y);
        x0[r] = m;
        peerbuf[r * 32 + lane] = o;
    }
    rows.release(h * L + 0);
    rows.release(h * L + 1);
    bar();   // partial sums delivered

    // ---- 3. output column h: add the partner's share, back to coefficients, accumulate
#pragma unroll
    for (int r = 0; r < 16; ++r) {
        const double2 o = mybuf[r * 32 + lane];
        x0[r].x += o.x;
        x0[r].y += o.y;
    }
    inverse16_head(x0, mybuf, lane, w1, w2);
    constexpr double kScale = 1.0 / kM;
    const double sgs = sg * kScale;
    int32_t *Q = ACC + h * kN;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        // the lower half warp finishes the even elements, the upper half the odd ones
        const double2 mine = sub ? x0[2 * k + 1] : x0[2 * k];
        const double2 send = sub ? x0[2 * k] : x0[2 * k + 1];
        double2 other;
        other.x = shfl_xor16(send.x);
        other.y = shfl_xor16(send.y);
        const double sx = mine.x + other.x, sy = mine.y + other.y;      // 512 z_e
        const double dx = mine.x - other.x, dy = mine.y - other.y;      // +-(U - V); z_{e+256} = conj(s0) (U - V)
        const int e = tt + 16 * (2 * k + sub);
        Q[e] += round_scaled_to_torus(sx, kScale);
        Q[e + 256] += round_scaled_to_torus(dx + dy, sgs);
        Q[e + 512] += round_scaled_to_torus(sy, kScale);
        Q[e + 768] += round_scaled_to_torus(dy - dx, sgs);
    }
    __syncwarp();   // polynomial h belongs to this warp alone: its next rotated reads need no group barrier
}

template <int BGBIT, int G>
__global__ void __launch_bounds__(64 * G, 1) k_blind_rotate5(const DeviceKey key, const BlindRotateArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int L = 2, KPL = 2 * L;
    unsigned char *ring = smem;
    RingState<kV5Stages> *st = reinterpret_cast<RingState<kV5Stages> *>(smem + kV5Stages * kStageBytes);
    const int g = threadIdx.x >> 6, h = (threadIdx.x >> 5) & 1, lane = threadIdx.x & 31, t = threadIdx.x & 63;
    unsigned char *mine = smem + kV5Stages * kStageBytes + 128 + (size_t)g * kV5GroupSmemBytes;
    int32_t *ACC = reinterpret_cast<int32_t *>(mine);
    double2 *buf0 = reinterpret_cast<double2 *>(mine + 2 * kN * 4);
    double2 *mybuf = buf0 + h * kBuf16Elems, *peerbuf = buf0 + (1 - h) * kBuf16Elems;
    uint16_t *bara = reinterpret_cast<uint16_t *>(mine + 2 * kN * 4 + 2 * kBuf16Elems * 16);

    const int n = key.n, n1 = n + 1;
    const int64_t item = (int64_t)blockIdx.x * G + g;
    const bool live = item < a.items;
    const int total = n * KPL;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kV5Stages; ++s) { mbar_init(&st->full[s], 1); st->done[s] = 0; }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        for (int s = 0; s < kV5Stages && s < total; ++s)
            tma_load_row_pair(ring + (size_t)s * kStageBytes, key.bkfft16 + (size_t)s * 2 * kM, &st->full[s]);
    }

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
        for (int i = t; i < n1; i += 64) {
            uint32_t x = (uint32_t)s0 * (uint32_t)r0[i] + (uint32_t)s1 * (uint32_t)r1[i];
            if (i == n) x += (uint32_t)cst;
            bara[i] = (uint16_t)((x + 0x00100000u) >> 21);
        }
    }
    __syncthreads();  // bara visible; barriers initialised

    // ---- ACC = (0, X^{2N-barb} * mu*(1+X+...+X^{N-1}))   (lwe-bootstrapping-functions-fft.cpp:146-148)
    if (live) {
        const int barb = bara[n];
        const int32_t mu = a.mu;
#pragma unroll
        for (int rr = 0; rr < 16; ++rr) {
            const int j = t + 64 * rr;
            ACC[j] = 0;
            ACC[kN + j] = ((j + barb) & kN) ? -mu : mu;
        }
    }
    const RootSet16 w1 = load_rootset16(c_tw16.p1[lane >> 4]), w2 = load_rootset16(c_tw16.p2[lane]);
    const NamedBarrier64 bar{1 + g};
    bar();

    // ---- blind rotation (lwe-bootstrapping-functions-fft.cpp:88-112): a row pair is read by one warp per group
    RingRows<kV5Stages> rows{ring, st, key.bkfft16, 0, total, G, (threadIdx.x & 31) == 0};
#pragma unroll 1
    for (int i = 0; i < n; ++i) {
        rows.seq0 = i * KPL;
        const int ai = live ? bara[i] : 0;
        if (ai == 0) {
            for (int p = 0; p < L; ++p) { rows.acquire(h * L + p); rows.release(h * L + p); }
            continue;
        }
        cmux_step5<BGBIT>(ACC, mybuf, peerbuf, ai, h, rows, lane, w1, w2, bar);
    }

    // ---- sample extract at index 0 (lwe.cpp:41-60)
    bar();   // both polynomials final
    if (live) {
        int32_t *u = a.u + (size_t)item * (kN + 1);
#pragma unroll
        for (int rr = 0; rr < 16; ++rr) {
            const int j = t + 64 * rr;
            u[j] = (j == 0) ? ACC[0] : -ACC[kN - j];
        }
        if (t == 0) u[kN] = ACC[kN];
    }
}

template <int BGBIT, int G>
static cudaError_t br5_launch(const DeviceKey &key, const BlindRotateArgs &a, cudaStream_t s) {
    static bool done[16] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    constexpr int smem = br5_smem_bytes(G);
    static_assert(smem <= 227 * 1024, "shared memory budget");
    if (!(dev < 16 && done[dev])) {
        cudaError_t e = cudaFuncSetAttribute(k_blind_rotate5<BGBIT, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        cudaFuncSetAttribute(k_blind_rotate5<BGBIT, G>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        if (dev < 16) done[dev] = true;
    }
    const unsigned grid = (unsigned)((a.items + G - 1) / G);
    k_blind_rotate5<BGBIT, G><<<grid, 64 * G, smem, s>>>(key, a);
    return cudaSuccess;
}

}  // namespace b200
