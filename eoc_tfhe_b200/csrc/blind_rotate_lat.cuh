// blind_rotate_lat.cuh -- latency kernel: ONE ciphertext per CTA, its CMux step spread over 2l groups of 64
// threads (included by kernels.cu).  Used for batches of at most one ciphertext per SM, i.e. the drop-in
// bootsNAND(LweSample*...) call and narrow circuit levels.
//
// A ciphertext's CMux chain is serial; in the throughput kernels one 64-thread group walks its 2l forward
// transforms, 2l spectral MACs and 2 inverse transforms one after the other (8,500 clk per step when it has
// the SM to itself).  Here group g = (q, p) owns ONE row of the external product
// (tgsw-fft-operations.cpp:89-114): it rotates and decomposes accumulator polynomial q, transforms digit
// p, multiplies with key row (q, p) -- fetched from L2 into registers at the top of the step, so the
// latency hides behind the transform -- and leaves its two partial spectra in shared memory; after one
// CTA barrier groups 0 and 1 add the 2l partials of their output polynomial and run the inverse transform.
// The critical path per step is one forward and one inverse transform instead of 2l + 2.
// Sums are formed in a different order than in cmux_step, but every external product is an exact integer
// convolution that FP64 resolves far inside the rounding margin: outputs are identical to k_blind_rotate.
#pragma once

namespace b200 {

template <int L> __host__ __device__ constexpr int lat_smem_bytes() {
    // ACC + bara + per group: two exchange buffers and two partial spectra (2 x 512 double2)
    return 2 * kN * 4 + 1280 + 2 * L * (2 * kBufElems * 16 + 2 * kM * 16);
}

template <int L, int BGBIT>
__global__ void __launch_bounds__(128 * L, 1) k_blind_rotate_lat(const DeviceKey key, const BlindRotateArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int G = 2 * L, KPL = 2 * L;
    constexpr uint32_t BG = 1u << BGBIT, HALF = BG >> 1;
    int32_t *ACC = reinterpret_cast<int32_t *>(smem);                                    // [2][1024]
    uint16_t *bara = reinterpret_cast<uint16_t *>(smem + 2 * kN * 4);                    // [n+1]
    const int g = threadIdx.x >> 6, t = threadIdx.x & 63;
    unsigned char *mine = smem + 2 * kN * 4 + 1280 + (size_t)g * (2 * kBufElems * 16 + 2 * kM * 16);
    double2 *bufs = reinterpret_cast<double2 *>(mine);                                   // [2][576]
    double2 *part = reinterpret_cast<double2 *>(mine + 2 * kBufElems * 16);              // [2 cols][512], pass-3 layout
    auto part_of = [&](int gg) {
        return reinterpret_cast<const double2 *>(smem + 2 * kN * 4 + 1280 + (size_t)gg * (2 * kBufElems * 16 + 2 * kM * 16) + 2 * kBufElems * 16);
    };
    const int q = g / L, p = g - q * L;

    const int n = key.n, n1 = n + 1;
    const int64_t item = blockIdx.x;

    // ---- gate prologue + modulus switch to 2N (numeric-functions.cpp:59-65)
    {
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
        for (int i = threadIdx.x; i < n1; i += 64 * G) {
            uint32_t x = (uint32_t)s0 * (uint32_t)r0[i] + (uint32_t)s1 * (uint32_t)r1[i];
            if (i == n) x += (uint32_t)cst;
            bara[i] = (uint16_t)((x + 0x00100000u) >> 21);
        }
    }
    __syncthreads();
    // ---- ACC = (0, X^{2N-barb} * mu*(1+X+...+X^{N-1}))   (lwe-bootstrapping-functions-fft.cpp:146-148)
    {
        const int barb = bara[n];
        const int32_t mu = a.mu;
        for (int j = threadIdx.x; j < kN; j += 64 * G) {
            ACC[j] = 0;
            ACC[kN + j] = ((j + barb) & kN) ? -mu : mu;
        }
    }
    const RootSet w2 = load_rootset(c_tw.p2[t >> 3]), w3 = load_rootset(c_tw.p3[t]);
    const NamedBarrier64 bar{1 + g};
    uint32_t offset = 0;  // tgsw.cpp:7-29
#pragma unroll
    for (int pp = 0; pp < L; ++pp) offset += HALF << (32 - (pp + 1) * BGBIT);
    const int shift = 32 - (p + 1) * BGBIT;
    __syncthreads();

    // ---- blind rotation (lwe-bootstrapping-functions-fft.cpp:88-112)
    int toggle = 0;
#pragma unroll 1
    for (int i = 0; i < n; ++i) {
        const int ai = bara[i];
        if (ai == 0) continue;                     // CTA-uniform
        // key row (q, p) of BK_i: both columns, 16 complex values per thread, in flight during the transform
        const double2 *row = key.bkfft + ((size_t)i * KPL + g) * 2 * kM + t;
        double2 b0[8], b1[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) { b0[r] = __ldg(row + r * 64); b1[r] = __ldg(row + kM + r * 64); }

        const int32_t *P = ACC + q * kN;
        double2 x[8];
        {
            uint32_t dv[16];   // (rotated difference) + decomposition offset, coefficient t + 64 rr
#pragma unroll
            for (int rr = 0; rr < 16; ++rr) {
                const int j = t + 64 * rr;
                const int m = (j - ai) & (2 * kN - 1);
                int32_t v = P[m & (kN - 1)];
                if (m & kN) v = -v;
                dv[rr] = (uint32_t)(v - P[j]) + offset;
            }
            digits_pretwisted<BGBIT>(x, dv, shift);   // level 0 of the transform on the integer side
        }
        negacyclic_forward<true>(x, bufs + toggle * kBufElems, t, w2, w3, bar);
        toggle ^= 1;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            double2 m0, m1;
            m0.x = x[r].x * b0[r].x;             m0.y = x[r].x * b0[r].y;
            m0.x = fma(-x[r].y, b0[r].y, m0.x);  m0.y = fma(x[r].y, b0[r].x, m0.y);
            m1.x = x[r].x * b1[r].x;             m1.y = x[r].x * b1[r].y;
            m1.x = fma(-x[r].y, b1[r].y, m1.x);  m1.y = fma(x[r].y, b1[r].x, m1.y);
            part[r * 64 + t] = m0;
            part[kM + r * 64 + t] = m1;
        }
        __syncthreads();   // all 2l partial spectra written; every group has finished reading ACC
        if (g < 2) {       // output polynomial g
            double2 s[8];
#pragma unroll
            for (int r = 0; r < 8; ++r) s[r] = part_of(0)[g * kM + r * 64 + t];
#pragma unroll
            for (int gg = 1; gg < G; ++gg) {
                const double2 *pg = part_of(gg) + g * kM + t;
#pragma unroll
                for (int r = 0; r < 8; ++r) { const double2 v = pg[r * 64]; s[r].x += v.x; s[r].y += v.y; }
            }
            negacyclic_inverse<true>(s, bufs + toggle * kBufElems, t, w2, w3, bar);
            toggle ^= 1;
            int32_t *Q = ACC + g * kN;
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const int j = t + 64 * r;
                const double sc = c_tw.p1_scale[r];   // 1/512 times the cos factors the last pass left out
                Q[j] += round_scaled_to_torus(s[r].x, sc);
                Q[j + kM] += round_scaled_to_torus(s[r].y, sc);
            }
        }
        __syncthreads();   // accumulator complete; partial buffers free
    }

    // ---- sample extract at index 0 (lwe.cpp:41-60)
    int32_t *u = a.u + (size_t)item * (kN + 1);
    for (int j = threadIdx.x; j < kN; j += 64 * G) u[j] = (j == 0) ? ACC[0] : -ACC[kN - j];
    if (threadIdx.x == 0) u[kN] = ACC[kN];
}

template <int L, int BGBIT>
static cudaError_t lat_launch(const DeviceKey &key, const BlindRotateArgs &a, cudaStream_t s) {
    static bool done[16] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    constexpr int smem = lat_smem_bytes<L>();
    static_assert(smem <= 227 * 1024, "shared memory budget");
    if (!(dev < 16 && done[dev])) {
        cudaError_t e = cudaFuncSetAttribute(k_blind_rotate_lat<L, BGBIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        if (dev < 16) done[dev] = true;
    }
    k_blind_rotate_lat<L, BGBIT><<<(unsigned)a.items, 128 * L, smem, s>>>(key, a);
    return cudaSuccess;
}

}  // namespace b200
