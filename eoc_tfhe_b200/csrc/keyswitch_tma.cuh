// keyswitch_tma.cuh -- key switch for large batches (included by kernels.cu).
//
// lweKeySwitch (libs/tfhe/src/libtfhe/lwe-keyswitch-functions.cpp:114-129,228-238), base 4, t = 8:
//   out = (0, u.b) - sum_{i<N} sum_{j<8, d_ij != 0} KS[i][j][d_ij],   d_ij = two bits of u.a[i] + 2^15
// The first kernel (k_keyswitch) gathers table rows with LDG per 8-sample tile: 82 % of its issued
// instructions were digit decode / row select, and every tile pulled the whole 49.5 MB table through L1
// (405 GB for 65,536 samples).  This one is organised around the table stream instead:
//   * a CTA owns NG x SG samples (64 at the 80-bit set); the table is walked ONCE per CTA in
//     half-coefficient slabs (12 rows = 24 KiB) that TMA bulk copies (cp.async.bulk) drop into a 4-stage
//     shared-memory ring completing on mbarriers -- L2 traffic falls 8x and rows are read with LDS.128;
//   * a thread owns two int4 column groups of SG = 8 samples (64 accumulator registers, 16 warps per SM:
//     the decode -> compare -> branch chain is latency bound, so warps matter more than tile height),
//     and one digit decode serves 8 integer subtractions;
//   * digits are warp-uniform: they are packed (16 bits per sample and coefficient) by a pre-pass per
//     chunk of 64 coefficients and moved to uniform registers with redux.sync, so the 3-way row choice is a
//     uniform branch, not a select.
// Integer wrap-around arithmetic only: bit-exact in any summation order.
#pragma once

namespace b200 {

constexpr int kKs2Chunk = 64;       // coefficients per digit chunk
constexpr int kKs2StagesMax = 6;
// ring depth: the eight-group CTA of the 80-bit set takes six stages (65,536 samples 19.7 -> 19.3 ms); the smaller
// shapes stay at four so that two of them fit an SM, and the 128-bit set's 30 KB stages gain nothing from more
__host__ __device__ constexpr int ks2_stages(int ng) { return ng >= 8 ? 6 : 4; }
constexpr int kKs2RowsPerStage = 12;  // half of the 24 rows of one coefficient (j = 0..3 or 4..7, h = 1..3)

template <int S> struct Ks2State {   // S = samples per CTA
    uint64_t full[kKs2StagesMax];
    uint64_t empty[kKs2StagesMax];    // consumer barriers: one arrival per warp and use of the stage
    int nz[S];
    int32_t bterm[S];
    int32_t outrow[S];
};

__host__ __device__ constexpr int ks2_state_bytes(int S) { return ((kKs2StagesMax * 16 + S * 12) + 127) / 128 * 128; }
__host__ __device__ constexpr int ks2_smem_bytes(int S, int stride, int stages) {
    return stages * kKs2RowsPerStage * stride * 4 + 2 * kKs2Chunk * S * 2 + ks2_state_bytes(S);
}

__device__ __forceinline__ void ks2_sub(int4 &a, const int4 &v) { a.x -= v.x; a.y -= v.y; a.z -= v.z; a.w -= v.w; }

// One decision for two samples at the same digit position (MODE 1): A, B = their digits.
template <int A, int B, int NC>
__device__ __forceinline__ void ks2_two_samples(int4 (&x)[NC], int4 (&y)[NC], const int4 (&r)[3][NC]) {
#pragma unroll
    for (int c = 0; c < NC; ++c) {
        if (A) ks2_sub(x[c], r[A ? A - 1 : 0][c]);
        if (B) ks2_sub(y[c], r[B ? B - 1 : 0][c]);
    }
}
// One decision for two adjacent digits of one sample (MODE 2): both rows leave in one three-input add.
template <int A, int B, int NC>
__device__ __forceinline__ void ks2_two_digits(int4 (&x)[NC], const int4 (&ra)[3][NC], const int4 (&rb)[3][NC]) {
#pragma unroll
    for (int c = 0; c < NC; ++c) {
        if (A && B) {
            const int4 &u = ra[A ? A - 1 : 0][c], &v = rb[B ? B - 1 : 0][c];
            x[c].x = x[c].x - u.x - v.x; x[c].y = x[c].y - u.y - v.y; x[c].z = x[c].z - u.z - v.z; x[c].w = x[c].w - u.w - v.w;
        } else if (A) ks2_sub(x[c], ra[A ? A - 1 : 0][c]);
        else if (B) ks2_sub(x[c], rb[B ? B - 1 : 0][c]);
    }
}
#include "keyswitch_brx.inc"
#define KS2_CASES16(F)                                                                                     \
    F(0, 1) F(0, 2) F(0, 3) F(1, 0) F(1, 1) F(1, 2) F(1, 3) F(2, 0) F(2, 1) F(2, 2) F(2, 3) F(3, 0) F(3, 1) \
    F(3, 2) F(3, 3)

// TG threads per group (each owns NC int4 column groups), NG groups per CTA, SG samples per group.
// MODE 0: per sample and digit a uniform 3-way branch; 1: one 16-way switch per sample pair; 2: one 16-way switch
// per digit pair; 3, 4: the same two with the choice made by one indexed jump (brx.idx, keyswitch_brx.inc).  STRIDE4 > 0: the row stride in int4 is a compile-time constant (immediate LDS offsets).
template <int TG, int NG, int SG, int NC, int MODE = 0, int STRIDE4 = 0>
__global__ void __launch_bounds__(TG * NG, (TG == 64 && NG <= 4) ? 2 : 1) k_keyswitch_tma(const DeviceKey key, const KeySwitchArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    static_assert(SG == 8 || SG == 16, "8 or 16 samples per group");
    constexpr int T = 8, S = NG * SG, WARPS = TG * NG / 32, kKs2Stages = ks2_stages(NG);
    const int Nin = key.N * key.k, n1 = key.n + 1;
    const int stride = STRIDE4 > 0 ? STRIDE4 * 4 : key.ks_stride, stride4 = stride >> 2;
    const int stage_bytes = kKs2RowsPerStage * stride * 4;
    unsigned char *ring = smem;
    uint16_t *dig = reinterpret_cast<uint16_t *>(smem + kKs2Stages * stage_bytes);     // [2][chunk][S]
    Ks2State<S> *st = reinterpret_cast<Ks2State<S> *>(smem + kKs2Stages * stage_bytes + 2 * kKs2Chunk * S * 2);

    const int grp = threadIdx.x / TG, t = threadIdx.x - grp * TG;
    const int64_t item0 = (int64_t)blockIdx.x * S;
    const uint32_t prec_offset = 1u << (32 - (1 + 2 * T));
    // gridDim.y > 1: the coefficient range is cut into gridDim.y pieces of whole 64-coefficient chunks, one CTA each;
    // the partial sums meet in the output rows through integer atomics (wrap-around addition: bit-exact in any order),
    // the rows are zeroed by the launcher.  Batches that cannot fill the SMs with whole-range CTAs use it.
    const int nchunks = (Nin + kKs2Chunk - 1) / kKs2Chunk;
    const int nsplit = (int)gridDim.y, split = (int)blockIdx.y;
    const int chunk0 = nchunks * split / nsplit, chunk1 = nchunks * (split + 1) / nsplit;
    const int q0 = chunk0 * kKs2Chunk * 2;                                   // first half-coefficient slab of this CTA
    const int total = min(chunk1 * kKs2Chunk, Nin) * 2;                     // one past its last
    const unsigned char *tab = reinterpret_cast<const unsigned char *>(key.ks);

    if (threadIdx.x == 0) {
        for (int s = 0; s < kKs2Stages; ++s) { mbar_init(&st->full[s], 1); mbar_init(&st->empty[s], WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        for (int s = 0; s < kKs2Stages && q0 + s < total; ++s) {
            const uint32_t b = smem_u32(&st->full[s]), d = smem_u32(ring + (size_t)s * stage_bytes);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(stage_bytes) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(d), "l"(tab + (size_t)(q0 + s) * stage_bytes), "r"(stage_bytes), "r"(b) : "memory");
        }
    }
    // per-sample constants: b term (lwe-keyswitch-functions.cpp:232) and destination row
    if (threadIdx.x < S) {
        const int g = threadIdx.x;
        st->nz[g] = 0;
        const int64_t item = item0 + g;
        int32_t b = 0, orow = -1;
        if (item < a.items) {
            int64_t u0, u1 = -1;
            int32_t add = 0;
            if (a.desc) { const KsDesc d = a.desc[item]; u0 = d.u0; u1 = d.u1; add = d.add_b; orow = d.out; }
            else if (a.pair) { u0 = 2 * item; u1 = 2 * item + 1; add = 1 << 29; orow = (int32_t)item; }
            else { u0 = item; orow = (int32_t)item; }
            b = a.u[u0 * (Nin + 1) + Nin] + add;
            if (u1 >= 0) b += a.u[u1 * (Nin + 1) + Nin];
        }
        st->bterm[g] = b;
        st->outrow[g] = orow;
    }
    __syncthreads();   // counters zeroed before the pre-pass adds to them; barriers initialised

    // digit pre-pass of one chunk (lwe-keyswitch-functions.cpp:122-125): top 2T bits of a_i + prec_offset
    auto prepass = [&](int chunk) {
        uint16_t *dst = dig + (size_t)((chunk - chunk0) & 1) * kKs2Chunk * S;
        const int i0 = chunk * kKs2Chunk;
        for (int idx = threadIdx.x; idx < S * kKs2Chunk; idx += TG * NG) {
            const int g = idx / kKs2Chunk, ii = idx - g * kKs2Chunk;
            const int64_t item = item0 + g;
            uint32_t packed = 0;
            if (item < a.items && i0 + ii < Nin) {
                int64_t u0, u1 = -1;
                if (a.desc) { const KsDesc d = a.desc[item]; u0 = d.u0; u1 = d.u1; }
                else if (a.pair) { u0 = 2 * item; u1 = 2 * item + 1; }
                else u0 = item;
                uint32_t v = (uint32_t)a.u[u0 * (Nin + 1) + i0 + ii];
                if (u1 >= 0) v += (uint32_t)a.u[u1 * (Nin + 1) + i0 + ii];
                packed = (v + prec_offset) >> (32 - 2 * T);
                const int c = __popc((packed | (packed >> 1)) & 0x5555u);
                if (c && a.nonzero) atomicAdd(&st->nz[g], c);
            }
            dst[ii * S + g] = (uint16_t)packed;
        }
    };
    prepass(chunk0);
    __syncthreads();

    int4 acc[SG][NC];
#pragma unroll
    for (int g = 0; g < SG; ++g)
#pragma unroll
        for (int c = 0; c < NC; ++c) acc[g][c] = make_int4(0, 0, 0, 0);
    int col[NC];
    bool has[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) { has[c] = t + TG * c < stride4; col[c] = has[c] ? t + TG * c : 0; }

#pragma unroll 1
    for (int chunk = chunk0; chunk < chunk1; ++chunk) {
        if (chunk + 1 < chunk1) prepass(chunk + 1);   // the other buffer; made visible by the barrier below
        const uint16_t *dg = dig + (size_t)((chunk - chunk0) & 1) * kKs2Chunk * S + grp * SG;
        const int icnt = min(kKs2Chunk, Nin - chunk * kKs2Chunk);
#pragma unroll 1
        for (int ii = 0; ii < icnt; ++ii) {
            // SG samples x 16 bits, warp uniform; redux.sync moves them to uniform registers
            uint32_t w[SG / 2];
            {
                const uint4 w0 = *reinterpret_cast<const uint4 *>(dg + ii * S);
                w[0] = w0.x; w[1] = w0.y; w[2] = w0.z; w[3] = w0.w;
                if (SG == 16) {
                    const uint4 w1 = *reinterpret_cast<const uint4 *>(dg + ii * S + 8);
                    w[SG / 2 - 4] = w1.x; w[SG / 2 - 3] = w1.y; w[SG / 2 - 2] = w1.z; w[SG / 2 - 1] = w1.w;
                }
            }
#pragma unroll
            for (int k = 0; k < SG / 2; ++k) w[k] = __reduce_or_sync(0xffffffffu, w[k]);
#pragma unroll 1
            for (int jh = 0; jh < 2; ++jh) {
                const int q = (chunk * kKs2Chunk + ii) * 2 + jh, s = (q - q0) % kKs2Stages;
                mbar_wait(&st->full[s], (uint32_t)((q - q0) / kKs2Stages) & 1u);
                const int4 *rows = reinterpret_cast<const int4 *>(ring + (size_t)s * stage_bytes);
                if (MODE == 0) {
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj) {
                        int4 r[3][NC];
#pragma unroll
                        for (int h = 0; h < 3; ++h)
#pragma unroll
                            for (int c = 0; c < NC; ++c) r[h][c] = rows[(jj * 3 + h) * stride4 + col[c]];
                        const int sh = 2 * (T - 1 - (jh * 4 + jj));
#pragma unroll
                        for (int g = 0; g < SG; ++g) {
                            const uint32_t d = ((w[g >> 1] >> (16 * (g & 1))) >> sh) & 3u;
                            if (d == 1) { _Pragma("unroll") for (int c = 0; c < NC; ++c) ks2_sub(acc[g][c], r[0][c]); }
                            else if (d == 2) { _Pragma("unroll") for (int c = 0; c < NC; ++c) ks2_sub(acc[g][c], r[1][c]); }
                            else if (d == 3) { _Pragma("unroll") for (int c = 0; c < NC; ++c) ks2_sub(acc[g][c], r[2][c]); }
                        }
                    }
                } else if (MODE == 1 || MODE == 3) {
#pragma unroll 1
                    for (int jj = 0; jj < 4; ++jj) {
                        int4 r[3][NC];
#pragma unroll
                        for (int h = 0; h < 3; ++h)
#pragma unroll
                            for (int c = 0; c < NC; ++c) r[h][c] = rows[(jj * 3 + h) * stride4 + col[c]];
                        const int sh = 2 * (T - 1 - (jh * 4 + jj));
#pragma unroll
                        for (int p = 0; p < SG / 2; ++p) {
                            const uint32_t wp = w[p] >> sh;           // sample 2p in bits 0-1, sample 2p+1 in bits 16-17
                            if constexpr (MODE == 3 && NC == 2) { ks2_brx_two_samples(acc[2 * p], acc[2 * p + 1], r, (wp & 3u) | ((wp >> 14) & 12u)); continue; }
                            switch ((wp & 3u) | ((wp >> 14) & 12u)) {
#define KS2_F(B, A) case (A) | ((B) << 2): ks2_two_samples<A, B, NC>(acc[2 * p], acc[2 * p + 1], r); break;
                                KS2_CASES16(KS2_F)
#undef KS2_F
                                default: break;
                            }
                        }
                    }
                } else {
#pragma unroll 1
                    for (int jp = 0; jp < 2; ++jp) {
                        int4 ra[3][NC], rb[3][NC];
#pragma unroll
                        for (int h = 0; h < 3; ++h)
#pragma unroll
                            for (int c = 0; c < NC; ++c) {
                                ra[h][c] = rows[(jp * 6 + h) * stride4 + col[c]];
                                rb[h][c] = rows[(jp * 6 + 3 + h) * stride4 + col[c]];
                            }
                        const int shb = 2 * (T - 1 - (jh * 4 + 2 * jp + 1));   // digit 2jp+1; digit 2jp sits two bits above
#pragma unroll
                        for (int g = 0; g < SG; ++g) {
                            if constexpr (MODE == 4 && NC == 2) { ks2_brx_two_digits(acc[g], ra, rb, (w[g >> 1] >> (16 * (g & 1) + shb)) & 15u); continue; }
                            switch ((w[g >> 1] >> (16 * (g & 1) + shb)) & 15u) {
#define KS2_F(A, B) case ((A) << 2) | (B): ks2_two_digits<A, B, NC>(acc[g], ra, rb); break;
                                KS2_CASES16(KS2_F)
#undef KS2_F
                                default: break;
                            }
                        }
                    }
                }
                // hand the stage back (one mbarrier arrival per warp); the arrival that completes the phase refills it
                __syncwarp();
                if ((threadIdx.x & 31) == 0 && mbar_arrive_is_last(&st->empty[s]) && q + kKs2Stages < total) {
                    const uint32_t b = smem_u32(&st->full[s]), d = smem_u32(ring + (size_t)s * stage_bytes);
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(stage_bytes) : "memory");
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 ::"r"(d), "l"(tab + (size_t)(q + kKs2Stages) * stage_bytes), "r"(stage_bytes), "r"(b) : "memory");
                }
            }
        }
        __syncthreads();   // next chunk's digits visible; this chunk's buffer free for the pre-pass after next
    }

#pragma unroll
    for (int g = 0; g < SG; ++g) {
        const int orow = st->outrow[grp * SG + g];
        if (orow < 0) continue;
        int32_t *o = a.out + (size_t)orow * n1;
        const int32_t bt = st->bterm[grp * SG + g];
#pragma unroll
        for (int c2 = 0; c2 < NC; ++c2) {
            if (!has[c2]) continue;
            const int cbase = (t + TG * c2) * 4;
            const int32_t vals[4] = {acc[g][c2].x, acc[g][c2].y, acc[g][c2].z, acc[g][c2].w};
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int colx = cbase + c;
                if (colx < n1) {
                    const int32_t v = vals[c] + ((colx == key.n && split == 0) ? bt : 0);
                    if (nsplit == 1) o[colx] = v;
                    else atomicAdd(o + colx, v);
                }
            }
        }
    }
    if (a.nonzero && threadIdx.x < S && item0 + threadIdx.x < a.items && st->nz[threadIdx.x])
        atomicAdd(a.nonzero + item0 + threadIdx.x, st->nz[threadIdx.x]);
}

template <int TG, int NG, int SG, int NC, int MODE = 0, int STRIDE4 = 0>
static cudaError_t ks2_launch(const DeviceKey &key, const KeySwitchArgs &a, cudaStream_t s, int nsplit = 1) {
    static bool done[16] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (STRIDE4 > 0 && key.ks_stride != STRIDE4 * 4) return cudaErrorInvalidValue;
    const int smem = ks2_smem_bytes(NG * SG, key.ks_stride, ks2_stages(NG));
    if (smem > 227 * 1024) return cudaErrorInvalidValue;
    if (!(dev < 16 && done[dev])) {
        cudaError_t e = cudaFuncSetAttribute(k_keyswitch_tma<TG, NG, SG, NC, MODE, STRIDE4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return e;
        if (dev < 16) done[dev] = true;
    }
    const dim3 grid((unsigned)((a.items + NG * SG - 1) / (NG * SG)), (unsigned)nsplit);
    k_keyswitch_tma<TG, NG, SG, NC, MODE, STRIDE4><<<grid, TG * NG, smem, s>>>(key, a);
    return cudaSuccess;
}

}  // namespace b200
