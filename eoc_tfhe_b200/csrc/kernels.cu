// kernels.cu -- sm_100a kernels of the gate-bootstrapping path.
//
//   k_bk_convert     one-time: integer bootstrapping key -> evaluation domain, coalesced layout
//                    (replaces init_LweBootstrappingKeyFFT / tGswToFFTConvert,
//                     libs/tfhe/src/libtfhe/lwe-bootstrapping-functions-fft.cpp:23-52,
//                     tgsw-fft-operations.cpp:49-54)
//   k_blind_rotate   gate prologue + modswitch + n CMux steps + sample extract, one ciphertext per
//                    64-thread CTA, everything on chip
//                    (boot-gates.cpp:35-370 prologues, lwe-bootstrapping-functions-fft.cpp:67-200,
//                     tgsw-fft-operations.cpp:89-114, tgsw-functions.cpp:298-348, lwe.cpp:41-60)
//   k_keyswitch      integer gather-accumulate over the key-switch table, batch tiled
//                    (lwe-keyswitch-functions.cpp:114-129,228-238)
//   k_cmux_test      a single CMux step / external product on caller accumulators (parity tests)
#include "engine.cuh"
#include "negacyclic_fft.cuh"
#include "negacyclic_fft_tm.cuh"
#ifdef B200_EXPERIMENTS
#include "negacyclic_fft16.cuh"   // round-1 experiments: built only with -DB200_EXPERIMENTS
#endif
#include "../../include/b200tfhe.h"

#include <algorithm>
#include <atomic>
#include <mutex>
#include <math.h>
#include <stdlib.h>
#ifdef B200_BR_PROFILE
#include <cstdio>
#endif

namespace b200 {

static std::atomic<int64_t> g_launches{0};
int64_t launch_count() { return g_launches.load(); }
#define COUNT_LAUNCH() g_launches.fetch_add(1, std::memory_order_relaxed)

// ------------------------------------------------------------------------------------------------
// twiddles: theta(level, b) = (pi/2 + 2 pi brv_level(b)) / 2^level is the argument of the root R of
// the block modulus X^n - R; a butterfly of that block multiplies by sqrt(R) = e^{i theta/2}.
// ------------------------------------------------------------------------------------------------
static unsigned bit_reverse(unsigned v, int bits) {
    unsigned r = 0;
    for (int i = 0; i < bits; ++i) r |= ((v >> i) & 1u) << (bits - 1 - i);
    return r;
}
static long double block_theta(int level, unsigned b) {
    const long double pi = 3.14159265358979323846264338327950288L;
    return (pi / 2 + 2 * pi * (long double)bit_reverse(b, level)) / (long double)(1u << level);
}
static double2 cos_tan(long double angle) {
    return make_double2((double)cosl(angle), (double)tanl(angle));
}
static RootSet make_rootset(int level, unsigned b) {
    const long double pi = 3.14159265358979323846264338327950288L;
    RootSet w;
    const long double c = block_theta(level + 2, 4 * b) / 2;
    w.A = cos_tan(block_theta(level, b) / 2);
    w.B = cos_tan(block_theta(level + 1, 2 * b) / 2);
    w.C = cos_tan(c);
    w.D = cos_tan(c + pi / 4);
    return w;
}
#ifdef B200_EXPERIMENTS
static RootSet16 make_rootset16(int level, unsigned b) {
    RootSet16 w;
    w.A = cos_tan(block_theta(level, b) / 2);
    w.B = cos_tan(block_theta(level + 1, 2 * b) / 2);
    w.C = cos_tan(block_theta(level + 2, 4 * b) / 2);
    w.D = cos_tan(block_theta(level + 2, 4 * b + 2) / 2);
    w.E = cos_tan(block_theta(level + 3, 8 * b) / 2);
    w.F = cos_tan(block_theta(level + 3, 8 * b + 2) / 2);
    w.G = cos_tan(block_theta(level + 3, 8 * b + 4) / 2);
    w.H = cos_tan(block_theta(level + 3, 8 * b + 6) / 2);
    return w;
}
#endif
cudaError_t upload_twiddles() {
    static TwiddleTables h;
    h.p1 = make_rootset(0, 0);
    for (unsigned b = 0; b < 8; ++b) h.p2[b] = make_rootset(3, b);
    for (unsigned b = 0; b < 64; ++b) h.p3[b] = make_rootset(6, b);
    {
        // deferred cos factors of the last inverse pass (negacyclic_fft.cuh, inv_radix8_last), in long double
        const long double pi = 3.14159265358979323846264338327950288L;
        const long double cA = cosl(block_theta(0, 0) / 2), cB = cosl(block_theta(1, 0) / 2);
        const long double cC = cosl(block_theta(2, 0) / 2), cD = cosl(block_theta(2, 0) / 2 + pi / 4);
        h.p1_rho = (double)(cD / cC);
        const long double sc[8] = {1, cC, cB, cC * cB, cA, cC * cA, cB * cA, cC * cB * cA};
        for (int r = 0; r < 8; ++r) h.p1_scale[r] = (double)(sc[r] / 512);
    }
    cudaError_t e = cudaMemcpyToSymbol(c_tw, &h, sizeof(h));
    if (e != cudaSuccess) return e;
    {
        // per-thread roots of the small passes of the tensor-memory transform (negacyclic_fft_tm.cuh): P3a is
        // indexed by b' = e >> 4, P3b by the thread (P3b layout: warp e8, lane (e6 e5 e4 e2 e3), registers (e0 e7 e1))
        static TmRootTables ht;
        for (unsigned b = 0; b < 32; ++b) {
            ht.p3a[b][0] = cos_tan(block_theta(6, 2 * b) / 2);
            ht.p3a[b][1] = cos_tan(block_theta(7, 4 * b) / 2);
            ht.p3a[b][2] = cos_tan(block_theta(7, 4 * b + 2) / 2);
        }
        for (unsigned t = 0; t < 64; ++t) {
            const unsigned lane = t & 31, e8 = t >> 5, e6 = (lane >> 4) & 1, e5 = (lane >> 3) & 1, e4 = (lane >> 2) & 1,
                           e2 = (lane >> 1) & 1, e3 = lane & 1;
            for (unsigned e7 = 0; e7 < 2; ++e7) {
                const unsigned e = (e8 << 8) | (e7 << 7) | (e6 << 6) | (e5 << 5) | (e4 << 4) | (e3 << 3) | (e2 << 2);  // e1 = e0 = 0
                ht.p3b[t][e7] = cos_tan(block_theta(8, e >> 1) / 2);
            }
        }
        e = cudaMemcpyToSymbol(c_tw_tm, &ht, sizeof(ht));
    }
#ifdef B200_EXPERIMENTS
    if (e != cudaSuccess) return e;
    static TwiddleTables16 h16;
    for (unsigned b = 0; b < 2; ++b) h16.p1[b] = make_rootset16(1, b);
    for (unsigned b = 0; b < 32; ++b) h16.p2[b] = make_rootset16(5, b);
    e = cudaMemcpyToSymbol(c_tw16, &h16, sizeof(h16));
#endif
    return e;
}

// ------------------------------------------------------------------------------------------------
// one-time key conversion
// ------------------------------------------------------------------------------------------------
// tensor memory of a CTA: all 512 columns, allocated by warp 0; every warp gets a 128-column window in its own lane
// quarter (a warp reaches lanes 32 (warp % 4) .. +31 only); warps that share a quarter get different columns
__device__ __forceinline__ uint32_t tm_alloc_cta(uint32_t *slot, const int window = kTmColsPerWarp, const int shared = 0) {
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t warp = threadIdx.x >> 5;
    return *slot + (((warp & 3u) * 32u) << 16) + (uint32_t)shared + (warp >> 2) * (uint32_t)window;
}
__device__ __forceinline__ void tm_free_cta(const uint32_t *slot) {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(*slot) : "memory");
}
template <class FFT> struct UsesTm { static constexpr bool value = false; };
template <> struct UsesTm<FftTm> { static constexpr bool value = true; };
template <> struct UsesTm<FftTmAcc> { static constexpr bool value = true; };
template <> struct UsesTm<FftTmS> { static constexpr bool value = true; };
template <class FFT> struct PairEvents { static constexpr bool value = false; };   // ring released per MAC event (release_n)
template <> struct PairEvents<FftTmAcc> { static constexpr bool value = true; };
template <class FFT> struct TmWindow { static constexpr int value = kTmColsPerWarp; };
template <> struct TmWindow<FftTmAcc> { static constexpr int value = kTmColsPerWarpAcc; };
template <> struct TmWindow<FftTmS> { static constexpr int value = kTmSColsPerWarp; };
template <class FFT> struct TmShared { static constexpr int value = 0; };       // columns shared by a lane quarter
template <> struct TmShared<FftTmS> { static constexpr int value = kTmSColsRoots; };

template <class FFT>
__global__ void __launch_bounds__(64) k_bk_convert(const int32_t *__restrict__ coefs, double2 *__restrict__ out, int polys) {
    __shared__ double2 buf[FFT::kBufElemsPerTransform];
    __shared__ uint32_t tm_slot_s;
    const int t = threadIdx.x;
    uint32_t tm = 0;
    if (UsesTm<FFT>::value) tm = tm_alloc_cta(&tm_slot_s);
    FFT fft;
    fft.init(t, tm);
    for (int poly = blockIdx.x; poly < polys; poly += gridDim.x) {
        const int32_t *p = coefs + (size_t)poly * kN;
        double2 x[8];
#pragma unroll
        for (int r = 0; r < 8; ++r)
            x[r] = make_double2((double)p[t + 64 * r], (double)p[t + 64 * r + kM]);
        fft.forward(x, buf, CtaBarrier());
        double2 *o = out + (size_t)poly * kM;
#pragma unroll
        for (int r = 0; r < 8; ++r) o[r * 64 + t] = x[r];
        __syncthreads();   // the exchange buffer is reused by the next polynomial
    }
    if (UsesTm<FFT>::value) tm_free_cta(&tm_slot_s);
}

__global__ void k_ks_pack(const int32_t *__restrict__ rows, int32_t *__restrict__ out, int n1, int stride,
                          int base, size_t nrows_out) {
    // rows: [N*t][base][n1]  ->  out: [N*t][base-1][stride]
    const size_t total = nrows_out * (size_t)stride;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total;
         idx += (size_t)gridDim.x * blockDim.x) {
        const size_t row = idx / stride;
        const int col = (int)(idx - row * stride);
        const size_t ij = row / (base - 1);
        const int h = (int)(row - ij * (base - 1)) + 1;
        out[idx] = col < n1 ? rows[(ij * base + h) * (size_t)n1 + col] : 0;
    }
}

// ------------------------------------------------------------------------------------------------
// gate prologue coefficients: x = (0, cst) + s0*in_a + s1*in_b   (boot-gates.cpp:42,68,94,120,146,
// 205,231,257,283,309 and :343-355 for the two halves of MUX)
// ------------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ void gate_coeffs(int op, int half, int &s0, int &s1, int32_t &cst, int &second) {
    const int32_t e8 = 1 << 29, e4 = 1 << 30;
    second = 1;
    switch (op) {
        case B200_NAND:  s0 = -1; s1 = -1; cst = e8;  break;
        case B200_OR:    s0 = 1;  s1 = 1;  cst = e8;  break;
        case B200_AND:   s0 = 1;  s1 = 1;  cst = -e8; break;
        case B200_XOR:   s0 = 2;  s1 = 2;  cst = e4;  break;
        case B200_XNOR:  s0 = -2; s1 = -2; cst = -e4; break;
        case B200_NOR:   s0 = -1; s1 = -1; cst = -e8; break;
        case B200_ANDNY: s0 = -1; s1 = 1;  cst = -e8; break;
        case B200_ANDYN: s0 = 1;  s1 = -1; cst = -e8; break;
        case B200_ORNY:  s0 = -1; s1 = 1;  cst = e8;  break;
        case B200_ORYN:  s0 = 1;  s1 = -1; cst = e8;  break;
        case B200_MUX:
            if (half == 0) { s0 = 1; s1 = 1; cst = -e8; second = 1; }
            else           { s0 = -1; s1 = 1; cst = -e8; second = 2; }
            break;
        default:         s0 = 1;  s1 = 0;  cst = 0;   break;  // kOpRaw
    }
}

// ------------------------------------------------------------------------------------------------
// Bootstrapping-key row providers for cmux_step.  A "row pair" is the two evaluation-domain
// polynomials (c = 0, 1) of TGSW row rp = q*l + p of BK_i: 2 x 512 double2 = 16 KiB, contiguous.
//   GlobalRows : read straight from global memory / L2 (single-step test kernel)
//   RingRows   : a shared-memory ring of row pairs filled by TMA bulk copies (cp.async.bulk) that all
//                ciphertext groups of the CTA consume in lockstep, so one L2 read serves the whole CTA
//                and the L2 latency is hidden behind the previous transforms.
// ------------------------------------------------------------------------------------------------
struct GlobalRows {
    const double2 *bk_i;
    __device__ __forceinline__ const double2 *acquire(int rp) const { return bk_i + (size_t)rp * 2 * kM; }
    __device__ __forceinline__ double2 load(const double2 *p) const { return __ldg(p); }
    __device__ __forceinline__ void release(int) const {}
    __device__ __forceinline__ int release_begin(int) const { return 0; }
    __device__ __forceinline__ void release_end(int, int) const {}
    __device__ __forceinline__ void release_n(int, int) const {}
    __device__ __forceinline__ uint64_t release_n_begin(int) const { return 0; }
    __device__ __forceinline__ void release_n_end(int, int, uint64_t) const {}
};

constexpr int kStageBytes = 2 * kM * 16;  // one row pair

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    const uint32_t a = smem_u32(bar);
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(a), "r"(parity) : "memory");
    } while (!done);
}
// one thread: announce `bytes` on the barrier and start the bulk copy global -> shared that completes it
__device__ __forceinline__ void tma_load_row_pair(void *dst, const void *src, uint64_t *bar) {
    const uint32_t b = smem_u32(bar), d = smem_u32(dst);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(kStageBytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(d), "l"(src), "r"(kStageBytes), "r"(b) : "memory");
}

template <int STAGES>
struct RingState {                 // lives in shared memory
    uint64_t full[STAGES];         // TMA completion barriers
    union {
        uint64_t empty[STAGES];    // consumer barriers: one arrival per consumer warp and use of the stage (RingRows<.., true>)
        int done[STAGES];          // the same as a plain counter (round-1 experiment kernels)
    };
};
// One arrival on a consumer barrier; true for the arrival that completes the phase (the barrier then re-arms itself).
// The returned state is the barrier BEFORE the arrival, so a pending count of 1 means "I am the last".  The arrival
// is a release at CTA scope: no separate fence, and nobody resets a counter.
__device__ __forceinline__ bool mbar_arrive_is_last(uint64_t *bar) {
    uint64_t state;
    uint32_t pending;
    asm volatile("mbarrier.arrive.shared::cta.b64 %0, [%1];" : "=l"(state) : "r"(smem_u32(bar)) : "memory");
    asm volatile("mbarrier.pending_count.b64 %0, %1;" : "=r"(pending) : "l"(state));
    return pending == 1u;
}

#ifdef B200_BR_PROFILE
__device__ long long g_br_prof[12];
#endif
template <int STAGES, bool MBAR = false>   // MBAR: stages handed back through mbarriers (st->empty) instead of fence + atomic counter
struct RingRows {
    unsigned char *ring;           // [STAGES][kStageBytes]
    RingState<STAGES> *st;
    const double2 *bk;             // key.bkfft
    int seq0;                      // sequence number of row pair 0 of the current CMux step (= i * kpl)
    int total;                     // n * kpl
    int warps;                     // consumer warps per CTA
    bool lane0;                    // this thread does the warp's ring bookkeeping
    __device__ __forceinline__ const double2 *acquire(int rp) const {
        const int q = seq0 + rp;
        mbar_wait(&st->full[q % STAGES], (uint32_t)(q / STAGES) & 1u);
        return reinterpret_cast<const double2 *>(ring + (size_t)(q % STAGES) * kStageBytes);
    }
    __device__ __forceinline__ double2 load(const double2 *p) const { return *p; }
    // the calling warp has consumed row pair rp; the last warp of the CTA to say so refills the stage
    __device__ __forceinline__ void release(int rp) const {
        const int q = seq0 + rp, s = q % STAGES;
        __syncwarp();
        if (MBAR) {
            if (lane0 && mbar_arrive_is_last(&st->empty[s]) && q + STAGES < total)
                tma_load_row_pair(ring + (size_t)s * kStageBytes, bk + (size_t)(q + STAGES) * 2 * kM, &st->full[s]);
            return;
        }
        if (lane0) {
            __threadfence_block();
            if (atomicAdd(&st->done[s], 1) == warps - 1) {
                st->done[s] = 0;
                if (q + STAGES < total)
                    tma_load_row_pair(ring + (size_t)s * kStageBytes, bk + (size_t)(q + STAGES) * 2 * kM, &st->full[s]);
            }
        }
    }
    // n consecutive row pairs that every warp consumes together (the pair-wise CMux step): ONE count for all of
    // them -- the counter of the first -- and the last warp refills the n stages
    __device__ __forceinline__ void release_n(int rp, int n) const {
        const int q = seq0 + rp, s = q % STAGES;
        if (MBAR) {
            __syncwarp();
            if (lane0 && mbar_arrive_is_last(&st->empty[s]))
                for (int i = 0; i < n; ++i)
                    if (q + i + STAGES < total)
                        tma_load_row_pair(ring + (size_t)((q + i) % STAGES) * kStageBytes, bk + (size_t)(q + i + STAGES) * 2 * kM,
                                          &st->full[(q + i) % STAGES]);
            return;
        }
        __syncwarp();
        if (lane0) {
            __threadfence_block();
            if (atomicAdd(&st->done[s], 1) == warps - 1) {
                st->done[s] = 0;
                for (int i = 0; i < n; ++i)
                    if (q + i + STAGES < total)
                        tma_load_row_pair(ring + (size_t)((q + i) % STAGES) * kStageBytes, bk + (size_t)(q + i + STAGES) * 2 * kM,
                                          &st->full[(q + i) % STAGES]);
            }
        }
    }
    // release_n in two halves (MBAR): the arrival is issued at once, its answer -- am I the last? -- is looked at later,
    // so the barrier's round trip (lane 0 needs the old state; the whole warp would wait for lane 0 at the
    // reconvergence point) overlaps with the caller's next piece of work.
    __device__ __forceinline__ uint64_t release_n_begin(int rp) const {
        const int q = seq0 + rp, s = q % STAGES;
        uint64_t state = 0;
        __syncwarp();
        if (lane0) asm volatile("mbarrier.arrive.shared::cta.b64 %0, [%1];" : "=l"(state) : "r"(smem_u32(&st->empty[s])) : "memory");
        return state;
    }
    __device__ __forceinline__ void release_n_end(int rp, int n, uint64_t state) const {
        const int q = seq0 + rp;
        if (lane0) {
            uint32_t pending;
            asm volatile("mbarrier.pending_count.b64 %0, %1;" : "=r"(pending) : "l"(state));
            if (pending == 1u)
                for (int i = 0; i < n; ++i)
                    if (q + i + STAGES < total)
                        tma_load_row_pair(ring + (size_t)((q + i) % STAGES) * kStageBytes, bk + (size_t)(q + i + STAGES) * 2 * kM,
                                          &st->full[(q + i) % STAGES]);
        }
    }
    // The same in two halves, so that the shared-memory atomic's round trip (lane 0 needs the old count to know
    // whether it is the last warp; the whole warp waits for lane 0 at the reconvergence point) overlaps with the
    // caller's next piece of work: release_begin issues the atomic, release_end looks at its result.
    __device__ __forceinline__ int release_begin(int rp) const {
        const int q = seq0 + rp, s = q % STAGES;
        int old = -1;
        __syncwarp();
        if (lane0) {
            __threadfence_block();
            old = atomicAdd(&st->done[s], 1);
        }
        return old;
    }
    __device__ __forceinline__ void release_end(int rp, int old) const {
        const int q = seq0 + rp, s = q % STAGES;
        if (lane0 && old == warps - 1) {
            st->done[s] = 0;
            if (q + STAGES < total)
                tma_load_row_pair(ring + (size_t)s * kStageBytes, bk + (size_t)(q + STAGES) * 2 * kM, &st->full[s]);
        }
    }
    // a group that does nothing this step (barai == 0, or no ciphertext) still has to hand the stages on
    __device__ __forceinline__ void skip_step(int kpl) const {
        for (int rp = 0; rp < kpl; ++rp) { acquire(rp); release(rp); }
    }
    // the same in the event structure of the pair-wise step (release_n): digits (p, p+1) of each polynomial together
    __device__ __forceinline__ void skip_step_pairs(int l) const {
        for (int q = 0; q < 2; ++q)
            for (int p = 0; p < l; p += 2) {
                const int n = p + 1 < l ? 2 : 1, rp = q * l + p;
                for (int i = 0; i < n; ++i) acquire(rp + i);
                release_n(rp, n);
            }
    }
};

// ------------------------------------------------------------------------------------------------
// one CMux step on a shared-memory accumulator, executed by one 64-thread group
//   MODE 0: ACC <- ACC + BK_i (x) ((X^a - 1) ACC)      (lwe-bootstrapping-functions-fft.cpp:67-76)
//   MODE 1: ACC <- BK_i (x) ACC                         (tgsw-fft-operations.cpp:89-114)
// ------------------------------------------------------------------------------------------------
template <int L, int BGBIT, int MODE, class FFT, class Barrier, class Rows>
__device__ __forceinline__ void cmux_step(int32_t *ACC, double2 *bufs, const int a, const Rows &rows,
                                          const FFT &fft, const Barrier &bar) {
    constexpr int kBuf = FFT::kBufElemsPerTransform;
    const int t = fft.t;
    constexpr uint32_t BG = 1u << BGBIT, HALF = BG >> 1;
    uint32_t offset = 0;  // tgsw.cpp:7-29
#pragma unroll
    for (int p = 0; p < L; ++p) offset += HALF << (32 - (p + 1) * BGBIT);

    double2 acc0[8], acc1[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) acc0[r] = acc1[r] = make_double2(0.0, 0.0);

    int toggle = 0;
#pragma unroll 1
    for (int q = 0; q < 2; ++q) {
        const int32_t *P = ACC + q * kN;
        uint32_t dv[16];  // (rotated difference) + decomposition offset, coefficient t + 64 rr
#pragma unroll
        for (int rr = 0; rr < 16; ++rr) {
            const int j = t + 64 * rr;
            if (MODE == 0) {
                const int m = (j - a) & (2 * kN - 1);
                int32_t v = P[m & (kN - 1)];
                if (m & kN) v = -v;
                dv[rr] = (uint32_t)(v - P[j]) + offset;
            } else {
                dv[rr] = (uint32_t)P[j] + offset;
            }
        }
#pragma unroll 1
        for (int p = 0; p < L; ++p) {
            const int shift = 32 - (p + 1) * BGBIT;
            double2 x[8];
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                x[r] = make_double2(digit_to_double<BGBIT>(dv[r], shift), digit_to_double<BGBIT>(dv[r + 8], shift));
            }
            fft.forward(x, bufs + toggle * kBuf, bar);
            toggle ^= 1;
            const double2 *row = rows.acquire(q * L + p) + t;
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const double2 b0 = rows.load(row + r * 64);
                const double2 b1 = rows.load(row + kM + r * 64);
                acc0[r].x = fma(x[r].x, b0.x, acc0[r].x);
                acc0[r].x = fma(-x[r].y, b0.y, acc0[r].x);
                acc0[r].y = fma(x[r].x, b0.y, acc0[r].y);
                acc0[r].y = fma(x[r].y, b0.x, acc0[r].y);
                acc1[r].x = fma(x[r].x, b1.x, acc1[r].x);
                acc1[r].x = fma(-x[r].y, b1.y, acc1[r].x);
                acc1[r].y = fma(x[r].x, b1.y, acc1[r].y);
                acc1[r].y = fma(x[r].y, b1.x, acc1[r].y);
            }
            rows.release(q * L + p);
        }
    }
    fft.template inverse<true>(acc0, bufs + toggle * kBuf, bar);
    toggle ^= 1;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const int j = t + 64 * r;
        const double sc = c_tw.p1_scale[r];   // 1/512 times the cos factors the last pass left out
        const int32_t lo = round_scaled_to_torus(acc0[r].x, sc), hi = round_scaled_to_torus(acc0[r].y, sc);
        if (MODE == 0) { ACC[j] += lo; ACC[j + kM] += hi; }
        else           { ACC[j] = lo;  ACC[j + kM] = hi; }
    }
    fft.template inverse<true>(acc1, bufs + toggle * kBuf, bar);
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const int j = kN + t + 64 * r;
        const double sc = c_tw.p1_scale[r];
        const int32_t lo = round_scaled_to_torus(acc1[r].x, sc), hi = round_scaled_to_torus(acc1[r].y, sc);
        if (MODE == 0) { ACC[j] += lo; ACC[j + kM] += hi; }
        else           { ACC[j] = lo;  ACC[j + kM] = hi; }
    }
    bar();  // accumulator complete before the next step's rotated reads
}

// The same step with the transforms taken two at a time (FftTm::forward2 / inverse2): the digits p, p+1 of one
// accumulator polynomial go through the forward transform together, the two output columns through the inverse.
template <int L, int BGBIT, int MODE, class Barrier, class Rows>
__device__ __forceinline__ void cmux_step_pairs(int32_t *ACC, double2 *bufs, const int a, const Rows &rows,
                                                const FftTm &fft, const Barrier &bar) {
    constexpr uint32_t BG = 1u << BGBIT, HALF = BG >> 1;
    constexpr int kBuf = FftTm::kBufElemsPerTransform;
    const int t = fft.t;
    uint32_t offset = 0;  // tgsw.cpp:7-29
#pragma unroll
    for (int p = 0; p < L; ++p) offset += HALF << (32 - (p + 1) * BGBIT);

    double2 acc0[8], acc1[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) acc0[r] = acc1[r] = make_double2(0.0, 0.0);

    auto mac = [&](const double2 (&x)[8], int rp) {
        const double2 *row = rows.acquire(rp) + t;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            const double2 b0 = rows.load(row + r * 64);
            const double2 b1 = rows.load(row + kM + r * 64);
            acc0[r].x = fma(x[r].x, b0.x, acc0[r].x);
            acc0[r].x = fma(-x[r].y, b0.y, acc0[r].x);
            acc0[r].y = fma(x[r].x, b0.y, acc0[r].y);
            acc0[r].y = fma(x[r].y, b0.x, acc0[r].y);
            acc1[r].x = fma(x[r].x, b1.x, acc1[r].x);
            acc1[r].x = fma(-x[r].y, b1.y, acc1[r].x);
            acc1[r].y = fma(x[r].x, b1.y, acc1[r].y);
            acc1[r].y = fma(x[r].y, b1.x, acc1[r].y);
        }
        rows.release(rp);
    };

    bool first = true;   // the barrier that ended the previous step also freed the exchange buffers
#pragma unroll 1
    for (int q = 0; q < 2; ++q) {
        const int32_t *P = ACC + q * kN;
        uint32_t dv[16];  // (rotated difference) + decomposition offset, coefficient t + 64 rr
#pragma unroll
        for (int rr = 0; rr < 16; ++rr) {
            const int j = t + 64 * rr;
            if (MODE == 0) {
                const int m = (j - a) & (2 * kN - 1);
                int32_t v = P[m & (kN - 1)];
                if (m & kN) v = -v;
                dv[rr] = (uint32_t)(v - P[j]) + offset;
            } else {
                dv[rr] = (uint32_t)P[j] + offset;
            }
        }
#pragma unroll
        for (int p = 0; p < L; p += 2) {
            if (!first) bar();   // every thread of the group has read the previous transforms' exchange buffers
            first = false;
            const int shift = 32 - (p + 1) * BGBIT;
            double2 xa[8];
#pragma unroll
            for (int r = 0; r < 8; ++r)
                xa[r] = make_double2(digit_to_double<BGBIT>(dv[r], shift), digit_to_double<BGBIT>(dv[r + 8], shift));
            if (p + 1 < L) {
                double2 xb[8];
#pragma unroll
                for (int r = 0; r < 8; ++r)
                    xb[r] = make_double2(digit_to_double<BGBIT>(dv[r], shift - BGBIT), digit_to_double<BGBIT>(dv[r + 8], shift - BGBIT));
                fft.forward2(xa, xb, bufs, bufs + kBuf, bar);
                mac(xa, q * L + p);
                mac(xb, q * L + p + 1);
            } else {
                fft.forward(xa, bufs, bar);
                mac(xa, q * L + p);
            }
        }
    }
    bar();
    fft.template inverse2<true>(acc0, acc1, bufs, bufs + kBuf, bar);
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const int j = t + 64 * r;
        const double sc = c_tw.p1_scale[r];   // 1/512 times the cos factors the last pass left out
        const int32_t lo0 = round_scaled_to_torus(acc0[r].x, sc), hi0 = round_scaled_to_torus(acc0[r].y, sc);
        const int32_t lo1 = round_scaled_to_torus(acc1[r].x, sc), hi1 = round_scaled_to_torus(acc1[r].y, sc);
        if (MODE == 0) { ACC[j] += lo0; ACC[j + kM] += hi0; ACC[kN + j] += lo1; ACC[kN + j + kM] += hi1; }
        else           { ACC[j] = lo0;  ACC[j + kM] = hi0;  ACC[kN + j] = lo1;  ACC[kN + j + kM] = hi1; }
    }
    bar();  // accumulator complete before the next step's rotated reads; exchange buffers free
}
template <int L, int BGBIT, int MODE, class Barrier, class Rows>
__device__ __forceinline__ void cmux_step(int32_t *ACC, double2 *bufs, const int a, const Rows &rows,
                                          const FftTm &fft, const Barrier &bar) {
    cmux_step_pairs<L, BGBIT, MODE>(ACC, bufs, a, rows, fft, bar);
}

// The pair-wise step with the spectral accumulators in TENSOR MEMORY while the next pair's forward transforms run:
// the 64 accumulator registers exist only inside a MAC (half of them at a time) and inside the inverse transforms,
// so six groups per SM have about 40 registers per thread more for the transforms than with resident accumulators.
// One tcgen05.st + one tcgen05.ld of 256 bytes per thread per step at l = 2 (three of each at l = 3).
#ifdef B200_BR_PROFILE   // phase clocks of thread 0 of CTA 0 (scratch builds only; scratch/prof_br.sh)
#define BR_T(k) do { if (prof_on) { const long long now_ = clock64(); g_br_prof[k] += now_ - last_; last_ = now_; } } while (0)
#define BR_T_BEGIN const bool prof_on = blockIdx.x == 0 && threadIdx.x == 0; long long last_ = clock64()
#else
#define BR_T(k) do { } while (0)
#define BR_T_BEGIN do { } while (0)
#endif
template <int L, int BGBIT, int MODE, class Barrier, class Rows>
__device__ __forceinline__ void cmux_step_pairs_tmacc(int32_t *ACC, double2 *bufs, const int a, const Rows &rows,
                                                      const FftTmAcc &fft, const Barrier &bar) {
    constexpr uint32_t BG = 1u << BGBIT, HALF = BG >> 1;
    constexpr int kBuf = FftTm::kBufElemsPerTransform;
    constexpr int kEvents = 2 * ((L + 1) / 2);      // MAC events of one step (a pair of rows, or a single last row)
    const int t = fft.t;
    const uint32_t tm_acc = fft.tm + kTmColAcc;
    uint32_t offset = 0;  // tgsw.cpp:7-29
#pragma unroll
    for (int p = 0; p < L; ++p) offset += HALF << (32 - (p + 1) * BGBIT);

    double2 acc0[8], acc1[8];   // defined by the last MAC event only
    uint64_t pending_release = 0;   // state of the ring's consumer barrier before this warp's last arrival (lane 0)
    int pending_rp = 0, pending_n = 0;
    BR_T_BEGIN;

    // one half (registers 4h .. 4h+3) of a MAC event over NX rows: c += x . key
    auto mac_half = [&](double2 (&c0)[4], double2 (&c1)[4], const double2 (&x)[8], const double2 *row, int h) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = 4 * h + i;
            const double2 b0 = rows.load(row + r * 64);
            const double2 b1 = rows.load(row + kM + r * 64);
            c0[i].x = fma(x[r].x, b0.x, c0[i].x);
            c0[i].x = fma(-x[r].y, b0.y, c0[i].x);
            c0[i].y = fma(x[r].x, b0.y, c0[i].y);
            c0[i].y = fma(x[r].y, b0.x, c0[i].y);
            c1[i].x = fma(x[r].x, b1.x, c1[i].x);
            c1[i].x = fma(-x[r].y, b1.y, c1[i].x);
            c1[i].y = fma(x[r].x, b1.y, c1[i].y);
            c1[i].y = fma(x[r].y, b1.x, c1[i].y);
        }
    };
    // MAC event ev over rows rp (and rp + 1 when xb is given): acc (tensor memory, or zero for the first event)
    // += x . rows; the last event of the step leaves the result in acc0 / acc1.  Half of the registers at a time.
    auto mac_event = [&](const int ev, const double2 (&xa)[8], const double2 (*xb)[8], const int rp) {
        const double2 *rowa = rows.acquire(rp) + t;
        const double2 *rowb = xb ? rows.acquire(rp + 1) + t : nullptr;
        BR_T(7);
        if (ev > 0) tm_wait_st();
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            double2 c0[4], c1[4];
            if (ev == 0) {
#pragma unroll
                for (int i = 0; i < 4; ++i) c0[i] = c1[i] = make_double2(0.0, 0.0);
            } else {
                double d[16], unused[8];
                tm_load16<0>(d, unused, tm_acc + 32 * h, 0);
#pragma unroll
                for (int i = 0; i < 4; ++i) { c0[i] = make_double2(d[4 * i], d[4 * i + 1]); c1[i] = make_double2(d[4 * i + 2], d[4 * i + 3]); }
            }
            mac_half(c0, c1, xa, rowa, h);
            if (xb) mac_half(c0, c1, *xb, rowb, h);
            if (ev == kEvents - 1) {
#pragma unroll
                for (int i = 0; i < 4; ++i) { acc0[4 * h + i] = c0[i]; acc1[4 * h + i] = c1[i]; }
            } else {
                tm_store16(tm_acc + 32 * h, c0[0].x, c0[0].y, c1[0].x, c1[0].y, c0[1].x, c0[1].y, c1[1].x, c1[1].y,
                           c0[2].x, c0[2].y, c1[2].x, c1[2].y, c0[3].x, c0[3].y, c1[3].x, c1[3].y);
            }
        }
        BR_T(8);
        // the arrival goes out now; whether it was the last one (-> refill) is looked at after the next piece of work
        pending_release = rows.release_n_begin(rp);
        pending_rp = rp;
        pending_n = xb ? 2 : 1;
    };

#pragma unroll
    for (int q = 0; q < 2; ++q) {
        const int32_t *P = ACC + q * kN;
        uint32_t dv[16];  // (rotated difference) + decomposition offset, coefficient t + 64 rr
#pragma unroll
        for (int rr = 0; rr < 16; ++rr) {
            const int j = t + 64 * rr;
            if (MODE == 0) {
                const int m = (j - a) & (2 * kN - 1);
                int32_t v = P[m & (kN - 1)];
                if (m & kN) v = -v;
                dv[rr] = (uint32_t)(v - P[j]) + offset;
            } else {
                dv[rr] = (uint32_t)P[j] + offset;
            }
        }
#pragma unroll
        for (int p = 0; p < L; p += 2) {
            const int ev = q * ((L + 1) / 2) + p / 2;
            if (ev > 0) rows.release_n_end(pending_rp, pending_n, pending_release);
            // One after the other, the transforms of a pair alternate between the two exchange buffers and each has a
            // group barrier between its writes and its reads: whoever writes a buffer again has passed the barrier of
            // the transform in between, which every thread reaches only after its reads of that buffer -- no extra
            // barrier.  Only an odd l leaves a lone transform last, and its buffer is the next one written.
#ifdef B200_SKEWED_PAIRS
            if (ev > 0) bar();   // every thread of the group has read the previous transforms' exchange buffers
#else
            if ((L & 1) && q == 1 && p == 0) bar();
#endif
            const int shift = 32 - (p + 1) * BGBIT;
            // digits as exact doubles, registers 4..7 pre-twisted by the level-0 root on the integer side
            double2 xa[8];
            digits_pretwisted<BGBIT>(xa, dv, shift);
            if (p + 1 < L) {
                double2 xb[8];
                digits_pretwisted<BGBIT>(xb, dv, shift - BGBIT);
                BR_T(0);
#ifndef B200_SKEWED_PAIRS
                // one after the other: with three warps per scheduler the other warps hide the exchange latency, and
                // the simpler instruction stream schedules better than FftTm::forward2's skewed pair (308.9 -> 303.3 ms
                // with the inverse below; the pair was worth 3 % when the kernel still had 168 registers and spills)
                fft.template forward<true>(xa, bufs, bar);
                fft.template forward<true>(xb, bufs + kBuf, bar);
#else
                fft.template forward2<true>(xa, xb, bufs, bufs + kBuf, bar);
#endif
                BR_T(1);
                mac_event(ev, xa, &xb, q * L + p);
                BR_T(2);
            } else {
                fft.template forward<true>(xa, bufs, bar);
                mac_event(ev, xa, nullptr, q * L + p);
            }
        }
    }
    rows.release_n_end(pending_rp, pending_n, pending_release);
#ifdef B200_SKEWED_PAIRS
    bar();
#else
    if (L & 1) bar();
#endif
    BR_T(3);
#ifndef B200_SKEWED_PAIRS
    fft.template inverse<true>(acc0, bufs, bar);
    fft.template inverse<true>(acc1, bufs + kBuf, bar);
#else
    fft.template inverse2<true>(acc0, acc1, bufs, bufs + kBuf, bar);
#endif
    BR_T(4);
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const int j = t + 64 * r;
        const double sc = c_tw.p1_scale[r];   // 1/512 times the cos factors the last pass left out
        const int32_t lo0 = round_scaled_to_torus(acc0[r].x, sc), hi0 = round_scaled_to_torus(acc0[r].y, sc);
        const int32_t lo1 = round_scaled_to_torus(acc1[r].x, sc), hi1 = round_scaled_to_torus(acc1[r].y, sc);
        if (MODE == 0) { ACC[j] += lo0; ACC[j + kM] += hi0; ACC[kN + j] += lo1; ACC[kN + j + kM] += hi1; }
        else           { ACC[j] = lo0;  ACC[j + kM] = hi0;  ACC[kN + j] = lo1;  ACC[kN + j + kM] = hi1; }
    }
    BR_T(5);
    bar();  // accumulator complete before the next step's rotated reads; exchange buffers free
    BR_T(6);
}
template <int L, int BGBIT, int MODE, class Barrier, class Rows>
__device__ __forceinline__ void cmux_step(int32_t *ACC, double2 *bufs, const int a, const Rows &rows,
                                          const FftTmAcc &fft, const Barrier &bar) {
    cmux_step_pairs_tmacc<L, BGBIT, MODE>(ACC, bufs, a, rows, fft, bar);
}

// Single transforms with the spectral accumulators in tensor memory across EVERY key row (FftTmS): per row the two
// halves of the accumulators are loaded (not for the first row), updated and stored; the inverse transforms load
// one output column at a time.  About 100 live registers, one exchange buffer: eight groups per SM.
template <int L, int BGBIT, int MODE, class Barrier, class Rows>
__device__ __forceinline__ void cmux_step_single_tmacc(int32_t *ACC, double2 *buf, const int a, const Rows &rows,
                                                       const FftTmS &fft, const Barrier &bar) {
    constexpr uint32_t BG = 1u << BGBIT, HALF = BG >> 1;
    const int t = fft.t;
    const uint32_t tm_acc = fft.tm + kTmSColAcc;     // per half h: [c0 x 4 | c1 x 4] = 32 columns
    uint32_t offset = 0;  // tgsw.cpp:7-29
#pragma unroll
    for (int p = 0; p < L; ++p) offset += HALF << (32 - (p + 1) * BGBIT);

#pragma unroll
    for (int q = 0; q < 2; ++q) {
        const int32_t *P = ACC + q * kN;
        uint32_t dv[16];  // (rotated difference) + decomposition offset, coefficient t + 64 rr
#pragma unroll
        for (int rr = 0; rr < 16; ++rr) {
            const int j = t + 64 * rr;
            if (MODE == 0) {
                const int m = (j - a) & (2 * kN - 1);
                int32_t v = P[m & (kN - 1)];
                if (m & kN) v = -v;
                dv[rr] = (uint32_t)(v - P[j]) + offset;
            } else {
                dv[rr] = (uint32_t)P[j] + offset;
            }
        }
#pragma unroll
        for (int p = 0; p < L; ++p) {
            const int ev = q * L + p;
            if (ev > 0) bar();   // the partner warp has read the previous transform's exchange buffer
            const int shift = 32 - (p + 1) * BGBIT;
            double2 x[8];
            digits_pretwisted<BGBIT>(x, dv, shift);   // level 0 of the transform on the integer side
            fft.template forward<true>(x, buf, bar);
            const double2 *row = rows.acquire(ev) + t;
            if (ev > 0) tm_wait_st();
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                double2 c0[4], c1[4];
                if (ev == 0) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) c0[i] = c1[i] = make_double2(0.0, 0.0);
                } else {
                    double d[16], unused[8];
                    tm_load16<0>(d, unused, tm_acc + 32 * h, 0);
#pragma unroll
                    for (int i = 0; i < 4; ++i) { c0[i] = make_double2(d[2 * i], d[2 * i + 1]); c1[i] = make_double2(d[8 + 2 * i], d[9 + 2 * i]); }
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int r = 4 * h + i;
                    const double2 b0 = rows.load(row + r * 64);
                    const double2 b1 = rows.load(row + kM + r * 64);
                    c0[i].x = fma(x[r].x, b0.x, c0[i].x);
                    c0[i].x = fma(-x[r].y, b0.y, c0[i].x);
                    c0[i].y = fma(x[r].x, b0.y, c0[i].y);
                    c0[i].y = fma(x[r].y, b0.x, c0[i].y);
                    c1[i].x = fma(x[r].x, b1.x, c1[i].x);
                    c1[i].x = fma(-x[r].y, b1.y, c1[i].x);
                    c1[i].y = fma(x[r].x, b1.y, c1[i].y);
                    c1[i].y = fma(x[r].y, b1.x, c1[i].y);
                }
                tm_store16(tm_acc + 32 * h, c0[0].x, c0[0].y, c0[1].x, c0[1].y, c0[2].x, c0[2].y, c0[3].x, c0[3].y,
                           c1[0].x, c1[0].y, c1[1].x, c1[1].y, c1[2].x, c1[2].y, c1[3].x, c1[3].y);
            }
            rows.release(ev);
        }
    }
    tm_wait_st();
#pragma unroll
    for (int c = 0; c < 2; ++c) {
        double2 acc[8];
        {
            double d[16], unused[8];
            // column c of both halves: 16 columns at +16 c of each half
            asm volatile("{\n\t.reg .b32 u<16>, v<16>;\n\t" TM_LD_L16("u", "%16") TM_LD_L16("v", "%17")
                         "tcgen05.wait::ld.sync.aligned;\n\t"
                         TM_PACK8("%0", "%1", "%2", "%3", "%4", "%5", "%6", "%7", "u") TM_PACK8("%8", "%9", "%10", "%11", "%12", "%13", "%14", "%15", "v") "}"
                         : "=d"(d[0]), "=d"(d[1]), "=d"(d[2]), "=d"(d[3]), "=d"(d[4]), "=d"(d[5]), "=d"(d[6]), "=d"(d[7]),
                           "=d"(d[8]), "=d"(d[9]), "=d"(d[10]), "=d"(d[11]), "=d"(d[12]), "=d"(d[13]), "=d"(d[14]), "=d"(d[15])
                         : "r"(tm_acc + 16 * c), "r"(tm_acc + 32 + 16 * c) : "memory");
            (void)unused;
#pragma unroll
            for (int r = 0; r < 8; ++r) acc[r] = make_double2(d[2 * r], d[2 * r + 1]);
        }
        bar();   // exchange buffer free
        fft.template inverse<true>(acc, buf, bar);
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            const int j = c * kN + t + 64 * r;
            const double sc = c_tw.p1_scale[r];   // 1/512 times the cos factors the last pass left out
            const int32_t lo = round_scaled_to_torus(acc[r].x, sc), hi = round_scaled_to_torus(acc[r].y, sc);
            if (MODE == 0) { ACC[j] += lo; ACC[j + kM] += hi; }
            else           { ACC[j] = lo;  ACC[j + kM] = hi; }
        }
    }
    bar();  // accumulator complete before the next step's rotated reads; exchange buffer free
}
template <int L, int BGBIT, int MODE, class Barrier, class Rows>
__device__ __forceinline__ void cmux_step(int32_t *ACC, double2 *bufs, const int a, const Rows &rows,
                                          const FftTmS &fft, const Barrier &bar) {
    cmux_step_single_tmacc<L, BGBIT, MODE>(ACC, bufs, a, rows, fft, bar);
}

// per ciphertext group: ACC + 2 exchange buffers + bara
template <class FFT> __host__ __device__ constexpr int group_smem_bytes() {
    return 2 * kN * 4 + FFT::kExchangeBuffers * FFT::kBufElemsPerTransform * 16 + 1280;
}
constexpr int kGroupSmemBytes = 2 * kN * 4 + 2 * kBufElems * 16 + 1280;   // FftSmem (also the experiments' kernels)
constexpr int kBrStages = 4;
template <class FFT> __host__ __device__ constexpr int br_smem_bytes(int groups, int stages) {
    return groups * group_smem_bytes<FFT>() + stages * kStageBytes + 128;
}

#ifndef BR_MAXREG
#define BR_MAXREG 255
#endif
// G ciphertexts per CTA (64 threads each), walking i = 0..n-1 in lockstep on one TMA-fed ring of
// bootstrapping-key row pairs.  Groups synchronise among themselves with named barriers only.
// FFT = FftSmem: both exchanges of a transform through shared memory (round 1, "v2");
//       FftTm:   the two warp-local exchanges and the per-thread roots in tensor memory ("tm", negacyclic_fft_tm.cuh).
template <int L, int BGBIT, int G, int STAGES, class FFT>
__global__ void __launch_bounds__(64 * G, 1) k_blind_rotate(const DeviceKey key, const BlindRotateArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int KPL = 2 * L;
    unsigned char *ring = smem;                                                       // [stages][16 KiB]
    RingState<STAGES> *st = reinterpret_cast<RingState<STAGES> *>(smem + STAGES * kStageBytes);
    // the group index is the same in every lane of a warp; the reduction tells the compiler, so that everything derived
    // from it (accumulator / buffer / tensor-memory bases, barrier ids, the group's item) is uniform-datapath work
    const int g = (int)__reduce_max_sync(0xffffffffu, threadIdx.x >> 6), t = threadIdx.x & 63;
    unsigned char *mine = smem + STAGES * kStageBytes + 128 + (size_t)g * group_smem_bytes<FFT>();
    int32_t *ACC = reinterpret_cast<int32_t *>(mine);                                // [2][1024]
    double2 *bufs = reinterpret_cast<double2 *>(mine + 2 * kN * 4);                  // [2][576 | 512]
    uint16_t *bara = reinterpret_cast<uint16_t *>(mine + 2 * kN * 4 + FFT::kExchangeBuffers * FFT::kBufElemsPerTransform * 16);  // [n+1]
    const double2 *bkfft = UsesTm<FFT>::value ? key.bkfft_tm : key.bkfft;
    uint32_t *tm_slot_p = reinterpret_cast<uint32_t *>(smem + STAGES * kStageBytes + 120);   // behind the ring state
    uint32_t tm = 0;
    static_assert(sizeof(RingState<STAGES>) <= 120, "ring state: 120 bytes before the tensor-memory slot");
    static_assert(!UsesTm<FFT>::value || TmShared<FFT>::value + ((2 * G + 3) / 4) * TmWindow<FFT>::value <= 512, "tensor-memory columns of the warps sharing a lane quarter");
    if (UsesTm<FFT>::value) tm = tm_alloc_cta(tm_slot_p, TmWindow<FFT>::value, TmShared<FFT>::value);

    const int n = key.n, n1 = n + 1;
    const int bmin = min((int)blockIdx.x, a.bal_r);
    const int64_t item = a.bal_q ? (int64_t)blockIdx.x * a.bal_q + bmin + g : (int64_t)blockIdx.x * G + g;
    const bool live = a.bal_q ? g < a.bal_q + ((int)blockIdx.x < a.bal_r) : item < a.items;
    const int total = n * KPL;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(&st->full[s], 1); mbar_init(&st->empty[s], 2 * G); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        for (int s = 0; s < STAGES && s < total; ++s)
            tma_load_row_pair(ring + (size_t)s * kStageBytes, bkfft + (size_t)s * 2 * kM, &st->full[s]);
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
    FFT fft;
    if constexpr (TmShared<FFT>::value != 0) fft.init(t, tm, (tm & 0xffff0000u) | (*tm_slot_p & 0xffffu));   // roots: columns 0.. of the lane quarter
    else fft.init(t, tm);
    const NamedBarrier64 bar{1 + g};
    bar();
    if (a.stagger > 0) {
        const long long until = clock64() + (long long)a.stagger * (g >> 1);
        while (clock64() < until) {}
    }

    // ---- blind rotation (lwe-bootstrapping-functions-fft.cpp:88-112)
    int lane_id = threadIdx.x & 31;
    asm volatile("" : "+r"(lane_id));   // keep it in a register: the compiler otherwise re-reads SR_TID in the loop
    RingRows<STAGES, true> rows{ring, st, bkfft, 0, total, 2 * G, lane_id == 0};
#pragma unroll 1
    for (int i = 0; i < n; ++i) {
        rows.seq0 = i * KPL;
        // the same value in every lane; the reduction makes that visible to the compiler, so the branch is uniform, the
        // step body is convergent code and warp-uniform constants (pass-1 roots) may live in uniform registers there
        const int ai = (int)__reduce_or_sync(0xffffffffu, (unsigned)(live ? bara[i] : 0));
        if (ai == 0) {
            if (PairEvents<FFT>::value) rows.skip_step_pairs(L);
            else rows.skip_step(KPL);
            continue;
        }
        cmux_step<L, BGBIT, 0>(ACC, bufs, ai, rows, fft, bar);
    }

    // ---- sample extract at index 0 (lwe.cpp:41-60)
    if (live) {
        int32_t *u = a.u + (size_t)item * (kN + 1);
#pragma unroll
        for (int rr = 0; rr < 16; ++rr) {
            const int j = t + 64 * rr;
            u[j] = (j == 0) ? ACC[0] : -ACC[kN - j];
        }
        if (t == 0) u[kN] = ACC[kN];
    }
    if (UsesTm<FFT>::value) tm_free_cta(tm_slot_p);
#ifdef B200_BR_PROFILE
    if (blockIdx.x == 0 && threadIdx.x == 0 && PairEvents<FFT>::value) {
        printf("tmacc group 0, clk per step (n = %d): decompose+digits %lld forward2 %lld mac event %lld | bar %lld inverse2 %lld round+update %lld bar %lld || inside mac event: acquire %lld compute %lld (release = rest: syncwarp %lld fence %lld atomic %lld)\n",
               n, g_br_prof[0] / n, g_br_prof[1] / n, g_br_prof[2] / n, g_br_prof[3] / n, g_br_prof[4] / n, g_br_prof[5] / n, g_br_prof[6] / n, g_br_prof[7] / n, g_br_prof[8] / n, g_br_prof[9] / n, g_br_prof[10] / n, g_br_prof[11] / n);
        for (int k = 0; k < 12; ++k) g_br_prof[k] = 0;
    }
#endif
}

}  // namespace b200
#ifdef B200_EXPERIMENTS
#include "blind_rotate16.cuh"
#include "blind_rotate3.cuh"
#include "blind_rotate4.cuh"
#include "blind_rotate5.cuh"
#endif
#include "blind_rotate_lat.cuh"
#include "blind_rotate_lat2.cuh"
namespace b200 {

// FftSmem: several CTAs per SM; FftTm: one CTA per SM owns the tensor memory (512 columns), so the CTA walks a slice
// of the accumulators
template <int L, int BGBIT, class FFT>
__global__ void __launch_bounds__(64, UsesTm<FFT>::value ? 1 : 6) k_cmux_test(const DeviceKey key, int32_t *acc, int bk_index,
                                                                            int barai, int mode, int64_t count) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint32_t tm_slot_s;
    int32_t *ACC = reinterpret_cast<int32_t *>(smem);
    double2 *bufs = reinterpret_cast<double2 *>(smem + 2 * kN * 4);
    const int t = threadIdx.x;
    uint32_t tm = 0;
    if (UsesTm<FFT>::value) tm = tm_alloc_cta(&tm_slot_s);
    FFT fft;
    fft.init(t, tm);
    const GlobalRows rows{(UsesTm<FFT>::value ? key.bkfft_tm : key.bkfft) + (size_t)bk_index * (2 * L * 2 * kM)};
    for (int64_t item = blockIdx.x; item < count; item += gridDim.x) {
        int32_t *g = acc + (size_t)item * 2 * kN;
        for (int j = t; j < 2 * kN; j += 64) ACC[j] = g[j];
        __syncthreads();
        if (mode == 0) {
            if (barai != 0) cmux_step<L, BGBIT, 0>(ACC, bufs, barai, rows, fft, CtaBarrier());
        } else {
            cmux_step<L, BGBIT, 1>(ACC, bufs, 0, rows, fft, CtaBarrier());
        }
        __syncthreads();
        for (int j = t; j < 2 * kN; j += 64) g[j] = ACC[j];
        __syncthreads();
    }
    if (UsesTm<FFT>::value) tm_free_cta(&tm_slot_s);
}

// The transform alone, for the parity tests that the reference keeps for its FFT processors
// (libs/tfhe/src/test/lagrangehalfc_test.cpp:16-36 inverse(forward) = identity, :130-218 product = Karatsuba):
//   mode 0: out = round(inverse(forward(a)) / 512)                      (int32 polynomials mod X^1024 + 1)
//   mode 1: out = round(inverse(forward(a) * forward(b)) / 512) mod 2^32 (a: small digits, b: Torus32)
template <class FFT>
__global__ void __launch_bounds__(64, UsesTm<FFT>::value ? 1 : 6) k_fft_test(const int32_t *__restrict__ a, const int32_t *__restrict__ b,
                                                                           int32_t *__restrict__ out, int mode, int64_t count) {
    __shared__ double2 buf[FFT::kBufElemsPerTransform];
    __shared__ uint32_t tm_slot_s;
    const int t = threadIdx.x;
    uint32_t tm = 0;
    if (UsesTm<FFT>::value) tm = tm_alloc_cta(&tm_slot_s);
    FFT fft;
    fft.init(t, tm);
    for (int64_t item = blockIdx.x; item < count; item += gridDim.x) {
        const int32_t *pa = a + (size_t)item * kN;
        double2 x[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) x[r] = make_double2(int_to_double(pa[t + 64 * r]), int_to_double(pa[t + 64 * r + kM]));
        fft.forward(x, buf, CtaBarrier());
        __syncthreads();
        if (mode == 1) {
            const int32_t *pb = b + (size_t)item * kN;
            double2 y[8];
#pragma unroll
            for (int r = 0; r < 8; ++r) y[r] = make_double2(int_to_double(pb[t + 64 * r]), int_to_double(pb[t + 64 * r + kM]));
            fft.forward(y, buf, CtaBarrier());
            __syncthreads();
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const double re = fma(x[r].x, y[r].x, -x[r].y * y[r].y), im = fma(x[r].x, y[r].y, x[r].y * y[r].x);
                x[r] = make_double2(re, im);
            }
        }
        fft.template inverse<false>(x, buf, CtaBarrier());
        __syncthreads();
        int32_t *po = out + (size_t)item * kN;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            // products reach 2^50: round to the nearest integer in 64 bits, then reduce mod 2^32
            po[t + 64 * r] = (int32_t)(uint32_t)(unsigned long long)__double2ll_rn(x[r].x * (1.0 / kM));
            po[t + 64 * r + kM] = (int32_t)(uint32_t)(unsigned long long)__double2ll_rn(x[r].y * (1.0 / kM));
        }
    }
    if (UsesTm<FFT>::value) tm_free_cta(&tm_slot_s);
}

// ------------------------------------------------------------------------------------------------
// key switch, base 4 (basebit 2): tile of TB ciphertexts per CTA, one int4 column group per thread,
// the three non-zero rows of (i,j) are loaded once per tile and selected by each ciphertext's digit.
// grid = (tiles, i-splits); with more than one split the partial sums are combined with integer
// atomics (wrap-around addition is associative, so the result is bit-exact in any order).
// ------------------------------------------------------------------------------------------------
constexpr int kKsTile = 8;

// MASKED (one or two tiles, i.e. the lone bootsNAND): only the rows some sample of the tile selects are loaded -- a
// lone sample needs a quarter of the 24 rows of a coefficient.  Full tiles need nearly all of them and the predicated
// loads cost more than they save (592 samples: 0.89 against 0.82 ms), so wider batches load unconditionally.
template <int T, bool MASKED>
__global__ void __launch_bounds__(160) k_keyswitch(const DeviceKey key, const KeySwitchArgs a, int i_per_split, int nsplit) {
    extern __shared__ __align__(128) unsigned char smem[];
    uint16_t *dig = reinterpret_cast<uint16_t *>(smem);  // [kKsTile][i_per_split]
    uint32_t *need = reinterpret_cast<uint32_t *>(smem + (((size_t)kKsTile * i_per_split * 2 + 15) & ~(size_t)15));  // [i_per_split]
    __shared__ int nz[kKsTile];
    __shared__ int32_t bterm[kKsTile];
    __shared__ int32_t outrow[kKsTile];

    const int Nin = key.N * key.k, n1 = key.n + 1, stride = key.ks_stride;
    const int64_t item0 = (int64_t)blockIdx.x * kKsTile;
    const int split = blockIdx.y;
    const int i0 = split * i_per_split;
    const int icnt = min(i_per_split, Nin - i0);
    const uint32_t prec_offset = 1u << (32 - (1 + 2 * T));

    if (threadIdx.x < kKsTile) {
        const int g = threadIdx.x;
        nz[g] = 0;
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
        bterm[g] = b;
        outrow[g] = orow;
    }
    if (MASKED)
        for (int ii = threadIdx.x; ii < icnt; ii += blockDim.x) need[ii] = 0;
    __syncthreads();
    // digit pre-pass (lwe-keyswitch-functions.cpp:122-125): top 2T bits of a_i + prec_offset
    for (int idx = threadIdx.x; idx < kKsTile * icnt; idx += blockDim.x) {
        const int g = idx / icnt, ii = idx - g * icnt;
        const int64_t item = item0 + g;
        uint32_t packed = 0;
        if (item < a.items) {
            int64_t u0, u1 = -1;
            if (a.desc) { const KsDesc d = a.desc[item]; u0 = d.u0; u1 = d.u1; }
            else if (a.pair) { u0 = 2 * item; u1 = 2 * item + 1; }
            else u0 = item;
            uint32_t v = (uint32_t)a.u[u0 * (Nin + 1) + i0 + ii];
            if (u1 >= 0) v += (uint32_t)a.u[u1 * (Nin + 1) + i0 + ii];
            packed = (v + prec_offset) >> (32 - 2 * T);
            const int c = __popc((packed | (packed >> 1)) & 0x5555u);
            if (c && a.nonzero) atomicAdd(&nz[g], c);
            if (MASKED) {   // rows of this coefficient that some sample of the tile selects: bit 3 j + (d - 1)
                uint32_t m = 0;
#pragma unroll
                for (int j = 0; j < T; ++j) {
                    const uint32_t d = (packed >> (2 * (T - 1 - j))) & 3u;
                    if (d) m |= 1u << (3 * j + d - 1);
                }
                if (m) atomicOr(&need[ii], m);
            }
        }
        dig[g * i_per_split + ii] = (uint16_t)packed;
    }
    __syncthreads();

    const int col4 = threadIdx.x;
    const int stride4 = stride >> 2;
    if (col4 < stride4) {
        int4 acc[kKsTile];
#pragma unroll
        for (int g = 0; g < kKsTile; ++g) acc[g] = make_int4(0, 0, 0, 0);
        const int4 *tab = reinterpret_cast<const int4 *>(key.ks) + (size_t)i0 * T * 3 * stride4 + col4;
#pragma unroll 1
        for (int ii = 0; ii < icnt; ++ii) {
            int4 r[T][3];
            const uint32_t nm = MASKED ? need[ii] : 0xffffffu;   // CTA-uniform
#pragma unroll
            for (int j = 0; j < T; ++j)
#pragma unroll
                for (int h = 0; h < 3; ++h)
                    if (nm & (1u << (3 * j + h))) r[j][h] = __ldg(tab + (size_t)((ii * T + j) * 3 + h) * stride4);
#pragma unroll
            for (int g = 0; g < kKsTile; ++g) {
                const uint32_t pk = dig[g * i_per_split + ii];
#pragma unroll
                for (int j = 0; j < T; ++j) {
                    const uint32_t d = (pk >> (2 * (T - 1 - j))) & 3u;
                    if (d) {  // warp-uniform
                        const int4 v = d == 1 ? r[j][0] : (d == 2 ? r[j][1] : r[j][2]);
                        acc[g].x -= v.x; acc[g].y -= v.y; acc[g].z -= v.z; acc[g].w -= v.w;
                    }
                }
            }
        }
#pragma unroll
        for (int g = 0; g < kKsTile; ++g) {
            const int orow = outrow[g];
            if (orow < 0) continue;
            int32_t *o = a.out + (size_t)orow * n1;
            const int32_t vals[4] = {acc[g].x, acc[g].y, acc[g].z, acc[g].w};
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int col = col4 * 4 + c;
                if (col < n1) {
                    int32_t v = vals[c];
                    if (col == key.n && split == 0) v += bterm[g];
                    if (nsplit == 1) o[col] = v;
                    else atomicAdd(o + col, v);
                }
            }
        }
    }
    if (a.nonzero && threadIdx.x < kKsTile && item0 + threadIdx.x < a.items && nz[threadIdx.x])
        atomicAdd(a.nonzero + item0 + threadIdx.x, nz[threadIdx.x]);
}

}  // namespace b200
#include "keyswitch_tma.cuh"
namespace b200 {

// zero the destination rows of a descriptor-driven key switch (needed when splits use atomics)
__global__ void k_zero_rows(int32_t *out, const KsDesc *desc, int64_t items, int n1) {
    const int64_t item = blockIdx.x;
    if (item >= items) return;
    const int64_t row = desc ? desc[item].out : item;
    for (int c = threadIdx.x; c < n1; c += blockDim.x) out[row * n1 + c] = 0;
}

// ------------------------------------------------------------------------------------------------
// linear gates (boot-gates.cpp:164-189): NOT = negate, COPY, CONSTANT = trivial +-1/8
// ------------------------------------------------------------------------------------------------
__global__ void k_linear(int32_t *out, const int32_t *in, int sign, int32_t cst, int n1, int64_t total) {
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int col = (int)(idx % n1);
        int32_t v = in ? sign * in[idx] : 0;
        if (col == n1 - 1) v += cst;
        out[idx] = v;
    }
}
__global__ void k_linear_desc(int32_t *pool, const int32_t *desc, int n1) {
    const int32_t *d = desc + (size_t)blockIdx.x * 4;
    const int op = d[0];
    int32_t *o = pool + (size_t)d[1] * n1;
    const int32_t *in = d[2] >= 0 ? pool + (size_t)d[2] * n1 : nullptr;
    for (int c = threadIdx.x; c < n1; c += blockDim.x) {
        int32_t v;
        if (op == B200_NOT) v = -in[c];
        else if (op == B200_COPY) v = in[c];
        else v = (c == n1 - 1) ? (op == B200_CONST1 ? (1 << 29) : -(1 << 29)) : 0;
        o[c] = v;
    }
}

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
#define DISPATCH_L_BG(FN, ...)                                                      \
    do {                                                                            \
        if (key.l == 2 && key.Bgbit == 10) { FN<2, 10> __VA_ARGS__; }               \
        else if (key.l == 3 && key.Bgbit == 7) { FN<3, 7> __VA_ARGS__; }            \
        else if (key.l == 2 && key.Bgbit == 8) { FN<2, 8> __VA_ARGS__; }            \
        else if (key.l == 1 && key.Bgbit == 16) { FN<1, 16> __VA_ARGS__; }          \
        else if (key.l == 4 && key.Bgbit == 6) { FN<4, 6> __VA_ARGS__; }            \
        else return cudaErrorInvalidValue;                                          \
    } while (0)

template <int L, int BGBIT, class FFT>
static cudaError_t set_br_attr() {
    static bool done[16] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 16 && done[dev]) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(k_cmux_test<L, BGBIT, FFT>, cudaFuncAttributeMaxDynamicSharedMemorySize, group_smem_bytes<FFT>());
    if (e != cudaSuccess) return e;
    if (dev < 16) done[dev] = true;
    return cudaSuccess;
}

static int sm_count_of(int dev);
cudaError_t launch_bk_convert(const DeviceKey &key, const int32_t *d_bk_coefs, cudaStream_t s) {
    const int polys = key.n * (key.k + 1) * key.l * (key.k + 1);
    k_bk_convert<FftSmem><<<polys, 64, 0, s>>>(d_bk_coefs, key.bkfft, polys);
    COUNT_LAUNCH();
    // the tensor-memory transform has its own evaluation order, hence its own copy of the key; a CTA of that
    // kernel owns an SM's tensor memory, so the grid is a few CTAs per SM walking the polynomials
    int dev = 0;
    cudaGetDevice(&dev);
    k_bk_convert<FftTm><<<sm_count_of(dev) * 4, 64, 0, s>>>(d_bk_coefs, key.bkfft_tm, polys);
    COUNT_LAUNCH();
#ifdef B200_EXPERIMENTS
    k_bk_convert16<<<polys, 32, 0, s>>>(d_bk_coefs, key.bkfft16);
    COUNT_LAUNCH();
#endif
    return cudaGetLastError();
}

cudaError_t launch_ks_pack(const DeviceKey &key, const int32_t *d_ks_rows, cudaStream_t s) {
    const int base = 1 << key.ks_basebit;
    const size_t nrows = (size_t)key.N * key.k * key.ks_t * (base - 1);
    k_ks_pack<<<148 * 8, 256, 0, s>>>(d_ks_rows, key.ks, key.n + 1, key.ks_stride, base, nrows);
    COUNT_LAUNCH();
    return cudaGetLastError();
}

// Ciphertexts per CTA: up to 4 or 6 (one CTA per SM, one key read per CTA) as soon as the batch has more than
// one ciphertext per SM -- spread evenly over whole waves of CTAs, see below; 1 only when the latency
// kernel is switched off (B200_LAT=0).
static int sm_count_of(int dev) {
    static int cached[16] = {0};
    if (dev < 16 && cached[dev]) return cached[dev];
    int v = 148;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    if (dev < 16) cached[dev] = v;
    return v;
}
static bool br_balance() {
    static const bool on = [] { const char *e = getenv("B200_BALANCE"); return !(e && e[0] == '0'); }();
    return on;
}
template <int L, int BGBIT, int G, int STAGES, class FFT = FftSmem>
static cudaError_t br_launch_st(const DeviceKey &key, const BlindRotateArgs &a, cudaStream_t s) {
    static bool done[16] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    constexpr int smem_bytes = br_smem_bytes<FFT>(G, STAGES);
    static_assert(smem_bytes <= 227 * 1024, "shared memory budget of one CTA");
    if (!(dev < 16 && done[dev])) {
        cudaError_t e = cudaFuncSetAttribute(k_blind_rotate<L, BGBIT, G, STAGES, FFT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
        if (e != cudaSuccess) return e;
        cudaFuncSetAttribute(k_blind_rotate<L, BGBIT, G, STAGES, FFT>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        if (dev < 16) done[dev] = true;
    }
    unsigned grid = (unsigned)((a.items + G - 1) / G);
    BlindRotateArgs b = a;
    {
        static int stagger = -1;
        if (stagger < 0) { const char *ev = getenv("B200_STAGGER"); stagger = ev ? atoi(ev) : 0; }
        b.stagger = stagger;
    }
    if (G > 1 && br_balance()) {
        // whole waves of one CTA per SM, items spread evenly over them: the last wave then runs CTAs
        // with fewer ciphertexts (which finish sooner) on every SM instead of full CTAs on a few
        const int64_t sms = sm_count_of(dev);
        const int64_t even = (grid + sms - 1) / sms * sms;
        if (even <= a.items) {
            grid = (unsigned)even;
            b.bal_q = (int)(a.items / even);
            b.bal_r = (int)(a.items % even);
        }
    }
    k_blind_rotate<L, BGBIT, G, STAGES, FFT><<<grid, 64 * G, smem_bytes, s>>>(key, b);
    return cudaSuccess;
}
// Ring depth: 3 stages for the 6-ciphertext CTA, 7 for the 4-ciphertext CTA (everything that fits in 227 KB),
// 2 stages for the one-ciphertext CTA of small batches so that three of them fit on an SM.
template <int L, int BGBIT, int G>
static cudaError_t br_launch(const DeviceKey &key, const BlindRotateArgs &a, cudaStream_t s) {
    if constexpr (G == 6) return br_launch_st<L, BGBIT, G, 3>(key, a, s);
    else if constexpr (G == 4) return br_launch_st<L, BGBIT, G, 7>(key, a, s);
    else return br_launch_st<L, BGBIT, G, 2>(key, a, s);
}
// the tensor-memory transform needs 25.9 KB per group instead of 27.9: one more ring stage with six groups
template <int L, int BGBIT, int G, class FFT = FftTm>
static cudaError_t br_launch_tm(const DeviceKey &key, const BlindRotateArgs &a, cudaStream_t s) {
    if constexpr (G == 6) return br_launch_st<L, BGBIT, G, 4, FFT>(key, a, s);
    else if constexpr (G == 5) return br_launch_st<L, BGBIT, G, 6, FFT>(key, a, s);
    else return br_launch_st<L, BGBIT, G, 7, FFT>(key, a, s);
}
// Kernel choice: 0 = 64 threads per ciphertext (default), 1 = warp per ciphertext with TMEM accumulators
// (B200_BR_KERNEL=w16 or b200_set_blind_rotate_kernel(1)).
static std::atomic<int> g_br_variant{-1};
void set_blind_rotate_variant(int v) { g_br_variant.store(v); }
bool experiments_built() {
#ifdef B200_EXPERIMENTS
    return true;
#else
    return false;
#endif
}
// 0 "v2": shared-memory transforms; 1..8: round-1 experiments; 9 "tm", 10 "tmacc", 11 "tms": tensor-memory
// transforms (round 2); 12 "auto" (default): tmacc where it is the fastest (l = 2, more than 4 ciphertexts per SM),
// v2 and the latency kernel elsewhere
constexpr int kVarAuto = 12;
static int br_variant_raw() {
    int v = g_br_variant.load();
    if (v < 0) {
        const char *e = getenv("B200_BR_KERNEL");
        v = (e && e[0] == 'w') ? 1 : (e && e[0] >= '0' && e[0] <= '9') ? atoi(e) : kVarAuto;
        g_br_variant.store(v);
    }
    return v;
}
static bool br_tm_g_forced() {
    static const bool forced = getenv("B200_TM_G") != nullptr;
    return forced;
}
// the variant that decides everything but the wide l = 2 launch, and whether that launch is the tmacc kernel
static int br_variant() {
    const int v = br_variant_raw();
    return (v == kVarAuto || (v == 10 && !br_tm_g_forced())) ? 0 : v;
}
static bool br_wide_tmacc() {
    const int v = br_variant_raw();
    return v == kVarAuto || (v == 10 && !br_tm_g_forced());
}
#ifdef B200_EXPERIMENTS
template <int L, int BGBIT, int G>
static cudaError_t br16_launch(const DeviceKey &key, const BlindRotateArgs &a, cudaStream_t s) {
    static bool done[16] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!(dev < 16 && done[dev])) {
        cudaError_t e = cudaFuncSetAttribute(k_blind_rotate16<L, BGBIT, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, br16_smem_bytes(G));
        if (e != cudaSuccess) return e;
        cudaFuncSetAttribute(k_blind_rotate16<L, BGBIT, G>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        if (dev < 16) done[dev] = true;
    }
    const unsigned grid = (unsigned)((a.items + G - 1) / G);
    k_blind_rotate16<L, BGBIT, G><<<grid, 32 * G, br16_smem_bytes(G), s>>>(key, a);
    return cudaSuccess;
}
#endif
cudaError_t launch_blind_rotate(const DeviceKey &key, const BlindRotateArgs &a, cudaStream_t s) {
    if (a.items <= 0) return cudaSuccess;
    if (a.items > 0x7fffffffLL || key.n + 1 > 640) return cudaErrorInvalidValue;
    int cur_dev = 0;
    cudaGetDevice(&cur_dev);
    const int64_t sms = sm_count_of(cur_dev);   // every threshold below is in units of "ciphertexts per SM"
    {
        // at most one ciphertext per SM: the latency kernel spreads each CMux step over 2l groups
        static int lat = -1;
        if (lat < 0) { const char *e = getenv("B200_LAT"); lat = e ? atoi(e) : 1; }
        if (lat && a.items <= sms && br_variant() == 0) {
            cudaError_t el = cudaErrorInvalidValue;
            // at most one ciphertext per SM pair: a cluster of two CTAs per ciphertext (blind_rotate_lat2.cuh).
            // B200_LAT=2: never; 3: with the tensor-memory transform (A/B: 0.918 ms against 0.896 ms per rotation at
            // n = 500 -- with two transforms per SM the shared-memory pipe is idle enough for the plain exchanges)
            if (lat != 2 && 2 * a.items <= sms) {
                if (lat == 3) {
                    if (key.l == 2 && key.Bgbit == 10) el = lat2_launch<2, 10, FftTm>(key, a, s);
                    else if (key.l == 3 && key.Bgbit == 7) el = lat2_launch<3, 7, FftTm>(key, a, s);
                } else {
                    if (key.l == 2 && key.Bgbit == 10) el = lat2_launch<2, 10, FftSmem>(key, a, s);
                    else if (key.l == 3 && key.Bgbit == 7) el = lat2_launch<3, 7, FftSmem>(key, a, s);
                    else if (key.l == 2 && key.Bgbit == 8) el = lat2_launch<2, 8, FftSmem>(key, a, s);
                }
                if (el == cudaSuccess) { COUNT_LAUNCH(); return cudaGetLastError(); }
                if (el != cudaErrorInvalidValue) return el;
            }
            if (key.l == 2 && key.Bgbit == 10) el = lat_launch<2, 10>(key, a, s);
            else if (key.l == 3 && key.Bgbit == 7) el = lat_launch<3, 7>(key, a, s);
            else if (key.l == 2 && key.Bgbit == 8) el = lat_launch<2, 8>(key, a, s);
            if (el == cudaSuccess) { COUNT_LAUNCH(); return cudaGetLastError(); }
            if (el != cudaErrorInvalidValue) return el;
        }
    }
#ifdef B200_EXPERIMENTS
    if (br_variant() == 8 && a.items >= sms * 4 && key.l == 2 && (key.Bgbit == 10 || key.Bgbit == 8)) {
        // v5: two warps per ciphertext, each a 16-point-per-thread transform engine owning one polynomial
        // (small batches stay on v2: a lone v5 ciphertext takes 5.6 us per step against 4.3 us)
        cudaError_t e5;
        if (key.Bgbit == 10) e5 = br5_launch<10, 4>(key, a, s);
        else e5 = br5_launch<8, 4>(key, a, s);
        if (e5 != cudaSuccess) return e5;
        COUNT_LAUNCH();
        return cudaGetLastError();
    }
    if (br_variant() == 7 && a.items >= sms * 4) {
        // v4: v2 with the key delivered through tensor memory (tcgen05.cp), one shared-memory read per CTA
        cudaError_t e4;
        if (key.l == 2 && key.Bgbit == 10) e4 = br4_launch<2, 10, 4>(key, a, s);
        else if (key.l == 3 && key.Bgbit == 7) e4 = br4_launch<3, 7, 4>(key, a, s);
        else return cudaErrorInvalidValue;
        if (e4 != cudaSuccess) return e4;
        COUNT_LAUNCH();
        return cudaGetLastError();
    }
    if (br_variant() >= 2 && br_variant() <= 6 && a.items >= sms * 4) {
        // v3: accumulators in tensor memory; variant = 2: 6 groups x 1 ciphertext, 3: 4 groups x 2 ciphertexts,
        // 4: 5 x 1, 5: 4 x 1, 6: 3 x 2  (small batches keep the one-group-per-CTA latency path below)
        const int v = br_variant();
        cudaError_t e3 = cudaErrorInvalidValue;
#define BR3_LAUNCH(L_, B_)                                                  \
        e3 = v == 2 ? br3_launch<L_, B_, 6, 1, 3>(key, a, s)                \
           : v == 3 ? br3_launch<L_, B_, 4, 2, 4>(key, a, s)                \
           : v == 4 ? br3_launch<L_, B_, 5, 1, 4>(key, a, s)                \
           : v == 5 ? br3_launch<L_, B_, 4, 1, 4>(key, a, s)                \
           :          br3_launch<L_, B_, 3, 2, 4>(key, a, s)
        if (key.l == 2 && key.Bgbit == 10) BR3_LAUNCH(2, 10);
        else if (key.l == 3 && key.Bgbit == 7) BR3_LAUNCH(3, 7);
        else return cudaErrorInvalidValue;
#undef BR3_LAUNCH
        if (e3 != cudaSuccess) return e3;
        COUNT_LAUNCH();
        return cudaGetLastError();
    }
    if (br_variant() == 1) {
        // 8 ciphertexts per CTA (one CTA per SM, one key read per 8 ciphertexts) once the batch fills
        // the GPU that way, otherwise one warp per CTA so that small batches spread over all SMs.
        const bool wide16 = a.items >= sms * 4;
        cudaError_t e16;
#define BR16_LAUNCH(L_, B_) e16 = wide16 ? br16_launch<L_, B_, 8>(key, a, s) : br16_launch<L_, B_, 1>(key, a, s)
        if (key.l == 2 && key.Bgbit == 10) BR16_LAUNCH(2, 10);
        else if (key.l == 3 && key.Bgbit == 7) BR16_LAUNCH(3, 7);
        else if (key.l == 2 && key.Bgbit == 8) BR16_LAUNCH(2, 8);
        else if (key.l == 1 && key.Bgbit == 16) BR16_LAUNCH(1, 16);
        else return cudaErrorInvalidValue;
#undef BR16_LAUNCH
        if (e16 != cudaSuccess) return e16;
        COUNT_LAUNCH();
        return cudaGetLastError();
    }
#endif
    if (br_variant() == 11 && a.items > sms * 4 && key.l == 2 && key.Bgbit == 10) {
        // "tms": single transforms, accumulators in tensor memory for every row, 6 / 7 / 8 groups per SM
        static int tmg = -1;
        if (tmg < 0) { const char *ev = getenv("B200_TM_G"); tmg = ev ? atoi(ev) : 8; }
        cudaError_t et = tmg == 6 ? br_launch_st<2, 10, 6, 7, FftTmS>(key, a, s)
                       : tmg == 7 ? br_launch_st<2, 10, 7, 6, FftTmS>(key, a, s)
                                  : br_launch_st<2, 10, 8, 5, FftTmS>(key, a, s);
        if (et != cudaSuccess) return et;
        COUNT_LAUNCH();
        return cudaGetLastError();
    }
    if (br_variant() == 10 && a.items > sms * 4) {
        // "tmacc": the tm kernel with the spectral accumulators parked in tensor memory between row pairs
        cudaError_t et;
        static int tmg = -1;
        if (tmg < 0) { const char *ev = getenv("B200_TM_G"); tmg = ev ? atoi(ev) : 6; }
        if (key.l == 2 && key.Bgbit == 10 && tmg == 4) et = br_launch_tm<2, 10, 4, FftTmAcc>(key, a, s);
        else if (key.l == 2 && key.Bgbit == 10 && tmg == 5) et = br_launch_tm<2, 10, 5, FftTmAcc>(key, a, s);
        else if (key.l == 2 && key.Bgbit == 10) et = br_launch_tm<2, 10, 6, FftTmAcc>(key, a, s);
        else if (key.l == 3 && key.Bgbit == 7 && tmg == 4) et = br_launch_tm<3, 7, 4, FftTmAcc>(key, a, s);
        else if (key.l == 3 && key.Bgbit == 7 && tmg == 5) et = br_launch_tm<3, 7, 5, FftTmAcc>(key, a, s);
        else if (key.l == 3 && key.Bgbit == 7) et = br_launch_tm<3, 7, 6, FftTmAcc>(key, a, s);
        else if (key.l == 2 && key.Bgbit == 8) et = br_launch_tm<2, 8, 6, FftTmAcc>(key, a, s);
        else return cudaErrorInvalidValue;
        if (et != cudaSuccess) return et;
        COUNT_LAUNCH();
        return cudaGetLastError();
    }
    if (br_variant() == 9 && a.items > sms * 4) {
        // "tm": the warp-local exchanges of every transform and the per-thread roots in tensor memory
        cudaError_t et;
        static int tmg = -1;   // ciphertexts per CTA at l = 2 (A/B knob: B200_TM_G=4|5|6)
        if (tmg < 0) { const char *ev = getenv("B200_TM_G"); tmg = ev ? atoi(ev) : 6; }
        if (key.l == 2 && key.Bgbit == 10 && tmg == 5) et = br_launch_tm<2, 10, 5>(key, a, s);
        else if (key.l == 2 && key.Bgbit == 10 && tmg == 4) et = br_launch_tm<2, 10, 4>(key, a, s);
        else if (key.l == 2 && key.Bgbit == 10) et = br_launch_tm<2, 10, 6>(key, a, s);
        else if (key.l == 3 && key.Bgbit == 7) et = br_launch_tm<3, 7, 4>(key, a, s);
        else if (key.l == 2 && key.Bgbit == 8) et = br_launch_tm<2, 8, 6>(key, a, s);
        else return cudaErrorInvalidValue;
        if (et != cudaSuccess) return et;
        COUNT_LAUNCH();
        return cudaGetLastError();
    }
    // More than one ciphertext per SM: multi-ciphertext CTAs on a balanced grid (br_launch_st).  Measured at
    // l = 2 (ms, G=1 x 3 CTAs per SM / G=4 / G=6): 296 items 2.39 / 2.29 / 2.66; 591: 4.69 / 3.44 / 3.73;
    // 600: - / 5.49 / 4.74; 887: - / 6.48 / 4.92.
    const bool wide = a.items > sms;
    cudaError_t e;
    // l = 2: six groups per CTA (12 warps, 168 registers, 3-stage ring) are 1.9 % faster than four (8 warps,
    // 246 registers, 7 stages); at l = 3 they are 3 % slower (636.6 vs 617.2 ms per 65,536).  B200_G6=0: four.
    static int g6 = -1;
    if (g6 < 0) { const char *ev = getenv("B200_G6"); g6 = ev ? atoi(ev) : 1; }
#define BR_LAUNCH(L_, B_) e = wide ? br_launch<L_, B_, 4>(key, a, s) : br_launch<L_, B_, 1>(key, a, s)
    if (g6 && a.items > sms * 4 && key.l == 2 && key.Bgbit == 10) {
        // A wave of the six-group kernel takes about as long with four groups per CTA as with six (4.74 vs
        // 4.92 ms), a wave of the four-group kernel 3.44 ms (2.3 ms at two per CTA) and the latency kernel
        // 1.35 ms: when what is left after whole six-group waves fits one narrower wave, it goes there.
        int dev = 0;
        cudaGetDevice(&dev);
        const int64_t wave6 = (int64_t)sm_count_of(dev) * 6, rest = a.items % wave6;
        if (br_balance() && a.items > wave6 && rest > 0 && rest <= (int64_t)sm_count_of(dev) * 4) {
            const int64_t head = a.items - rest;
            BlindRotateArgs h = a, r = a;
            h.items = head;
            r.items = rest;
            r.u += (size_t)head * (kN + 1);
            if (a.desc) r.desc += head;
            else {
                const size_t rows = (size_t)(a.op == B200_MUX ? head / 2 : head) * (key.n + 1);   // wave6 is even
                if (r.in0) r.in0 += rows;
                if (r.in1) r.in1 += rows;
                if (r.in2) r.in2 += rows;
            }
            e = br_wide_tmacc() ? br_launch_tm<2, 10, 6, FftTmAcc>(key, h, s) : br_launch<2, 10, 6>(key, h, s);
            if (e != cudaSuccess) return e;
            COUNT_LAUNCH();
            return launch_blind_rotate(key, r, s);
        }
        e = br_wide_tmacc() ? br_launch_tm<2, 10, 6, FftTmAcc>(key, a, s) : br_launch<2, 10, 6>(key, a, s);
    } else
    if (key.l == 2 && key.Bgbit == 10) BR_LAUNCH(2, 10);
    else if (key.l == 3 && key.Bgbit == 7) {
        // l = 3 in 'auto': four groups per CTA with the tensor-memory transform and parked accumulators (259.0 ms per
        // 29,600 rotations; without parking, "tm", 261.5; shared-memory transform 265.1; six groups 282)
        if (br_variant_raw() == kVarAuto && a.items > sms * 4) e = br_launch_tm<3, 7, 4, FftTmAcc>(key, a, s);
        else BR_LAUNCH(3, 7);
    }
    else if (key.l == 2 && key.Bgbit == 8) BR_LAUNCH(2, 8);
    else if (key.l == 1 && key.Bgbit == 16) BR_LAUNCH(1, 16);
    else if (key.l == 4 && key.Bgbit == 6) BR_LAUNCH(4, 6);
    else return cudaErrorInvalidValue;
#undef BR_LAUNCH
    if (e != cudaSuccess) return e;
    COUNT_LAUNCH();
    return cudaGetLastError();
}

cudaError_t launch_cmux_test(const DeviceKey &key, int32_t *d_acc, int bk_index, int barai, int mode,
                             int64_t count, cudaStream_t s) {
    if (count <= 0) return cudaSuccess;
#ifdef B200_EXPERIMENTS
    if (br_variant() == 1) {
#define CT16_LAUNCH(L_, B_)                                                                                      \
    do {                                                                                                         \
        cudaError_t e = cudaFuncSetAttribute(k_cmux_test16<L_, B_>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                             2 * kN * 4 + kBuf16Elems * 16);                                    \
        if (e != cudaSuccess) return e;                                                                          \
        k_cmux_test16<L_, B_><<<(unsigned)count, 32, 2 * kN * 4 + kBuf16Elems * 16, s>>>(key, d_acc, bk_index, barai, mode); \
    } while (0)
        if (key.l == 2 && key.Bgbit == 10) CT16_LAUNCH(2, 10);
        else if (key.l == 3 && key.Bgbit == 7) CT16_LAUNCH(3, 7);
        else if (key.l == 2 && key.Bgbit == 8) CT16_LAUNCH(2, 8);
        else if (key.l == 1 && key.Bgbit == 16) CT16_LAUNCH(1, 16);
        else return cudaErrorInvalidValue;
#undef CT16_LAUNCH
        COUNT_LAUNCH();
        return cudaGetLastError();
    }
#endif
    if ((br_variant_raw() >= 9 && br_variant_raw() <= 11) ||
        (br_variant_raw() == kVarAuto && ((key.l == 2 && key.Bgbit == 10) || (key.l == 3 && key.Bgbit == 7)))) {
        int dev = 0;
        cudaGetDevice(&dev);
        const unsigned grid = (unsigned)std::min<int64_t>(count, sm_count_of(dev));
#define CTT_LAUNCH(L_, B_)                                                                \
    do {                                                                                  \
        cudaError_t e = set_br_attr<L_, B_, FftTm>();                                     \
        if (e != cudaSuccess) return e;                                                   \
        k_cmux_test<L_, B_, FftTm><<<grid, 64, group_smem_bytes<FftTm>(), s>>>(key, d_acc, bk_index, barai, mode, count); \
    } while (0)
        if (key.l == 2 && key.Bgbit == 10) CTT_LAUNCH(2, 10);
        else if (key.l == 3 && key.Bgbit == 7) CTT_LAUNCH(3, 7);
        else if (key.l == 2 && key.Bgbit == 8) CTT_LAUNCH(2, 8);
        else return cudaErrorInvalidValue;
#undef CTT_LAUNCH
        COUNT_LAUNCH();
        return cudaGetLastError();
    }
#define CT_LAUNCH(L_, B_)                                                                 \
    do {                                                                                  \
        cudaError_t e = set_br_attr<L_, B_, FftSmem>();                                   \
        if (e != cudaSuccess) return e;                                                   \
        k_cmux_test<L_, B_, FftSmem><<<(unsigned)count, 64, group_smem_bytes<FftSmem>(), s>>>(key, d_acc, bk_index, barai, mode, count); \
    } while (0)
    if (key.l == 2 && key.Bgbit == 10) CT_LAUNCH(2, 10);
    else if (key.l == 3 && key.Bgbit == 7) CT_LAUNCH(3, 7);
    else if (key.l == 2 && key.Bgbit == 8) CT_LAUNCH(2, 8);
    else if (key.l == 1 && key.Bgbit == 16) CT_LAUNCH(1, 16);
    else if (key.l == 4 && key.Bgbit == 6) CT_LAUNCH(4, 6);
    else return cudaErrorInvalidValue;
#undef CT_LAUNCH
    COUNT_LAUNCH();
    return cudaGetLastError();
}

cudaError_t launch_fft_test(const int32_t *d_a, const int32_t *d_b, int32_t *d_out, int mode, int64_t count, cudaStream_t s) {
    if (count <= 0) return cudaSuccess;
    int dev = 0;
    cudaGetDevice(&dev);
    const int v = br_variant_raw();
    if ((v >= 9 && v <= 11) || v == kVarAuto)
        k_fft_test<FftTm><<<(unsigned)std::min<int64_t>(count, sm_count_of(dev)), 64, 0, s>>>(d_a, d_b, d_out, mode, count);
    else
        k_fft_test<FftSmem><<<(unsigned)std::min<int64_t>(count, 148 * 6), 64, 0, s>>>(d_a, d_b, d_out, mode, count);
    COUNT_LAUNCH();
    return cudaGetLastError();
}

static std::atomic<int> g_ks_variant{-1};
void set_keyswitch_variant(int v) { g_ks_variant.store(v); }
cudaError_t launch_keyswitch(const DeviceKey &key, const KeySwitchArgs &a, cudaStream_t s) {
    if (a.items <= 0) return cudaSuccess;
    if (key.ks_basebit != 2 || key.ks_t != 8) return cudaErrorInvalidValue;  // both reference sets: t=8, base 4
    const int Nin = key.N * key.k;
    int cur_dev = 0;
    cudaGetDevice(&cur_dev);
    const int64_t sms = sm_count_of(cur_dev);
    // large batches: table streamed once per 64 (48) samples through a TMA-fed shared-memory ring
    int ks_sel = g_ks_variant.load();   // 0: by batch size, 1: gather kernel, 2: streaming kernel
    if (ks_sel < 0) { const char *e = getenv("B200_KS"); ks_sel = e ? atoi(e) : 0; g_ks_variant.store(ks_sel); }
    const int stride4 = key.ks_stride / 4;
    if (ks_sel != 1 && Nin % kKs2Chunk == 0 && stride4 <= 192) {
        const bool small = stride4 <= 128;
        // 16 warps per SM, 8 samples x 2 int4 column groups per thread (measured best of the shapes tried:
        // 16 samples per group -> 8 warps, 43 ms; digit pairs with a 16-way branch -> 34-49 ms; this -> 26 ms).
        // Batches that cannot fill 148 CTAs of 64 samples use CTAs of 32 (two per SM) down to 2,960 samples (below that the gather kernel is as fast).
        const int64_t per_cta = small ? 64 : 40;
        const bool full = a.items >= sms * per_cta;
        // (with the coefficient range split up to 16 ways the streaming kernel beats the gather kernel from about a
        // hundred samples on: 592 samples 0.82 -> 0.2 ms; below that the gather kernel's 256-way split wins)
        static const int64_t ks_min = [] { const char *m = getenv("B200_KS_MIN"); return (int64_t)(m ? atoi(m) : 96); }();
        if (ks_sel == 2 || a.items >= ks_min) {
            // Row choice (B200_KS_MODE): 4 (default) = one brx.idx jump table per sample and DIGIT PAIR, both rows leaving
            // in one three-input add; 0 = uniform 3-way branch per sample and digit; 1 / 2 = one 16-way switch per sample
            // pair / digit pair (nvcc lowers it to a compare tree); 3 = sample pairs through a jump table.  All
            // bit-identical; 65,536 samples at the 80-bit set: 23.5 / 24.8 / 28.1 / 26.7 / 24.7 ms.
            static const int ks_mode = [] { const char *m = getenv("B200_KS_MODE"); return m ? atoi(m) : 4; }();
            cudaError_t e;
            // Batches that leave SMs (or second CTA slots) empty cut the coefficient range instead: nsplit CTAs per
            // sample tile, partial sums combined with integer atomics in zeroed rows (4,096 samples: 2.94 -> 1.6 ms)
            int nsplit = 1;
            if (!full) {
                const int64_t ctas = (a.items + (small ? 32 : 24) - 1) / (small ? 32 : 24), slots = small ? 2 * sms : sms;
                static const int ks_split = [] { const char *m = getenv("B200_KS_SPLIT"); return m ? atoi(m) : 16; }();
                nsplit = (int)std::max<int64_t>(1, std::min<int64_t>(ks_split, slots / ctas));
                if (nsplit > 1) {
                    k_zero_rows<<<(unsigned)a.items, 128, 0, s>>>(a.out, a.desc, a.items, key.n + 1);
                    COUNT_LAUNCH();
                }
            }
            if (small && stride4 == 126) {                // 80-bit set: row stride known at compile time
                if (!full) e = ks_mode == 4 ? ks2_launch<64, 4, 8, 2, 4, 126>(key, a, s, nsplit) : ks2_launch<64, 4, 8, 2, 0, 126>(key, a, s, nsplit);
                else
                e = ks_mode == 1 ? ks2_launch<64, 8, 8, 2, 1, 126>(key, a, s)
                  : ks_mode == 2 ? ks2_launch<64, 8, 8, 2, 2, 126>(key, a, s)
                  : ks_mode == 3 ? ks2_launch<64, 8, 8, 2, 3, 126>(key, a, s)
                  : ks_mode == 4 ? ks2_launch<64, 8, 8, 2, 4, 126>(key, a, s)
                                 : ks2_launch<64, 8, 8, 2, 0, 126>(key, a, s);
            } else if (small)
                e = full ? ks2_launch<64, 8, 8, 2>(key, a, s) : ks2_launch<64, 4, 8, 2>(key, a, s, nsplit);
            else if (stride4 == 158) {                    // 128-bit set
                if (ks_mode == 4) e = full ? ks2_launch<96, 5, 8, 2, 4, 158>(key, a, s) : ks2_launch<96, 3, 8, 2, 4, 158>(key, a, s, nsplit);
                else e = full ? ks2_launch<96, 5, 8, 2, 0, 158>(key, a, s) : ks2_launch<96, 3, 8, 2, 0, 158>(key, a, s, nsplit);
            } else
                e = full ? ks2_launch<96, 5, 8, 2>(key, a, s) : ks2_launch<96, 3, 8, 2>(key, a, s, nsplit);
            if (e != cudaSuccess) return e;
            COUNT_LAUNCH();
            return cudaGetLastError();
        }
    }
    const int64_t tiles = (a.items + kKsTile - 1) / kKsTile;
    // enough CTAs to fill 148 SMs several times over; small batches split the i range instead
    int nsplit = 1;
    while (tiles * nsplit < sms * 4 && nsplit < (tiles <= 2 ? 256 : 64)) nsplit *= 2;   // a lone gate: 4 coefficients per CTA
    const int i_per_split = (Nin + nsplit - 1) / nsplit;
    const int threads = ((key.ks_stride / 4) + 31) / 32 * 32;
    if (threads > 160) return cudaErrorInvalidValue;
    if (nsplit > 1) {
        k_zero_rows<<<(unsigned)a.items, 128, 0, s>>>(a.out, a.desc, a.items, key.n + 1);
        COUNT_LAUNCH();
    }
    const size_t smem = (((size_t)kKsTile * i_per_split * sizeof(uint16_t) + 15) & ~(size_t)15) + (size_t)i_per_split * sizeof(uint32_t);
    dim3 grid((unsigned)tiles, nsplit);
    if (tiles <= 2) k_keyswitch<8, true><<<grid, threads, smem, s>>>(key, a, i_per_split, nsplit);
    else k_keyswitch<8, false><<<grid, threads, smem, s>>>(key, a, i_per_split, nsplit);
    COUNT_LAUNCH();
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// Gate batch without any allocation (b200_gate_batch_dev_noalloc; no counterpart in the reference, added
// for callers that cannot call cudaMalloc: graph capture, tight serving loops).
// The extracted samples that travel from the blind rotation to the key switch -- the workspace b200_gate_batch_dev
// keeps per key and grows with cudaMalloc -- live in a file-scope device array of fixed size instead: every GPU
// that loads the library holds kNoAllocMaxRotations x (N + 1) words (16.8 MB) until exit.
// In host code `g_noalloc_u` is only a stand-in for the symbol, not a device address: the address each device
// gives it is looked up once per device with cudaGetSymbolAddress (that device current) and passed to the kernels.
// Every call shares the array and it keeps the previous call's contents.  Nothing in it has to start at zero:
// the blind rotation writes all N + 1 words of rows 0..rotations-1 before the key switch, later on the same
// stream, reads exactly those rows.  Calls that overlap on the GPU (different streams, no ordering) race on it.
// ------------------------------------------------------------------------------------------------
static __device__ __align__(256) int32_t g_noalloc_u[(size_t)kNoAllocMaxRotations * (kN + 1)];

static cudaError_t noalloc_scratch(int device, int32_t **out) {
    static std::mutex mu;
    static int32_t *addr[64] = {nullptr};
    if (device < 0 || device >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> g(mu);
    if (!addr[device]) {
        int cur = -1;
        cudaError_t e = cudaGetDevice(&cur);
        if (e != cudaSuccess) return e;
        if (cur != device) return cudaErrorInvalidDevice;     // the symbol resolves on the CURRENT device
        int32_t *p = nullptr;
        e = cudaGetSymbolAddress((void **)&p, g_noalloc_u);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        addr[device] = p;
    }
    *out = addr[device];
    return cudaSuccess;
}

cudaError_t launch_gates_noalloc(const DeviceKey &key, int op, int32_t *d_out, const int32_t *d_a, const int32_t *d_b,
                                 const int32_t *d_c, int64_t count, cudaStream_t s) {
    const int per = op == B200_MUX ? 2 : 1;
    if (count <= 0) return cudaSuccess;
    if (key.N * key.k != kN || count * per > kNoAllocMaxRotations) return cudaErrorInvalidValue;
    int32_t *u = nullptr;
    cudaError_t e = noalloc_scratch(key.device, &u);
    if (e != cudaSuccess) return e;
    BlindRotateArgs br{};
    br.in0 = d_a;
    br.in1 = d_b;
    br.in2 = d_c;
    br.op = op;
    br.mu = 1 << 29;
    br.u = u;
    br.items = count * per;
    e = launch_blind_rotate(key, br, s);
    if (e != cudaSuccess) return e;
    KeySwitchArgs ks{};
    ks.u = u;
    ks.pair = per == 2;
    ks.out = d_out;
    ks.items = count;
    return launch_keyswitch(key, ks, s);
}

cudaError_t launch_linear(const DeviceKey &key, int32_t *d_out, const int32_t *d_in, int sign, int32_t cst,
                          int64_t count, cudaStream_t s) {
    if (count <= 0) return cudaSuccess;
    const int64_t total = count * (key.n + 1);
    const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
    k_linear<<<blocks, 256, 0, s>>>(d_out, d_in, sign, cst, key.n + 1, total);
    COUNT_LAUNCH();
    return cudaGetLastError();
}

cudaError_t launch_linear_desc(const DeviceKey &key, int32_t *pool, const int32_t *d_desc, int64_t items, cudaStream_t s) {
    if (items <= 0) return cudaSuccess;
    k_linear_desc<<<(unsigned)items, 128, 0, s>>>(pool, d_desc, key.n + 1);
    COUNT_LAUNCH();
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// FP64 FMA peak of this GPU (roofline denominator of the blind rotation): 8 independent DFMA
// chains per thread, 256 threads x 8 CTAs per SM, timed with CUDA events.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_fp64_peak(double *sink, int iters, double seed) {
    double a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double m = 1.0 - 1e-9, c = 1e-9;
#pragma unroll 1
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
            a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
        }
    }
    const double r = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
    if (r == 123.456) sink[0] = r;
}
cudaError_t measure_fp64_peak(double *tflops) {
    double *sink = nullptr;
    cudaError_t e = cudaMalloc(&sink, 8);
    if (e != cudaSuccess) return e;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int blocks = 148 * 8, iters = 4096;
    double best = 0;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        k_fp64_peak<<<blocks, 256>>>(sink, iters, 1.0);
        COUNT_LAUNCH();
        cudaEventRecord(e1);
        e = cudaEventSynchronize(e1);
        if (e != cudaSuccess) break;
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flop = 2.0 * 64.0 * iters * 256.0 * blocks;
        if (rep > 0 && ms > 0) best = fmax(best, flop / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    *tflops = best;
    return e;
}

// int32 add issue peak (roofline denominator of the key switch, which is a stream of 32-bit subtractions):
// 16 accumulators per thread, each subtraction taking another accumulator as its operand so that nothing folds
// (the first version subtracted a loop-invariant value four times and the compiler turned that into one IMAD).
__global__ void __launch_bounds__(256) k_int_add_peak(int32_t *sink, int iters, int32_t seed) {
    int32_t a[16];
#pragma unroll
    for (int r = 0; r < 16; ++r) a[r] = seed + threadIdx.x * 17 + r;
#pragma unroll 1
    for (int i = 0; i < iters; ++i) {
        // every subtrahend is another accumulator, so no two subtractions can be folded into one multiply-add
#pragma unroll
        for (int k = 0; k < 4; ++k) {
#pragma unroll
            for (int r = 0; r < 16; ++r) a[r] -= a[(r + 5) & 15];
        }
    }
    int32_t r = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) r ^= a[k];
    if (r == 0x7fffffff) sink[0] = r;
}
cudaError_t measure_int_add_peak(double *tera_adds_per_s) {
    int32_t *sink = nullptr;
    cudaError_t e = cudaMalloc(&sink, 4);
    if (e != cudaSuccess) return e;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int blocks = 148 * 8, iters = 4096;
    double best = 0;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        k_int_add_peak<<<blocks, 256>>>(sink, iters, 12345 + rep);
        COUNT_LAUNCH();
        cudaEventRecord(e1);
        e = cudaEventSynchronize(e1);
        if (e != cudaSuccess) break;
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        const double adds = 64.0 * iters * 256.0 * blocks;
        if (rep > 0 && ms > 0) best = fmax(best, adds / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    *tera_adds_per_s = best;
    return e;
}

}  // namespace b200

namespace b200 {
void b200_gate_coeffs(int op, int half, int *s0, int *s1, int32_t *cst, int *second) {
    gate_coeffs(op, half, *s0, *s1, *cst, *second);
}
}  // namespace b200
