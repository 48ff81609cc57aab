// blind_rotate_lat2.cuh -- latency kernel for very small batches: ONE ciphertext per CLUSTER OF TWO CTAs (two SMs),
// included by kernels.cu.  Used for at most one ciphertext per SM pair: the drop-in bootsNAND(LweSample*...) call.
//
// k_blind_rotate_lat (one CTA per ciphertext) is bound by one SM's shared-memory pipe (68 % busy, 2l + 2 transforms
// plus 2l partial spectra per step through it) and by the 2l forward transforms sharing one FP64 pipe.  Here the
// two accumulator polynomials of the TLWE sample live on different SMs:
//   * CTA q (cluster rank) owns ACC[q].  Its l groups of 64 threads rotate and decompose ACC[q] (digit p in group p),
//     transform the digit and multiply it with key row (q, p) of BK_i (tgsw-fft-operations.cpp:89-114): one partial
//     spectrum for its own output polynomial q and one for the other CTA's;
//   * the l partials for the other polynomial are added locally and pushed into the peer's shared memory with
//     st.async (8 KiB per step and direction, completion counted on the peer's mbarrier): the only traffic between
//     the two SMs, and no cluster barrier inside the loop -- the receive buffer is double buffered and the CTA
//     barriers of the two sides order its reuse (see the comment at the send);
//   * group 0 adds its own l partials and the received one, runs ONE inverse transform and updates ACC[q], which only
//     this CTA reads in the next step.
// Per step and SM: l forward transforms side by side, one inverse, two CTA barriers, one mbarrier wait.
// The transform is a policy (negacyclic_fft_tm.cuh); with l transforms per SM the shared-memory pipe is nearly idle
// and the plain shared-memory exchanges are as fast as the tensor-memory ones (0.896 against 0.918 ms per rotation
// at n = 500), so FftSmem is the default.  Measured per step at n = 500 (clock64 build, -DB200_LAT2_PROFILE):
// decomposition 350, forward 1,000, MAC + hand-off 300, peer's partial in flight 500, inverse 900, update 150 clk;
// one CTA per ciphertext (k_blind_rotate_lat) takes 5,150.  Sums are formed in a different order than in cmux_step, but
// every external product is an exact integer convolution that FP64 resolves far inside the rounding margin:
// outputs are identical to k_blind_rotate (tests/test_gpu_workloads.py).
#pragma once
#ifdef B200_LAT2_PROFILE
#include <cstdio>
#endif

namespace b200 {

constexpr int kLat2PartBytes = kM * 16;      // one partial spectrum: 512 double2
constexpr int kLat2HandOffBarrier = 8;       // named barrier (1..l are the groups' own)

template <int L, class FFT> __host__ __device__ constexpr int lat2_smem_used() {
    return kN * 4 + 1280 + L * FFT::kBufElemsPerTransform * 16 + 2 * (L - 1) * kLat2PartBytes + 2 * kLat2PartBytes + 64;
}
// more than half an SM's shared memory: two of these CTAs never share an SM, so the whole-SM tensor-memory
// allocation of one cannot wait for a CTA whose peer waits for ours
template <int L, class FFT> __host__ __device__ constexpr int lat2_smem_bytes() {
    return lat2_smem_used<L, FFT>() > 116 * 1024 ? lat2_smem_used<L, FFT>() : 116 * 1024;
}

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_peer(uint32_t local_smem_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(rank));
    return r;
}
// 16 bytes into the peer's shared memory; the peer's mbarrier counts them
__device__ __forceinline__ void st_async_peer(uint32_t peer_addr, const double2 v, uint32_t peer_mbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.b64 [%0], {%1, %2}, [%3];"
                 ::"r"(peer_addr), "l"(__double_as_longlong(v.x)), "l"(__double_as_longlong(v.y)), "r"(peer_mbar) : "memory");
}

#ifdef B200_LAT2_PROFILE
// phase clocks of one thread per group, printed at the end (scratch builds only)
#define LAT2_T(k) do { const long long now_ = clock64(); prof[k] += now_ - last_; last_ = now_; } while (0)
#else
#define LAT2_T(k) do { } while (0)
#endif

template <int L, int BGBIT, class FFT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(64 * L, 1) k_blind_rotate_lat2(const DeviceKey key, const BlindRotateArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int KPL = 2 * L, THREADS = 64 * L;
    constexpr uint32_t BG = 1u << BGBIT, HALF = BG >> 1;
    constexpr int BUF = FFT::kBufElemsPerTransform * 16;
    int32_t *ACC = reinterpret_cast<int32_t *>(smem);                                    // [1024]: polynomial q
    uint16_t *bara = reinterpret_cast<uint16_t *>(smem + kN * 4);                        // [n+1]
    unsigned char *base = smem + kN * 4 + 1280;
    const int p = threadIdx.x >> 6, t = threadIdx.x & 63;
    double2 *buf = reinterpret_cast<double2 *>(base + (size_t)p * BUF);
    double2 *s_own = reinterpret_cast<double2 *>(base + (size_t)L * BUF);                // [L-1][512], from groups 1..
    double2 *s_oth = s_own + (size_t)(L - 1) * kM;                                       // [L-1][512], from groups ..L-2
    double2 *rx = s_oth + (size_t)(L - 1) * kM;                                          // [2][512], written by the peer
    uint64_t *rx_bar = reinterpret_cast<uint64_t *>(rx + 2 * kM);                        // [2]
    uint32_t *tm_slot_p = reinterpret_cast<uint32_t *>(rx_bar + 2);

    const uint32_t q = cluster_ctarank();
    const int n = key.n, n1 = n + 1;
    const int64_t item = blockIdx.x >> 1;

    if (threadIdx.x == 0) {
        mbar_init(&rx_bar[0], 1);
        mbar_init(&rx_bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    uint32_t tm = 0;
    if (UsesTm<FFT>::value) tm = tm_alloc_cta(tm_slot_p);
    FFT fft;
    fft.init(t, tm);

    // ---- gate prologue + modulus switch to 2N (numeric-functions.cpp:59-65); both CTAs compute all of bara
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
        for (int i = threadIdx.x; i < n1; i += THREADS) {
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
        for (int j = threadIdx.x; j < kN; j += THREADS) ACC[j] = q ? (((j + barb) & kN) ? -mu : mu) : 0;
    }
    const NamedBarrier64 bar{1 + p};
    uint32_t offset = 0;  // tgsw.cpp:7-29
#pragma unroll
    for (int pp = 0; pp < L; ++pp) offset += HALF << (32 - (pp + 1) * BGBIT);
    const int shift = 32 - (p + 1) * BGBIT;
    const double2 *bk = UsesTm<FFT>::value ? key.bkfft_tm : key.bkfft;
    const uint32_t peer = q ^ 1u;
    const uint32_t peer_rx = map_to_peer(smem_u32(rx), peer), peer_bar = map_to_peer(smem_u32(rx_bar), peer);
    cluster_sync_all();   // the peer's barriers exist before anything is sent to them; ACC and bara complete

    // ---- blind rotation (lwe-bootstrapping-functions-fft.cpp:88-112)
    uint32_t cnt = 0;     // steps taken (a_i != 0): the same on both CTAs
#ifdef B200_LAT2_PROFILE
    long long prof[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, last_ = clock64();
#endif
    // Key row (q, p) of BK_i, both columns, 16 complex values per thread.  The registers are free from the spectral
    // MAC of one step to that of the next, so the rows of the NEXT step are requested right after the MAC, in time
    // every group would otherwise spend waiting (for the peer's partial, or for group 0's inverse transform): neither
    // the L2 latency nor the issue of the 16 loads is on the step's critical path.
    double2 b_own[8], b_oth[8];
    auto next_step = [&](int from) { while (from < n && bara[from] == 0) ++from; return from; };   // cluster-uniform
    auto request_rows = [&](int step) {
        const double2 *row = bk + ((size_t)step * KPL + (q * L + p)) * 2 * kM + t;
#pragma unroll
        for (int r = 0; r < 8; ++r) { b_own[r] = __ldg(row + q * kM + r * 64); b_oth[r] = __ldg(row + peer * kM + r * 64); }
    };
    int i = next_step(0);
    if (i < n) request_rows(i);
#pragma unroll 1
    while (i < n) {
        const int ai = bara[i], inext = next_step(i + 1);
        const uint32_t par = cnt & 1u, phase = (cnt >> 1) & 1u;
        ++cnt;
        if (threadIdx.x == 0)
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&rx_bar[par])), "r"(kLat2PartBytes) : "memory");
        LAT2_T(9);

        double2 x[8];
        {
            uint32_t dv[16];   // (rotated difference) + decomposition offset, coefficient t + 64 rr
#pragma unroll
            for (int rr = 0; rr < 16; ++rr) {
                const int j = t + 64 * rr;
                const int m = (j - ai) & (2 * kN - 1);
                int32_t v = ACC[m & (kN - 1)];
                if (m & kN) v = -v;
                dv[rr] = (uint32_t)(v - ACC[j]) + offset;
            }
            digits_pretwisted<BGBIT>(x, dv, shift);   // level 0 of the transform on the integer side
        }
        LAT2_T(0);
        fft.template forward<true>(x, buf, bar);
        LAT2_T(1);
        // the partial for the peer first: it has the longer way.  Groups 0..l-2 hand theirs to group l-1 through
        // shared memory and a named barrier they only arrive at; group l-1 adds and pushes while the others go on
        double2 m_own[8], m_oth[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            m_oth[r].x = x[r].x * b_oth[r].x;                  m_oth[r].y = x[r].x * b_oth[r].y;
            m_oth[r].x = fma(-x[r].y, b_oth[r].y, m_oth[r].x); m_oth[r].y = fma(x[r].y, b_oth[r].x, m_oth[r].y);
        }
        if (p <= L - 2) {
#pragma unroll
            for (int r = 0; r < 8; ++r) s_oth[p * kM + r * 64 + t] = m_oth[r];
            asm volatile("bar.arrive %0, %1;" ::"r"(kLat2HandOffBarrier), "r"(THREADS) : "memory");
        } else {
            asm volatile("bar.sync %0, %1;" ::"r"(kLat2HandOffBarrier), "r"(THREADS) : "memory");
#pragma unroll
            for (int pp = 0; pp < L - 1; ++pp)
#pragma unroll
                for (int r = 0; r < 8; ++r) { const double2 v = s_oth[pp * kM + r * 64 + t]; m_oth[r].x += v.x; m_oth[r].y += v.y; }
            // Push into the peer's rx[par].  The peer read rx[par] two steps ago, before its end-of-step CTA
            // barrier; it then had to receive OUR previous partial, which we sent after our hand-off barrier of that
            // step, which came after our end-of-step barrier of the step before, which our group 0 passed only after
            // reading the peer's partial of that step: so the peer's readers of rx[par] are done whenever we get
            // here, without a cluster barrier.
            const uint32_t dst = peer_rx + (par * kM + t) * 16u, mb = peer_bar + par * 8u;
#pragma unroll
            for (int r = 0; r < 8; ++r) st_async_peer(dst + r * 64 * 16, m_oth[r], mb);
        }
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            m_own[r].x = x[r].x * b_own[r].x;                  m_own[r].y = x[r].x * b_own[r].y;
            m_own[r].x = fma(-x[r].y, b_own[r].y, m_own[r].x); m_own[r].y = fma(x[r].y, b_own[r].x, m_own[r].y);
        }
        if (p >= 1) {
#pragma unroll
            for (int r = 0; r < 8; ++r) s_own[(p - 1) * kM + r * 64 + t] = m_own[r];
        }
        if (p != L - 1 && inext < n) request_rows(inext);   // group l-1 is awaited at the barrier: it asks after it
        LAT2_T(2);
        __syncthreads();   // partial spectra written; every group has finished reading ACC
        LAT2_T(3);
        if (p == L - 1 && inext < n) request_rows(inext);
        if (p == 0) {      // output polynomial q
#pragma unroll
            for (int pp = 0; pp < L - 1; ++pp)
#pragma unroll
                for (int r = 0; r < 8; ++r) { const double2 v = s_own[pp * kM + r * 64 + t]; m_own[r].x += v.x; m_own[r].y += v.y; }
            LAT2_T(4);
            mbar_wait(&rx_bar[par], phase);
            LAT2_T(5);
#pragma unroll
            for (int r = 0; r < 8; ++r) { const double2 v = rx[par * kM + r * 64 + t]; m_own[r].x += v.x; m_own[r].y += v.y; }
            fft.template inverse<true>(m_own, buf, bar);
            LAT2_T(6);
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const int j = t + 64 * r;
                const double sc = c_tw.p1_scale[r];   // 1/512 times the cos factors the last pass left out
                ACC[j] += round_scaled_to_torus(m_own[r].x, sc);
                ACC[j + kM] += round_scaled_to_torus(m_own[r].y, sc);
            }
        }
        LAT2_T(7);
        __syncthreads();   // accumulator complete; partial buffers free
        LAT2_T(8);
        i = inext;
    }
#ifdef B200_LAT2_PROFILE
    if (t == 0 && item == 0)
        printf("lat2 cta %u group %d, clk per step: decompose %lld forward %lld mac %lld bar1 %lld sum(+send) %lld wait_rx %lld inverse %lld update %lld bar2 %lld key-ldg-issue %lld (steps %u)\n",
               q, p, prof[0] / cnt, prof[1] / cnt, prof[2] / cnt, prof[3] / cnt, prof[4] / cnt, prof[5] / cnt, prof[6] / cnt, prof[7] / cnt, prof[8] / cnt, prof[9] / cnt, cnt);
#endif

    // ---- sample extract at index 0 (lwe.cpp:41-60): a from polynomial 0, b from polynomial 1
    int32_t *u = a.u + (size_t)item * (kN + 1);
    if (q == 0) {
        for (int j = threadIdx.x; j < kN; j += THREADS) u[j] = (j == 0) ? ACC[0] : -ACC[kN - j];
    } else if (threadIdx.x == 0) u[kN] = ACC[0];
    cluster_sync_all();   // nobody leaves while the peer may still write into its shared memory
    if (UsesTm<FFT>::value) tm_free_cta(tm_slot_p);
}

template <int L, int BGBIT, class FFT>
static cudaError_t lat2_launch(const DeviceKey &key, const BlindRotateArgs &a, cudaStream_t s) {
    static bool done[16] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    constexpr int smem = lat2_smem_bytes<L, FFT>();
    static_assert(smem <= 227 * 1024, "shared memory budget");
    if (!(dev < 16 && done[dev])) {
        cudaError_t e = cudaFuncSetAttribute(k_blind_rotate_lat2<L, BGBIT, FFT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        if (dev < 16) done[dev] = true;
    }
    k_blind_rotate_lat2<L, BGBIT, FFT><<<(unsigned)(2 * a.items), 64 * L, smem, s>>>(key, a);
    return cudaSuccess;
}

}  // namespace b200
