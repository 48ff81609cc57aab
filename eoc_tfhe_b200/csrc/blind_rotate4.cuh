// blind_rotate4.cuh -- blind rotation "v4": the v2 kernel with the bootstrapping key delivered through
// TENSOR MEMORY instead of being read from shared memory by every ciphertext group.
//
// Why: v2 is co-limited by the shared-memory data pipe (72 %) and the FP64 pipe (51 %); of its 2426
// shared-memory wavefronts per ciphertext and CMux step, 512 are the four groups of a CTA each reading
// the same 64 KiB of key (profiles/r1_v2_ncu.txt, scratch/mb_fft.cu: forward transform + MAC runs the
// shared-memory pipe at 91 %).  Here a key row pair still arrives in shared memory by TMA
// (cp.async.bulk), but is then copied ONCE per CTA into tensor memory by 16 tcgen05.cp instructions of
// shape 64x128b.warpx2::02_13: rows 0-31 of the source land in the lanes of warps 0 and 2 (mod 4), rows
// 32-63 in those of warps 1 and 3 (measured with scratch/mb_utccp.cu), which is exactly "element t of a
// 64-element key column to thread t of every 64-thread group".  The groups read their operands with
// tcgen05.ld (the tensor-memory path is separate from the shared-memory pipe,
// profiles/r1_microbench_pipes.txt), so key traffic on the shared-memory pipe falls from 512 to 32
// wavefronts per ciphertext-step.
//
// Pipeline per key row pair q (sequence number over all n*kpl row pairs):
//   TMA(q)  -> shared stage q % S   (mbarrier full[s], transaction bytes)
//   cp(q)   -> tensor-memory slot q % T, 64 columns (tcgen05.commit -> tfull[slot], and -> sfree[s])
//   MAC(q)  by all 2G warps (tcgen05.ld), then one arrival per warp on a counter; the LAST warp to arrive
//           runs the "event" of row q:  cp(q + T) into the slot just freed, TMA(q + T + S - 1) into the stage
//           whose copy was issued one event earlier.  Events are totally ordered (a warp releases rows in
//           order), so exactly one thread at a time drives the pipeline and nobody else ever waits on it.
// Arithmetic is that of cmux_step, FMA for FMA: results are bit-identical to k_blind_rotate.
#pragma once

namespace b200 {

constexpr int kV4Stages = 3;        // S: shared-memory stages (16 KiB each)
constexpr int kV4Slots = 4;         // T: tensor-memory slots (64 columns each)
constexpr int kV4SlotCols = 64;     // 16 double2 per lane: b0[r] at column 4r, b1[r] at 32 + 4r

struct V4State {                    // shared memory
    uint64_t full[kV4Stages];       // TMA landed
    uint64_t sfree[kV4Stages];      // tcgen05.cp out of the stage finished
    uint64_t tfull[kV4Slots];       // tcgen05.cp into the slot finished
    int done[kV4Slots];             // warps that finished reading the slot
    uint32_t tm_base;
};

__device__ __forceinline__ uint64_t v4_smem_desc(const void *p) {
    // K-major, no swizzle: 8 rows x 16 bytes per core matrix, core matrices 128 bytes apart
    uint64_t d = (uint64_t)((smem_u32(p) >> 4) & 0x3fff);
    d |= (uint64_t)(128 >> 4) << 16;   // leading byte offset (unused: one 16-byte chunk per row)
    d |= (uint64_t)(128 >> 4) << 32;   // stride byte offset between 8-row groups
    d |= (uint64_t)1 << 46;            // descriptor version of sm_100
    return d;
}

struct V4Rows {
    unsigned char *ring;            // [S][kStageBytes]
    V4State *st;
    const double2 *bk;
    uint32_t tm;                    // tensor-memory base of the calling warp's lane quarter (slot 0)
    int seq0, total, warps;

    __device__ __forceinline__ void tma(int q) const {
        const int s = q % kV4Stages;
        tma_load_row_pair(ring + (size_t)s * kStageBytes, bk + (size_t)q * 2 * kM, &st->full[s]);
    }
    // one thread: copy row pair q from its shared stage into its tensor-memory slot
    __device__ __forceinline__ void cp(int q) const {
        const int s = q % kV4Stages, slot = q % kV4Slots;
        mbar_wait(&st->full[s], (uint32_t)(q / kV4Stages) & 1u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const unsigned char *src = ring + (size_t)s * kStageBytes;
        const uint32_t dst = st->tm_base + (uint32_t)slot * kV4SlotCols;
#pragma unroll
        for (int k = 0; k < 16; ++k) {   // k = col * 8 + r : 64 double2 = 1 KiB each
            const uint64_t d = v4_smem_desc(src + k * 1024);
            asm volatile("tcgen05.cp.cta_group::1.64x128b.warpx2::02_13 [%0], %1;" ::"r"(dst + 4 * k), "l"(d) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&st->tfull[slot])) : "memory");
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&st->sfree[s])) : "memory");
    }
    // one thread: refill the stage that held row pair q - S with row pair q, once its copy has drained
    __device__ __forceinline__ void tma_after_copy(int q) const {
        const int p = q - kV4Stages;
        mbar_wait(&st->sfree[p % kV4Stages], (uint32_t)(p / kV4Stages) & 1u);
        tma(q);
    }
    // thread 0, before the first use: fill the pipeline (cp 0..T-1, TMA 0..T+S-2)
    __device__ __forceinline__ void prologue() const {
        int issued = 0;
        for (; issued < kV4Stages && issued < total; ++issued) tma(issued);
        for (int c = 0; c < kV4Slots && c < total; ++c) {
            cp(c);
            for (; issued <= c + kV4Stages - 1 && issued < total; ++issued) tma_after_copy(issued);
        }
    }
    // all threads of a warp: wait until row pair rp of the current step is readable; returns the slot address
    __device__ __forceinline__ uint32_t acquire(int rp) const {
        const int q = seq0 + rp;
        mbar_wait(&st->tfull[q % kV4Slots], (uint32_t)(q / kV4Slots) & 1u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        return tm + (uint32_t)(q % kV4Slots) * kV4SlotCols;
    }
    // the calling warp's tensor-memory loads of row pair rp have completed (tcgen05.wait::ld)
    __device__ __forceinline__ void release(int rp) const {
        const int q = seq0 + rp, slot = q % kV4Slots;
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if ((threadIdx.x & 31) == 0) {
            __threadfence_block();
            if (atomicAdd(&st->done[slot], 1) == warps - 1) {
                st->done[slot] = 0;
                if (q + kV4Slots < total) cp(q + kV4Slots);
                if (q + kV4Slots + kV4Stages - 1 < total) tma_after_copy(q + kV4Slots + kV4Stages - 1);
            }
        }
    }
    __device__ __forceinline__ void skip_step(int kpl) const {
        for (int rp = 0; rp < kpl; ++rp) { acquire(rp); release(rp); }
    }
};

// four complex doubles (16 columns) of the calling lane in one tcgen05.ld
__device__ __forceinline__ void tmem_ld16(uint32_t addr, double2 (&v)[4]) {
    int w[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]),
                   "=r"(w[8]), "=r"(w[9]), "=r"(w[10]), "=r"(w[11]), "=r"(w[12]), "=r"(w[13]), "=r"(w[14]), "=r"(w[15])
                 : "r"(addr) : "memory");
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        v[k].x = __hiloint2double(w[4 * k + 1], w[4 * k]);
        v[k].y = __hiloint2double(w[4 * k + 3], w[4 * k + 2]);
    }
}

// one CMux step (see cmux_step in kernels.cu); key operands come from tensor memory
template <int L, int BGBIT, class Barrier>
__device__ __forceinline__ void cmux_step4(int32_t *ACC, double2 *bufs, const int a, const V4Rows &rows,
                                           const int t, const RootSet &w2, const RootSet &w3, const Barrier &bar) {
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
        uint32_t dv[16];
#pragma unroll
        for (int rr = 0; rr < 16; ++rr) {
            const int j = t + 64 * rr;
            const int m = (j - a) & (2 * kN - 1);
            int32_t v = P[m & (kN - 1)];
            if (m & kN) v = -v;
            dv[rr] = (uint32_t)(v - P[j]) + offset;
        }
#pragma unroll 1
        for (int p = 0; p < L; ++p) {
            const int shift = 32 - (p + 1) * BGBIT;
            double2 x[8];
#pragma unroll
            for (int r = 0; r < 8; ++r)
                x[r] = make_double2(digit_to_double<BGBIT>(dv[r], shift), digit_to_double<BGBIT>(dv[r + 8], shift));
            negacyclic_forward(x, bufs + toggle * kBufElems, t, w2, w3, bar);
            toggle ^= 1;
            const uint32_t slot = rows.acquire(q * L + p);
#pragma unroll
            for (int r0 = 0; r0 < 8; r0 += 4) {
                double2 b0[4], b1[4];
                tmem_ld16(slot + 4 * r0, b0);
                tmem_ld16(slot + 32 + 4 * r0, b1);
                tmem_wait_ld();
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int r = r0 + k;
                    acc0[r].x = fma(x[r].x, b0[k].x, acc0[r].x);
                    acc0[r].x = fma(-x[r].y, b0[k].y, acc0[r].x);
                    acc0[r].y = fma(x[r].x, b0[k].y, acc0[r].y);
                    acc0[r].y = fma(x[r].y, b0[k].x, acc0[r].y);
                    acc1[r].x = fma(x[r].x, b1[k].x, acc1[r].x);
                    acc1[r].x = fma(-x[r].y, b1[k].y, acc1[r].x);
                    acc1[r].y = fma(x[r].x, b1[k].y, acc1[r].y);
                    acc1[r].y = fma(x[r].y, b1[k].x, acc1[r].y);
                }
            }
            rows.release(q * L + p);
        }
    }
    constexpr double kScale = 1.0 / kM;
    negacyclic_inverse(acc0, bufs + toggle * kBufElems, t, w2, w3, bar);
    toggle ^= 1;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const int j = t + 64 * r;
        ACC[j] += round_scaled_to_torus(acc0[r].x, kScale);
        ACC[j + kM] += round_scaled_to_torus(acc0[r].y, kScale);
    }
    negacyclic_inverse(acc1, bufs + toggle * kBufElems, t, w2, w3, bar);
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const int j = kN + t + 64 * r;
        ACC[j] += round_scaled_to_torus(acc1[r].x, kScale);
        ACC[j + kM] += round_scaled_to_torus(acc1[r].y, kScale);
    }
    bar();  // accumulator complete before the next step's rotated reads
}

template <int G> __host__ __device__ constexpr int br4_smem_bytes() { return kV4Stages * kStageBytes + 256 + G * kGroupSmemBytes; }

template <int L, int BGBIT, int G>
__global__ void __maxnreg__(255) k_blind_rotate4(const DeviceKey key, const BlindRotateArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int KPL = 2 * L;
    constexpr int TM_COLS = kV4Slots * kV4SlotCols;   // 256
    unsigned char *ring = smem;
    V4State *st = reinterpret_cast<V4State *>(smem + kV4Stages * kStageBytes);
    const int g = threadIdx.x >> 6, t = threadIdx.x & 63, warp = threadIdx.x >> 5;
    unsigned char *mine = smem + kV4Stages * kStageBytes + 256 + (size_t)g * kGroupSmemBytes;
    int32_t *ACC = reinterpret_cast<int32_t *>(mine);
    double2 *bufs = reinterpret_cast<double2 *>(mine + 2 * kN * 4);
    uint16_t *bara = reinterpret_cast<uint16_t *>(mine + 2 * kN * 4 + 2 * kBufElems * 16);

    const int n = key.n, n1 = n + 1;
    const int64_t item = (int64_t)blockIdx.x * G + g;
    const bool live = item < a.items;
    const int total = n * KPL;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kV4Stages; ++s) { mbar_init(&st->full[s], 1); mbar_init(&st->sfree[s], 1); }
        for (int s = 0; s < kV4Slots; ++s) { mbar_init(&st->tfull[s], 1); st->done[s] = 0; }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&st->tm_base)), "n"(TM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
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
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();  // bara visible; barriers initialised; tensor-memory base published
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    V4Rows rows{ring, st, key.bkfft, st->tm_base + (((uint32_t)(warp & 3) * 32u) << 16), 0, total, 2 * G};
    if (threadIdx.x == 0) rows.prologue();

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
    const RootSet w2 = load_rootset(c_tw.p2[t >> 3]), w3 = load_rootset(c_tw.p3[t]);
    const NamedBarrier64 bar{1 + g};
    bar();

    // ---- blind rotation (lwe-bootstrapping-functions-fft.cpp:88-112)
#pragma unroll 1
    for (int i = 0; i < n; ++i) {
        rows.seq0 = i * KPL;
        const int ai = live ? bara[i] : 0;
        if (ai == 0) { rows.skip_step(KPL); continue; }
        cmux_step4<L, BGBIT>(ACC, bufs, ai, rows, t, w2, w3, bar);
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
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(st->tm_base), "n"(TM_COLS) : "memory");
}

template <int L, int BGBIT, int G>
static cudaError_t br4_launch(const DeviceKey &key, const BlindRotateArgs &a, cudaStream_t s) {
    static bool done[16] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    constexpr int smem = br4_smem_bytes<G>();
    if (!(dev < 16 && done[dev])) {
        cudaError_t e = cudaFuncSetAttribute(k_blind_rotate4<L, BGBIT, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        cudaFuncSetAttribute(k_blind_rotate4<L, BGBIT, G>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        if (dev < 16) done[dev] = true;
    }
    const unsigned grid = (unsigned)((a.items + G - 1) / G);
    k_blind_rotate4<L, BGBIT, G><<<grid, 64 * G, smem, s>>>(key, a);
    return cudaSuccess;
}

}  // namespace b200
