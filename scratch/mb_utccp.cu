// mb_utccp.cu -- what lands where when tcgen05.cp copies shared memory into tensor memory?
// (no-swizzle K-major descriptor, rows of 16 bytes; shapes 64x128b.warpx2::02_13 / ::01_23, 32x128b.warpx4, 128x128b)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scratch/mb_utccp scratch/mb_utccp.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(const void *p, uint32_t sbo_bytes, uint32_t lbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_u32(p) >> 4) & 0x3fff);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;   // descriptor version for sm_100
    return d;                 // base offset 0, layout type 0 = no swizzle
}

__global__ void __launch_bounds__(128, 1) k_test(uint32_t *out) {
    __shared__ __align__(1024) uint32_t rows[128 * 4];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tm_slot;
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    for (int i = t; i < 128 * 4; i += 128) rows[i] = (i & 3) * 1000 + (i >> 2);
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 32;" ::"r"(smem_u32(&tm_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (t == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = tm_slot;
    // zero the 32 columns first so that untouched cells are visible
    {
        const uint32_t a = tm + (((uint32_t)warp * 32u) << 16);
        for (int c = 0; c < 32; c += 4)
            asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%1,%1,%1};" ::"r"(a + c), "r"(0xdeadu) : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (t == 0) {
        const uint64_t d = make_desc(rows, 128, 128);
        asm volatile("tcgen05.cp.cta_group::1.64x128b.warpx2::02_13 [%0], %1;" ::"r"(tm + 0), "l"(d) : "memory");
        asm volatile("tcgen05.cp.cta_group::1.64x128b.warpx2::01_23 [%0], %1;" ::"r"(tm + 4), "l"(d) : "memory");
        asm volatile("tcgen05.cp.cta_group::1.32x128b.warpx4 [%0], %1;" ::"r"(tm + 8), "l"(d) : "memory");
        asm volatile("tcgen05.cp.cta_group::1.128x128b [%0], %1;" ::"r"(tm + 12), "l"(d) : "memory");
        const uint64_t d2 = make_desc(rows + 64 * 4, 128, 0);   // second half of the rows, LBO 0
        asm volatile("tcgen05.cp.cta_group::1.64x128b.warpx2::02_13 [%0], %1;" ::"r"(tm + 16), "l"(d2) : "memory");
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    {
        uint32_t done;
        do {
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done) : "r"(smem_u32(&bar)), "r"(0) : "memory");
        } while (!done);
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t a = tm + (((uint32_t)warp * 32u) << 16);
    for (int c = 0; c < 20; c += 4) {
        uint32_t v0, v1, v2, v3;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v0), "=r"(v1), "=r"(v2), "=r"(v3) : "r"(a + c) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        uint32_t *o = out + (size_t)t * 20 + c;
        o[0] = v0; o[1] = v1; o[2] = v2; o[3] = v3;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 32;" ::"r"(tm) : "memory");
}

int main() {
    uint32_t *d, h[128 * 20];
    cudaMalloc(&d, sizeof(h));
    cudaMemset(d, 0xff, sizeof(h));
    k_test<<<1, 128>>>(d);
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    const char *names[5] = {"64x128b.warpx2::02_13", "64x128b.warpx2::01_23", "32x128b.warpx4", "128x128b", "64x128b 02_13, rows 64.."};
    for (int s = 0; s < 5; ++s) {
        printf("%s: lane -> source row (word0), words 1..3 minus 1000/2000/3000\n", names[s]);
        for (int t = 0; t < 128; ++t) {
            const uint32_t *v = h + t * 20 + s * 4;
            if (t % 16 == 0) printf("  lane %3d:", t);
            if (v[0] == 0xdead) printf("    -"); else printf(" %4u", v[0]);
            if (v[0] != 0xdead && (v[1] != v[0] + 1000 || v[2] != v[0] + 2000 || v[3] != v[0] + 3000)) printf("!");
            if (t % 16 == 15) printf("\n");
        }
    }
    return 0;
}
