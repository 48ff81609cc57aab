// negacyclic_fft16.cuh -- N=1024 negacyclic FP64 transform for one polynomial per WARP (32 threads x
// 16 complex points), the layout the wide blind-rotation kernel uses.
//
// Same mathematics as negacyclic_fft.cuh (recursion on the modulus X^n - R, evaluations at the 512
// roots of X^512 = i, replacing libs/tfhe/src/libtfhe/fft_processors/spqlios/fft_processor_spqlios.cpp:38-115),
// re-cut for the measured B200 pipe rates: the v2 kernel (64 threads x 8 points, two shared-memory
// exchanges per transform) ran the shared-memory data pipe at 72 % while FP64 sat at 51 %
// (profiles/r1_v2_blind_rotate_ncu.txt), i.e. it was bound by exchange wavefronts.  Here
//   * level 0 is folded into the operand load (forward: each half warp builds its own half-butterfly
//     from the four integer digits; inverse: one 8-point SHFL exchange with lane^16),
//   * levels 1-4 and 5-8 are two radix-16 passes in registers with ONE padded shared-memory exchange
//     between them (16x16 transposes inside each half warp, 128-bit accesses, conflict free),
//   * a polynomial belongs to one warp, so every synchronisation is a __syncwarp.
// Exchange traffic per transform drops from 2 x 16 KiB to 16 KiB (+ 32 SHFL on the inverse).
//
// Layouts (lane = 16*sub + tt, r = register 0..15):
//   pass 1  element e = tt + 16 r of sub-block `sub` (global complex index 256*sub + e)
//   pass 2  element e = 16 tt + r  -> evaluation index g = 16*lane + r  ("bit-reversed-root order")
// Evaluation-domain arrays in memory are stored [r][lane] (position r*32 + lane) so that a warp's
// 128-bit loads of one r are contiguous.
#pragma once
#include "negacyclic_fft.cuh"

namespace b200 {

struct RootSet16 {                // (cos, tan) of the eight stored roots of one radix-16 pass
    double2 A, B, C, D, E, F, G, H;
};

struct TwiddleTables16 {
    RootSet16 p1[2];              // level 1, block = sub
    RootSet16 p2[32];             // level 5, block = lane
};

__constant__ TwiddleTables16 c_tw16;

constexpr int kSubStride = 272;   // 256 + 256/16 padding (double2 elements) per sub-block
constexpr int kBuf16Elems = 2 * kSubStride;

__device__ __forceinline__ RootSet16 load_rootset16(const RootSet16 &src) {
    RootSet16 w = src;
    pin(w.A); pin(w.B); pin(w.C); pin(w.D); pin(w.E); pin(w.F); pin(w.G); pin(w.H);
    return w;
}

__device__ __forceinline__ void fwd_radix16(double2 (&x)[16], const RootSet16 &w) {
#pragma unroll
    for (int r = 0; r < 8; ++r) ct_bfly(x[r], x[r + 8], w.A);
#pragma unroll
    for (int r = 0; r < 4; ++r) ct_bfly(x[r], x[r + 4], w.B);
#pragma unroll
    for (int r = 8; r < 12; ++r) ct_bfly_i(x[r], x[r + 4], w.B);
    ct_bfly(x[0], x[2], w.C);    ct_bfly(x[1], x[3], w.C);
    ct_bfly_i(x[4], x[6], w.C);  ct_bfly_i(x[5], x[7], w.C);
    ct_bfly(x[8], x[10], w.D);   ct_bfly(x[9], x[11], w.D);
    ct_bfly_i(x[12], x[14], w.D); ct_bfly_i(x[13], x[15], w.D);
    ct_bfly(x[0], x[1], w.E);    ct_bfly_i(x[2], x[3], w.E);
    ct_bfly(x[4], x[5], w.F);    ct_bfly_i(x[6], x[7], w.F);
    ct_bfly(x[8], x[9], w.G);    ct_bfly_i(x[10], x[11], w.G);
    ct_bfly(x[12], x[13], w.H);  ct_bfly_i(x[14], x[15], w.H);
}

__device__ __forceinline__ void inv_radix16(double2 (&x)[16], const RootSet16 &w) {
    gs_bfly(x[0], x[1], w.E);    gs_bfly_i(x[2], x[3], w.E);
    gs_bfly(x[4], x[5], w.F);    gs_bfly_i(x[6], x[7], w.F);
    gs_bfly(x[8], x[9], w.G);    gs_bfly_i(x[10], x[11], w.G);
    gs_bfly(x[12], x[13], w.H);  gs_bfly_i(x[14], x[15], w.H);
    gs_bfly(x[0], x[2], w.C);    gs_bfly(x[1], x[3], w.C);
    gs_bfly_i(x[4], x[6], w.C);  gs_bfly_i(x[5], x[7], w.C);
    gs_bfly(x[8], x[10], w.D);   gs_bfly(x[9], x[11], w.D);
    gs_bfly_i(x[12], x[14], w.D); gs_bfly_i(x[13], x[15], w.D);
#pragma unroll
    for (int r = 0; r < 4; ++r) gs_bfly(x[r], x[r + 4], w.B);
#pragma unroll
    for (int r = 8; r < 12; ++r) gs_bfly_i(x[r], x[r + 4], w.B);
#pragma unroll
    for (int r = 0; r < 8; ++r) gs_bfly(x[r], x[r + 8], w.A);
}

// Levels 1..8 of the forward transform.  In: x[r] = level-0 output, element tt + 16 r of the lane's
// sub-block.  Out: x[r] = evaluation 16*lane + r.  `buf` is the warp's exchange buffer.
__device__ __forceinline__ void forward16_tail(double2 (&x)[16], double2 *buf, const int lane,
                                               const RootSet16 &w1, const RootSet16 &w2) {
    fwd_radix16(x, w1);
    double2 *sb = buf + (lane >> 4) * kSubStride;
    const int tt = lane & 15;
    __syncwarp();  // previous readers of the buffer are done
#pragma unroll
    for (int r = 0; r < 16; ++r) sb[tt + 17 * r] = x[r];
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 16; ++r) x[r] = sb[17 * tt + r];
    fwd_radix16(x, w2);
}

// Levels 8..1 of the inverse transform (no 1/2 factors: the result carries a factor 256 here and 512
// after level 0).  In: evaluations 16*lane + r.  Out: x[r] = element tt + 16 r of the lane's sub-block.
__device__ __forceinline__ void inverse16_head(double2 (&x)[16], double2 *buf, const int lane,
                                               const RootSet16 &w1, const RootSet16 &w2) {
    inv_radix16(x, w2);
    double2 *sb = buf + (lane >> 4) * kSubStride;
    const int tt = lane & 15;
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 16; ++r) sb[17 * tt + r] = x[r];
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 16; ++r) x[r] = sb[tt + 17 * r];
    inv_radix16(x, w1);
}

__device__ __forceinline__ double shfl_xor16(double v) {
    int lo = __double2loint(v), hi = __double2hiint(v);
    lo = __shfl_xor_sync(0xffffffffu, lo, 16);
    hi = __shfl_xor_sync(0xffffffffu, hi, 16);
    return __hiloint2double(hi, lo);
}

constexpr double kCos45 = 0.70710678118654752440;  // level-0 root sqrt(i) = cos45 (1 + i)

}  // namespace b200
