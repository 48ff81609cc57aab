// negacyclic_fft.cuh -- N=1024 negacyclic FP64 transform for one polynomial per 64-thread group.
//
// Replaces the reference FFT processor for this path
// (libs/tfhe/src/libtfhe/fft_processors/spqlios/fft_processor_spqlios.cpp:38-115,
//  algorithm model spqlios-fft-impl.cpp:204-397,469-641): a real polynomial a(X) mod X^1024+1
// is folded to z_j = a_j + i*a_{j+512} (j<512) and evaluated at the 512 roots of X^512 = i
// (the "Lagrange half-complex" domain, nayuki/lagrangehalfc_impl.h:36-41).  The ordering and
// scaling of that domain are private to each reference processor (include/polynomials.h:36-47);
// this engine converts the bootstrapping key itself, so it uses its own order (below).
//
// B200 design (not a translation of the x86 code):
//   * Recursion on the modulus: z mod (X^n - R) -> z mod (X^{n/2} - sqrt R), z mod (X^{n/2} + sqrt R).
//     All butterflies of a block share one root, the twist by w^j is absorbed, and the 512
//     evaluations come out in bit-reversed-root order with no reordering pass.
//   * 64 threads x 8 points, three radix-8 passes (9 levels) held in registers, two exchanges
//     through a padded shared-memory buffer with 128-bit accesses (conflict free).  Only the exchange
//     next to pass 1 crosses the two warps (one group barrier per transform); the other one stays inside
//     octets of threads and needs a __syncwarp only.
//   * Butterflies in (cos, tan) form: u +- c*(v + i*t*v) costs 6 DFMA instead of 8 FP64 ops.
//   * Per-thread roots (4 per pass) live in registers for the whole blind rotation; pass-1 roots
//     are warp-uniform and come from the constant bank.
//
// Layouts:  pass 1 holds element e = t + 64 r, pass 2 e = 64 (t>>3) + (t&7) + 8 r, pass 3 e = 8 t + r
// (t = thread 0..63, r = register 0..7).  Evaluation-domain arrays are always in pass-3 layout.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200 {

constexpr int kN = 1024;          // ring degree
constexpr int kM = 512;           // complex points
constexpr int kGroup = 64;        // threads per polynomial
constexpr int kBufElems = 576;    // 512 + 512/8 padding

struct RootSet {                  // (cos, tan) of the four stored roots of one radix-8 pass
    double2 A, B, C, D;
};

struct TwiddleTables {
    RootSet p1;                   // level 0, block 0            (warp uniform)
    RootSet p2[8];                // level 3, block = e >> 6
    RootSet p3[64];               // level 6, block = e >> 3
    // last inverse pass with its cos factors deferred (inv_radix8_last): ratio cos D / cos C and, per register,
    // the product of the cos factors it skipped times 1/512 -- folded into the rounding FMA
    double p1_rho;
    double p1_scale[8];
};

__constant__ TwiddleTables c_tw;

__device__ __forceinline__ int pad_idx(int e) { return e + (e >> 3); }

// Per-thread root sets are loaded once and must stay in registers: without the opaque asm the
// compiler rematerialises them with lane-divergent LDC loads inside the CMux loop (measured:
// 75% of all stall samples on the first profile).
__device__ __forceinline__ void pin(double2 &v) { asm volatile("" : "+d"(v.x), "+d"(v.y)); }
__device__ __forceinline__ RootSet load_rootset(const RootSet &src) {
    RootSet w = src;
    pin(w.A); pin(w.B); pin(w.C); pin(w.D);
    return w;
}

// (u, v) <- (u + s v, u - s v),  s = c (1 + i t)
__device__ __forceinline__ void ct_bfly(double2 &u, double2 &v, const double2 ct) {
    const double wr = fma(-ct.y, v.y, v.x);
    const double wi = fma(ct.y, v.x, v.y);
    const double ur = u.x, ui = u.y;
    u.x = fma(ct.x, wr, ur);
    u.y = fma(ct.x, wi, ui);
    v.x = fma(-ct.x, wr, ur);
    v.y = fma(-ct.x, wi, ui);
}
// same with root i*s
__device__ __forceinline__ void ct_bfly_i(double2 &u, double2 &v, const double2 ct) {
    const double wr = fma(-ct.y, v.y, v.x);
    const double wi = fma(ct.y, v.x, v.y);
    const double ur = u.x, ui = u.y;
    u.x = fma(-ct.x, wi, ur);
    u.y = fma(ct.x, wr, ui);
    v.x = fma(ct.x, wi, ur);
    v.y = fma(-ct.x, wr, ui);
}
// inverse butterflies (without the 1/2): (a, b) <- (a + b, (a - b) * conj(s))
__device__ __forceinline__ void gs_bfly(double2 &u, double2 &v, const double2 ct) {
    const double dr = u.x - v.x, di = u.y - v.y;
    u.x += v.x;
    u.y += v.y;
    const double wr = fma(ct.y, di, dr);
    const double wi = fma(-ct.y, dr, di);
    v.x = ct.x * wr;
    v.y = ct.x * wi;
}
// root i*s: conj(i s) = -i conj(s)
__device__ __forceinline__ void gs_bfly_i(double2 &u, double2 &v, const double2 ct) {
    const double dr = u.x - v.x, di = u.y - v.y;
    u.x += v.x;
    u.y += v.y;
    const double wr = fma(ct.y, di, dr);
    const double wi = fma(-ct.y, dr, di);
    v.x = ct.x * wi;
    v.y = -ct.x * wr;
}

__device__ __forceinline__ void fwd_radix8(double2 (&x)[8], const RootSet &w) {
    ct_bfly(x[0], x[4], w.A);
    ct_bfly(x[1], x[5], w.A);
    ct_bfly(x[2], x[6], w.A);
    ct_bfly(x[3], x[7], w.A);
    ct_bfly(x[0], x[2], w.B);
    ct_bfly(x[1], x[3], w.B);
    ct_bfly_i(x[4], x[6], w.B);
    ct_bfly_i(x[5], x[7], w.B);
    ct_bfly(x[0], x[1], w.C);
    ct_bfly_i(x[2], x[3], w.C);
    ct_bfly(x[4], x[5], w.D);
    ct_bfly_i(x[6], x[7], w.D);
}

// First pass of a forward transform whose inputs are gadget digits: the root of level 0 is e^{i pi/4} (tan = 1), so
// its twist w = v (1 + i) = (v.x - v.y, v.x + v.y) is an INTEGER sum that the caller forms before converting to
// double (digits_pretwisted): x[4..7] arrive as w, and the level costs 4 DFMA per butterfly instead of 6.
__device__ __forceinline__ void fwd_radix8_first_pretwisted(double2 (&x)[8]) {
    const double c = c_tw.p1.A.x;
    const double2 B = c_tw.p1.B, C = c_tw.p1.C, D = c_tw.p1.D;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const double ur = x[r].x, ui = x[r].y, wr = x[r + 4].x, wi = x[r + 4].y;
        x[r].x = fma(c, wr, ur);
        x[r].y = fma(c, wi, ui);
        x[r + 4].x = fma(-c, wr, ur);
        x[r + 4].y = fma(-c, wi, ui);
    }
    ct_bfly(x[0], x[2], B);
    ct_bfly(x[1], x[3], B);
    ct_bfly_i(x[4], x[6], B);
    ct_bfly_i(x[5], x[7], B);
    ct_bfly(x[0], x[1], C);
    ct_bfly_i(x[2], x[3], C);
    ct_bfly(x[4], x[5], D);
    ct_bfly_i(x[6], x[7], D);
}

__device__ __forceinline__ void inv_radix8(double2 (&x)[8], const RootSet &w) {
    gs_bfly(x[0], x[1], w.C);
    gs_bfly_i(x[2], x[3], w.C);
    gs_bfly(x[4], x[5], w.D);
    gs_bfly_i(x[6], x[7], w.D);
    gs_bfly(x[0], x[2], w.B);
    gs_bfly(x[1], x[3], w.B);
    gs_bfly_i(x[4], x[6], w.B);
    gs_bfly_i(x[5], x[7], w.B);
    gs_bfly(x[0], x[4], w.A);
    gs_bfly(x[1], x[5], w.A);
    gs_bfly(x[2], x[6], w.A);
    gs_bfly(x[3], x[7], w.A);
}

// Last inverse pass (warp-uniform roots from the constant bank): the butterflies leave out the multiplication by
// cos (6 FP64 operations instead of 8); what each register is short of is a constant, c_tw.p1_scale[r], which the
// caller folds into the rounding FMA.  Where a level adds two registers that are short of different factors
// (cos C against cos D) the second enters with their ratio: u +- rho v is two DFMA instead of two DADD.
// (tools/fft_model.py checks the algebra: inv_radix8 == inv_radix8_last followed by the per-register scale.)
__device__ __forceinline__ void gsu_bfly(double2 &u, double2 &v, const double t) {
    const double dr = u.x - v.x, di = u.y - v.y;
    u.x += v.x;
    u.y += v.y;
    v.x = fma(t, di, dr);
    v.y = fma(-t, dr, di);
}
__device__ __forceinline__ void gsu_bfly_i(double2 &u, double2 &v, const double t) {
    const double ndr = v.x - u.x, di = u.y - v.y;   // (wr, wi) * (-i) = (wi, -wr), written without a negation
    u.x += v.x;
    u.y += v.y;
    v.x = fma(t, ndr, di);
    v.y = fma(-t, di, ndr);
}
__device__ __forceinline__ void gsu_bfly_rho(double2 &u, double2 &v, const double t, const double rho) {
    const double dr = fma(-rho, v.x, u.x), di = fma(-rho, v.y, u.y);
    u.x = fma(rho, v.x, u.x);
    u.y = fma(rho, v.y, u.y);
    v.x = fma(t, di, dr);
    v.y = fma(-t, dr, di);
}
__device__ __forceinline__ void inv_radix8_last(double2 (&x)[8]) {
    const double tA = c_tw.p1.A.y, tB = c_tw.p1.B.y, tC = c_tw.p1.C.y, tD = c_tw.p1.D.y, rho = c_tw.p1_rho;
    gsu_bfly(x[0], x[1], tC);
    gsu_bfly_i(x[2], x[3], tC);
    gsu_bfly(x[4], x[5], tD);
    gsu_bfly_i(x[6], x[7], tD);
    gsu_bfly(x[0], x[2], tB);
    gsu_bfly(x[1], x[3], tB);
    gsu_bfly_i(x[4], x[6], tB);
    gsu_bfly_i(x[5], x[7], tB);
    gsu_bfly(x[0], x[4], tA);
    gsu_bfly_rho(x[1], x[5], tA, rho);
    gsu_bfly(x[2], x[6], tA);
    gsu_bfly_rho(x[3], x[7], tA, rho);
}

// Group-wide barrier hook: the kernels pass a functor so that a CTA may hold one or several
// 64-thread groups (named barriers) without changing the transform code.
struct CtaBarrier {
    __device__ __forceinline__ void operator()() const { __syncthreads(); }
};
struct NamedBarrier64 {
    int id;  // 1..15
    __device__ __forceinline__ void operator()() const {
        asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory");
    }
};

// coefficients (pass-1 layout, x[r] = z_{t+64r}) -> evaluations (pass-3 layout, e = 8t + r)
// PRE: x[4..7] arrive twisted by the level-0 root (digits_pretwisted)
template <bool PRE = false, class Barrier>
__device__ __forceinline__ void negacyclic_forward(double2 (&x)[8], double2 *buf, const int t,
                                                   const RootSet &w2, const RootSet &w3,
                                                   const Barrier &bar) {
    if (PRE) fwd_radix8_first_pretwisted(x); else fwd_radix8(x, c_tw.p1);
#pragma unroll
    for (int r = 0; r < 8; ++r) buf[pad_idx(t + 64 * r)] = x[r];
    bar();
    const int base2 = pad_idx(64 * (t >> 3) + (t & 7));  // + 9 r  (8 r + r padding)
#pragma unroll
    for (int r = 0; r < 8; ++r) x[r] = buf[base2 + 9 * r];
    fwd_radix8(x, w2);
#pragma unroll
    for (int r = 0; r < 8; ++r) buf[base2 + 9 * r] = x[r];
    // The second exchange is an 8x8 transpose inside each aligned octet of threads (thread t reads what
    // threads 8(t>>3)+r wrote, all within elements 72(t>>3)..72(t>>3)+71 of the padded buffer), i.e. it
    // never leaves the warp: a warp-level sync replaces the group barrier.
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 8; ++r) x[r] = buf[9 * t + r];
    fwd_radix8(x, w3);
}

// evaluations (pass-3 layout) -> 512 * coefficients (pass-1 layout)
// DEFER_LAST: the result in register r is short of the factor 512 * c_tw.p1_scale[r] (see inv_radix8_last): the
// caller rounds with round_scaled_to_torus(x[r], c_tw.p1_scale[r]) instead of (x[r], 1/512).
template <bool DEFER_LAST = false, class Barrier>
__device__ __forceinline__ void negacyclic_inverse(double2 (&x)[8], double2 *buf, const int t,
                                                   const RootSet &w2, const RootSet &w3,
                                                   const Barrier &bar) {
    inv_radix8(x, w3);
#pragma unroll
    for (int r = 0; r < 8; ++r) buf[9 * t + r] = x[r];
    __syncwarp();  // octet-local exchange, see negacyclic_forward
    const int base2 = pad_idx(64 * (t >> 3) + (t & 7));
#pragma unroll
    for (int r = 0; r < 8; ++r) x[r] = buf[base2 + 9 * r];
    inv_radix8(x, w2);
#pragma unroll
    for (int r = 0; r < 8; ++r) buf[base2 + 9 * r] = x[r];
    bar();
#pragma unroll
    for (int r = 0; r < 8; ++r) x[r] = buf[pad_idx(t + 64 * r)];
    if (DEFER_LAST) inv_radix8_last(x);
    else inv_radix8(x, c_tw.p1);
}

// exact int32 -> double for |v| < 2^31 with one integer op and one DADD (no I2F on the slow pipe)
__device__ __forceinline__ double int_to_double(int32_t v) {
#ifdef B200_I2F
    return (double)v;
#endif
    return __hiloint2double(0x43300000, (int)(0x80000000u ^ (uint32_t)v)) - 4503601774854144.0;  // 2^52 + 2^31
}
// gadget digit p of (x + offset) as an exact double, centred: ((v >> shift) & (Bg-1)) - Bg/2
// (tgsw-functions.cpp:298-348).  The unsigned field goes straight into the mantissa of 2^52 and the
// centring rides on the same DADD: two integer instructions and one DADD per digit.
template <int BGBIT>
__device__ __forceinline__ double digit_to_double(uint32_t v, int shift) {
    constexpr uint32_t MASK = (1u << BGBIT) - 1;
    constexpr double kBias = 4503599627370496.0 + (double)(1u << (BGBIT - 1));  // 2^52 + Bg/2
    return __hiloint2double(0x43300000, (int)((v >> shift) & MASK)) - kBias;
}
// The eight complex inputs of a forward transform from the 16 decomposition words of a thread (coefficients
// t + 64 rr, rr = 0..15; x[r] = (digit rr = r, digit rr = r + 8)), registers 4..7 already twisted by the level-0
// root: (d1 - d2, d1 + d2) with d = field - Bg/2, i.e. (f1 - f2, f1 + f2 - Bg) -- exact small integers, converted
// like int_to_double.  For fwd_radix8_first_pretwisted.
template <int BGBIT>
__device__ __forceinline__ void digits_pretwisted(double2 (&x)[8], const uint32_t (&dv)[16], const int shift) {
    constexpr uint32_t MASK = (1u << BGBIT) - 1;
#pragma unroll
    for (int r = 0; r < 4; ++r)
        x[r] = make_double2(digit_to_double<BGBIT>(dv[r], shift), digit_to_double<BGBIT>(dv[r + 8], shift));
#pragma unroll
    for (int r = 4; r < 8; ++r) {
        const int32_t f1 = (int32_t)((dv[r] >> shift) & MASK), f2 = (int32_t)((dv[r + 8] >> shift) & MASK);
        x[r] = make_double2(int_to_double(f1 - f2), int_to_double(f1 + f2 - (int32_t)(1u << BGBIT)));
    }
}

// round-to-nearest of x*scale to an integer, reduced mod 2^32 (|x*scale| < 2^51)
__device__ __forceinline__ int32_t round_scaled_to_torus(double x, double scale) {
    return __double2loint(fma(x, scale, 6755399441055744.0));  // 1.5 * 2^52
}

}  // namespace b200
