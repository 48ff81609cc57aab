// negacyclic_fft_tm.cuh -- the N=1024 negacyclic FP64 transform with its two warp-local exchanges in TENSOR MEMORY.
//
// Same transform as negacyclic_fft.cuh (recursion on the modulus, in place on element index e, (cos, tan)
// butterflies; replaces fft_processors/spqlios/fft_processor_spqlios.cpp:38-115 of the reference), cut differently:
// the shared-memory data pipe (one 128-byte wavefront per clock) is what bounds k_blind_rotate
// (profiles/r1_final_ncu.txt: 82 %), and 63 % of its wavefronts are the two exchanges per transform.  Here only
// the exchange that crosses the two warps of a group goes through shared memory; the others are done by the
// tensor-memory load/store path, which is a separate datapath (2-4x the LDS/STS rate, runs concurrently with
// it: profiles/r2_microbench_tmem.txt).  tcgen05.st 32x32b followed by tcgen05.ld 16x256b is a fixed permutation
// between the lanes of a warp -- a 4x4 transpose of (lane bits 4,3) against two register-index bits, in 64-bit
// units -- so each such exchange brings TWO new element bits into registers and the passes become
//
//     P1   levels 0-2  registers (e8 e7 e6)   thread t = e & 63                       roots: constant bank
//     S    shared memory, slot(e) = e ^ (((e >> 6) & 3) << 1): conflict free both ways, no padding
//     P2   levels 3-5  registers (e5 e4 e3)   warp e8, lane (e2 e1 e0 e7 e6)          roots: 8 sets
//     X1   tensor memory: registers (e5 e4) <-> lane bits (e2 e1); e3 stays
//     P3a  levels 6-7  registers (e2 e1 e3)   warp e8, lane (e0 e7 e6 e5 e4)          roots: 32 sets of 3
//     X2   tensor memory: registers (e2 e3) <-> lane bits (e0 e7); e1 stays
//     P3b  level 8     registers (e0 e7 e1)   warp e8, lane (e6 e5 e4 e2 e3)          roots: 2 per thread
//
// (tools/fft_tm_model.py is the bit-level model of exactly this plan on the measured lane map: it reproduces the
// in-place reference transform, the inverse round trip and the bank-conflict count.)  Evaluation-domain arrays are
// in the P3b layout; the bootstrapping key is converted by the same code, so no order is ever undone.
//
// The per-thread roots (18 doubles) also live in tensor memory, 72 columns of the thread's own lane, and are
// fetched by the load that brings the exchanged data: they cost no registers between passes.
#pragma once
#include "negacyclic_fft.cuh"

namespace b200 {

constexpr int kTmBufElems = 512;          // one exchange buffer: 8 KiB, no padding
constexpr int kTmColsPerWarp = 128;       // 0-31, 32-63 exchange windows A, B | 64-79 P2 roots | 80-91 P3a roots | 96-103 P3b roots
constexpr int kTmColExB = 32, kTmColRootsBase = 64;
constexpr int kTmColRoots2 = 0, kTmColRoots3a = 16, kTmColRoots3b = 32;   // relative to the roots base (FftTm::tmr)

struct TmRootTables {
    double2 p3a[32][3];                   // A = s(6, 2b'), C = s(7, 4b'), D = s(7, 4b'+2);  b' = e >> 4
    double2 p3b[64][2];                   // E0, E1 = s(8, e >> 1) at (e7, e1) = (0, 0), (1, 0); index = thread
};
__constant__ TmRootTables c_tw_tm;

__device__ __forceinline__ int tm_slot(int e) { return e ^ (((e >> 6) & 3) << 1); }

// ---- tensor-memory primitives ---------------------------------------------------------------------------------
// tcgen05.ld is asynchronous: its destination registers are defined only after tcgen05.wait::ld.  Every load below
// is therefore ONE asm statement -- loads, wait, then the 32-bit halves packed into the double outputs -- so the
// compiler cannot place a use (or a register move) between a load and its wait.  Stores unpack inside the statement
// likewise.  ptxas keeps packed / unpacked pairs in place when they are adjacent registers.
#define TM_UNPACK4(a, b, c, d) "mov.b64 {t0,t1}, " a ";\n\tmov.b64 {t2,t3}, " b ";\n\tmov.b64 {t4,t5}, " c ";\n\tmov.b64 {t6,t7}, " d ";\n\t"
#define TM_ST8(addr, off) "tcgen05.st.sync.aligned.32x32b.x8.b32 [" addr "+" off "], {t0,t1,t2,t3,t4,t5,t6,t7};\n\t"

// 16 doubles of one thread -> 32 columns of its lane; the caller issues tm_wait_st() before the loads that read them.
// ONE st 32x32b.x2 PER DOUBLE: a double is an aligned register pair wherever it lives, so nothing has to be gathered.
// Four st .x8 -- a quarter of the store instructions -- need the doubles in 8-register blocks in column order, and
// with the register file full ptxas assembles those blocks with moves: 206 IMAD.MOV per CMux step, 318.2 -> 309.7 ms
// per 65,536 rotations without them (B200_TM_ST_X8 restores the wide stores for A/B).
__device__ __forceinline__ void tm_store16(uint32_t addr, double a0, double a1, double a2, double a3, double b0, double b1,
                                           double b2, double b3, double c0, double c1, double c2, double c3, double d0,
                                           double d1, double d2, double d3) {
#ifndef B200_TM_ST_X8
#define TM_ST2(a, off) "mov.b64 {t0,t1}, " a ";\n\ttcgen05.st.sync.aligned.32x32b.x2.b32 [%0+" off "], {t0,t1};\n\t"
    asm volatile("{\n\t.reg .b32 t<2>;\n\t"
                 TM_ST2("%1", "0") TM_ST2("%2", "2") TM_ST2("%3", "4") TM_ST2("%4", "6")
                 TM_ST2("%5", "8") TM_ST2("%6", "10") TM_ST2("%7", "12") TM_ST2("%8", "14")
                 TM_ST2("%9", "16") TM_ST2("%10", "18") TM_ST2("%11", "20") TM_ST2("%12", "22")
                 TM_ST2("%13", "24") TM_ST2("%14", "26") TM_ST2("%15", "28") TM_ST2("%16", "30")
                 "}"
                 ::"r"(addr), "d"(a0), "d"(a1), "d"(a2), "d"(a3), "d"(b0), "d"(b1), "d"(b2), "d"(b3), "d"(c0), "d"(c1), "d"(c2),
                   "d"(c3), "d"(d0), "d"(d1), "d"(d2), "d"(d3) : "memory");
#undef TM_ST2
#else
    asm volatile("{\n\t.reg .b32 t<8>;\n\t"
                 TM_UNPACK4("%1", "%2", "%3", "%4") TM_ST8("%0", "0")
                 TM_UNPACK4("%5", "%6", "%7", "%8") TM_ST8("%0", "8")
                 TM_UNPACK4("%9", "%10", "%11", "%12") TM_ST8("%0", "16")
                 TM_UNPACK4("%13", "%14", "%15", "%16") TM_ST8("%0", "24")
                 "}"
                 ::"r"(addr), "d"(a0), "d"(a1), "d"(a2), "d"(a3), "d"(b0), "d"(b1), "d"(b2), "d"(b3), "d"(c0), "d"(c1), "d"(c2),
                   "d"(c3), "d"(d0), "d"(d1), "d"(d2), "d"(d3) : "memory");
#endif
}
__device__ __forceinline__ void tm_store4(uint32_t addr, double a0, double a1, double a2, double a3) {
    asm volatile("{\n\t.reg .b32 t<8>;\n\t" TM_UNPACK4("%1", "%2", "%3", "%4") TM_ST8("%0", "0") "}"
                 ::"r"(addr), "d"(a0), "d"(a1), "d"(a2), "d"(a3) : "memory");
}
__device__ __forceinline__ void tm_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

struct Tm8 { double v[8]; };
#define TM_LD_X16(regs, addr) "tcgen05.ld.sync.aligned.16x256b.x4.b32 {" regs "0," regs "1," regs "2," regs "3," regs "4," regs "5," regs "6," regs "7," regs "8," regs "9," regs "10," regs "11," regs "12," regs "13," regs "14," regs "15}, [" addr "];\n\t"
#define TM_LD_L16(regs, addr) "tcgen05.ld.sync.aligned.32x32b.x16.b32 {" regs "0," regs "1," regs "2," regs "3," regs "4," regs "5," regs "6," regs "7," regs "8," regs "9," regs "10," regs "11," regs "12," regs "13," regs "14," regs "15}, [" addr "];\n\t"
#define TM_PACK8(o0, o1, o2, o3, o4, o5, o6, o7, r) \
    "mov.b64 " o0 ", {" r "0," r "1};\n\tmov.b64 " o1 ", {" r "2," r "3};\n\tmov.b64 " o2 ", {" r "4," r "5};\n\tmov.b64 " o3 ", {" r "6," r "7};\n\t" \
    "mov.b64 " o4 ", {" r "8," r "9};\n\tmov.b64 " o5 ", {" r "10," r "11};\n\tmov.b64 " o6 ", {" r "12," r "13};\n\tmov.b64 " o7 ", {" r "14," r "15};\n\t"
#define TM_OUT8(a) "=d"(a[0]), "=d"(a[1]), "=d"(a[2]), "=d"(a[3]), "=d"(a[4]), "=d"(a[5]), "=d"(a[6]), "=d"(a[7])

// Reader side of a forward exchange (2 x ld 16x256b.x4) plus NR (0 | 4 | 6) root doubles from columns tm_roots of the
// thread's own lane:   lo.v[2j+k] / hi.v[2j+k] = double of lane (0 | 16) + 8k + T/4, column pair 4j + T%4
template <int NR>
__device__ __forceinline__ void tm_exchange_fwd_load(Tm8 &lo, Tm8 &hi, double (&rt)[8], const uint32_t tm, const uint32_t tm_hi,
                                                     const uint32_t tm_roots) {
    if constexpr (NR == 0) {
        asm volatile("{\n\t.reg .b32 u<16>, v<16>;\n\t" TM_LD_X16("u", "%16") TM_LD_X16("v", "%17")
                     "tcgen05.wait::ld.sync.aligned;\n\t"
                     TM_PACK8("%0", "%1", "%2", "%3", "%4", "%5", "%6", "%7", "u") TM_PACK8("%8", "%9", "%10", "%11", "%12", "%13", "%14", "%15", "v") "}"
                     : TM_OUT8(lo.v), TM_OUT8(hi.v) : "r"(tm), "r"(tm_hi) : "memory");
    } else if constexpr (NR == 4) {
        asm volatile("{\n\t.reg .b32 u<16>, v<16>, w<8>;\n\t" TM_LD_X16("u", "%20") TM_LD_X16("v", "%21")
                     "tcgen05.ld.sync.aligned.32x32b.x8.b32 {w0,w1,w2,w3,w4,w5,w6,w7}, [%22];\n\t"
                     "tcgen05.wait::ld.sync.aligned;\n\t"
                     TM_PACK8("%0", "%1", "%2", "%3", "%4", "%5", "%6", "%7", "u") TM_PACK8("%8", "%9", "%10", "%11", "%12", "%13", "%14", "%15", "v")
                     "mov.b64 %16, {w0,w1};\n\tmov.b64 %17, {w2,w3};\n\tmov.b64 %18, {w4,w5};\n\tmov.b64 %19, {w6,w7};\n\t}"
                     : TM_OUT8(lo.v), TM_OUT8(hi.v), "=d"(rt[0]), "=d"(rt[1]), "=d"(rt[2]), "=d"(rt[3])
                     : "r"(tm), "r"(tm_hi), "r"(tm_roots) : "memory");
    } else {
        asm volatile("{\n\t.reg .b32 u<16>, v<16>, w<12>;\n\t" TM_LD_X16("u", "%22") TM_LD_X16("v", "%23")
                     "tcgen05.ld.sync.aligned.32x32b.x8.b32 {w0,w1,w2,w3,w4,w5,w6,w7}, [%24];\n\t"
                     "tcgen05.ld.sync.aligned.32x32b.x4.b32 {w8,w9,w10,w11}, [%24+8];\n\t"
                     "tcgen05.wait::ld.sync.aligned;\n\t"
                     TM_PACK8("%0", "%1", "%2", "%3", "%4", "%5", "%6", "%7", "u") TM_PACK8("%8", "%9", "%10", "%11", "%12", "%13", "%14", "%15", "v")
                     "mov.b64 %16, {w0,w1};\n\tmov.b64 %17, {w2,w3};\n\tmov.b64 %18, {w4,w5};\n\tmov.b64 %19, {w6,w7};\n\t"
                     "mov.b64 %20, {w8,w9};\n\tmov.b64 %21, {w10,w11};\n\t}"
                     : TM_OUT8(lo.v), TM_OUT8(hi.v), "=d"(rt[0]), "=d"(rt[1]), "=d"(rt[2]), "=d"(rt[3]), "=d"(rt[4]), "=d"(rt[5])
                     : "r"(tm), "r"(tm_hi), "r"(tm_roots) : "memory");
    }
}
// 16 doubles of one thread from 32 columns of its lane (two ld 32x32b.x16) plus NR (0 | 6 | 8) root doubles, then wait::ld
template <int NR>
__device__ __forceinline__ void tm_load16(double (&d)[16], double (&rt)[8], const uint32_t tm, const uint32_t tm_roots) {
    double (&d0)[8] = reinterpret_cast<double (&)[8]>(d[0]);
    double (&d1)[8] = reinterpret_cast<double (&)[8]>(d[8]);
    if constexpr (NR == 0) {
        asm volatile("{\n\t.reg .b32 u<16>, v<16>;\n\t" TM_LD_L16("u", "%16") TM_LD_L16("v", "%16+16")
                     "tcgen05.wait::ld.sync.aligned;\n\t"
                     TM_PACK8("%0", "%1", "%2", "%3", "%4", "%5", "%6", "%7", "u") TM_PACK8("%8", "%9", "%10", "%11", "%12", "%13", "%14", "%15", "v") "}"
                     : TM_OUT8(d0), TM_OUT8(d1) : "r"(tm) : "memory");
    } else if constexpr (NR == 6) {
        asm volatile("{\n\t.reg .b32 u<16>, v<16>, w<12>;\n\t" TM_LD_L16("u", "%22") TM_LD_L16("v", "%22+16")
                     "tcgen05.ld.sync.aligned.32x32b.x8.b32 {w0,w1,w2,w3,w4,w5,w6,w7}, [%23];\n\t"
                     "tcgen05.ld.sync.aligned.32x32b.x4.b32 {w8,w9,w10,w11}, [%23+8];\n\t"
                     "tcgen05.wait::ld.sync.aligned;\n\t"
                     TM_PACK8("%0", "%1", "%2", "%3", "%4", "%5", "%6", "%7", "u") TM_PACK8("%8", "%9", "%10", "%11", "%12", "%13", "%14", "%15", "v")
                     "mov.b64 %16, {w0,w1};\n\tmov.b64 %17, {w2,w3};\n\tmov.b64 %18, {w4,w5};\n\tmov.b64 %19, {w6,w7};\n\t"
                     "mov.b64 %20, {w8,w9};\n\tmov.b64 %21, {w10,w11};\n\t}"
                     : TM_OUT8(d0), TM_OUT8(d1), "=d"(rt[0]), "=d"(rt[1]), "=d"(rt[2]), "=d"(rt[3]), "=d"(rt[4]), "=d"(rt[5])
                     : "r"(tm), "r"(tm_roots) : "memory");
    } else {
        asm volatile("{\n\t.reg .b32 u<16>, v<16>, w<16>;\n\t" TM_LD_L16("u", "%24") TM_LD_L16("v", "%24+16") TM_LD_L16("w", "%25")
                     "tcgen05.wait::ld.sync.aligned;\n\t"
                     TM_PACK8("%0", "%1", "%2", "%3", "%4", "%5", "%6", "%7", "u") TM_PACK8("%8", "%9", "%10", "%11", "%12", "%13", "%14", "%15", "v")
                     TM_PACK8("%16", "%17", "%18", "%19", "%20", "%21", "%22", "%23", "w") "}"
                     : TM_OUT8(d0), TM_OUT8(d1), TM_OUT8(rt) : "r"(tm), "r"(tm_roots) : "memory");
    }
}
// roots only (8 doubles), then wait::ld
__device__ __forceinline__ void tm_load_roots8(double (&rt)[8], const uint32_t tm_roots) {
    asm volatile("{\n\t.reg .b32 w<16>;\n\t"
                 "tcgen05.ld.sync.aligned.32x32b.x16.b32 {w0,w1,w2,w3,w4,w5,w6,w7,w8,w9,w10,w11,w12,w13,w14,w15}, [%8];\n\t"
                 "tcgen05.wait::ld.sync.aligned;\n\t"
                 "mov.b64 %0, {w0,w1};\n\tmov.b64 %1, {w2,w3};\n\tmov.b64 %2, {w4,w5};\n\tmov.b64 %3, {w6,w7};\n\t"
                 "mov.b64 %4, {w8,w9};\n\tmov.b64 %5, {w10,w11};\n\tmov.b64 %6, {w12,w13};\n\tmov.b64 %7, {w14,w15};\n\t}"
                 : "=d"(rt[0]), "=d"(rt[1]), "=d"(rt[2]), "=d"(rt[3]), "=d"(rt[4]), "=d"(rt[5]), "=d"(rt[6]), "=d"(rt[7])
                 : "r"(tm_roots) : "memory");
}
// inverse direction, writer side: st 16x256b.x4 on both lane halves (tm_wait_st() before the loads that follow)
__device__ __forceinline__ void tm_exchange_inv_store16(const Tm8 &lo, const Tm8 &hi, const uint32_t tm, const uint32_t tm_hi) {
    asm volatile("{\n\t.reg .b32 u<16>, v<16>;\n\t"
                 "mov.b64 {u0,u1}, %2;\n\tmov.b64 {u2,u3}, %3;\n\tmov.b64 {u4,u5}, %4;\n\tmov.b64 {u6,u7}, %5;\n\t"
                 "mov.b64 {u8,u9}, %6;\n\tmov.b64 {u10,u11}, %7;\n\tmov.b64 {u12,u13}, %8;\n\tmov.b64 {u14,u15}, %9;\n\t"
                 "mov.b64 {v0,v1}, %10;\n\tmov.b64 {v2,v3}, %11;\n\tmov.b64 {v4,v5}, %12;\n\tmov.b64 {v6,v7}, %13;\n\t"
                 "mov.b64 {v8,v9}, %14;\n\tmov.b64 {v10,v11}, %15;\n\tmov.b64 {v12,v13}, %16;\n\tmov.b64 {v14,v15}, %17;\n\t"
                 "tcgen05.st.sync.aligned.16x256b.x4.b32 [%0], {u0,u1,u2,u3,u4,u5,u6,u7,u8,u9,u10,u11,u12,u13,u14,u15};\n\t"
                 "tcgen05.st.sync.aligned.16x256b.x4.b32 [%1], {v0,v1,v2,v3,v4,v5,v6,v7,v8,v9,v10,v11,v12,v13,v14,v15};\n\t"
                 "}"
                 ::"r"(tm), "r"(tm_hi), "d"(lo.v[0]), "d"(lo.v[1]), "d"(lo.v[2]), "d"(lo.v[3]), "d"(lo.v[4]), "d"(lo.v[5]), "d"(lo.v[6]),
                   "d"(lo.v[7]), "d"(hi.v[0]), "d"(hi.v[1]), "d"(hi.v[2]), "d"(hi.v[3]), "d"(hi.v[4]), "d"(hi.v[5]), "d"(hi.v[6]), "d"(hi.v[7])
                 : "memory");
}

// ---- the exchanges ------------------------------------------------------------------------------------------------
// Writer side of X1 / X2 (st 32x32b): column pair d = 8 s + 4 part + q, where s is the register bit that stays and q
// the two register bits that become lane bits.  X1: registers r = 4 e5 + 2 e4 + e3, s = e3, q = (e5 e4) -> r = 2 q + s.
//                                              X2: registers r = 4 e2 + 2 e1 + e3, s = e1, q = (e2 e3) -> r = 4 (q>>1) + 2 s + (q&1).
template <int WHICH>
__host__ __device__ constexpr int tm_writer_reg(int s, int q) { return WHICH == 1 ? 2 * q + s : 4 * (q >> 1) + 2 * s + (q & 1); }

template <int WHICH>
__device__ __forceinline__ void tm_exchange_fwd_store(const double2 (&x)[8], const uint32_t tm) {
#define XW(s, q) x[tm_writer_reg<WHICH>(s, q)]
    tm_store16(tm, XW(0, 0).x, XW(0, 1).x, XW(0, 2).x, XW(0, 3).x, XW(0, 0).y, XW(0, 1).y, XW(0, 2).y, XW(0, 3).y,
                   XW(1, 0).x, XW(1, 1).x, XW(1, 2).x, XW(1, 3).x, XW(1, 0).y, XW(1, 1).y, XW(1, 2).y, XW(1, 3).y);
#undef XW
}
// Reader side: register r = 4 B + 2 k + s comes from load B (lo | hi), double 2 j + k with j = 2 s + part.
__device__ __forceinline__ void tm_unpack(double2 (&x)[8], const Tm8 &lo, const Tm8 &hi) {
#pragma unroll
    for (int k = 0; k < 2; ++k)
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            x[2 * k + s]     = make_double2(lo.v[2 * (2 * s) + k], lo.v[2 * (2 * s + 1) + k]);
            x[4 + 2 * k + s] = make_double2(hi.v[2 * (2 * s) + k], hi.v[2 * (2 * s + 1) + k]);
        }
}
__device__ __forceinline__ void tm_pack(Tm8 &lo, Tm8 &hi, const double2 (&x)[8]) {
#pragma unroll
    for (int k = 0; k < 2; ++k)
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            lo.v[2 * (2 * s) + k] = x[2 * k + s].x;     lo.v[2 * (2 * s + 1) + k] = x[2 * k + s].y;
            hi.v[2 * (2 * s) + k] = x[4 + 2 * k + s].x; hi.v[2 * (2 * s + 1) + k] = x[4 + 2 * k + s].y;
        }
}
__device__ __forceinline__ void tm_exchange_inv_store(const double2 (&x)[8], const uint32_t tm) {
    Tm8 lo, hi;
    tm_pack(lo, hi, x);
    tm_exchange_inv_store16(lo, hi, tm, tm + (16u << 16));
}
// inverse direction, reader side: the 16 doubles of the thread's lane back into the writer-side registers
template <int WHICH>
__device__ __forceinline__ void tm_unpack_inv(double2 (&x)[8], const double (&d)[16]) {
#pragma unroll
    for (int s = 0; s < 2; ++s)
#pragma unroll
        for (int q = 0; q < 4; ++q) x[tm_writer_reg<WHICH>(s, q)] = make_double2(d[8 * s + q], d[8 * s + 4 + q]);
}

// ---- small passes -------------------------------------------------------------------------------------------------
// P3a: registers r = 4 e2 + 2 e1 + e3; level 6 pairs (r, r+4) with root A (e3 = 0) or iA (e3 = 1);
// level 7 pairs (r, r+2) with root C, iC (e3 = 0; e2 = 0, 1) or D, iD (e3 = 1).
__device__ __forceinline__ void fwd_p3a(double2 (&x)[8], const double2 A, const double2 C, const double2 D) {
    ct_bfly(x[0], x[4], A);
    ct_bfly(x[2], x[6], A);
    ct_bfly_i(x[1], x[5], A);
    ct_bfly_i(x[3], x[7], A);
    ct_bfly(x[0], x[2], C);
    ct_bfly_i(x[4], x[6], C);
    ct_bfly(x[1], x[3], D);
    ct_bfly_i(x[5], x[7], D);
}
__device__ __forceinline__ void inv_p3a(double2 (&x)[8], const double2 A, const double2 C, const double2 D) {
    gs_bfly(x[0], x[2], C);
    gs_bfly_i(x[4], x[6], C);
    gs_bfly(x[1], x[3], D);
    gs_bfly_i(x[5], x[7], D);
    gs_bfly(x[0], x[4], A);
    gs_bfly(x[2], x[6], A);
    gs_bfly_i(x[1], x[5], A);
    gs_bfly_i(x[3], x[7], A);
}
// P3b: registers r = 4 e0 + 2 e7 + e1; level 8 pairs (r, r+4) with root E0, iE0 (e7 = 0; e1 = 0, 1), E1, iE1 (e7 = 1).
__device__ __forceinline__ void fwd_p3b(double2 (&x)[8], const double2 E0, const double2 E1) {
    ct_bfly(x[0], x[4], E0);
    ct_bfly_i(x[1], x[5], E0);
    ct_bfly(x[2], x[6], E1);
    ct_bfly_i(x[3], x[7], E1);
}
__device__ __forceinline__ void inv_p3b(double2 (&x)[8], const double2 E0, const double2 E1) {
    gs_bfly(x[0], x[4], E0);
    gs_bfly_i(x[1], x[5], E0);
    gs_bfly(x[2], x[6], E1);
    gs_bfly_i(x[3], x[7], E1);
}

// ---- the transform object of one thread -----------------------------------------------------------------------
// t = thread in the 64-thread group; tm = tensor-memory address of this warp's 128-column window (lane field = the
// warp's lane quarter, column field = window base).
struct FftTm {
    static constexpr int kBufElemsPerTransform = kTmBufElems;
    static constexpr int kExchangeBuffers = 2;
    int t;
    uint32_t tm;   // exchange window(s) of this warp
    uint32_t tmr;  // roots of this thread: 40 columns of its lane (per warp, or shared by the warps of a lane quarter)
    int rd0;       // element index of register 0 in the P2 layout (registers add 8 r): shared-memory read base

    __device__ __forceinline__ static int elem_p2(int t) {
        const int lane = t & 31;
        return ((t >> 5) << 8) | ((lane & 3) << 6) | (((lane >> 4) & 1) << 2) | (((lane >> 3) & 1) << 1) | ((lane >> 2) & 1);
    }
    // one-time: this thread's roots into its tensor-memory lane
    __device__ __forceinline__ void init(int t_, uint32_t tm_) { init(t_, tm_, tm_ + kTmColRootsBase); }
    __device__ __forceinline__ void init(int t_, uint32_t tm_, uint32_t tmr_) {
        t = t_;
        tm = tm_;
        tmr = tmr_;
        rd0 = elem_p2(t);
        const int lane = t & 31, w = t >> 5;
        const RootSet &p2 = c_tw.p2[4 * w + (lane & 3)];
        tm_store4(tmr + kTmColRoots2, p2.A.x, p2.A.y, p2.B.x, p2.B.y);
        tm_store4(tmr + kTmColRoots2 + 8, p2.C.x, p2.C.y, p2.D.x, p2.D.y);
        const double2 *a = c_tw_tm.p3a[16 * w + (lane & 15)];
        tm_store4(tmr + kTmColRoots3a, a[0].x, a[0].y, a[1].x, a[1].y);
        tm_store4(tmr + kTmColRoots3a + 8, a[2].x, a[2].y, 0.0, 0.0);
        const double2 *b = c_tw_tm.p3b[t];
        tm_store4(tmr + kTmColRoots3b, b[0].x, b[0].y, b[1].x, b[1].y);
        tm_wait_st();
    }

    // coefficients (x[r] = z_{t+64r}) -> evaluations in the P3b layout
    template <bool PRE = false, class Barrier>
    __device__ __forceinline__ void forward(double2 (&x)[8], double2 *buf, const Barrier &bar) const {
        if (PRE) fwd_radix8_first_pretwisted(x); else fwd_radix8(x, c_tw.p1);
#pragma unroll
        for (int r = 0; r < 8; ++r) buf[tm_slot(t + 64 * r)] = x[r];
        double rt[8];
        tm_load_roots8(rt, tmr + kTmColRoots2);     // x is dead here: the roots cost no registers before the barrier
        bar();
        const int base = tm_slot(rd0);             // + 8 r: the swizzle depends on element bits 7,6 only
#pragma unroll
        for (int r = 0; r < 8; ++r) x[r] = buf[base + 8 * r];
        {
            RootSet w2;
            w2.A = make_double2(rt[0], rt[1]); w2.B = make_double2(rt[2], rt[3]);
            w2.C = make_double2(rt[4], rt[5]); w2.D = make_double2(rt[6], rt[7]);
            fwd_radix8(x, w2);
        }
        Tm8 lo, hi;
        tm_exchange_fwd_store<1>(x, tm);
        tm_wait_st();
        tm_exchange_fwd_load<6>(lo, hi, rt, tm, tm + (16u << 16), tmr + kTmColRoots3a);
        tm_unpack(x, lo, hi);
        fwd_p3a(x, make_double2(rt[0], rt[1]), make_double2(rt[2], rt[3]), make_double2(rt[4], rt[5]));
        tm_exchange_fwd_store<2>(x, tm);
        tm_wait_st();
        tm_exchange_fwd_load<4>(lo, hi, rt, tm, tm + (16u << 16), tmr + kTmColRoots3b);
        tm_unpack(x, lo, hi);
        fwd_p3b(x, make_double2(rt[0], rt[1]), make_double2(rt[2], rt[3]));
    }

    // TWO independent transforms of one thread, skewed by half a stage: while the tensor-memory round trip of one
    // is in flight the butterflies of the other run, so the exchange latency (about 350 clk per round trip,
    // profiles/r2_microbench_tmem.txt) is hidden inside the warp instead of by other warps, and every butterfly
    // level has twice the independent DFMA chains.  bufa / bufb: two exchange buffers; the caller guarantees that
    // nobody still reads them (group barrier after the previous transform's reads).
    struct NoMid { __device__ __forceinline__ void operator()() const {} };
    // mid(): called when xa is complete and xb's last exchange is in flight (the caller's spectral MAC of xa)
    // PRE: xa[4..7], xb[4..7] arrive twisted by the level-0 root (digits_pretwisted)
    template <bool PRE = false, class Barrier, class Mid = NoMid>
    __device__ __forceinline__ void forward2(double2 (&xa)[8], double2 (&xb)[8], double2 *bufa, double2 *bufb,
                                             const Barrier &bar, const Mid &mid = Mid()) const {
        const uint32_t tma = tm, tmb = tm + kTmColExB;
        if (PRE) fwd_radix8_first_pretwisted(xa); else fwd_radix8(xa, c_tw.p1);
#pragma unroll
        for (int r = 0; r < 8; ++r) bufa[tm_slot(t + 64 * r)] = xa[r];
        if (PRE) fwd_radix8_first_pretwisted(xb); else fwd_radix8(xb, c_tw.p1);
#pragma unroll
        for (int r = 0; r < 8; ++r) bufb[tm_slot(t + 64 * r)] = xb[r];
        double rt[8], r3[8];
        tm_load_roots8(rt, tmr + kTmColRoots2);
        bar();
        const int base = tm_slot(rd0);
#pragma unroll
        for (int r = 0; r < 8; ++r) xa[r] = bufa[base + 8 * r];
#pragma unroll
        for (int r = 0; r < 8; ++r) xb[r] = bufb[base + 8 * r];
        RootSet w2;
        w2.A = make_double2(rt[0], rt[1]); w2.B = make_double2(rt[2], rt[3]);
        w2.C = make_double2(rt[4], rt[5]); w2.D = make_double2(rt[6], rt[7]);
        Tm8 lo, hi;
        fwd_radix8(xa, w2);
        tm_exchange_fwd_store<1>(xa, tma);
        fwd_radix8(xb, w2);                     // hides a's store
        tm_wait_st();
        tm_exchange_fwd_load<6>(lo, hi, r3, tma, tma + (16u << 16), tmr + kTmColRoots3a);
        tm_exchange_fwd_store<1>(xb, tmb);
        tm_unpack(xa, lo, hi);
        const double2 A = make_double2(r3[0], r3[1]), C = make_double2(r3[2], r3[3]), D = make_double2(r3[4], r3[5]);
        fwd_p3a(xa, A, C, D);                   // hides b's store
        tm_wait_st();
        tm_exchange_fwd_load<0>(lo, hi, rt, tmb, tmb + (16u << 16), 0);
        tm_exchange_fwd_store<2>(xa, tma);
        tm_unpack(xb, lo, hi);
        fwd_p3a(xb, A, C, D);
        tm_wait_st();
        tm_exchange_fwd_load<4>(lo, hi, r3, tma, tma + (16u << 16), tmr + kTmColRoots3b);
        tm_exchange_fwd_store<2>(xb, tmb);
        tm_unpack(xa, lo, hi);
        const double2 E0 = make_double2(r3[0], r3[1]), E1 = make_double2(r3[2], r3[3]);
        fwd_p3b(xa, E0, E1);
        tm_wait_st();
        tm_exchange_fwd_load<0>(lo, hi, rt, tmb, tmb + (16u << 16), 0);
        mid();
        tm_unpack(xb, lo, hi);
        fwd_p3b(xb, E0, E1);
    }

    template <bool DEFER_LAST, class Barrier>
    __device__ __forceinline__ void inverse2(double2 (&xa)[8], double2 (&xb)[8], double2 *bufa, double2 *bufb,
                                             const Barrier &bar) const {
        const uint32_t tma = tm, tmb = tm + kTmColExB;
        double rt[8], r3[8], d[16];
        tm_load_roots8(r3, tmr + kTmColRoots3b);
        {
            const double2 E0 = make_double2(r3[0], r3[1]), E1 = make_double2(r3[2], r3[3]);
            inv_p3b(xa, E0, E1);
            tm_exchange_inv_store(xa, tma);
            inv_p3b(xb, E0, E1);
        }
        tm_wait_st();
        tm_load16<6>(d, r3, tma, tmr + kTmColRoots3a);
        tm_exchange_inv_store(xb, tmb);
        tm_unpack_inv<2>(xa, d);
        const double2 A = make_double2(r3[0], r3[1]), C = make_double2(r3[2], r3[3]), D = make_double2(r3[4], r3[5]);
        inv_p3a(xa, A, C, D);
        tm_wait_st();
        tm_load16<0>(d, rt, tmb, 0);
        tm_exchange_inv_store(xa, tma);
        tm_unpack_inv<2>(xb, d);
        inv_p3a(xb, A, C, D);
        tm_wait_st();
        tm_load16<8>(d, rt, tma, tmr + kTmColRoots2);
        tm_exchange_inv_store(xb, tmb);
        tm_unpack_inv<1>(xa, d);
        RootSet w2;
        w2.A = make_double2(rt[0], rt[1]); w2.B = make_double2(rt[2], rt[3]);
        w2.C = make_double2(rt[4], rt[5]); w2.D = make_double2(rt[6], rt[7]);
        inv_radix8(xa, w2);
        const int base = tm_slot(rd0);
#pragma unroll
        for (int r = 0; r < 8; ++r) bufa[base + 8 * r] = xa[r];
        tm_wait_st();
        tm_load16<0>(d, r3, tmb, 0);
        tm_unpack_inv<1>(xb, d);
        inv_radix8(xb, w2);
#pragma unroll
        for (int r = 0; r < 8; ++r) bufb[base + 8 * r] = xb[r];
        bar();
#pragma unroll
        for (int r = 0; r < 8; ++r) xa[r] = bufa[tm_slot(t + 64 * r)];
#pragma unroll
        for (int r = 0; r < 8; ++r) xb[r] = bufb[tm_slot(t + 64 * r)];
        if (DEFER_LAST) { inv_radix8_last(xa); inv_radix8_last(xb); }
        else { inv_radix8(xa, c_tw.p1); inv_radix8(xb, c_tw.p1); }
    }

    // evaluations (P3b layout) -> 512 * coefficients (x[r] = 512 z_{t+64r}); DEFER_LAST as in negacyclic_inverse
    template <bool DEFER_LAST, class Barrier>
    __device__ __forceinline__ void inverse(double2 (&x)[8], double2 *buf, const Barrier &bar) const {
        double rt[8], d[16];
        tm_load_roots8(rt, tmr + kTmColRoots3b);    // (E0, E1 and 8 unused columns)
        inv_p3b(x, make_double2(rt[0], rt[1]), make_double2(rt[2], rt[3]));
        tm_exchange_inv_store(x, tm);
        tm_wait_st();
        tm_load16<6>(d, rt, tm, tmr + kTmColRoots3a);
        tm_unpack_inv<2>(x, d);
        inv_p3a(x, make_double2(rt[0], rt[1]), make_double2(rt[2], rt[3]), make_double2(rt[4], rt[5]));
        tm_exchange_inv_store(x, tm);
        tm_wait_st();
        tm_load16<8>(d, rt, tm, tmr + kTmColRoots2);
        tm_unpack_inv<1>(x, d);
        {
            RootSet w2;
            w2.A = make_double2(rt[0], rt[1]); w2.B = make_double2(rt[2], rt[3]);
            w2.C = make_double2(rt[4], rt[5]); w2.D = make_double2(rt[6], rt[7]);
            inv_radix8(x, w2);
        }
        const int base = tm_slot(rd0);
#pragma unroll
        for (int r = 0; r < 8; ++r) buf[base + 8 * r] = x[r];
        bar();
#pragma unroll
        for (int r = 0; r < 8; ++r) x[r] = buf[tm_slot(t + 64 * r)];
        if (DEFER_LAST) inv_radix8_last(x);
        else inv_radix8(x, c_tw.p1);
    }
};

// Same transform; selects the CMux step that parks the two spectral accumulators in tensor memory between the
// row pairs of an external product (kernels.cu, cmux_step_pairs_tmacc): 64 columns per thread behind the roots.
constexpr int kTmColAcc = 104;            // 104-167: acc0, acc1 (32 doubles)
constexpr int kTmColsPerWarpAcc = 168;
struct FftTmAcc : FftTm {};
// Single transforms, accumulators in tensor memory for EVERY key row, one exchange buffer: what fits eight groups
// (16 warps, 128 registers per thread) on an SM.  Tensor-memory columns of a lane quarter: 0-39 the roots, shared
// by the four warps of the quarter (they hold the same half of the 64 threads); then per warp 32 exchange + 64
// accumulator columns.
constexpr int kTmSColsRoots = 64, kTmSColsPerWarp = 96, kTmSColAcc = 32;
struct FftTmS : FftTm {
    static constexpr int kExchangeBuffers = 1;
};

// the shared-memory transform of negacyclic_fft.cuh behind the same interface
struct FftSmem {
    static constexpr int kBufElemsPerTransform = kBufElems;
    static constexpr int kExchangeBuffers = 2;
    int t;
    RootSet w2, w3;
    __device__ __forceinline__ void init(int t_, uint32_t) {
        t = t_;
        w2 = load_rootset(c_tw.p2[t >> 3]);
        w3 = load_rootset(c_tw.p3[t]);
    }
    template <bool PRE = false, class Barrier>
    __device__ __forceinline__ void forward(double2 (&x)[8], double2 *buf, const Barrier &bar) const {
        negacyclic_forward<PRE>(x, buf, t, w2, w3, bar);
    }
    template <bool DEFER_LAST, class Barrier>
    __device__ __forceinline__ void inverse(double2 (&x)[8], double2 *buf, const Barrier &bar) const {
        negacyclic_inverse<DEFER_LAST>(x, buf, t, w2, w3, bar);
    }
};

}  // namespace b200
