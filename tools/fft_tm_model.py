"""fft_tm_model.py -- bit-level model of the transform whose last two exchanges go through tensor memory
(csrc/negacyclic_fft_tm.cuh).  Design aid and generator of the operand orders; not product code.

A group of 64 threads (2 warps) x 8 double2 registers holds the 512 elements e.  Passes and exchanges:

    P1   levels 0-2  regs (e8 e7 e6)   thread t = e & 63                          roots: warp uniform
    S    shared memory, slot(e) = e ^ (((e >> 6) & 3) << 1)   (conflict free both ways, no padding)
    P2   levels 3-5  regs (e5 e4 e3)   warp e8, lane (e2 e1 e0 e7 e6)              roots: 8 sets  (e8 e7 e6)
    X1   tensor memory: regs (e5 e4) <-> lane bits (4,3) = (e2 e1); e3 stays
    P3a  levels 6-7  regs (e2 e1 e3)   warp e8, lane (e0 e7 e6 e5 e4)              roots: 32 sets (e8..e4)
    X2   tensor memory: regs (e2 e3) <-> lane bits (4,3) = (e0 e7); e1 stays
    P3b  level 8     regs (e0 e7 e1)   warp e8, lane (e6 e5 e4 e2 e3)              roots: 2 per thread

The tensor-memory exchange is the hardware map measured by scratch/mb_r2 `map` (profiles/r2_microbench_tmem.txt):
st 32x32b writes lane = thread, column pair d (one double); ld 16x256b.x4 at lane offset 16 B returns to thread T,
for j = 0..3 and k = 0, 1, the double of lane 16 B + 8 k + T/4, column pair 4 j + T%4.
"""
import cmath
import math

import numpy as np

from fft_model import M, forward_reference, root, theta


def bits(v, n):
    return [(v >> i) & 1 for i in range(n)]          # little endian list


def slot(e):
    return e ^ (((e >> 6) & 3) << 1)


# ---- layouts: (t, r) -> e -------------------------------------------------------------------------
def L1(t, r):
    return t + 64 * r


def L2(t, r):
    w, lane = t >> 5, t & 31
    l = bits(lane, 5)
    e8, e2, e1, e0, e7, e6 = w, l[4], l[3], l[2], l[1], l[0]
    e5, e4, e3 = (r >> 2) & 1, (r >> 1) & 1, r & 1
    return (e8 << 8) | (e7 << 7) | (e6 << 6) | (e5 << 5) | (e4 << 4) | (e3 << 3) | (e2 << 2) | (e1 << 1) | e0


def L3(t, r):
    w, lane = t >> 5, t & 31
    l = bits(lane, 5)
    e8, e0, e7, e6, e5, e4 = w, l[4], l[3], l[2], l[1], l[0]
    e2, e1, e3 = (r >> 2) & 1, (r >> 1) & 1, r & 1
    return (e8 << 8) | (e7 << 7) | (e6 << 6) | (e5 << 5) | (e4 << 4) | (e3 << 3) | (e2 << 2) | (e1 << 1) | e0


def L4(t, r):
    w, lane = t >> 5, t & 31
    l = bits(lane, 5)
    e8, e6, e5, e4, e2, e3 = w, l[4], l[3], l[2], l[1], l[0]
    e0, e7, e1 = (r >> 2) & 1, (r >> 1) & 1, r & 1
    return (e8 << 8) | (e7 << 7) | (e6 << 6) | (e5 << 5) | (e4 << 4) | (e3 << 3) | (e2 << 2) | (e1 << 1) | e0


def check_bijection(L):
    seen = sorted(L(t, r) for t in range(64) for r in range(8))
    assert seen == list(range(M))


# ---- hardware tensor-memory exchange (one warp) -----------------------------------------------------
def tmem_exchange(regs, order_st, order_ld):
    """regs[lane][r] complex.  order_st(d) -> (r, part) says which double a writer puts in column pair d;
    order_ld(B, j, k) -> (r, part) says where the reader files the double of ld at lane offset 16 B, repeat j,
    lane half k.  Returns the readers' registers."""
    tm = np.zeros((32, 16))                       # [lane][column pair]
    for lane in range(32):
        for d in range(16):
            r, part = order_st(d)
            v = regs[lane][r]
            tm[lane][d] = v.imag if part else v.real
    out = [[[None, None] for _ in range(8)] for _ in range(32)]
    for T in range(32):
        for B in range(2):
            for j in range(4):
                for k in range(2):
                    r, part = order_ld(B, j, k)
                    out[T][r][part] = tm[16 * B + 8 * k + (T >> 2)][4 * j + (T & 3)]
    return [[complex(out[T][r][0], out[T][r][1]) for r in range(8)] for T in range(32)]


def tmem_exchange_inv(regs, order_st, order_ld):
    """the opposite direction: st 16x256b.x4 (both lane halves) -> ld 32x32b.x32; same two order functions"""
    tm = np.zeros((32, 16))
    for T in range(32):
        for B in range(2):
            for j in range(4):
                for k in range(2):
                    r, part = order_ld(B, j, k)
                    v = regs[T][r]
                    tm[16 * B + 8 * k + (T >> 2)][4 * j + (T & 3)] = v.imag if part else v.real
    out = [[[None, None] for _ in range(8)] for _ in range(32)]
    for lane in range(32):
        for d in range(16):
            r, part = order_st(d)
            out[lane][r][part] = tm[lane][d]
    return [[complex(out[l][r][0], out[l][r][1]) for r in range(8)] for l in range(32)]


# X1: writer regs r = (e5 e4 e3); column pair d = 8 e3 + 4 part + 2 e5 + e4
def x1_st(d):
    e3, part, e5, e4 = (d >> 3) & 1, (d >> 2) & 1, (d >> 1) & 1, d & 1
    return (e5 << 2) | (e4 << 1) | e3, part


# reader regs r' = (e2 e1 e3): B = e2 (writer lane bit 4), k = e1 (writer lane bit 3), j = (e3, part)
def x1_ld(B, j, k):
    return (B << 2) | (k << 1) | (j >> 1), j & 1


# X2: writer regs r' = (e2 e1 e3); column pair d = 8 e1 + 4 part + 2 e2 + e3
def x2_st(d):
    e1, part, e2, e3 = (d >> 3) & 1, (d >> 2) & 1, (d >> 1) & 1, d & 1
    return (e2 << 2) | (e1 << 1) | e3, part


# reader regs r'' = (e0 e7 e1): B = e0, k = e7, j = (e1, part)
def x2_ld(B, j, k):
    return (B << 2) | (k << 1) | (j >> 1), j & 1


# ---- butterflies on registers -------------------------------------------------------------------------
def bfly_level(x, layout, t, level, rbit):
    """in-place level on the 8 registers of thread t; rbit = register bit that carries element bit 8-level"""
    for r in range(8):
        if r & (1 << rbit):
            continue
        r2 = r | (1 << rbit)
        e = layout(t, r)
        assert layout(t, r2) == e + (256 >> level), (t, r, level)
        s = root(level, e >> (9 - level))
        u, v = x[r], x[r2]
        x[r], x[r2] = u + s * v, u - s * v


def inv_bfly_level(x, layout, t, level, rbit):
    for r in range(8):
        if r & (1 << rbit):
            continue
        r2 = r | (1 << rbit)
        e = layout(t, r)
        s = root(level, e >> (9 - level))
        a, b = x[r], x[r2]
        x[r], x[r2] = a + b, (a - b) * s.conjugate()


def forward(z):
    R = [[z[L1(t, r)] for r in range(8)] for t in range(64)]
    for t in range(64):
        for level, rbit in ((0, 2), (1, 1), (2, 0)):
            bfly_level(R[t], L1, t, level, rbit)
    buf = [None] * M
    for t in range(64):
        for r in range(8):
            buf[slot(L1(t, r))] = R[t][r]
    R = [[buf[slot(L2(t, r))] for r in range(8)] for t in range(64)]
    for t in range(64):
        for level, rbit in ((3, 2), (4, 1), (5, 0)):
            bfly_level(R[t], L2, t, level, rbit)
    R = tmem_exchange(R[:32], x1_st, x1_ld) + tmem_exchange(R[32:], x1_st, x1_ld)
    for t in range(64):
        for level, rbit in ((6, 2), (7, 1)):
            bfly_level(R[t], L3, t, level, rbit)
    R = tmem_exchange(R[:32], x2_st, x2_ld) + tmem_exchange(R[32:], x2_st, x2_ld)
    for t in range(64):
        bfly_level(R[t], L4, t, 8, 2)
    return R


def inverse(R):
    R = [list(x) for x in R]
    for t in range(64):
        inv_bfly_level(R[t], L4, t, 8, 2)
    R = tmem_exchange_inv(R[:32], x2_st, x2_ld) + tmem_exchange_inv(R[32:], x2_st, x2_ld)
    for t in range(64):
        for level, rbit in ((7, 1), (6, 2)):
            inv_bfly_level(R[t], L3, t, level, rbit)
    R = tmem_exchange_inv(R[:32], x1_st, x1_ld) + tmem_exchange_inv(R[32:], x1_st, x1_ld)
    for t in range(64):
        for level, rbit in ((5, 0), (4, 1), (3, 2)):
            inv_bfly_level(R[t], L2, t, level, rbit)
    buf = [None] * M
    for t in range(64):
        for r in range(8):
            buf[slot(L2(t, r))] = R[t][r]
    R = [[buf[slot(L1(t, r))] for r in range(8)] for t in range(64)]
    for t in range(64):
        for level, rbit in ((2, 0), (1, 1), (0, 2)):
            inv_bfly_level(R[t], L1, t, level, rbit)
    return R


def conflicts(layout, r):
    """worst number of lanes of a quarter warp that hit the same 16-byte bank group for register r"""
    worst = 1
    for t0 in range(0, 64, 8):
        groups = {}
        for t in range(t0, t0 + 8):
            groups.setdefault(slot(layout(t, r)) & 7, []).append(t)
        worst = max(worst, max(len(v) for v in groups.values()))
    return worst


if __name__ == "__main__":
    for L in (L1, L2, L3, L4):
        check_bijection(L)
    rng = np.random.default_rng(3)
    z = rng.integers(-512, 512, M) + 1j * rng.integers(-512, 512, M)
    want = forward_reference(z)
    R = forward(z)
    err = max(abs(R[t][r] - want[L4(t, r)]) for t in range(64) for r in range(8))
    print("forward through the modelled exchanges vs in-place reference: max err %.2e" % err)
    back = inverse(R)
    err = max(abs(back[t][r] / 512 - z[L1(t, r)]) for t in range(64) for r in range(8))
    print("inverse(forward(z)) / 512 - z: max err %.2e" % err)
    print("shared-memory exchange, worst bank multiplicity per quarter warp: write (L1) %d, read (L2) %d"
          % (max(conflicts(L1, r) for r in range(8)), max(conflicts(L2, r) for r in range(8))))
    # roots per thread
    for name, L, levels in (("P2", L2, (3, 4, 5)), ("P3a", L3, (6, 7)), ("P3b", L4, (8,))):
        sets = set()
        for t in range(64):
            sets.add(frozenset((lv, L(t, r) >> (9 - lv)) for lv in levels for r in range(8)))
        print(name, "distinct root sets over the 64 threads:", len(sets), "roots per thread:", len(next(iter(sets))))
