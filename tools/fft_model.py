"""fft_model.py -- CPU model of the negacyclic transform's data movement (design aid, not product code).

The transform (csrc/negacyclic_fft.cuh) is in place on 512 complex elements e = 0..511: level L pairs e with
e ^ (256 >> L), root s(L, e >> (9-L)), and leaves z(rho_e) in element e.  What a kernel variant chooses is only
WHICH (thread, register) holds e during each pass and how elements move between passes.  This model checks a
layout plan before it is written in CUDA:
  * every butterfly of a pass has both inputs in one thread,
  * which distinct roots a thread needs per pass,
  * the exchange between two passes is the permutation the hardware path offers (shared memory: any;
    tensor memory st 32x32b -> ld 16x256b: the fixed 4x4 lane transpose measured by scratch/mb_r2 `map`),
  * the result equals direct evaluation at the 512 roots of X^512 = i.
"""
import cmath
import math

import numpy as np

M = 512


def brv(v, bits):
    r = 0
    for i in range(bits):
        r |= ((v >> i) & 1) << (bits - 1 - i)
    return r


def theta(level, b):
    return (math.pi / 2 + 2 * math.pi * brv(b, level)) / (1 << level)


def root(level, b):
    return cmath.exp(1j * theta(level, b) / 2)


def level_inplace(x, level):
    h = 256 >> level
    for e in range(M):
        if e & h:
            continue
        s = root(level, e >> (9 - level))
        u, v = x[e], x[e + h]
        x[e], x[e + h] = u + s * v, u - s * v


def forward_reference(z):
    x = np.array(z, dtype=complex)
    for level in range(9):
        level_inplace(x, level)
    return x


def direct(z):
    out = np.zeros(M, dtype=complex)
    j = np.arange(M)
    for e in range(M):
        rho = cmath.exp(1j * theta(9, e))
        out[e] = np.sum(np.asarray(z) * rho ** j)
    return out


class Layout:
    """element e <-> (thread t in 0..63, register r in 0..7), given as a list of 9 (kind, bit) entries from
    element bit 8 down to element bit 0, kind in {'t', 'r'}: where that element bit lives."""

    def __init__(self, spec):
        assert len(spec) == 9
        self.spec = spec
        assert sorted(b for k, b in spec if k == 't') == list(range(6))
        assert sorted(b for k, b in spec if k == 'r') == list(range(3))

    def where(self, e):
        t = r = 0
        for i, (k, b) in enumerate(self.spec):
            bit = (e >> (8 - i)) & 1
            if k == 't':
                t |= bit << b
            else:
                r |= bit << b
        return t, r

    def elem(self, t, r):
        e = 0
        for i, (k, b) in enumerate(self.spec):
            bit = ((t if k == 't' else r) >> b) & 1
            e |= bit << (8 - i)
        return e

    def reg_bit_of_level(self, level):
        k, b = self.spec[level]      # element bit 8-level is spec[level]
        return b if k == 'r' else None


def check_pass(layout, levels):
    """all butterflies of `levels` are thread local; returns {thread: sorted list of (level, block)} roots used"""
    need = {}
    for level in levels:
        assert layout.reg_bit_of_level(level) is not None, "level %d is not register resident" % level
        for e in range(M):
            t, _ = layout.where(e)
            need.setdefault(t, set()).add((level, e >> (9 - level)))
    return need


def tmem_exchange(lay_w):
    """st 32x32b (lane = thread within the warp, columns = registers) -> ld 16x256b on both lane halves.
    Hardware map (mb_r2 `map`): reader thread T gets, per 16-lane half B in {0, 16} and 8-column group j,
        regs (4j+0, 4j+1) <- lane B + T/4,     columns 8j + 2(T%4) + {0,1}
        regs (4j+2, 4j+3) <- lane B + T/4 + 8, columns 8j + 2(T%4) + {0,1}
    With a double2 register x[r] stored in columns 4r..4r+3 (re.lo, re.hi, im.lo, im.hi) a 64-bit unit is one
    double; the kernel stores double d (d = 2r + part) in column pair d.  Returns the function
    (reader lane T, slot) -> (writer lane, writer double index d)."""
    raise NotImplementedError


if __name__ == "__main__":
    rng = np.random.default_rng(1)
    z = rng.integers(-512, 512, M) + 1j * rng.integers(-512, 512, M)
    a = forward_reference(z)
    b = direct(z)
    print("in-place levels vs direct evaluation: max err %.3e" % np.max(np.abs(a - b)))
    # the shipped layouts
    p1 = Layout([('r', 2), ('r', 1), ('r', 0), ('t', 5), ('t', 4), ('t', 3), ('t', 2), ('t', 1), ('t', 0)])
    p2 = Layout([('t', 5), ('t', 4), ('t', 3), ('r', 2), ('r', 1), ('r', 0), ('t', 2), ('t', 1), ('t', 0)])
    p3 = Layout([('t', 5), ('t', 4), ('t', 3), ('t', 2), ('t', 1), ('t', 0), ('r', 2), ('r', 1), ('r', 0)])
    for name, lay, lv in (("pass 1", p1, (0, 1, 2)), ("pass 2", p2, (3, 4, 5)), ("pass 3", p3, (6, 7, 8))):
        need = check_pass(lay, lv)
        print(name, "thread local; roots per thread:", len(need[0]), "distinct root sets:", len({frozenset(v) for v in need.values()}))
    assert p1.elem(5, 3) == 5 + 64 * 3 and p2.elem(13, 2) == 64 * 1 + 5 + 8 * 2 and p3.elem(7, 4) == 8 * 7 + 4


# ---------------------------------------------------------------------------------------------------
# inverse radix-8 pass with the cos factors deferred (negacyclic_fft.cuh inv_radix8_deferred): algebra check
# ---------------------------------------------------------------------------------------------------
def ct(level, b):
    a = theta(level, b) / 2
    return math.cos(a), math.tan(a)


def gs(u, v, c, t):          # (a, b) <- (a + b, (a - b) conj(s)),  s = c (1 + i t)
    d = u - v
    return u + v, c * complex(d.real + t * d.imag, d.imag - t * d.real)


def gs_i(u, v, c, t):        # root i s: conj(i s) = -i conj(s)
    d = u - v
    w = complex(d.real + t * d.imag, d.imag - t * d.real)
    return u + v, c * complex(w.imag, -w.real)


def inv_radix8(x, A, B, C, D):
    x = list(x)
    x[0], x[1] = gs(x[0], x[1], *C); x[2], x[3] = gs_i(x[2], x[3], *C)
    x[4], x[5] = gs(x[4], x[5], *D); x[6], x[7] = gs_i(x[6], x[7], *D)
    x[0], x[2] = gs(x[0], x[2], *B); x[1], x[3] = gs(x[1], x[3], *B)
    x[4], x[6] = gs_i(x[4], x[6], *B); x[5], x[7] = gs_i(x[5], x[7], *B)
    for r in range(4):
        x[r], x[r + 4] = gs(x[r], x[r + 4], *A)
    return x


def gsu(u, v, t, rho=None, rot=False):     # unscaled: (a + rho b, (a - rho b)(1 - i t)) [then * -i if rot]
    if rho is None:
        s, d = u + v, u - v
    else:
        s, d = u + rho * v, u - rho * v     # two FMA per component instead of two ADD
    w = complex(d.real + t * d.imag, d.imag - t * d.real)
    return s, (complex(w.imag, -w.real) if rot else w)


def inv_radix8_deferred(x, A, B, C, D):
    """same result as inv_radix8 after multiplying x[r] by scale[r]: 6 FP64 ops per butterfly + 14 scalings"""
    x = list(x)
    cA, tA = A; cB, tB = B; cC, tC = C; cD, tD = D
    x[0], x[1] = gsu(x[0], x[1], tC); x[2], x[3] = gsu(x[2], x[3], tC, rot=True)
    x[4], x[5] = gsu(x[4], x[5], tD); x[6], x[7] = gsu(x[6], x[7], tD, rot=True)
    # pending: [1, cC, 1, cC, 1, cD, 1, cD]
    x[0], x[2] = gsu(x[0], x[2], tB); x[1], x[3] = gsu(x[1], x[3], tB)
    x[4], x[6] = gsu(x[4], x[6], tB, rot=True); x[5], x[7] = gsu(x[5], x[7], tB, rot=True)
    # pending: [1, cC, cB, cC cB, 1, cD, cB, cD cB]
    rho = cD / cC
    x[0], x[4] = gsu(x[0], x[4], tA)
    x[1], x[5] = gsu(x[1], x[5], tA, rho=rho)
    x[2], x[6] = gsu(x[2], x[6], tA)
    x[3], x[7] = gsu(x[3], x[7], tA, rho=rho)
    scale = [1, cC, cB, cC * cB, cA, cC * cA, cB * cA, cC * cB * cA]
    return x, scale


if __name__ == "__main__":
    rng = np.random.default_rng(2)
    worst = 0.0
    for level, b in ((0, 0), (3, 5), (6, 37)):
        A = ct(level, b); B = ct(level + 1, 2 * b); C = ct(level + 2, 4 * b); D = ct(level + 2, 4 * b + 2)
        x = rng.normal(size=8) + 1j * rng.normal(size=8)
        want = inv_radix8(x, A, B, C, D)
        got, scale = inv_radix8_deferred(x, A, B, C, D)
        worst = max(worst, max(abs(w - g * s) for w, g, s in zip(want, got, scale)))
    print("deferred-cos inverse radix-8 vs plain: max err %.2e" % worst)
