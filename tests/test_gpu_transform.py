"""The negacyclic transform alone, restating the checks the reference keeps for its FFT
processors (libs/tfhe/src/test/lagrangehalfc_test.cpp:16-36: inverse(forward) = identity; :130-218: product through
the transform = Karatsuba product) for BOTH transforms of this engine: the shared-memory one ("v2") and the one with
its warp-local exchanges in tensor memory (default).  Integer results, so the bar is exact equality."""
import ctypes as C

import numpy as np
import pytest

import reflib

pytestmark = [pytest.mark.gpu]
N = 1024


def negacyclic_mul_exact(a, b):
    """a * b mod (X^N + 1) mod 2^32 in exact int64 arithmetic (|a| <= 2^9, |b| < 2^31, N = 2^10: < 2^51)."""
    out = np.empty_like(b)
    for i in range(a.shape[0]):
        full = np.convolve(a[i].astype(np.int64), b[i].astype(np.int64))          # length 2N - 1, exact
        c = full[:N].copy()
        c[: N - 1] -= full[N:]
        out[i] = (c & 0xFFFFFFFF).astype(np.uint32).astype(np.int32)
    return out


@pytest.fixture(params=["v2", "auto"])
def transform(request):
    from eoc_tfhe_b200 import set_blind_rotate_kernel
    set_blind_rotate_kernel(request.param)
    yield request.param
    set_blind_rotate_kernel("auto")


def test_inverse_of_forward_is_identity(transform):
    from eoc_tfhe_b200 import fft_roundtrip
    rng = np.random.default_rng(11)
    a = rng.integers(-2**31, 2**31, size=(300, N), dtype=np.int64).astype(np.int32)
    a[0] = 0
    a[1] = -2**31
    a[2] = 2**31 - 1
    a[3, ::2] = 2**31 - 1
    a[3, 1::2] = -2**31
    a[4] = 0
    a[4, 0] = 1                      # the constant 1: every evaluation equals 1
    a[5] = 0
    a[5, N - 1] = -1                 # X^(N-1), the last coefficient
    assert np.array_equal(fft_roundtrip(a), a)
    assert np.array_equal(fft_roundtrip(a[:1]), a[:1])     # a single polynomial, fewer CTAs than SMs


def test_product_through_the_transform_is_exact(transform):
    from eoc_tfhe_b200 import poly_mul
    rng = np.random.default_rng(12)
    count = 96
    a = rng.integers(-512, 513, size=(count, N), dtype=np.int64).astype(np.int32)     # gadget digits, Bg = 1024
    b = rng.integers(-2**31, 2**31, size=(count, N), dtype=np.int64).astype(np.int32)
    a[0] = 512
    b[0] = -2**31                    # the largest products the blind rotation can form
    a[1] = 0
    a[1, 1] = 1                      # multiplication by X: a negacyclic rotation
    a[2] = -512
    b[2] = 2**31 - 1
    got = poly_mul(a, b)
    assert np.array_equal(got, negacyclic_mul_exact(a, b))
    assert np.array_equal(got[1, 1:], b[1, :-1]) and got[1, 0] == -b[1, N - 1]


@pytest.mark.skipif(not reflib.available(), reason="oracle/_ref not built (needs /root/reference)")
def test_product_equals_reference_karatsuba(transform):
    """torusPolynomialMultKaratsuba of the unmodified reference (polynomials.cpp / multiplication.cpp:146-160)."""
    from eoc_tfhe_b200 import poly_mul
    L = reflib.RefLib().L
    rng = np.random.default_rng(13)
    count = 8
    a = rng.integers(-64, 64, size=(count, N), dtype=np.int64).astype(np.int32)       # 7-bit digits (128-bit set)
    b = rng.integers(-2**31, 2**31, size=(count, N), dtype=np.int64).astype(np.int32)

    class Poly(C.Structure):
        _fields_ = [("N", C.c_int32), ("coefs", C.POINTER(C.c_int32))]
    L.torusPolynomialMultKaratsuba.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    want = np.empty_like(b)
    for i in range(count):
        ai, bi, ri = np.ascontiguousarray(a[i]), np.ascontiguousarray(b[i]), np.zeros(N, dtype=np.int32)
        pa = Poly(N, ai.ctypes.data_as(C.POINTER(C.c_int32)))
        pb = Poly(N, bi.ctypes.data_as(C.POINTER(C.c_int32)))
        pr = Poly(N, ri.ctypes.data_as(C.POINTER(C.c_int32)))
        L.torusPolynomialMultKaratsuba(C.byref(pr), C.byref(pa), C.byref(pb))
        want[i] = ri
    assert np.array_equal(poly_mul(a, b), want)
    assert np.array_equal(want, negacyclic_mul_exact(a, b))
