"""eoc's evaluation-side entry points (csrc/eoc_abi.cpp) against the UNMODIFIED reference.

The cases follow tests/tfhe.test.js of the reference (add 15 + 27, subtract; :142-186) with libtfhe called
directly where the JS test goes through WASM + Lua: integers are encrypted as eoc-tfhe-run.cpp:282-300 does
(modSwitchToTorus32(value, Msize), alpha = 1/(10 Msize)), serialized with libtfhe's own exporter, base64'd,
and the result strings are compared BYTE FOR BYTE with what lweCopy + lweAddTo / lweSubTo + export give.
"""
import base64
import ctypes as C
import tempfile

import numpy as np
import pytest

import reflib
from eoc_tfhe_b200 import eoc, lib

needs_ref = pytest.mark.skipif(not reflib.available(), reason="oracle/_ref not built (needs /root/reference)")

MSIZE = (1 << 31) - 1            # eoc-tfhe-run.cpp:35
ALPHA = 1.0 / (10.0 * MSIZE)     # eoc-tfhe-run.cpp:36


class EocRef:
    """The reference side of the eoc calls: libtfhe functions eoc-tfhe-run.cpp uses, on native samples."""

    def __init__(self, rk):
        self.rk, self.L = rk, rk.L
        L, VP = self.L, C.c_void_p
        L.modSwitchToTorus32.restype = C.c_int32
        L.modSwitchToTorus32.argtypes = [C.c_int32, C.c_int32]
        L.modSwitchFromTorus32.restype = C.c_int32
        L.modSwitchFromTorus32.argtypes = [C.c_int32, C.c_int32]
        L.lweSymEncrypt.argtypes = [VP, C.c_int32, C.c_double, VP]
        L.lweCopy.argtypes = L.lweAddTo.argtypes = L.lweSubTo.argtypes = [VP, VP, VP]
        L.export_gate_bootstrapping_ciphertext_toFile.argtypes = [VP, VP, VP]
        L.import_gate_bootstrapping_ciphertext_fromFile.argtypes = [VP, VP, VP]

    def encrypt(self, value):
        s = self.rk._new(1)
        self.L.lweSymEncrypt(C.byref(s[0]), self.L.modSwitchToTorus32(value, MSIZE), ALPHA, self.rk.sk.contents.lwe_key)
        return s

    def combine(self, s1, s2, sub):
        out = self.rk._new(1)
        self.L.lweCopy(C.byref(out[0]), C.byref(s1[0]), self.rk.in_out_params)
        (self.L.lweSubTo if sub else self.L.lweAddTo)(C.byref(out[0]), C.byref(s2[0]), self.rk.in_out_params)
        return out

    def to_b64(self, s):
        with tempfile.NamedTemporaryFile(suffix=".ct") as tf:
            f = self.rk.ref.libc.fopen(tf.name.encode(), b"wb")
            self.L.export_gate_bootstrapping_ciphertext_toFile(f, C.byref(s[0]), self.rk.sk.contents.params)
            self.rk.ref.libc.fclose(f)
            return base64.b64encode(open(tf.name, "rb").read()).decode()

    def decrypt_b64(self, text):
        """decryptInteger (eoc-tfhe-run.cpp:393-424): phase -> modSwitchFromTorus32(phase, Msize)."""
        s = self.rk._new(1)
        with tempfile.NamedTemporaryFile(suffix=".ct") as tf:
            tf.write(base64.b64decode(text))
            tf.flush()
            f = self.rk.ref.libc.fopen(tf.name.encode(), b"rb")
            self.L.import_gate_bootstrapping_ciphertext_fromFile(f, C.byref(s[0]), self.rk.sk.contents.params)
            self.rk.ref.libc.fclose(f)
        return self.L.modSwitchFromTorus32(self.L.lwePhase(C.byref(s[0]), self.rk.sk.contents.lwe_key), MSIZE)


@pytest.fixture(scope="module")
def eoc_ref():
    ref = reflib.RefLib()
    return EocRef(ref.keygen_custom(n=630, l=3, Bgbit=7, seed=11))   # the shape of eoc's lambda = 128 samples


@needs_ref
@pytest.mark.parametrize("x,y", [(27, 15), (0, 0), (654321, 123456), (-5, 12), (MSIZE // 2, MSIZE // 2 + 7)])
def test_add_and_subtract_ciphertexts_byte_exact(eoc_ref, x, y):
    r = eoc_ref
    s1, s2 = r.encrypt(x), r.encrypt(y)
    b1, b2 = r.to_b64(s1), r.to_b64(s2)
    got_add, got_sub = eoc.add_ciphertexts(b1, b2, "ignored"), eoc.subtract_ciphertexts(b1, b2)
    assert got_add == r.to_b64(r.combine(s1, s2, sub=False))      # a[], b AND the variance field
    assert got_sub == r.to_b64(r.combine(s1, s2, sub=True))
    # and they decrypt as the JS test expects; Msize = 2^31 - 1 does not divide 2^32, so the reference's own
    # encoding is off by one unit for negative values and wrapped sums (modSwitchToTorus32,
    # numeric-functions.cpp:71-76) -- exact for the JS test's small non-negative cases
    def off(got, want):
        return abs((got - want + MSIZE // 2) % MSIZE - MSIZE // 2)
    tol = 0 if (x >= 0 and y >= 0 and x - y >= 0 and x + y < MSIZE // 2) else 2
    assert off(r.decrypt_b64(got_add), x + y) <= tol
    assert off(r.decrypt_b64(got_sub), x - y) <= tol


@needs_ref
def test_add_ciphertexts_accepts_what_the_reference_decoder_accepts(eoc_ref):
    """base64_decode (eoc-tfhe-run.cpp:70-92) stops at the first character outside the alphabet."""
    r = eoc_ref
    s1, s2 = r.encrypt(40), r.encrypt(2)
    b1, b2 = r.to_b64(s1), r.to_b64(s2)
    assert eoc.add_ciphertexts(b1 + "\n trailing", b2) == r.to_b64(r.combine(s1, s2, sub=False))


def test_malformed_input_returns_null_with_message():
    L = lib.load()
    good = base64.b64encode(np.array([42] + [0] * 9 + [0, 0], dtype=np.int32).tobytes()).decode()   # n = 8 record
    assert eoc.add_ciphertexts(good, good)          # a zero sample of dimension 8 adds fine
    for a, b in [("", good), (good, good[:-8]), ("AAAA", "AAAA")]:
        assert not L.addCiphertexts(a.encode(), b.encode(), b"")
        assert b"addCiphertexts" in L.b200_last_error()
    bad_uid = base64.b64encode(np.array([41] + [0] * 11, dtype=np.int32).tobytes()).decode()
    assert not L.subtractCiphertexts(bad_uid.encode(), bad_uid.encode(), b"")
    assert b"uid" in L.b200_last_error()
    # bit-sliced calls need an evaluation key first (eoc-tfhe-run.cpp:464-468)
    eoc.set_key(None)
    assert not L.addCiphertextBits(good.encode(), good.encode(), 1)
    assert b"Public key not initialized" in L.b200_last_error()


@pytest.mark.gpu
@needs_ref
def test_bit_sliced_add_and_subtract_on_gpu():
    """Integers encrypted bit by bit, added / subtracted by the bootstrapped ripple-carry adder on the GPU:
    48 pairs through b200_add_bits_host, one pair through the base64 entry points with the key loaded
    from libtfhe's serialized cloud key."""
    from eoc_tfhe_b200 import CloudKey, circuits
    ref = reflib.RefLib()
    rk = ref.keygen(80, seed=4321)
    ck = CloudKey.from_libtfhe(rk.cloud_ptr)
    nb, count = 16, 48
    rng = np.random.default_rng(5)
    x, y = rng.integers(0, 1 << nb, size=count), rng.integers(0, 1 << nb, size=count)
    x[0], y[0] = 0xFFFF, 0xFFFF
    x[1], y[1] = 0, 1
    x[2], y[2] = 1234, 1234
    cx = rk.encrypt_bits(circuits.int_to_bits(x, nb).ravel()).reshape(count, nb, -1)
    cy = rk.encrypt_bits(circuits.int_to_bits(y, nb).ravel()).reshape(count, nb, -1)

    def value(out):
        return circuits.bits_to_int(rk.decrypt_bits(out.reshape(-1, out.shape[-1])).reshape(out.shape[0], nb + 1))

    assert np.array_equal(value(eoc.add_bits(ck, cx, cy)), x + y)
    assert np.array_equal(value(eoc.add_bits(ck, cx, cy, subtract=True)), (x - y) % (1 << nb) + ((x >= y).astype(np.int64) << nb))

    with tempfile.NamedTemporaryFile(suffix=".key") as tf:
        rk.export_cloud(tf.name)
        eoc.set_cloud_key(base64.b64encode(open(tf.name, "rb").read()).decode())
    bx = base64.b64encode(rk.export_samples(cx[3])).decode()
    by = base64.b64encode(rk.export_samples(cy[3])).decode()
    s = rk.import_samples(base64.b64decode(eoc.add_ciphertext_bits(bx, by, nb)), nb + 1)
    assert circuits.bits_to_int(rk.decrypt_bits(s)[None, :])[0] == x[3] + y[3]
    d = rk.import_samples(base64.b64decode(eoc.add_ciphertext_bits(bx, by, nb, subtract=True)), nb + 1)
    assert circuits.bits_to_int(rk.decrypt_bits(d)[None, :])[0] == (x[3] - y[3]) % (1 << nb) + (int(x[3] >= y[3]) << nb)
    # identical ciphertexts from both routes (same kernels, same key bits)
    assert np.array_equal(s, eoc.add_bits(ck, cx[3:4], cy[3:4])[0])
    eoc.set_key(None)
    ck.close()
