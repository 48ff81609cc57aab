"""GPU path vs the UNMODIFIED reference (oracle/_ref/libtfhe-spqlios-fma.so) on the same keys.

Parity protocol of SURVEY.md section 8c: (ii) key switch bit-exact, (iii) one external product /
one CMux step within 2 LSB per coefficient of the reference's FP64 FFT result (and of the exact
integer product), (i) decrypted gate outputs identical.
"""
import numpy as np
import pytest

import reflib

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not reflib.available(), reason="oracle/_ref not built (needs /root/reference)")]


# every parity check runs against both blind-rotation kernels (64 threads per ciphertext / warp per
# ciphertext with TMEM accumulators)
@pytest.fixture(scope="module", params=[(80, "v2"), (128, "v2"), (80, "tm"), (128, "tm"), (80, "w16"), (128, "w16")],
                ids=["P80-v2", "P128-v2", "P80-tm", "P128-tm", "P80-w16", "P128-w16"])
def keys(request):
    from eoc_tfhe_b200 import CloudKey, set_blind_rotate_kernel, experiments_built
    lam, kernel = request.param
    if kernel not in ("v2", "tm") and not experiments_built():
        pytest.skip("round-1 variant: built only with B200_EXPERIMENTS=1")
    set_blind_rotate_kernel(kernel)
    ref = reflib.RefLib()
    rk = ref.keygen(lam, seed=42)
    ck = CloudKey.from_libtfhe(rk.cloud_ptr)
    yield rk, ck
    ck.close()
    set_blind_rotate_kernel("auto")


def test_params_roundtrip(keys):
    rk, ck = keys
    p = ck.params
    assert (p.n, p.N, p.k, p.l, p.Bgbit, p.ks_t, p.ks_basebit) == (rk.n, rk.N, rk.k, rk.l, rk.Bgbit, rk.ks_t, rk.ks_basebit)


def test_keyswitch_bit_exact(keys):
    rk, ck = keys
    rng = np.random.default_rng(1)
    u = rng.integers(-2**31, 2**31, size=(24, rk.N + 1), dtype=np.int64).astype(np.int32)
    u[0] = 0                      # all digits hit the rounding offset only
    u[1, :rk.N] = -1              # carries through every digit
    u[2, :rk.N] = 0x7FFFFFFF
    want, var = rk.keyswitch(u)
    got, nz = ck.keyswitch(u, with_counts=True)
    assert np.array_equal(got, want)
    assert np.allclose(nz * rk.ks_variance(), var, rtol=1e-12, atol=0)


def test_keyswitch_streaming_kernel_bit_exact(keys):
    """The table-streaming kernel (TMA-fed ring, 64/40 samples per CTA) against the reference and against the
    gather kernel: edge rows, a ragged count that leaves a partly filled CTA, non-zero-digit counts."""
    from eoc_tfhe_b200 import set_keyswitch_kernel
    rk, ck = keys
    rng = np.random.default_rng(5)
    u = rng.integers(-2**31, 2**31, size=(3 * 64 + 37, rk.N + 1), dtype=np.int64).astype(np.int32)
    u[0] = 0
    u[1, :rk.N] = -1
    u[2, :rk.N] = 0x7FFFFFFF
    u[3, :rk.N] = 0x00007FFF      # one below the rounding offset: every digit 0
    try:
        set_keyswitch_kernel("stream")
        got, nz = ck.keyswitch(u, with_counts=True)
        set_keyswitch_kernel("gather")
        got1, nz1 = ck.keyswitch(u, with_counts=True)
    finally:
        set_keyswitch_kernel("auto")
    assert np.array_equal(got, got1) and np.array_equal(nz, nz1)
    want, var = rk.keyswitch(u[:40])
    assert np.array_equal(got[:40], want)
    assert np.allclose(nz[:40] * rk.ks_variance(), var, rtol=1e-12, atol=0)
    assert nz[3] == 0 and np.array_equal(got[3, :rk.n], np.zeros(rk.n, dtype=np.int32))
    # a batch large enough for the 64-sample CTAs (the call above ran the 32-sample shape), ragged again
    big = rng.integers(-2**31, 2**31, size=(148 * 64 + 21, rk.N + 1), dtype=np.int64).astype(np.int32)
    try:
        set_keyswitch_kernel("stream")
        g2, n2 = ck.keyswitch(big, with_counts=True)
        set_keyswitch_kernel("gather")
        g1, n1 = ck.keyswitch(big, with_counts=True)
    finally:
        set_keyswitch_kernel("auto")
    assert np.array_equal(g1, g2) and np.array_equal(n1, n2)


def test_keyswitch_large_batch_matches_small(keys):
    rk, ck = keys
    rng = np.random.default_rng(2)
    u = rng.integers(-2**31, 2**31, size=(700, rk.N + 1), dtype=np.int64).astype(np.int32)
    big = ck.keyswitch(u)                      # tiles only
    small = np.concatenate([ck.keyswitch(u[i:i + 7]) for i in range(0, 700, 7)])  # split-i + atomics
    assert np.array_equal(big, small)
    want, _ = rk.keyswitch(u[:8])
    assert np.array_equal(big[:8], want)


def test_extern_product_within_2lsb(keys):
    rk, ck = keys
    rng = np.random.default_rng(3)
    acc = rng.integers(-2**31, 2**31, size=(4, 2, rk.N), dtype=np.int64).astype(np.int32)
    for idx in (0, 1, rk.n - 1):
        got = ck.extern_product(acc, idx)
        for g in range(acc.shape[0]):
            fft = rk.extern_product_fft(acc[g], idx)
            exact = rk.extern_product_exact(acc[g], idx)
            d_fft = np.abs((got[g].astype(np.int64) - fft.astype(np.int64) + 2**31) % 2**32 - 2**31)
            d_exact = np.abs((got[g].astype(np.int64) - exact.astype(np.int64) + 2**31) % 2**32 - 2**31)
            assert d_fft.max() <= 2, d_fft.max()      # bound on |GPU - reference FP64 FFT|, in 2^-32 units
            assert d_exact.max() <= 2, d_exact.max()  # bound on |GPU - exact product|


def test_cmux_step_within_2lsb(keys):
    rk, ck = keys
    rng = np.random.default_rng(4)
    acc = rng.integers(-2**31, 2**31, size=(3, 2, rk.N), dtype=np.int64).astype(np.int32)
    for barai in (1, 517, 1024, 1531, 2047):
        got = ck.cmux_step(acc, 5, barai)
        for g in range(acc.shape[0]):
            want = rk.cmux_step(acc[g], 5, barai)
            d = np.abs((got[g].astype(np.int64) - want.astype(np.int64) + 2**31) % 2**32 - 2**31)
            assert d.max() <= 2, (barai, d.max())
    assert np.array_equal(ck.cmux_step(acc, 5, 0), acc)  # barai == 0 is skipped (fft.cpp:101)


def test_bootstrap_woks_phase(keys):
    rk, ck = keys
    rng = np.random.default_rng(5)
    bits = rng.integers(0, 2, size=16)
    x = rk.encrypt_bits(bits)
    mu = 1 << 29
    got = ck.bootstrap_woks(x, mu)
    want = rk.bootstrap_woks(x, mu)
    pg, pw = rk.phase_extracted(got), rk.phase_extracted(want)
    sign = np.where(bits > 0, 1, -1)
    assert np.all(np.abs(pg.astype(np.int64) - sign * mu) < 2**26)   # well inside the +-1/16 margin
    assert np.all((pg > 0) == (pw > 0))


GATE_TRUTH = {
    "NAND": lambda a, b: 1 - (a & b), "OR": lambda a, b: a | b, "AND": lambda a, b: a & b,
    "XOR": lambda a, b: a ^ b, "XNOR": lambda a, b: 1 - (a ^ b), "NOR": lambda a, b: 1 - (a | b),
    "ANDNY": lambda a, b: (1 - a) & b, "ANDYN": lambda a, b: a & (1 - b),
    "ORNY": lambda a, b: (1 - a) | b, "ORYN": lambda a, b: a | (1 - b),
}


def test_all_gates_decrypt_like_reference(keys):
    from eoc_tfhe_b200 import GateOp
    rk, ck = keys
    rng = np.random.default_rng(6)
    count = 8
    a_bits, b_bits, c_bits = (rng.integers(0, 2, size=count) for _ in range(3))
    a, b, c = rk.encrypt_bits(a_bits), rk.encrypt_bits(b_bits), rk.encrypt_bits(c_bits)
    for name, fn in GATE_TRUTH.items():
        got = ck.gate(GateOp[name], a, b)
        want, _ = rk.gate(name, a[:2], b[:2])
        assert np.array_equal(rk.decrypt_bits(got), fn(a_bits, b_bits)), name
        assert np.array_equal(rk.decrypt_bits(want), fn(a_bits, b_bits)[:2]), name
    got = ck.gate(GateOp.MUX, a, b, c)
    assert np.array_equal(rk.decrypt_bits(got), np.where(a_bits > 0, b_bits, c_bits))
    want, _ = rk.gate("MUX", a[:2], b[:2], c[:2])
    assert np.array_equal(rk.decrypt_bits(want), np.where(a_bits > 0, b_bits, c_bits)[:2])
    # linear gates are bit-exact
    assert np.array_equal(ck.gate(GateOp.NOT, a), -a)
    assert np.array_equal(ck.gate(GateOp.COPY, a), a)
    one = ck.gate(GateOp.CONST1, out=np.empty_like(a))
    assert np.all(one[:, :-1] == 0) and np.all(one[:, -1] == 1 << 29)


@pytest.mark.skipif(not reflib.available("nayuki-portable"), reason="oracle/_ref/libtfhe-nayuki-portable.so not built")
def test_against_nayuki_portable_processor():
    """The FFT processor eoc actually ships (config.yml:21-25, fft_processors/nayuki/fft_processor_nayuki.cpp:34-86) as
    the second oracle: one external product within 2 LSB of its FP64 result and of the exact product, one CMux step
    within 2 LSB, gates decrypt to the same bits -- on a key generated by that build."""
    from eoc_tfhe_b200 import CloudKey, GateOp
    ref = reflib.RefLib("nayuki-portable")
    rk = ref.keygen(80, seed=77)
    ck = CloudKey.from_libtfhe(rk.cloud_ptr)
    try:
        rng = np.random.default_rng(9)
        acc = rng.integers(-2**31, 2**31, size=(3, 2, rk.N), dtype=np.int64).astype(np.int32)
        got = ck.extern_product(acc, 4)
        for g in range(acc.shape[0]):
            for want in (rk.extern_product_fft(acc[g], 4), rk.extern_product_exact(acc[g], 4)):
                d = np.abs((got[g].astype(np.int64) - want.astype(np.int64) + 2**31) % 2**32 - 2**31)
                assert d.max() <= 2, d.max()
        got = ck.cmux_step(acc, 7, 1531)
        for g in range(acc.shape[0]):
            want = rk.cmux_step(acc[g], 7, 1531)
            d = np.abs((got[g].astype(np.int64) - want.astype(np.int64) + 2**31) % 2**32 - 2**31)
            assert d.max() <= 2, d.max()
        a_bits, b_bits = rng.integers(0, 2, size=16), rng.integers(0, 2, size=16)
        a, b = rk.encrypt_bits(a_bits), rk.encrypt_bits(b_bits)
        out = ck.gate(GateOp.NAND, a, b)
        ref_out, _ = rk.gate("NAND", a, b)
        assert np.array_equal(rk.decrypt_bits(out), 1 - (a_bits & b_bits))
        assert np.array_equal(rk.decrypt_bits(ref_out), rk.decrypt_bits(out))
        assert np.array_equal(ck.keyswitch(rk.bootstrap_woks(a)), rk.keyswitch(rk.bootstrap_woks(a))[0])   # bit-exact
    finally:
        ck.close()
