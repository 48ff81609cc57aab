"""Generate tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libtfhe-spqlios-fma.so,
compiled by oracle/Makefile from /root/reference/libs/tfhe/src).  Run in the build container:

    python tests/golden/make_golden.py

The reference holds no golden vectors for this path, so these fixtures are
outputs of the reference itself on seeded inputs.  The default parameter sets have keys of
~100 MB; to keep the fixtures small they use the reference's own code with a SMALL LWE dimension
(n = 8; n = 4 trips an out-of-bounds AVX tail in the reference) and otherwise the 80-bit / 128-bit shapes (N=1024, k=1, (l,Bgbit) = (2,10) / (3,7), t=8,
basebit=2), built like default_80bit_gate_bootstrapping_parameters (tfhe_gate_bootstrapping.cpp:23-48).
Every array is what libtfhe computed:
    bk, ks                      cloud key in wire-format order (tfhe_io.cpp:757-930)
    lwe_key, tlwe_key           secret keys (to decrypt)
    bits_a/b/c, ct_a/b/c        bootsSymEncrypt inputs
    gate_<NAME>                 boots<NAME>(ct_a, ct_b[, ct_c]) outputs (spqlios-fma)
    woks_x, woks_u              tfhe_bootstrap_woKS_FFT input / output
    ks_u, ks_out, ks_var        lweKeySwitch input / output / current_variance
    ep_acc, ep_fft, ep_exact    tGswFFTExternMulToTLwe and tGswExternMulToTLwe on bk[ep_index]
    cmux_acc, cmux_barai, cmux_out   tfhe_MuxRotate_FFT on bk[cmux_index]
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import reflib  # noqa: E402

SETS = {"small80": dict(n=8, l=2, Bgbit=10, ks_stdev=2.44e-5, bk_stdev=7.18e-9),
        "small128": dict(n=8, l=3, Bgbit=7, ks_stdev=2.0 ** -15, bk_stdev=2.0 ** -25)}


def make(name, cfg):
    ref = reflib.RefLib("spqlios-fma")
    rk = ref.keygen_custom(seed=20240 + len(name), **cfg)
    rng = np.random.default_rng(77)
    out = dict(params=np.array([rk.n, rk.N, rk.k, rk.l, rk.Bgbit, rk.ks_t, rk.ks_basebit], dtype=np.int32),
               ks_variance=np.float64(rk.ks_variance()), bk=rk.bk_coefs(), ks=rk.ks_rows(),
               lwe_key=rk.lwe_key(), tlwe_key=rk.tlwe_key())
    count = 6
    for tag in "abc":
        bits = rng.integers(0, 2, size=count).astype(np.int32)
        out["bits_" + tag] = bits
        out["ct_" + tag] = rk.encrypt_bits(bits)
    for g in reflib.GATES:
        out["gate_" + g], _ = rk.gate(g, out["ct_a"], out["ct_b"])
    out["gate_MUX"], _ = rk.gate("MUX", out["ct_a"], out["ct_b"], out["ct_c"])
    out["woks_x"] = out["ct_a"][:3]
    out["woks_u"] = rk.bootstrap_woks(out["woks_x"], 1 << 29)
    u = rng.integers(-2 ** 31, 2 ** 31, size=(5, rk.N + 1), dtype=np.int64).astype(np.int32)
    u[0] = 0
    u[1, :rk.N] = -1
    out["ks_u"] = u
    out["ks_out"], out["ks_var"] = rk.keyswitch(u)
    acc = rng.integers(-2 ** 31, 2 ** 31, size=(3, 2, rk.N), dtype=np.int64).astype(np.int32)
    out["ep_acc"], out["ep_index"] = acc, np.int32(1)
    out["ep_fft"] = np.stack([rk.extern_product_fft(a, 1) for a in acc])
    out["ep_exact"] = np.stack([rk.extern_product_exact(a, 1) for a in acc])
    barai = np.array([1, 700, 1024, 2047], dtype=np.int32)
    out["cmux_acc"], out["cmux_index"], out["cmux_barai"] = acc[0], np.int32(2), barai
    out["cmux_out"] = np.stack([rk.cmux_step(acc[0], 2, int(a)) for a in barai])
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    for name, cfg in SETS.items():
        make(name, cfg)
