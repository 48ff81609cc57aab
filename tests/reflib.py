"""ctypes binding of the UNMODIFIED reference (oracle/_ref/libtfhe-*.so, built by oracle/Makefile).

Test infrastructure only.  Gives the tests the reference's own keygen / encrypt / decrypt and its
CPU implementation of the path (bootsNAND..., tfhe_bootstrap_woKS_FFT, lweKeySwitch,
tGswFFTExternMulToTLwe, tfhe_MuxRotate_FFT) on numpy arrays.
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")


def ref_path(variant="spqlios-fma"):
    return os.path.join(REF_DIR, "libtfhe-%s.so" % variant)


def available(variant="spqlios-fma"):
    return os.path.exists(ref_path(variant))


class LweParams(C.Structure):
    _fields_ = [("n", C.c_int32), ("alpha_min", C.c_double), ("alpha_max", C.c_double)]


class LweSample(C.Structure):
    _fields_ = [("a", C.POINTER(C.c_int32)), ("b", C.c_int32), ("current_variance", C.c_double)]


class TLweParams(C.Structure):
    _fields_ = [("N", C.c_int32), ("k", C.c_int32), ("alpha_min", C.c_double), ("alpha_max", C.c_double),
                ("extracted_lweparams", LweParams)]


class TGswParams(C.Structure):
    _fields_ = [("l", C.c_int32), ("Bgbit", C.c_int32), ("Bg", C.c_int32), ("halfBg", C.c_int32),
                ("maskMod", C.c_uint32), ("tlwe_params", C.POINTER(TLweParams)), ("kpl", C.c_int32),
                ("h", C.POINTER(C.c_int32)), ("offset", C.c_uint32)]


class TorusPolynomial(C.Structure):
    _fields_ = [("N", C.c_int32), ("coefsT", C.POINTER(C.c_int32))]


class TLweSample(C.Structure):
    _fields_ = [("a", C.POINTER(TorusPolynomial)), ("b", C.POINTER(TorusPolynomial)),
                ("current_variance", C.c_double), ("k", C.c_int32)]


class TGswSample(C.Structure):
    _fields_ = [("all_sample", C.POINTER(TLweSample)), ("bloc_sample", C.c_void_p), ("k", C.c_int32), ("l", C.c_int32)]


class LweKeySwitchKey(C.Structure):
    _fields_ = [("n", C.c_int32), ("t", C.c_int32), ("basebit", C.c_int32), ("base", C.c_int32),
                ("out_params", C.POINTER(LweParams)), ("ks0_raw", C.POINTER(LweSample)),
                ("ks1_raw", C.c_void_p), ("ks", C.POINTER(C.POINTER(C.POINTER(LweSample))))]


class LweBootstrappingKey(C.Structure):
    _fields_ = [("in_out_params", C.POINTER(LweParams)), ("bk_params", C.POINTER(TGswParams)),
                ("accum_params", C.POINTER(TLweParams)), ("extract_params", C.POINTER(LweParams)),
                ("bk", C.POINTER(TGswSample)), ("ks", C.POINTER(LweKeySwitchKey))]


class LweBootstrappingKeyFFT(C.Structure):
    _fields_ = [("in_out_params", C.POINTER(LweParams)), ("bk_params", C.POINTER(TGswParams)),
                ("accum_params", C.POINTER(TLweParams)), ("extract_params", C.POINTER(LweParams)),
                ("bkFFT", C.c_void_p), ("ks", C.POINTER(LweKeySwitchKey))]


class ParameterSet(C.Structure):
    _fields_ = [("ks_t", C.c_int32), ("ks_basebit", C.c_int32), ("in_out_params", C.POINTER(LweParams)),
                ("tgsw_params", C.POINTER(TGswParams))]


class CloudKeySet(C.Structure):
    _fields_ = [("params", C.POINTER(ParameterSet)), ("bk", C.POINTER(LweBootstrappingKey)),
                ("bkFFT", C.POINTER(LweBootstrappingKeyFFT))]


class LweKey(C.Structure):
    _fields_ = [("params", C.POINTER(LweParams)), ("key", C.POINTER(C.c_int32))]


class IntPolynomial(C.Structure):
    _fields_ = [("N", C.c_int32), ("coefs", C.POINTER(C.c_int32))]


class TGswKey(C.Structure):
    _fields_ = [("params", C.POINTER(TGswParams)), ("tlwe_params", C.POINTER(TLweParams)),
                ("key", C.POINTER(IntPolynomial))]


class SecretKeySet(C.Structure):
    _fields_ = [("params", C.POINTER(ParameterSet)), ("lwe_key", C.POINTER(LweKey)),
                ("tgsw_key", C.POINTER(TGswKey)), ("cloud", CloudKeySet)]


GATES = ["NAND", "OR", "AND", "XOR", "XNOR", "NOR", "ANDNY", "ANDYN", "ORNY", "ORYN"]


class RefLib:
    def __init__(self, variant="spqlios-fma"):
        self.L = L = C.CDLL(ref_path(variant), mode=C.RTLD_LOCAL)
        VP = C.c_void_p
        L.new_default_gate_bootstrapping_parameters.restype = C.POINTER(ParameterSet)
        L.new_default_gate_bootstrapping_parameters.argtypes = [C.c_int32]
        L.new_random_gate_bootstrapping_secret_keyset.restype = C.POINTER(SecretKeySet)
        L.new_random_gate_bootstrapping_secret_keyset.argtypes = [VP]
        L.new_LweSample_array.restype = C.POINTER(LweSample)
        L.new_LweSample_array.argtypes = [C.c_int32, VP]
        L.delete_LweSample_array.argtypes = [C.c_int32, VP]
        L.bootsSymEncrypt.argtypes = [VP, C.c_int32, VP]
        L.bootsSymDecrypt.argtypes = [VP, VP]
        L.bootsSymDecrypt.restype = C.c_int32
        L.lwePhase.argtypes = [VP, VP]
        L.lwePhase.restype = C.c_int32
        for g in GATES:
            getattr(L, "boots" + g).argtypes = [VP] * 4
            getattr(L, "boots" + g).restype = None
        L.bootsMUX.argtypes = [VP] * 5
        L.bootsMUX.restype = None
        L.tfhe_bootstrap_woKS_FFT.argtypes = [VP, VP, C.c_int32, VP]
        L.tfhe_bootstrap_FFT.argtypes = [VP, VP, C.c_int32, VP]
        L.lweKeySwitch.argtypes = [VP, VP, VP]
        L.new_TLweSample.restype = C.POINTER(TLweSample)
        L.new_TLweSample.argtypes = [VP]
        L.delete_TLweSample.argtypes = [VP]
        L.tGswFFTExternMulToTLwe.argtypes = [VP, VP, VP]
        L.tGswExternMulToTLwe.argtypes = [VP, VP, VP]
        self._muxrot = getattr(L, "_Z18tfhe_MuxRotate_FFTP10TLweSamplePKS_PK13TGswSampleFFTiPK10TGswParams")
        self._muxrot.argtypes = [VP, VP, VP, C.c_int32, VP]
        L.tfhe_random_generator_setSeed.argtypes = [C.POINTER(C.c_uint32), C.c_int32]
        L.export_tfheGateBootstrappingCloudKeySet_toFile.argtypes = [VP, VP]
        L.export_tfheGateBootstrappingSecretKeySet_toFile.argtypes = [VP, VP]
        L.new_tfheGateBootstrappingCloudKeySet_fromFile.restype = C.POINTER(CloudKeySet)
        L.new_tfheGateBootstrappingCloudKeySet_fromFile.argtypes = [VP]
        L.new_tfheGateBootstrappingSecretKeySet_fromFile.restype = C.POINTER(SecretKeySet)
        L.new_tfheGateBootstrappingSecretKeySet_fromFile.argtypes = [VP]
        self.libc = C.CDLL(None)
        self.libc.fopen.restype = VP
        self.libc.fopen.argtypes = [C.c_char_p, C.c_char_p]
        self.libc.fclose.argtypes = [VP]

    def seed(self, *values):
        arr = (C.c_uint32 * len(values))(*values)
        self.L.tfhe_random_generator_setSeed(arr, len(values))

    def keygen(self, minimum_lambda=80, seed=42):
        self.seed(seed)
        params = self.L.new_default_gate_bootstrapping_parameters(minimum_lambda)
        sk = self.L.new_random_gate_bootstrapping_secret_keyset(params)
        return RefKeys(self, sk)

    def keygen_custom(self, n, l, Bgbit, ks_t=8, ks_basebit=2, ks_stdev=2.44e-5, bk_stdev=7.18e-9,
                      max_stdev=0.012467, seed=42):
        """Keys for a non-default parameter set, built the way default_80bit_gate_bootstrapping_parameters
        does (tfhe_gate_bootstrapping.cpp:23-48) -- used for SMALL golden fixtures (small n)."""
        L = self.L
        L.new_LweParams.restype = C.POINTER(LweParams)
        L.new_LweParams.argtypes = [C.c_int32, C.c_double, C.c_double]
        L.new_TLweParams.restype = C.POINTER(TLweParams)
        L.new_TLweParams.argtypes = [C.c_int32, C.c_int32, C.c_double, C.c_double]
        L.new_TGswParams.restype = C.POINTER(TGswParams)
        L.new_TGswParams.argtypes = [C.c_int32, C.c_int32, C.c_void_p]
        self.seed(seed)
        p_in = L.new_LweParams(n, ks_stdev, max_stdev)
        p_acc = L.new_TLweParams(1024, 1, bk_stdev, max_stdev)
        p_bk = L.new_TGswParams(l, Bgbit, p_acc)
        ps = ParameterSet(ks_t, ks_basebit, p_in, p_bk)
        self._keep = getattr(self, "_keep", []) + [ps]
        sk = L.new_random_gate_bootstrapping_secret_keyset(C.byref(ps))
        return RefKeys(self, sk)

    def load_secret_keyset(self, path):
        f = self.libc.fopen(path.encode(), b"rb")
        sk = self.L.new_tfheGateBootstrappingSecretKeySet_fromFile(f)
        self.libc.fclose(f)
        return RefKeys(self, sk)


class RefKeys:
    """A reference secret keyset + views of its cloud key as numpy arrays."""

    def __init__(self, ref, sk):
        self.ref, self.L, self.sk = ref, ref.L, sk
        s = sk.contents
        self.cloud = s.cloud
        self.cloud_ptr = C.addressof(s.cloud)
        ps = s.params.contents
        gp = ps.tgsw_params.contents
        tp = gp.tlwe_params.contents
        self.n = ps.in_out_params.contents.n
        self.N, self.k, self.l, self.Bgbit = tp.N, tp.k, gp.l, gp.Bgbit
        self.kpl = gp.kpl
        self.ks_t, self.ks_basebit = ps.ks_t, ps.ks_basebit
        self.in_out_params = ps.in_out_params
        self.extract_params = C.pointer(tp.extracted_lweparams)
        self.tgsw_params = ps.tgsw_params
        self.tlwe_params = gp.tlwe_params

    # ---- key material as arrays ----------------------------------------------------------
    def lwe_key(self):
        return np.ctypeslib.as_array(self.sk.contents.lwe_key.contents.key, shape=(self.n,)).copy()

    def tlwe_key(self):
        return np.ctypeslib.as_array(self.sk.contents.tgsw_key.contents.key[0].coefs, shape=(self.N,)).copy()

    def bk_coefs(self):
        bk = self.cloud.bk.contents
        out = np.empty((self.n, self.kpl, self.k + 1, self.N), dtype=np.int32)
        for i in range(self.n):
            rows = bk.bk[i].all_sample
            for r in range(self.kpl):
                for c in range(self.k + 1):
                    out[i, r, c] = np.ctypeslib.as_array(rows[r].a[c].coefsT, shape=(self.N,))
        return out

    def ks_rows(self):
        ks = self.cloud.bkFFT.contents.ks.contents
        base = 1 << self.ks_basebit
        Nin = ks.n
        # ks0_raw is one contiguous array of N*t*base samples in [i][j][h] order (lwekeyswitch.cpp)
        out = np.empty((Nin, self.ks_t, base, self.n + 1), dtype=np.int32)
        flat = out.reshape(-1, self.n + 1)
        raw = ks.ks0_raw
        for idx in range(Nin * self.ks_t * base):
            s = raw[idx]
            flat[idx, :self.n] = np.ctypeslib.as_array(s.a, shape=(self.n,))
            flat[idx, self.n] = s.b
        return out

    def ks_variance(self):
        return self.cloud.bkFFT.contents.ks.contents.ks[0][0][1].current_variance

    # ---- samples ---------------------------------------------------------------------------
    def _new(self, count, extracted=False):
        params = self.extract_params if extracted else self.in_out_params
        return self.L.new_LweSample_array(count, params)

    def _fill(self, samples, arr):
        dim = arr.shape[1] - 1
        for i in range(arr.shape[0]):
            C.memmove(samples[i].a, arr[i].ctypes.data, dim * 4)
            samples[i].b = int(arr[i, dim])
            samples[i].current_variance = 0.0

    def _read(self, samples, count, dim):
        out = np.empty((count, dim + 1), dtype=np.int32)
        for i in range(count):
            out[i, :dim] = np.ctypeslib.as_array(samples[i].a, shape=(dim,))
            out[i, dim] = samples[i].b
        return out

    def encrypt_bits(self, bits):
        bits = np.asarray(bits).astype(np.int32).ravel()
        s = self._new(len(bits))
        for i, b in enumerate(bits):
            self.L.bootsSymEncrypt(C.byref(s[i]), int(b), self.sk)
        out = self._read(s, len(bits), self.n)
        self.L.delete_LweSample_array(len(bits), s)
        return out

    def phase(self, cts):
        """b - <a, s> for in/out-dimension samples (lwe-functions.cpp:72-81), exact int32."""
        key = self.lwe_key().astype(np.int64)
        cts = np.asarray(cts)
        ph = cts[:, self.n].astype(np.int64) - cts[:, :self.n].astype(np.int64) @ key
        return ph.astype(np.uint64).astype(np.uint32).astype(np.int32) if False else (ph & 0xFFFFFFFF).astype(np.uint32).astype(np.int32)

    def phase_extracted(self, us):
        key = self.tlwe_key().astype(np.int64)
        us = np.asarray(us)
        ph = us[:, self.N].astype(np.int64) - us[:, :self.N].astype(np.int64) @ key
        return (ph & 0xFFFFFFFF).astype(np.uint32).astype(np.int32)

    def decrypt_bits(self, cts):
        return (self.phase(cts) > 0).astype(np.int32)

    # ---- the reference path on arrays ------------------------------------------------------
    def gate(self, name, a, b, c=None):
        count = a.shape[0]
        sa, sb, so = self._new(count), self._new(count), self._new(count)
        self._fill(sa, a)
        self._fill(sb, b)
        sc = None
        if c is not None:
            sc = self._new(count)
            self._fill(sc, c)
        for i in range(count):
            if name == "MUX":
                self.L.bootsMUX(C.byref(so[i]), C.byref(sa[i]), C.byref(sb[i]), C.byref(sc[i]), self.cloud_ptr)
            else:
                getattr(self.L, "boots" + name)(C.byref(so[i]), C.byref(sa[i]), C.byref(sb[i]), self.cloud_ptr)
        out = self._read(so, count, self.n)
        var = np.array([so[i].current_variance for i in range(count)])
        for s in (sa, sb, so) + ((sc,) if sc is not None else ()):
            self.L.delete_LweSample_array(count, s)
        return out, var

    def gate_mt(self, name, a, b, c=None, threads=None):
        """gate() with one gate stream per host thread (ctypes releases the GIL; the reference's gate
        evaluation is re-entrant: per-call scratch, thread_local FFT state)."""
        import os
        from concurrent.futures import ThreadPoolExecutor
        threads = threads or len(os.sched_getaffinity(0))
        count = a.shape[0]
        cuts = np.linspace(0, count, threads + 1).astype(int)
        parts = [(cuts[i], cuts[i + 1]) for i in range(threads) if cuts[i + 1] > cuts[i]]
        with ThreadPoolExecutor(max_workers=len(parts)) as ex:
            res = list(ex.map(lambda se: self.gate(name, a[se[0]:se[1]], b[se[0]:se[1]],
                                                   None if c is None else c[se[0]:se[1]]), parts))
        return np.concatenate([r[0] for r in res]), np.concatenate([r[1] for r in res])

    def bootstrap_woks(self, x, mu=1 << 29):
        count = x.shape[0]
        sx, su = self._new(count), self._new(count, extracted=True)
        self._fill(sx, x)
        for i in range(count):
            self.L.tfhe_bootstrap_woKS_FFT(C.byref(su[i]), self.cloud.bkFFT, mu, C.byref(sx[i]))
        out = self._read(su, count, self.N)
        self.L.delete_LweSample_array(count, sx)
        self.L.delete_LweSample_array(count, su)
        return out

    def keyswitch(self, u):
        count = u.shape[0]
        su, so = self._new(count, extracted=True), self._new(count)
        self._fill(su, u)
        for i in range(count):
            self.L.lweKeySwitch(C.byref(so[i]), self.cloud.bkFFT.contents.ks, C.byref(su[i]))
        out = self._read(so, count, self.n)
        var = np.array([so[i].current_variance for i in range(count)])
        self.L.delete_LweSample_array(count, su)
        self.L.delete_LweSample_array(count, so)
        return out, var

    def _tlwe_from(self, acc):
        t = self.L.new_TLweSample(self.tlwe_params)
        for c in range(self.k + 1):
            C.memmove(t.contents.a[c].coefsT, acc[c].ctypes.data, self.N * 4)
        return t

    def _tlwe_read(self, t):
        out = np.empty((self.k + 1, self.N), dtype=np.int32)
        for c in range(self.k + 1):
            out[c] = np.ctypeslib.as_array(t.contents.a[c].coefsT, shape=(self.N,))
        return out

    def extern_product_fft(self, acc, bk_index):
        """tGswFFTExternMulToTLwe (tgsw-fft-operations.cpp:89-114) on acc[k+1][N]."""
        t = self._tlwe_from(np.ascontiguousarray(acc, dtype=np.int32))
        bkfft = self.cloud.bkFFT.contents.bkFFT  # TGswSampleFFT array; sizeof = 24
        self.L.tGswFFTExternMulToTLwe(t, C.c_void_p(bkfft + 24 * bk_index), self.tgsw_params)
        out = self._tlwe_read(t)
        self.L.delete_TLweSample(t)
        return out

    def extern_product_exact(self, acc, bk_index):
        """External product in EXACT integer arithmetic from the reference's own pieces:
        tGswTLweDecompH (tgsw-functions.cpp:353-360) + torusPolynomialAddMulRKaratsuba
        (multiplication.cpp:146-160), summed like tGswExternMulToTLwe (tgsw-functions.cpp:156-170).
        (tGswExternMulToTLwe itself multiplies through the FFT: polynomials_arithmetic.h:113.)"""
        L = self.L
        L.new_IntPolynomial_array.restype = C.POINTER(IntPolynomial)
        L.new_IntPolynomial_array.argtypes = [C.c_int32, C.c_int32]
        L.delete_IntPolynomial_array.argtypes = [C.c_int32, C.c_void_p]
        L.tGswTLweDecompH.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.tLweClear.argtypes = [C.c_void_p, C.c_void_p]
        L.torusPolynomialAddMulRKaratsuba.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        t = self._tlwe_from(np.ascontiguousarray(acc, dtype=np.int32))
        dec = L.new_IntPolynomial_array(self.kpl, self.N)
        L.tGswTLweDecompH(dec, t, self.tgsw_params)
        L.tLweClear(t, self.tlwe_params)
        rows = self.cloud.bk.contents.bk[bk_index].all_sample
        for i in range(self.kpl):
            for c in range(self.k + 1):
                L.torusPolynomialAddMulRKaratsuba(C.byref(t.contents.a[c]), C.byref(dec[i]), C.byref(rows[i].a[c]))
        out = self._tlwe_read(t)
        L.delete_IntPolynomial_array(self.kpl, dec)
        L.delete_TLweSample(t)
        return out

    def cmux_step(self, acc, bk_index, barai):
        """tfhe_MuxRotate_FFT (lwe-bootstrapping-functions-fft.cpp:67-76)."""
        src = self._tlwe_from(np.ascontiguousarray(acc, dtype=np.int32))
        dst = self.L.new_TLweSample(self.tlwe_params)
        bkfft = self.cloud.bkFFT.contents.bkFFT
        self.ref._muxrot(dst, src, C.c_void_p(bkfft + 24 * bk_index), int(barai), self.tgsw_params)
        out = self._tlwe_read(dst)
        self.L.delete_TLweSample(src)
        self.L.delete_TLweSample(dst)
        return out

    def export_samples(self, cts):
        """Bytes written by export_gate_bootstrapping_ciphertext_toFile (tfhe_io.cpp:1214) for each sample."""
        import tempfile
        cts = np.asarray(cts, dtype=np.int32)
        smp = self._new(len(cts))
        self._fill(smp, cts)
        L = self.L
        L.export_gate_bootstrapping_ciphertext_toFile.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        with tempfile.NamedTemporaryFile(suffix=".ct") as tf:
            f = self.ref.libc.fopen(tf.name.encode(), b"wb")
            for i in range(len(cts)):
                L.export_gate_bootstrapping_ciphertext_toFile(f, C.byref(smp[i]), self.sk.contents.params)
            self.ref.libc.fclose(f)
            data = open(tf.name, "rb").read()
        L.delete_LweSample_array(len(cts), smp)
        return data

    def import_samples(self, data, count):
        """What import_gate_bootstrapping_ciphertext_fromFile (tfhe_io.cpp:1220) reads back from `data`."""
        import tempfile
        smp = self._new(count)
        L = self.L
        L.import_gate_bootstrapping_ciphertext_fromFile.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        with tempfile.NamedTemporaryFile(suffix=".ct") as tf:
            tf.write(data)
            tf.flush()
            f = self.ref.libc.fopen(tf.name.encode(), b"rb")
            for i in range(count):
                L.import_gate_bootstrapping_ciphertext_fromFile(f, C.byref(smp[i]), self.sk.contents.params)
            self.ref.libc.fclose(f)
        out = self._read(smp, count, self.n)
        L.delete_LweSample_array(count, smp)
        return out

    def export_secret(self, path):
        f = self.ref.libc.fopen(path.encode(), b"wb")
        self.L.export_tfheGateBootstrappingSecretKeySet_toFile(f, self.sk)
        self.ref.libc.fclose(f)

    def export_cloud(self, path):
        f = self.ref.libc.fopen(path.encode(), b"wb")
        self.L.export_tfheGateBootstrappingCloudKeySet_toFile(f, self.cloud_ptr)
        self.ref.libc.fclose(f)
