"""ctypes binding of oracle/liboracle.so (the CPU restatement, oracle/tfhe_port.c). Test infrastructure only."""
import ctypes as C
import os
import subprocess

import numpy as np

ORACLE_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(ORACLE_DIR, "liboracle.so")

OPS = {"NAND": 0, "OR": 1, "AND": 2, "XOR": 3, "XNOR": 4, "NOR": 5, "ANDNY": 6, "ANDYN": 7, "ORNY": 8,
       "ORYN": 9, "MUX": 10, "NOT": 11, "COPY": 12, "CONST0": 13, "CONST1": 14}


class port_params(C.Structure):
    _fields_ = [(f, C.c_int32) for f in ("n", "N", "k", "l", "Bgbit", "ks_t", "ks_basebit")]


_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    src = [os.path.join(ORACLE_DIR, f) for f in ("tfhe_port.c", "tfhe_port.h")]
    if not os.path.exists(LIB_PATH) or any(os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in src):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "port"], stdout=subprocess.DEVNULL)
    L = C.CDLL(LIB_PATH, mode=C.RTLD_LOCAL)
    VP, I32 = C.c_void_p, C.c_int32
    L.port_key_new.restype = VP
    L.port_key_new.argtypes = [C.POINTER(port_params), VP, VP]
    L.port_key_free.argtypes = [VP]
    L.port_modswitch_from_torus32.restype = I32
    L.port_modswitch_from_torus32.argtypes = [I32, I32]
    L.port_modswitch_to_torus32.restype = I32
    L.port_modswitch_to_torus32.argtypes = [I32, I32]
    L.port_decomp_offset.restype = C.c_uint32
    L.port_decomp_offset.argtypes = [C.POINTER(port_params)]
    L.port_mul_by_xai.argtypes = [VP, I32, VP, I32]
    L.port_mul_by_xai_minus_one.argtypes = [VP, I32, VP, I32]
    L.port_decompose.argtypes = [VP, VP, C.POINTER(port_params)]
    L.port_ifft_int.argtypes = [VP, VP]
    L.port_fft_torus.argtypes = [VP, VP]
    L.port_extern_product_fft.argtypes = [VP, VP, I32]
    L.port_extern_product_exact.argtypes = [VP, VP, I32]
    L.port_cmux_step.argtypes = [VP, VP, I32, I32, C.c_int]
    L.port_bootstrap_woks.argtypes = [VP, VP, VP, I32]
    L.port_keyswitch.restype = I32
    L.port_keyswitch.argtypes = [VP, VP, VP]
    L.port_gate_prologue.argtypes = [C.c_int, C.c_int, VP, VP, VP, VP, I32]
    L.port_gate.argtypes = [VP, C.c_int, VP, VP, VP, VP]
    L.port_gate_batch.argtypes = [VP, C.c_int, VP, VP, VP, VP, C.c_int64, C.c_int]
    _lib = L
    return L


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


class PortKey:
    """CPU oracle key from wire-format arrays: bk [n][kpl][k+1][N], ks [N][t][base][n+1]."""

    def __init__(self, n, l, Bgbit, bk, ks, N=1024, k=1, ks_t=8, ks_basebit=2):
        self.L = load()
        self.p = port_params(n, N, k, l, Bgbit, ks_t, ks_basebit)
        self.n, self.N, self.k, self.l, self.Bgbit = n, N, k, l, Bgbit
        bk, ks = _i32(bk), _i32(ks)
        assert bk.shape == (n, (k + 1) * l, k + 1, N) and ks.shape == (N * k, ks_t, 1 << ks_basebit, n + 1)
        self._h = self.L.port_key_new(C.byref(self.p), bk.ctypes.data, ks.ctypes.data)
        assert self._h

    def close(self):
        if self._h:
            self.L.port_key_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def extern_product(self, acc, idx, exact=False):
        acc = _i32(acc).copy()
        (self.L.port_extern_product_exact if exact else self.L.port_extern_product_fft)(self._h, acc.ctypes.data, idx)
        return acc

    def cmux_step(self, acc, idx, barai, exact=False):
        acc = _i32(acc).copy()
        self.L.port_cmux_step(self._h, acc.ctypes.data, idx, barai, int(exact))
        return acc

    def bootstrap_woks(self, x, mu=1 << 29):
        x = _i32(x)
        u = np.empty((x.shape[0], self.N + 1), dtype=np.int32)
        for i in range(x.shape[0]):
            self.L.port_bootstrap_woks(self._h, u[i].ctypes.data, x[i].ctypes.data, mu)
        return u

    def keyswitch(self, u):
        u = _i32(u)
        out = np.empty((u.shape[0], self.n + 1), dtype=np.int32)
        nz = np.empty(u.shape[0], dtype=np.int32)
        for i in range(u.shape[0]):
            nz[i] = self.L.port_keyswitch(self._h, out[i].ctypes.data, u[i].ctypes.data)
        return out, nz

    def gate(self, name, a=None, b=None, c=None, threads=1, count=None):
        arrs = [None if x is None else _i32(x) for x in (a, b, c)]
        count = count if count is not None else arrs[0].shape[0]
        out = np.empty((count, self.n + 1), dtype=np.int32)
        ptr = [None if x is None else x.ctypes.data for x in arrs]
        self.L.port_gate_batch(self._h, OPS[name], out.ctypes.data, ptr[0], ptr[1], ptr[2], count, threads)
        return out


def prologue(name, half, a, b, c=None):
    L = load()
    a, b = _i32(a), _i32(b)
    c = None if c is None else _i32(c)
    x = np.empty_like(a)
    L.port_gate_prologue(OPS[name], half, x.ctypes.data, a.ctypes.data, b.ctypes.data,
                         None if c is None else c.ctypes.data, a.shape[0] - 1)
    return x
