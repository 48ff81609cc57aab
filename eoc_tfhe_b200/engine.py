"""Host-side mirror of the reference's gate API for the GPU engine.

Names follow libtfhe: a *cloud key* (TFheGateBootstrappingCloudKeySet,
libs/tfhe/src/include/tfhe_gate_bootstrapping_structures.h:26-42) evaluates *gates*
(libs/tfhe/src/libtfhe/boot-gates.cpp) on *LWE samples* -- here batches `int32[count][n+1]`
(mask a[0..n) then b), as numpy arrays on the host or torch CUDA tensors resident in HBM.
"""
import ctypes as C
import enum
from dataclasses import dataclass

import numpy as np

from . import lib as _lib


class GateOp(enum.IntEnum):
    NAND = 0
    OR = 1
    AND = 2
    XOR = 3
    XNOR = 4
    NOR = 5
    ANDNY = 6
    ANDYN = 7
    ORNY = 8
    ORYN = 9
    MUX = 10
    NOT = 11
    COPY = 12
    CONST0 = 13
    CONST1 = 14


@dataclass(frozen=True)
class Params:
    """Gate-bootstrapping parameter set (tfhe_gate_bootstrapping.cpp:23-75)."""
    n: int
    N: int = 1024
    k: int = 1
    l: int = 2
    Bgbit: int = 10
    ks_t: int = 8
    ks_basebit: int = 2
    alpha_lwe: float = 2.44e-5   # in_out_params->alpha_min (ciphertext / KS noise)
    alpha_bk: float = 7.18e-9    # tlwe_params->alpha_min (bootstrapping-key noise)

    @property
    def kpl(self):
        return (self.k + 1) * self.l

    @property
    def bk_shape(self):
        return (self.n, self.kpl, self.k + 1, self.N)

    @property
    def ks_shape(self):
        return (self.N * self.k, self.ks_t, 1 << self.ks_basebit, self.n + 1)


# new_default_gate_bootstrapping_parameters(lambda<=80) / (80<lambda<=128)
PARAMS_80 = Params(n=500, l=2, Bgbit=10, alpha_lwe=2.44e-5, alpha_bk=7.18e-9)
PARAMS_128 = Params(n=630, l=3, Bgbit=7, alpha_lwe=2.0 ** -15, alpha_bk=2.0 ** -25)


def _np_i32(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.int32)
    if shape is not None and tuple(a.shape) != tuple(shape):
        raise ValueError("expected shape %s, got %s" % (tuple(shape), tuple(a.shape)))
    return a


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return a.data_ptr()  # torch tensor


class CloudKey:
    """Device-resident cloud key: FFT-domain bootstrapping key + key-switch table in HBM."""

    def __init__(self, handle, params, owns=True):
        self._h = C.c_void_p(handle)
        self.params = params
        self._owns = owns

    # ---- construction --------------------------------------------------------------------
    @staticmethod
    def _cparams(p, ks_variance):
        return _lib.b200_params(p.n, p.N, p.k, p.l, p.Bgbit, p.ks_t, p.ks_basebit, float(ks_variance))

    @classmethod
    def from_arrays(cls, params, bk_coefs, ks_rows, ks_variance=None, device=None):
        """bk_coefs int32 [n][kpl][k+1][N], ks_rows int32 [N][t][base][n+1] (wire-format order)."""
        L = _lib.load()
        if device is not None:
            _lib.check(L.b200_set_device(int(device)))
        bk = _np_i32(bk_coefs, params.bk_shape)
        ks = _np_i32(ks_rows, params.ks_shape)
        if ks_variance is None:
            ks_variance = params.alpha_lwe ** 2
        cp = cls._cparams(params, ks_variance)
        h = C.c_void_p()
        _lib.check(L.b200_key_create(C.byref(cp), bk.ctypes.data, ks.ctypes.data, C.byref(h)))
        return cls(h.value, params)

    @classmethod
    def empty(cls, params, ks_variance=None, device=None):
        """Allocated but unfilled key (non-root ranks before the broadcast)."""
        L = _lib.load()
        if device is not None:
            _lib.check(L.b200_set_device(int(device)))
        cp = cls._cparams(params, params.alpha_lwe ** 2 if ks_variance is None else ks_variance)
        h = C.c_void_p()
        _lib.check(L.b200_key_create_empty(C.byref(cp), C.byref(h)))
        return cls(h.value, params)

    @classmethod
    def from_libtfhe(cls, cloud_key_ptr, device=None):
        """From a live `TFheGateBootstrappingCloudKeySet*` (an int address)."""
        L = _lib.load()
        if device is not None:
            _lib.check(L.b200_set_device(int(device)))
        h = C.c_void_p()
        _lib.check(L.b200_cloudkey_upload(C.c_void_p(cloud_key_ptr), C.byref(h)))
        cp = _lib.b200_params()
        _lib.check(L.b200_key_params(h, C.byref(cp)))
        p = Params(n=cp.n, N=cp.N, k=cp.k, l=cp.l, Bgbit=cp.Bgbit, ks_t=cp.ks_t, ks_basebit=cp.ks_basebit)
        return cls(h.value, p, owns=False)

    def close(self):
        if self._h and self._owns:
            _lib.load().b200_key_free(self._h)
        self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- resident blobs (for the NCCL broadcast) -----------------------------------------
    def device_blobs(self):
        """[(ptr, nbytes), (ptr, nbytes)] of the FFT-domain BK and the KS table."""
        a, b = C.c_void_p(), C.c_void_p()
        na, nb = C.c_size_t(), C.c_size_t()
        _lib.check(_lib.load().b200_key_device_blobs(self._h, C.byref(a), C.byref(na), C.byref(b), C.byref(nb)))
        return [(a.value, na.value), (b.value, nb.value)]

    def device_blob_tensors(self):
        """The two resident blobs as zero-copy torch uint8 CUDA tensors (for dist.broadcast)."""
        import torch

        class _Raw:
            def __init__(self, ptr, nbytes):
                self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}

        return [torch.as_tensor(_Raw(p, nb), device="cuda") for p, nb in self.device_blobs()]

    def timing_drain(self):
        """Summed kernel times since the previous drain: {"blind_rotate": ms, "keyswitch": ms, "calls": n}."""
        arr = (C.c_double * 2)()
        calls = C.c_int32()
        _lib.check(_lib.load().b200_timing_drain(self._h, C.byref(arr), C.byref(calls)))
        return {"blind_rotate": arr[0], "keyswitch": arr[1], "calls": calls.value}

    # ---- gates -----------------------------------------------------------------------------
    def _rowlen(self):
        return self.params.n + 1

    def gate(self, op, a=None, b=None, c=None, out=None, stream=None):
        """out = op(a, b, c) on a batch. numpy in -> numpy out (copies inside the call);
        torch CUDA tensors in -> torch CUDA tensor out (asynchronous on `stream`)."""
        op = GateOp(op)
        L = _lib.load()
        first = a if a is not None else out
        if first is None:
            raise ValueError("CONST gates need `out`")
        if isinstance(first, np.ndarray) or (a is None and isinstance(out, np.ndarray)):
            a_ = None if a is None else _np_i32(a)
            b_ = None if b is None else _np_i32(b)
            c_ = None if c is None else _np_i32(c)
            count = (a_ if a_ is not None else out).shape[0]
            if out is None:
                out = np.empty((count, self._rowlen()), dtype=np.int32)
            for x in (a_, b_, c_, out):
                if x is not None and tuple(x.shape) != (count, self._rowlen()):
                    raise ValueError("batch shape must be (count, n+1)")
            _lib.check(L.b200_gate_batch_host(self._h, int(op), _ptr(out), _ptr(a_), _ptr(b_), _ptr(c_), count))
            return out
        import torch
        ts = [x for x in (a, b, c) if x is not None]
        count = first.shape[0]
        for x in ts:
            if not (x.is_cuda and x.dtype == torch.int32 and x.is_contiguous() and tuple(x.shape) == (count, self._rowlen())):
                raise ValueError("device batches must be contiguous int32 CUDA tensors of shape (count, n+1)")
        if out is None:
            out = torch.empty((count, self._rowlen()), dtype=torch.int32, device=first.device)
        s = stream if stream is not None else torch.cuda.current_stream(first.device).cuda_stream
        _lib.check(L.b200_gate_batch_dev(self._h, int(op), _ptr(out), _ptr(a), _ptr(b), _ptr(c), count, C.c_void_p(s)))
        return out

    def gate_noalloc(self, op, a=None, b=None, c=None, out=None, stream=None):
        """`gate` on device tensors through b200_gate_batch_dev_noalloc: the library allocates nothing on the call
        (static scratch, at most NOALLOC_MAX_ROTATIONS blind rotations; MUX takes two per gate).  `out` is required:
        allocating it here would defeat the purpose."""
        import torch
        op = GateOp(op)
        if out is None:
            raise ValueError("gate_noalloc needs a preallocated `out`")
        count = out.shape[0]
        for x in (a, b, c, out):
            if x is not None and not (x.is_cuda and x.dtype == torch.int32 and x.is_contiguous()
                                      and tuple(x.shape) == (count, self._rowlen())):
                raise ValueError("device batches must be contiguous int32 CUDA tensors of shape (count, n+1)")
        s = stream if stream is not None else torch.cuda.current_stream(out.device).cuda_stream
        _lib.check(_lib.load().b200_gate_batch_dev_noalloc(self._h, int(op), _ptr(out), _ptr(a), _ptr(b), _ptr(c), count,
                                                           C.c_void_p(s)))
        return out

    def nand(self, a, b, **kw):
        return self.gate(GateOp.NAND, a, b, **kw)

    def mux(self, a, b, c, **kw):
        return self.gate(GateOp.MUX, a, b, c, **kw)

    # ---- pieces of the path (parity tests) -------------------------------------------------
    def bootstrap_woks(self, x, mu=1 << 29):
        x = _np_i32(x)
        u = np.empty((x.shape[0], self.params.N * self.params.k + 1), dtype=np.int32)
        _lib.check(_lib.load().b200_bootstrap_woks_host(self._h, _ptr(u), _ptr(x), int(mu), x.shape[0]))
        return u

    def keyswitch(self, u, with_counts=False):
        u = _np_i32(u)
        out = np.empty((u.shape[0], self._rowlen()), dtype=np.int32)
        nz = np.zeros(u.shape[0], dtype=np.int32)
        _lib.check(_lib.load().b200_keyswitch_host(self._h, _ptr(out), _ptr(u), _ptr(nz), u.shape[0]))
        return (out, nz) if with_counts else out

    def cmux_step(self, acc, bk_index, barai):
        acc = _np_i32(acc).copy()
        _lib.check(_lib.load().b200_cmux_step_host(self._h, _ptr(acc), int(bk_index), int(barai), acc.shape[0]))
        return acc

    def extern_product(self, acc, bk_index):
        acc = _np_i32(acc).copy()
        _lib.check(_lib.load().b200_extern_product_host(self._h, _ptr(acc), int(bk_index), acc.shape[0]))
        return acc

    # ---- circuits ----------------------------------------------------------------------------
    def run_circuit(self, gates, wires, stream=None):
        """gates int32 [G][5] = (op, out, a, b, c); wires [W][n+1] pool (numpy or torch CUDA).
        Returns the number of bootstrapped levels; outputs are written into `wires`."""
        gates = _np_i32(gates)
        L = _lib.load()
        lv = C.c_int32()
        if isinstance(wires, np.ndarray):
            if wires.dtype != np.int32 or not wires.flags.c_contiguous:
                raise ValueError("wires must be a contiguous int32 array")
            _lib.check(L.b200_circuit_run_host(self._h, gates.ctypes.data, gates.shape[0], wires.ctypes.data,
                                               wires.shape[0], C.byref(lv)))
        else:
            import torch
            s = stream if stream is not None else torch.cuda.current_stream(wires.device).cuda_stream
            _lib.check(L.b200_circuit_run_dev(self._h, gates.ctypes.data, gates.shape[0], wires.data_ptr(),
                                              wires.shape[0], C.byref(lv), C.c_void_p(s)))
        return lv.value

    def run_circuit_io(self, gates, num_wires, in_wires, in_rows, out_wires):
        """Circuit with packed host I/O (b200_circuit_run_host_io): in_rows [len(in_wires)][n+1] numpy int32 go
        to the wires in_wires; returns (rows of out_wires, bootstrapped levels)."""
        gates = _np_i32(gates)
        iw, ow = _np_i32(in_wires).ravel(), _np_i32(out_wires).ravel()
        rows = _np_i32(in_rows, (iw.size, self.params.n + 1))
        out = np.empty((ow.size, self.params.n + 1), dtype=np.int32)
        lv = C.c_int32()
        _lib.check(_lib.load().b200_circuit_run_host_io(self._h, gates.ctypes.data, gates.shape[0], int(num_wires),
                                                        iw.ctypes.data, iw.size, rows.ctypes.data,
                                                        ow.ctypes.data, ow.size, out.ctypes.data, C.byref(lv)))
        return out, lv.value

    def run_level(self, wires, rot_desc, ks_desc, stream=None):
        """One level from launch descriptors (see b200_expand_gates) over a device wire pool."""
        import torch
        rot = _np_i32(rot_desc)
        ks = _np_i32(ks_desc)
        s = stream if stream is not None else torch.cuda.current_stream(wires.device).cuda_stream
        _lib.check(_lib.load().b200_level_run_dev(self._h, wires.data_ptr(), rot.ctypes.data, rot.shape[0],
                                                  ks.ctypes.data, ks.shape[0], C.c_void_p(s)))

    def last_timing_ms(self):
        arr = (C.c_float * 3)()
        _lib.check(_lib.load().b200_last_timing_ms(self._h, C.byref(arr)))
        return {"blind_rotate": arr[0], "keyswitch": arr[1], "total": arr[2]}


def fp64_peak_tflops(device=0):
    """FP64 FMA peak of the device measured by the library's DFMA-chain kernel (TFLOP/s)."""
    v = C.c_double()
    _lib.check(_lib.load().b200_fp64_peak_tflops(int(device), C.byref(v)))
    return v.value


def int32_add_peak(device=0):
    """32-bit integer add throughput of the device measured by the library's chain kernel (10^12 adds/s)."""
    v = C.c_double()
    _lib.check(_lib.load().b200_int32_add_peak(int(device), C.byref(v)))
    return v.value


def set_blind_rotate_kernel(variant):
    """12 / "auto" (default): the fastest kernel per launch ("tmacc" for wide batches at l = 2, else "v2" and the
    latency kernel); 0 / "v2": shared-memory transforms; 9 / "tm", 10 / "tmacc", 11 / "tms": tensor-memory transforms
    (exchanges and roots in tensor memory; + accumulators; single transforms); 1..8: round-1 experiments (only in a
    B200_EXPERIMENTS build)."""
    v = {"v2": 0, "w16": 1, "v3g6": 2, "v3c2": 3, "v3g5": 4, "v3g4": 5, "v3g3c2": 6, "v4": 7, "v5": 8, "tm": 9, "tmacc": 10, "tms": 11, "auto": 12}.get(variant, variant)
    _lib.check(_lib.load().b200_set_blind_rotate_kernel(int(v)))


def fft_roundtrip(a):
    """inverse(forward(a)) of the selected kernel's transform on int32 polynomials [count][1024]: the identity."""
    a = _np_i32(a)
    out = np.empty_like(a)
    _lib.check(_lib.load().b200_fft_roundtrip_host(_ptr(out), _ptr(a), a.shape[0]))
    return out


def poly_mul(a, b):
    """a * b mod (X^1024 + 1, 2^32) through the transform; a small (gadget digits), b any Torus32."""
    a, b = _np_i32(a), _np_i32(b)
    out = np.empty_like(a)
    _lib.check(_lib.load().b200_poly_mul_host(_ptr(out), _ptr(a), _ptr(b), a.shape[0]))
    return out


def experiments_built():
    """True when the library carries the round-1 blind-rotation variants (-DB200_EXPERIMENTS)."""
    return bool(_lib.load().b200_experiments_built())


def set_keyswitch_kernel(variant):
    """0 / "auto": by batch size; 1 / "gather": 8-sample tiles; 2 / "stream": table streamed by TMA, 64 samples per CTA."""
    v = {"auto": 0, "gather": 1, "stream": 2}.get(variant, variant)
    _lib.check(_lib.load().b200_set_keyswitch_kernel(int(v)))


def launch_count():
    return int(_lib.load().b200_launch_count())
