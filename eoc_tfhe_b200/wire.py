"""libtfhe wire format for the objects that cross the gate-bootstrapping path (host only).

Thin mirror of csrc/wire.cpp (the parsing and formatting happen in C++): serialized cloud keys as
written by `export_tfheGateBootstrappingCloudKeySet_toFile` (libs/tfhe/src/libtfhe/tfhe_io.cpp:1099-1112)
and LWE sample records as written by `export_gate_bootstrapping_ciphertext_toFile` (:90-107, :1214).
"""
import ctypes as C

import numpy as np

from . import lib as _lib
from .engine import CloudKey, Params


def parse_cloud_key(data):
    """bytes -> (Params, ks_variance, bk int32 [n][kpl][k+1][N], ks int32 [N][t][base][n+1], info dict).
    The arrays are copies (the payloads sit at unaligned offsets after the text sections)."""
    L = _lib.load()
    buf = (C.c_char * len(data)).from_buffer_copy(data)
    p, info = _lib.b200_params(), _lib.b200_wire_cloudkey_info()
    _lib.check(L.b200_wire_parse_cloudkey(C.addressof(buf), len(data), C.byref(p), C.byref(info)))
    params = Params(n=p.n, N=p.N, k=p.k, l=p.l, Bgbit=p.Bgbit, ks_t=p.ks_t, ks_basebit=p.ks_basebit,
                    alpha_lwe=info.lwe_alpha_min, alpha_bk=info.tlwe_alpha_min)
    raw = np.frombuffer(data, dtype=np.uint8)
    ks = raw[info.ks_offset:info.ks_offset + info.ks_bytes].copy().view(np.int32).reshape(params.ks_shape)
    bk = raw[info.bk_offset:info.bk_offset + info.bk_bytes].copy().view(np.int32).reshape(params.bk_shape)
    meta = {"lwe_alpha": (info.lwe_alpha_min, info.lwe_alpha_max), "tlwe_alpha": (info.tlwe_alpha_min, info.tlwe_alpha_max),
            "bk_variance": info.bk_variance, "total_bytes": info.total_bytes}
    return params, p.ks_variance, bk, ks, meta


def write_cloud_key(params, bk, ks, ks_variance=0.0, bk_variance=0.0, lwe_alpha=(0.0, 0.0), tlwe_alpha=(0.0, 0.0)):
    """Serialize raw key arrays the way libtfhe does; the result loads with
    new_tfheGateBootstrappingCloudKeySet_fromFile."""
    L = _lib.load()
    bk = np.ascontiguousarray(bk, dtype=np.int32).reshape(params.bk_shape)
    ks = np.ascontiguousarray(ks, dtype=np.int32).reshape(params.ks_shape)
    p = _lib.b200_params(params.n, params.N, params.k, params.l, params.Bgbit, params.ks_t, params.ks_basebit, float(ks_variance))
    info = _lib.b200_wire_cloudkey_info()
    info.lwe_alpha_min, info.lwe_alpha_max = lwe_alpha
    info.tlwe_alpha_min, info.tlwe_alpha_max = tlwe_alpha
    info.bk_variance = bk_variance
    cap = int(L.b200_wire_cloudkey_size(C.byref(p))) + 64
    out = np.empty(cap, dtype=np.uint8)
    written = C.c_size_t()
    _lib.check(L.b200_wire_write_cloudkey(out.ctypes.data, cap, C.byref(written), C.byref(p), C.byref(info),
                                          bk.ctypes.data, ks.ctypes.data))
    return out[:written.value].tobytes()


def cloud_key_from_bytes(data, device=None):
    """Device-resident CloudKey from a serialized libtfhe cloud key (needs a GPU)."""
    L = _lib.load()
    if device is not None:
        _lib.check(L.b200_set_device(int(device)))
    buf = (C.c_char * len(data)).from_buffer_copy(data)
    h = C.c_void_p()
    _lib.check(L.b200_key_from_cloudkey_bytes(C.addressof(buf), len(data), C.byref(h)))
    cp = _lib.b200_params()
    _lib.check(L.b200_key_params(h, C.byref(cp)))
    return CloudKey(h.value, Params(n=cp.n, N=cp.N, k=cp.k, l=cp.l, Bgbit=cp.Bgbit, ks_t=cp.ks_t, ks_basebit=cp.ks_basebit))


def cloud_key_from_file(path, device=None):
    L = _lib.load()
    if device is not None:
        _lib.check(L.b200_set_device(int(device)))
    h = C.c_void_p()
    _lib.check(L.b200_key_from_cloudkey_file(str(path).encode(), C.byref(h)))
    cp = _lib.b200_params()
    _lib.check(L.b200_key_params(h, C.byref(cp)))
    return CloudKey(h.value, Params(n=cp.n, N=cp.N, k=cp.k, l=cp.l, Bgbit=cp.Bgbit, ks_t=cp.ks_t, ks_basebit=cp.ks_basebit))


def read_samples(data, n, count=None):
    """bytes of consecutive LWE sample records -> (int32 [count][n+1], float64 [count] variances)."""
    L = _lib.load()
    rec = int(L.b200_wire_sample_bytes(n))
    if count is None:
        count = len(data) // rec
    buf = (C.c_char * len(data)).from_buffer_copy(data)
    cts = np.empty((count, n + 1), dtype=np.int32)
    var = np.empty(count, dtype=np.float64)
    _lib.check(L.b200_wire_read_samples(C.addressof(buf), len(data), n, count, cts.ctypes.data, var.ctypes.data))
    return cts, var


def write_samples(cts, variances=None):
    L = _lib.load()
    cts = np.ascontiguousarray(cts, dtype=np.int32)
    count, n1 = cts.shape
    rec = int(L.b200_wire_sample_bytes(n1 - 1))
    out = np.empty(count * rec, dtype=np.uint8)
    var = None if variances is None else np.ascontiguousarray(variances, dtype=np.float64)
    _lib.check(L.b200_wire_write_samples(out.ctypes.data, out.size, n1 - 1, count, cts.ctypes.data,
                                         None if var is None else var.ctypes.data))
    return out.tobytes()
