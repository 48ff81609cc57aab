"""Evaluation side of eoc's entry points (csrc/eoc_abi.cpp; ao-tfhe/eoc-tfhe-run.cpp:427-513).

`add_ciphertexts` / `subtract_ciphertexts` are the reference's `addCiphertexts` / `subtractCiphertexts`:
base64 strings of one LweSample wire record each, linear, bit-exact, no device.  The `*_bits` functions are
the bootstrapped counterpart for integers encrypted bit by bit: the ripple-carry adder of
libs/tfhe/src/test/test-addition-boot.cpp:70-94 evaluated level by level on the GPU.
"""
import ctypes as C

import numpy as np

from . import lib as _lib


def _take(ptr):
    if not ptr:
        raise _lib.B200Error(_lib.load().b200_last_error().decode("utf-8", "replace"))
    try:
        return C.string_at(ptr).decode("ascii")
    finally:
        _lib.load().b200_eoc_free(ptr)


def add_ciphertexts(b64_1, b64_2, b64_public_key=""):
    """eoc-tfhe-run.cpp:427-469 (the third argument is ignored there too)."""
    return _take(_lib.load().addCiphertexts(b64_1.encode(), b64_2.encode(), b64_public_key.encode()))


def subtract_ciphertexts(b64_1, b64_2, b64_public_key=""):
    """eoc-tfhe-run.cpp:472-513."""
    return _take(_lib.load().subtractCiphertexts(b64_1.encode(), b64_2.encode(), b64_public_key.encode()))


def set_cloud_key(b64_cloud_key):
    """Evaluation key of the base64 bit-sliced calls from base64(export_tfheGateBootstrappingCloudKeySet_toStream)."""
    _lib.check(_lib.load().b200_eoc_set_cloud_key(b64_cloud_key.encode()))


def set_key(cloud_key):
    """Use an existing CloudKey (kept alive by the caller); None clears it."""
    _lib.check(_lib.load().b200_eoc_set_key(cloud_key._h if cloud_key is not None else None))


def add_ciphertext_bits(b64_x, b64_y, nb_bits, subtract=False):
    """x + y (or x - y) of two nb_bits-bit integers, each base64 of nb_bits concatenated sample records
    (LSB first); returns base64 of nb_bits + 1 records, the last one the carry out."""
    L = _lib.load()
    f = L.subtractCiphertextBits if subtract else L.addCiphertextBits
    return _take(f(b64_x.encode(), b64_y.encode(), int(nb_bits)))


def add_bits(cloud_key, x, y, subtract=False):
    """x, y: int32 [count][nb_bits][n+1] -> [count][nb_bits+1][n+1] through b200_add_bits_host."""
    x = np.ascontiguousarray(x, dtype=np.int32)
    y = np.ascontiguousarray(y, dtype=np.int32)
    if x.ndim != 3 or x.shape != y.shape or x.shape[2] != cloud_key.params.n + 1:
        raise ValueError("operands must both be [count][nb_bits][n+1]")
    count, nb, n1 = x.shape
    out = np.empty((count, nb + 1, n1), dtype=np.int32)
    _lib.check(_lib.load().b200_add_bits_host(cloud_key._h, out.ctypes.data, x.ctypes.data, y.ctypes.data, nb, count,
                                              1 if subtract else 0))
    return out
