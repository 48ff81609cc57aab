"""The Node-API addon (bindings/node/b200tfhe_napi.cc) loaded and called through the test host
bindings/node/napi_host.cc -- node itself is not installed on these boxes.  The calls mirror
bindings/node/b200tfhe.test.js, which is the same test for a machine that has node.

not gpu: the addon links, registers its twelve exports and fails loudly without a device.
gpu:     cloud key from libtfhe's serialization, gateBatch, runCircuit, readSamples, the eoc string entry points."""
import ctypes as C
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "bindings", "node", "build")
EXPORTS = {"info", "uploadCloudKey", "freeKey", "keyParams", "gateBatch", "runCircuit", "readSamples", "addCiphertexts",
           "subtractCiphertexts", "setCloudKeyBase64", "addCiphertextBits", "subtractCiphertextBits"}
UNDEFINED, NUL, INT32, DOUBLE, STRING, EXTERNAL, ARRAYBUFFER, TYPEDARRAY, BUFFER, OBJECT, FUNCTION = range(11)


class Host:
    def __init__(self):
        self.h = h = C.CDLL(os.path.join(BUILD, "libnapi_host.so"), mode=C.RTLD_GLOBAL)   # the addon's napi_* bind here
        VP = C.c_void_p
        h.host_load.restype = VP
        h.host_load.argtypes = [C.c_char_p, C.c_char_p, C.c_size_t]
        for name, res, args in (("host_export_count", C.c_int, [VP]), ("host_export_name", C.c_int, [VP, C.c_int, C.c_char_p, C.c_size_t]),
                                ("host_int32", VP, [VP, C.c_int32]), ("host_string", VP, [VP, C.c_char_p]),
                                ("host_buffer", VP, [VP, VP, C.c_size_t]), ("host_int32array", VP, [VP, VP, C.c_size_t]),
                                ("host_call", VP, [VP, C.c_char_p, C.c_int, C.POINTER(VP)]), ("host_error", C.c_char_p, [VP]),
                                ("host_kind", C.c_int, [VP]), ("host_get_int32", C.c_int32, [VP]), ("host_get_double", C.c_double, [VP]),
                                ("host_get_string", C.c_char_p, [VP]), ("host_get_property", VP, [VP, C.c_char_p]),
                                ("host_typedarray_length", C.c_size_t, [VP]), ("host_typedarray_data", VP, [VP]),
                                ("host_reset", None, [VP]), ("host_unload", None, [VP])):
            getattr(h, name).restype = res
            getattr(h, name).argtypes = args
        err = C.create_string_buffer(512)
        self.env = h.host_load(os.path.join(BUILD, "b200tfhe.node").encode(), err, 512)
        assert self.env, err.value.decode()

    def exports(self):
        names = {}
        for i in range(self.h.host_export_count(self.env)):
            buf = C.create_string_buffer(64)
            is_fn = self.h.host_export_name(self.env, i, buf, 64)
            names[buf.value.decode()] = bool(is_fn)
        return names

    def val(self, x):
        if isinstance(x, C.c_void_p) or x is None:
            return x
        if isinstance(x, (int, np.integer)):
            return C.c_void_p(self.h.host_int32(self.env, int(x)))
        if isinstance(x, str):
            return C.c_void_p(self.h.host_string(self.env, x.encode()))
        if isinstance(x, bytes):
            return C.c_void_p(self.h.host_buffer(self.env, C.cast(C.c_char_p(x), C.c_void_p), len(x)))
        x = np.ascontiguousarray(x, dtype=np.int32)
        return C.c_void_p(self.h.host_int32array(self.env, x.ctypes.data, x.size))

    def call(self, name, *args):
        """-> the result converted to Python, or raises RuntimeError with the JavaScript exception's message"""
        argv = (C.c_void_p * max(1, len(args)))(*[self.val(a) for a in args])
        r = self.h.host_call(self.env, name.encode(), len(args), argv)
        if not r:
            raise RuntimeError((self.h.host_error(self.env) or b"no result").decode())
        return self.py(C.c_void_p(r))

    def py(self, v):
        k = self.h.host_kind(v)
        if k in (UNDEFINED, NUL):
            return None
        if k == INT32:
            return self.h.host_get_int32(v)
        if k == DOUBLE:
            return self.h.host_get_double(v)
        if k == STRING:
            return self.h.host_get_string(v).decode()
        if k == EXTERNAL:
            return v
        if k == TYPEDARRAY:
            n = self.h.host_typedarray_length(v)
            return np.ctypeslib.as_array(C.cast(self.h.host_typedarray_data(v), C.POINTER(C.c_int32)), shape=(n,)).copy()
        if k == OBJECT:
            return v
        raise AssertionError("unexpected value kind %d" % k)

    def prop(self, obj, name):
        p = self.h.host_get_property(obj, name.encode())
        return None if not p else self.py(C.c_void_p(p))


@pytest.fixture(scope="module")
def host():
    if not os.path.exists(os.path.join(BUILD, "b200tfhe.node")):
        pytest.skip("addon not built (python __graft_entry__.py)")
    return Host()


def test_addon_links_and_registers_its_exports(host):
    names = host.exports()
    assert set(names) == EXPORTS and all(names.values())
    info = host.call("info")
    assert host.prop(info, "devices") >= 0 and host.prop(info, "launches") >= 0


def test_addon_argument_errors_become_exceptions(host):
    with pytest.raises(RuntimeError, match="uploadCloudKey"):
        host.call("uploadCloudKey", 5)                        # not a Buffer
    with pytest.raises(RuntimeError, match="cloud key"):
        host.call("uploadCloudKey", b"not a key file")        # parse error text from the library
    with pytest.raises(RuntimeError, match="gateBatch"):
        host.call("gateBatch")
    with pytest.raises(RuntimeError, match="readSamples"):
        host.call("readSamples", b"", 0)
    assert host.call("addCiphertexts", "!!not base64!!", "AAAA") is None     # the reference returns nil here


@pytest.mark.gpu
def test_addon_on_the_gpu(host, tmp_path):
    import reflib
    if not reflib.available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    from eoc_tfhe_b200 import circuits
    rk = reflib.RefLib().keygen(80, seed=5)
    path = str(tmp_path / "cloud.key")
    rk.export_cloud(path)
    key = host.call("uploadCloudKey", open(path, "rb").read())
    try:
        p = host.call("keyParams", key)
        assert (host.prop(p, "n"), host.prop(p, "N"), host.prop(p, "l"), host.prop(p, "Bgbit")) == (rk.n, rk.N, rk.l, rk.Bgbit)
        rng = np.random.default_rng(3)
        ab, bb, cb = (rng.integers(0, 2, size=40).astype(np.int32) for _ in range(3))
        a, b, c = rk.encrypt_bits(ab), rk.encrypt_bits(bb), rk.encrypt_bits(cb)
        n1 = rk.n + 1
        out = host.call("gateBatch", key, 0, a.ravel(), b.ravel()).reshape(-1, n1)          # NAND
        assert np.array_equal(rk.decrypt_bits(out), 1 - (ab & bb))
        out = host.call("gateBatch", key, 10, a.ravel(), b.ravel(), c.ravel()).reshape(-1, n1)   # MUX
        assert np.array_equal(rk.decrypt_bits(out), np.where(ab > 0, bb, cb))
        # readSamples on libtfhe's own export of the same ciphertexts
        assert np.array_equal(host.call("readSamples", rk.export_samples(a), rk.n).reshape(-1, n1), a)
        # a 4-bit ripple-carry adder netlist through runCircuit
        nb = 4
        nl = circuits.full_adder(nb)
        x, y = 11, 6
        wires = np.zeros((nl.num_wires, n1), dtype=np.int32)
        wires[nl.inputs["x"]] = rk.encrypt_bits(circuits.int_to_bits([x], nb)[0])
        wires[nl.inputs["y"]] = rk.encrypt_bits(circuits.int_to_bits([y], nb)[0])
        flat = host.val(wires.ravel())
        levels = host.call("runCircuit", key, nl.to_array().ravel(), flat)
        assert levels >= nb
        res = host.py(flat).reshape(-1, n1)
        assert int(circuits.bits_to_int(rk.decrypt_bits(res[nl.outputs["sum"]]))) == x + y
        with pytest.raises(RuntimeError, match="batches differ"):
            host.call("gateBatch", key, 0, a.ravel(), b.ravel()[:-n1])
    finally:
        host.call("freeKey", key)
