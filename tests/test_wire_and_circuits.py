"""Host-side logic that needs no GPU: the C ABI exports every declared symbol, the libtfhe wire format
reader/writer (csrc/wire.cpp) agrees with the UNMODIFIED reference's own export/import, and the netlist
builders + levelizer (circuits.py, csrc/circuit.cpp) describe the circuits of
libs/tfhe/src/test/test-addition-boot.cpp."""
import ctypes as C
import os
import re
import tempfile

import numpy as np
import pytest

import reflib
from eoc_tfhe_b200 import GateOp, circuits, lib, wire

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
needs_ref = pytest.mark.skipif(not reflib.available(), reason="oracle/_ref not built (needs /root/reference)")


def test_library_exports_every_declared_symbol():
    """Every function declared in include/b200tfhe.h is exported by libb200tfhe.so and bound in lib.py."""
    text = open(os.path.join(ROOT, "include", "b200tfhe.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = set(re.findall(r"\b((?:b200_|boots|tfhe_bootstrap|lweKeySwitch|addCiphertext|subtractCiphertext)\w*)\s*\(", text))
    names -= {"b200_gate_op"}
    L = lib.load()
    missing = [n for n in sorted(names) if not hasattr(L, n)]
    assert not missing, missing
    unbound = [n for n in sorted(names) if n not in lib.SIGNATURES]
    assert not unbound, unbound
    assert len(names) >= 53


def test_no_device_fails_loudly():
    """Without a CUDA device every compute entry returns an error (there is no CPU fallback)."""
    L = lib.load()
    if L.b200_device_count() > 0:
        pytest.skip("a CUDA device is present")
    p = lib.b200_params(8, 1024, 1, 2, 10, 8, 2, 0.0)
    h = C.c_void_p()
    assert L.b200_key_create_empty(C.byref(p), C.byref(h)) != 0
    assert b"no CUDA device" in L.b200_last_error()


# ---- wire format ---------------------------------------------------------------------------------
@needs_ref
@pytest.mark.parametrize("cfg", [dict(n=8, l=2, Bgbit=10), dict(n=12, l=3, Bgbit=7)], ids=["l2", "l3"])
def test_cloud_key_parse_matches_reference_export(cfg):
    ref = reflib.RefLib()
    rk = ref.keygen_custom(seed=7, **cfg)
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "cloud.key")
        rk.export_cloud(path)
        data = open(path, "rb").read()
    params, ks_var, bk, ks, meta = wire.parse_cloud_key(data)
    assert (params.n, params.N, params.k, params.l, params.Bgbit, params.ks_t, params.ks_basebit) == \
           (rk.n, rk.N, rk.k, rk.l, rk.Bgbit, rk.ks_t, rk.ks_basebit)
    assert meta["total_bytes"] == len(data)
    assert np.array_equal(bk, rk.bk_coefs())
    assert np.array_equal(ks, rk.ks_rows())
    assert ks_var == pytest.approx(rk.ks_variance(), rel=1e-12)
    # our writer reproduces the reference's file byte for byte
    again = wire.write_cloud_key(params, bk, ks, ks_variance=ks_var, bk_variance=meta["bk_variance"],
                                 lwe_alpha=meta["lwe_alpha"], tlwe_alpha=meta["tlwe_alpha"])
    assert again == data


@needs_ref
def test_cloud_key_written_here_loads_in_reference():
    ref = reflib.RefLib()
    rk = ref.keygen_custom(seed=9, n=8, l=2, Bgbit=10)
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "cloud.key")
        rk.export_cloud(path)
        params, ks_var, bk, ks, meta = wire.parse_cloud_key(open(path, "rb").read())
        mine = os.path.join(d, "mine.key")
        with open(mine, "wb") as f:
            f.write(wire.write_cloud_key(params, bk, ks, ks_variance=ks_var, bk_variance=meta["bk_variance"],
                                         lwe_alpha=meta["lwe_alpha"], tlwe_alpha=meta["tlwe_alpha"]))
        fh = ref.libc.fopen(mine.encode(), b"rb")
        ck = ref.L.new_tfheGateBootstrappingCloudKeySet_fromFile(fh)
        ref.libc.fclose(fh)
    assert ck and ck.contents.params.contents.in_out_params.contents.n == 8
    first = ck.contents.bk.contents.bk[0].all_sample[0].a[0].coefsT
    assert [first[i] for i in range(5)] == [int(v) for v in bk[0, 0, 0, :5]]


def test_cloud_key_parse_rejects_garbage():
    with pytest.raises(lib.B200Error):
        wire.parse_cloud_key(b"not a key at all\n" * 10)
    with pytest.raises(lib.B200Error):
        wire.parse_cloud_key(b"")


@needs_ref
def test_sample_records_match_reference_export():
    ref = reflib.RefLib()
    rk = ref.keygen_custom(seed=11, n=8, l=2, Bgbit=10)
    bits = np.array([0, 1, 1, 0, 1], dtype=np.int32)
    cts = rk.encrypt_bits(bits)
    data = rk.export_samples(cts)                      # export_gate_bootstrapping_ciphertext_toFile per sample
    got, var = wire.read_samples(data, rk.n)
    assert np.array_equal(got, cts)
    assert np.all(var >= 0)
    assert wire.write_samples(got, var) == data
    back = rk.import_samples(wire.write_samples(cts, var), len(bits))   # the reference reads what we write
    assert np.array_equal(back, cts)


def test_sample_records_reject_wrong_uid():
    rec = wire.write_samples(np.zeros((2, 9), dtype=np.int32))
    bad = bytearray(rec)
    bad[0] = 43
    with pytest.raises(lib.B200Error):
        wire.read_samples(bytes(bad), 8)
    with pytest.raises(lib.B200Error):
        wire.read_samples(rec[:-1], 8, count=2)


# ---- netlists ------------------------------------------------------------------------------------
@pytest.mark.parametrize("builder,gates_per_bit,rot_per_bit", [(circuits.full_adder, 5, 5), (circuits.full_adder_mux, 4, 5)])
def test_adders_compute_sums(builder, gates_per_bit, rot_per_bit):
    nb = 16
    nl = builder(nb)
    assert nl.bootstrapped_gates() == gates_per_bit * nb
    assert nl.rotations() == rot_per_bit * nb
    rng = np.random.default_rng(99)
    x = rng.integers(0, 1 << nb, size=200)
    y = rng.integers(0, 1 << nb, size=200)
    xb, yb = circuits.int_to_bits(x, nb), circuits.int_to_bits(y, nb)
    vals = {w: xb[:, i] for i, w in enumerate(nl.inputs["x"])}
    vals.update({w: yb[:, i] for i, w in enumerate(nl.inputs["y"])})
    out = nl.evaluate_plain(vals)
    s = np.stack([np.broadcast_to(out[w], (200,)) for w in nl.outputs["sum"]], axis=1)
    assert np.array_equal(circuits.bits_to_int(s), x + y)


def test_comparison_netlist():
    nb = 16
    nl = circuits.comparison_mux(nb)
    rng = np.random.default_rng(5)
    x = rng.integers(0, 1 << nb, size=300)
    y = rng.integers(0, 1 << nb, size=300)
    y[:20] = x[:20]
    xb, yb = circuits.int_to_bits(x, nb), circuits.int_to_bits(y, nb)
    vals = {w: xb[:, i] for i, w in enumerate(nl.inputs["x"])}
    vals.update({w: yb[:, i] for i, w in enumerate(nl.inputs["y"])})
    out = nl.evaluate_plain(vals)
    assert np.array_equal(np.broadcast_to(out[nl.outputs["comp"][0]], (300,)), (x <= y).astype(np.int32))


def test_levelizer_depths():
    """All independent gates of a depth share a level: the ripple-carry adder needs 2 levels per bit
    (AND(carry, x^y) then XOR) after one level for x^y / x&y; linear gates add no level."""
    nl = circuits.full_adder(16)
    lv, n = nl.levelize()
    assert n == 1 + 2 * 16
    g = nl.to_array()
    assert np.all(lv[g[:, 0] == int(GateOp.CONST0)] == 0)
    assert (lv == 1).sum() == 32                # x^y and x&y of every bit position run in the first launch
    nlm = circuits.full_adder_mux(16)
    assert nlm.levelize()[1] == 1 + 16
    assert circuits.comparison_mux(16).levelize()[1] == 1 + 16


def test_replicate_offsets_wires():
    nl = circuits.full_adder(4)
    g, wires = nl.replicate(3)
    base = nl.to_array()
    assert wires == 3 * nl.num_wires and g.shape == (3 * len(base), 5)
    second = g[len(base):2 * len(base)]
    assert np.array_equal(second[:, 0], base[:, 0])
    assert np.array_equal(second[:, 1], base[:, 1] + nl.num_wires)
    assert np.array_equal(second[:, 2] >= 0, base[:, 2] >= 0)
    # levels are unchanged by replication
    lv = np.zeros(len(g), dtype=np.int32)
    n = C.c_int32()
    lib.check(lib.load().b200_circuit_levelize(g.ctypes.data, len(g), wires, lv.ctypes.data, C.byref(n)))
    assert n.value == nl.levelize()[1]


def test_levelizer_rejects_bad_netlists():
    L = lib.load()
    n = C.c_int32()

    def run(g, wires):
        g = np.asarray(g, dtype=np.int32)
        return L.b200_circuit_levelize(g.ctypes.data, len(g), wires, None, C.byref(n))

    assert run([[0, 2, 0, 1, -1]], 3) == 0
    assert run([[0, 2, 0, 5, -1]], 3) != 0            # input wire out of range
    assert run([[0, 2, 0, 1, -1], [0, 2, 0, 1, -1]], 3) != 0   # wire written twice
    assert run([[0, 3, 2, 1, -1], [0, 2, 0, 1, -1]], 4) != 0   # not topological
    assert run([[99, 2, 0, 1, -1]], 3) != 0           # invalid op


def test_expand_gates_descriptors():
    """MUX expands to two rotations (-1/8 + a + b, -1/8 - a + c) and one key switch with +1/8
    (boot-gates.cpp:341-363); NAND to 1/8 - a - b (boot-gates.cpp:42)."""
    g = np.array([[int(GateOp.NAND), 3, 0, 1, -1], [int(GateOp.MUX), 4, 0, 1, 2]], dtype=np.int32)
    rot = np.zeros((3, 4), dtype=np.int32)
    ks = np.zeros((2, 4), dtype=np.int32)
    nr, nk = C.c_int64(), C.c_int64()
    lib.check(lib.load().b200_expand_gates(g.ctypes.data, 2, rot.ctypes.data, C.byref(nr), ks.ctypes.data, C.byref(nk)))
    assert (nr.value, nk.value) == (3, 2)

    def s(v):
        lo, hi = v & 0xffff, (v >> 16) & 0xffff
        return (lo - 65536 if lo >= 32768 else lo, hi - 65536 if hi >= 32768 else hi)

    assert list(rot[0, :2]) == [0, 1] and s(int(rot[0, 2])) == (-1, -1) and rot[0, 3] == 1 << 29
    assert list(rot[1, :2]) == [0, 1] and s(int(rot[1, 2])) == (1, 1) and rot[1, 3] == -(1 << 29)
    assert list(rot[2, :2]) == [0, 2] and s(int(rot[2, 2])) == (-1, 1) and rot[2, 3] == -(1 << 29)
    assert list(ks[0]) == [0, -1, 0, 3]
    assert list(ks[1]) == [1, 2, 1 << 29, 4]
