"""b200_gate_batch_dev_noalloc (include/b200tfhe.h): the allocation-free variant of b200_gate_batch_dev, whose extracted
samples go through a static device array instead of the key's cudaMalloc-grown workspace.  No counterpart in the
reference, so the oracle here is the existing entry point: same kernels,
same dispatch for the same batch size, hence the SAME words -- integer outputs, compared exactly.

The file sorts last on purpose (a failure here must not hide other tests) and needs nothing
but the library and a GPU: random key material, no reference build, no state from other tests.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_ROTATIONS = 4096          # B200_NOALLOC_MAX_ROTATIONS


@pytest.fixture(scope="module")
def setup():
    import torch
    from eoc_tfhe_b200 import CloudKey, PARAMS_80
    P = PARAMS_80
    rng = np.random.default_rng(20261105)
    bk = rng.integers(-2**31, 2**31, size=P.bk_shape, dtype=np.int64).astype(np.int32)
    ks = rng.integers(-2**31, 2**31, size=P.ks_shape, dtype=np.int64).astype(np.int32)
    ck = CloudKey.from_arrays(P, bk, ks, device=0)
    g = torch.Generator(device="cuda")
    g.manual_seed(7)

    def batch(count):
        return torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda", generator=g)
    yield ck, P, batch
    ck.close()


def test_two_calls_in_a_row_equal_the_allocating_entry(setup):
    """Two calls back to back on one stream share the static array: the second (smaller, other inputs, other gate)
    must not see anything of the first.  Sizes cover the wide kernels + streaming key switch (3,300 NAND), the
    four-per-CTA kernel + gather key switch (500 XOR), MUX (two rotations per gate, paired key switch) and the
    cluster latency kernel (1 gate)."""
    import torch
    from eoc_tfhe_b200 import GateOp
    ck, P, batch = setup
    a, b, c = batch(3300), batch(3300), batch(3300)
    plan = [(GateOp.NAND, 3300), (GateOp.XOR, 500), (GateOp.MUX, 1200), (GateOp.NAND, 1), (GateOp.NAND, 3300)]
    want = [ck.gate(op, a[:n], b[:n], c[:n] if op == GateOp.MUX else None) for op, n in plan]
    torch.cuda.synchronize()
    outs = [torch.full((n, P.n + 1), 0x5A5A5A5A, dtype=torch.int32, device="cuda") for _, n in plan]
    before, free_before = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
    for (op, n), out in zip(plan, outs):        # enqueued back to back, no synchronisation in between
        ck.gate_noalloc(op, a[:n], b[:n], c[:n] if op == GateOp.MUX else None, out=out)
    torch.cuda.synchronize()
    # neither torch's allocator nor the driver handed out memory during the calls (kernels already loaded above)
    assert torch.cuda.memory_allocated() == before and torch.cuda.mem_get_info()[0] >= free_before
    for (op, n), out, w in zip(plan, outs, want):
        assert torch.equal(out, w), (op, n)


def test_linear_gates_and_limits(setup):
    import torch
    from eoc_tfhe_b200 import GateOp, B200Error
    ck, P, batch = setup
    a = batch(64)
    out = torch.empty_like(a)
    ck.gate_noalloc(GateOp.NOT, a, out=out)
    assert torch.equal(out, ck.gate(GateOp.NOT, a))
    # exactly the maximum is accepted ...
    big = batch(MAX_ROTATIONS)
    out = torch.empty_like(big)
    ck.gate_noalloc(GateOp.AND, big, big, out=out)
    assert torch.equal(out, ck.gate(GateOp.AND, big, big))
    # ... one more is rejected with an error that names the limit, and nothing is written
    over = batch(MAX_ROTATIONS + 1)
    out = torch.full_like(over, 7)
    with pytest.raises(B200Error, match="4096"):
        ck.gate_noalloc(GateOp.AND, over, over, out=out)
    half = batch(MAX_ROTATIONS // 2 + 1)          # MUX: two rotations per gate
    with pytest.raises(B200Error, match="B200_NOALLOC_MAX_ROTATIONS"):
        ck.gate_noalloc(GateOp.MUX, half, half, half, out=torch.empty_like(half))
    torch.cuda.synchronize()
    assert int((out != 7).sum()) == 0


def test_raw_c_abi_call(setup):
    """Straight through ctypes, as a C caller would: plain pointers and sizes, NULL stream."""
    import torch
    from eoc_tfhe_b200 import GateOp, lib
    ck, P, batch = setup
    L = lib.load()
    a, b = batch(300), batch(300)
    out = torch.zeros_like(a)
    rc = L.b200_gate_batch_dev_noalloc(ck._h, int(GateOp.OR), C.c_void_p(out.data_ptr()), C.c_void_p(a.data_ptr()),
                                       C.c_void_p(b.data_ptr()), None, 300, None)
    assert rc == 0, L.b200_last_error()
    torch.cuda.synchronize()
    assert torch.equal(out, ck.gate(GateOp.OR, a, b))
    assert L.b200_gate_batch_dev_noalloc(ck._h, int(GateOp.OR), None, None, None, None, 5, None) == 1
