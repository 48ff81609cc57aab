"""The N>1 host logic on CPU: world_size 2 over gloo.  Ranks slice a gate level, evaluate their slice
with the CPU oracle standing in for the GPU (tests may use oracle/ as the checker), all-gather, and the
result equals the unsharded evaluation bit for bit; independent circuit instances shard with no per-level
traffic."""
import os
import socket
import sys

import numpy as np
import pytest

from eoc_tfhe_b200 import shard

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_slices_cover_and_balance():
    for count in (0, 1, 7, 64, 65536, 2 ** 20 + 3):
        for world in (1, 2, 3, 4, 8):
            cuts = [shard.slice_for_rank(count, world, r) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == count
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
            sizes = [e - s for s, e in cuts]
            assert max(sizes) - min(sizes) <= 1 and max(sizes) == shard.padded_slice_len(count, world)
    with pytest.raises(ValueError):
        shard.slice_for_rank(4, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, result_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch
    import torch.distributed as dist
    import portlib
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = dict(np.load(os.path.join(ROOT, "tests", "golden", "small80.npz")))
    n, N, k, l, Bgbit, t, bb = (int(v) for v in g["params"])
    pk = portlib.PortKey(n, l, Bgbit, g["bk"], g["ks"])
    rng = np.random.default_rng(5)                       # same inputs on every rank
    count = 7                                            # ragged: 4 + 3
    a = torch.from_numpy(np.tile(g["ct_a"], (2, 1))[:count].copy())
    b = torch.from_numpy(np.tile(g["ct_b"], (2, 1))[:count].copy())
    c = torch.from_numpy(np.tile(g["ct_c"], (2, 1))[:count].copy())
    calls = []

    def evaluate(op, xa, xb, xc):
        calls.append(xa.shape[0])
        out = pk.gate(op, xa.numpy(), xb.numpy(), None if xc is None else xc.numpy())
        return torch.from_numpy(out)

    nand = shard.sharded_gate(evaluate, "NAND", a, b)
    mux = shard.sharded_gate(evaluate, "MUX", a, b, c)
    s, e = shard.slice_for_rank(count, world, rank)
    assert calls == [e - s, e - s]
    full_nand = pk.gate("NAND", a.numpy(), b.numpy())
    full_mux = pk.gate("MUX", a.numpy(), b.numpy(), c.numpy())
    ok = np.array_equal(nand.numpy(), full_nand) and np.array_equal(mux.numpy(), full_mux)

    # two-level dependency: level 2 consumes level 1's gathered outputs on every rank
    lvl2 = shard.sharded_gate(evaluate, "XOR", nand, mux)
    ok = ok and np.array_equal(lvl2.numpy(), pk.gate("XOR", full_nand, full_mux))

    # independent instances: 5 one-bit "circuits" x AND y, sharded by instance, gathered once
    def run_local(inp, n_local):
        return {"z": torch.from_numpy(pk.gate("AND", inp["x"][:, 0].numpy(), inp["y"][:, 0].numpy()))[:, None, :]}

    inst = shard.sharded_instances(run_local, {"x": a[:5, None, :], "y": b[:5, None, :]}, 5)
    ok = ok and np.array_equal(inst["z"][:, 0].numpy(), pk.gate("AND", a[:5].numpy(), b[:5].numpy()))
    pk.close()
    dist.barrier()
    dist.destroy_process_group()
    with open(os.path.join(result_dir, "rank%d" % rank), "w") as f:
        f.write("ok" if ok else "mismatch")


def test_sharded_levels_world2_gloo(tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert [open(tmp_path / ("rank%d" % r)).read() for r in range(2)] == ["ok", "ok"]
