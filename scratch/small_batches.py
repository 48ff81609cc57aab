"""Latency / throughput of small gate batches (the one-ciphertext-per-CTA path and the drop-in bootsNAND).
usage: python scratch/small_batches.py"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from eoc_tfhe_b200 import CloudKey, PARAMS_80, GateOp, set_blind_rotate_kernel
if len(sys.argv) > 1: set_blind_rotate_kernel(int(sys.argv[1]))
P = PARAMS_80
rng = np.random.default_rng(0)
bk = rng.integers(-2**31, 2**31, size=P.bk_shape, dtype=np.int64).astype(np.int32)
ks = rng.integers(-2**31, 2**31, size=P.ks_shape, dtype=np.int64).astype(np.int32)
ck = CloudKey.from_arrays(P, bk, ks)
for count in (1, 16, 148, 296, 444, 591, 592, 1184, 4096, 8192):
    a = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda")
    out = torch.empty_like(a)
    best = (1e9, 0, 0)
    for it in range(4):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        ck.gate(GateOp.NAND, a, a, out=out)
        torch.cuda.synchronize(); wall = (time.perf_counter() - t0) * 1e3
        t = ck.timing_drain()
        if wall < best[0]: best = (wall, t["blind_rotate"], t["keyswitch"])
    print("%5d gates: wall %.3f ms, blind rotate %.3f ms, key switch %.3f ms -> %.0f gates/s" % (count, best[0], best[1], best[2], count / best[0] * 1e3), flush=True)
