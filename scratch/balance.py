"""Blind-rotation time of batch sizes that do not fill whole waves, balanced grid on/off (B200_BALANCE=0|1).
usage: B200_BALANCE=0 python scratch/balance.py"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from eoc_tfhe_b200 import CloudKey, PARAMS_80, GateOp
P = PARAMS_80
rng = np.random.default_rng(0)
bk = rng.integers(-2**31, 2**31, size=P.bk_shape, dtype=np.int64).astype(np.int32)
ks = rng.integers(-2**31, 2**31, size=P.ks_shape, dtype=np.int64).astype(np.int32)
ck = CloudKey.from_arrays(P, bk, ks)
ref = {}
sizes = [int(v) for v in sys.argv[1].split(',')] if len(sys.argv) > 1 else (600, 889, 1000, 1332, 1500, 2000, 4096, 8192, 12288, 65536)
torch.manual_seed(1)
for count in sizes:
    a = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda")
    out = torch.empty_like(a)
    best = (1e9, 0)
    for it in range(3):
        ck.gate(GateOp.NAND, a, a, out=out)
        torch.cuda.synchronize()
        t = ck.timing_drain()
        if t["blind_rotate"] < best[0]: best = (t["blind_rotate"], t["keyswitch"])
    print("%6d gates: blind rotate %.3f ms, key switch %.3f ms; checksum %d" % (count, best[0], best[1], int(out.to(torch.int64).sum().item())), flush=True)
