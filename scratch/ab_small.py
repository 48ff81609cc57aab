"""small/medium batches: blind-rotation time per kernel choice"""
import os, sys
import numpy as np, torch
sys.path.insert(0, ".")
from eoc_tfhe_b200 import CloudKey, PARAMS_80, GateOp, set_blind_rotate_kernel
P = PARAMS_80
rng = np.random.default_rng(0)
bk = rng.integers(-2**31, 2**31, size=P.bk_shape, dtype=np.int64).astype(np.int32)
ks = rng.integers(-2**31, 2**31, size=P.ks_shape, dtype=np.int64).astype(np.int32)
ck = CloudKey.from_arrays(P, bk, ks, device=0)
for count in (296, 444, 592, 600, 888, 1184):
    a = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda")
    out = torch.empty_like(a)
    res = []
    for v in ("auto", "v2", "tm", "tmacc"):
        set_blind_rotate_kernel(v)
        for _ in range(2): ck.gate(GateOp.NAND, a, a, out=out)
        torch.cuda.synchronize(); ck.timing_drain()
        for _ in range(5): ck.gate(GateOp.NAND, a, a, out=out)
        torch.cuda.synchronize(); t = ck.timing_drain()
        res.append("%s %.3f" % (v, t["blind_rotate"] / t["calls"]))
    print(count, "gates, B200_TM_G=%s:" % os.environ.get("B200_TM_G", "-"), " | ".join(res), flush=True)
