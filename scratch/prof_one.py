"""One wave-sized NAND batch on random data: kernel times of the last iterations (for ncu and quick A/B runs).
usage: python scratch/prof_one.py [count] [80|128] [iters]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from eoc_tfhe_b200 import CloudKey, PARAMS_80, PARAMS_128, GateOp
P = PARAMS_128 if (len(sys.argv) > 2 and sys.argv[2] == "128") else PARAMS_80
count = int(sys.argv[1]) if len(sys.argv) > 1 else 2368
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 3
rng = np.random.default_rng(0)
bk = rng.integers(-2**31, 2**31, size=P.bk_shape, dtype=np.int64).astype(np.int32)
ks = rng.integers(-2**31, 2**31, size=P.ks_shape, dtype=np.int64).astype(np.int32)
ck = CloudKey.from_arrays(P, bk, ks)
a = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda")
b = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda")
out = torch.empty_like(a)
for it in range(iters):
    ck.gate(GateOp.NAND, a, b, out=out)
    torch.cuda.synchronize()
    t = ck.timing_drain()
    print(os.environ.get("B200_BR_KERNEL", "v2"), count, "items: blind_rotate %.3f ms, keyswitch %.3f ms" % (t["blind_rotate"], t["keyswitch"]), flush=True)
print("checksum", int(out.to(torch.int64).sum().item()))
