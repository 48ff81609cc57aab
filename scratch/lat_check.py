"""Latency kernel vs the throughput kernel on small batches: extracted samples identical, timings."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from eoc_tfhe_b200 import CloudKey, PARAMS_80, PARAMS_128, GateOp
P = PARAMS_128 if (len(sys.argv) > 1 and sys.argv[1] == "128") else PARAMS_80
rng = np.random.default_rng(0)
bk = rng.integers(-2**31, 2**31, size=P.bk_shape, dtype=np.int64).astype(np.int32)
ks = rng.integers(-2**31, 2**31, size=P.ks_shape, dtype=np.int64).astype(np.int32)
ck = CloudKey.from_arrays(P, bk, ks)
x = rng.integers(-2**31, 2**31, size=(300, P.n + 1), dtype=np.int64).astype(np.int32)
x[3] = 0
big = ck.bootstrap_woks(x)            # 300 items: one-group-per-CTA throughput path
small = np.concatenate([ck.bootstrap_woks(x[i:i + 100]) for i in range(0, 300, 100)])   # <= 148: latency kernel
print("identical:", np.array_equal(big, small), "lat kernel env:", os.environ.get("B200_LAT", "1"))
for count in (1, 16, 148):
    a = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda")
    out = torch.empty_like(a)
    best = (1e9, 0, 0)
    for it in range(5):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        ck.gate(GateOp.NAND, a, a, out=out)
        torch.cuda.synchronize(); wall = (time.perf_counter() - t0) * 1e3
        t = ck.timing_drain()
        if wall < best[0]: best = (wall, t["blind_rotate"], t["keyswitch"])
    print("%5d gates: wall %.3f ms, blind rotate %.3f ms, key switch %.3f ms" % (count, best[0], best[1], best[2]), flush=True)
