import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from eoc_tfhe_b200 import CloudKey, PARAMS_80, PARAMS_128, GateOp
for P, name in ((PARAMS_80, "P80"), (PARAMS_128, "P128")):
    rng = np.random.default_rng(0)
    bk = rng.integers(-2**31, 2**31, size=P.bk_shape, dtype=np.int64).astype(np.int32)
    ks = rng.integers(-2**31, 2**31, size=P.ks_shape, dtype=np.int64).astype(np.int32)
    t = time.time(); ck = CloudKey.from_arrays(P, bk, ks); print(name, "key upload+convert s", time.time() - t)
    for count in (1, 148*6, 148*6*4, 16384):
        a = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda")
        b = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda")
        out = torch.empty_like(a)
        for it in range(2):
            torch.cuda.synchronize(); t = time.time()
            ck.gate(GateOp.NAND, a, b, out=out)
            torch.cuda.synchronize(); dt = time.time() - t
        tm = ck.last_timing_ms()
        print(name, "count", count, "wall ms %.2f" % (dt * 1e3), "gates/s %.0f" % (count / dt), tm)
    ck.close()
