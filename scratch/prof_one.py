import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from eoc_tfhe_b200 import CloudKey, PARAMS_80, PARAMS_128, GateOp
P = PARAMS_128 if (len(sys.argv) > 2 and sys.argv[2] == "128") else PARAMS_80
count = int(sys.argv[1]) if len(sys.argv) > 1 else 888
rng = np.random.default_rng(0)
bk = rng.integers(-2**31, 2**31, size=P.bk_shape, dtype=np.int64).astype(np.int32)
ks = rng.integers(-2**31, 2**31, size=P.ks_shape, dtype=np.int64).astype(np.int32)
ck = CloudKey.from_arrays(P, bk, ks)
a = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda")
b = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda")
out = torch.empty_like(a)
for it in range(2):
    ck.gate(GateOp.NAND, a, b, out=out)
    torch.cuda.synchronize()
print(ck.last_timing_ms())
