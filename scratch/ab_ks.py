"""A/B of the streaming key switch's row-choice modes (B200_KS_MODE, read once per process): time (library CUDA
events inside a NAND batch) and output hash."""
import hashlib
import os
import sys
import numpy as np, torch
sys.path.insert(0, ".")
from eoc_tfhe_b200 import CloudKey, PARAMS_80, PARAMS_128, GateOp
P = PARAMS_128 if (len(sys.argv) > 2 and sys.argv[2] == "128") else PARAMS_80
count = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
rng = np.random.default_rng(0)
bk = rng.integers(-2**31, 2**31, size=P.bk_shape, dtype=np.int64).astype(np.int32)
ks = rng.integers(-2**31, 2**31, size=P.ks_shape, dtype=np.int64).astype(np.int32)
ck = CloudKey.from_arrays(P, bk, ks, device=0)
g = torch.Generator(device="cuda"); g.manual_seed(1)
a = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda", generator=g)
b = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda", generator=g)
out = torch.empty_like(a)
ck.gate(GateOp.NAND, a, b, out=out)
torch.cuda.synchronize(); ck.timing_drain()
for _ in range(3): ck.gate(GateOp.NAND, a, b, out=out)
torch.cuda.synchronize()
tm = ck.timing_drain()
print("mode %s: %d gates, keyswitch %.3f ms (blind rotation %.2f ms), sha %s" % (os.environ.get("B200_KS_MODE", "0"), count,
      tm["keyswitch"] / tm["calls"], tm["blind_rotate"] / tm["calls"], hashlib.sha256(out.cpu().numpy().tobytes()).hexdigest()[:16]), flush=True)
