"""A/B of the small-batch blind-rotation kernels (B200_LAT: 0 wide kernel, 2 one CTA per ciphertext, 1 cluster of
two CTAs where it fits, 3 the same with the shared-memory transform): output hashes and times."""
import hashlib
import os
import sys
import time
import numpy as np, torch
sys.path.insert(0, ".")
from eoc_tfhe_b200 import CloudKey, PARAMS_80, PARAMS_128, GateOp
lam = int(sys.argv[1]) if len(sys.argv) > 1 else 80
P = PARAMS_80 if lam == 80 else PARAMS_128
rng = np.random.default_rng(0)
bk = rng.integers(-2**31, 2**31, size=P.bk_shape, dtype=np.int64).astype(np.int32)
ks = rng.integers(-2**31, 2**31, size=P.ks_shape, dtype=np.int64).astype(np.int32)
ck = CloudKey.from_arrays(P, bk, ks, device=0)
res = []
for count in (1, 5, 74, 75, 148):
    g = torch.Generator(device="cuda"); g.manual_seed(count)
    a = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda", generator=g)
    b = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda", generator=g)
    out = torch.empty_like(a)
    ck.gate(GateOp.NAND, a, b, out=out)
    torch.cuda.synchronize(); ck.timing_drain()
    for _ in range(5): ck.gate(GateOp.NAND, a, b, out=out)
    torch.cuda.synchronize()
    tm = ck.timing_drain()
    res.append("%d: %.3f ms %s" % (count, tm["blind_rotate"] / tm["calls"], hashlib.sha256(out.cpu().numpy().tobytes()).hexdigest()[:8]))
# host path, one gate
ah = rng.integers(-2**31, 2**31, size=(1, P.n + 1), dtype=np.int64).astype(np.int32)
bh = rng.integers(-2**31, 2**31, size=(1, P.n + 1), dtype=np.int64).astype(np.int32)
for _ in range(5): ck.gate(GateOp.NAND, ah, bh)
t0 = time.perf_counter()
for _ in range(50): o = ck.gate(GateOp.NAND, ah, bh)
dt = (time.perf_counter() - t0) / 50 * 1e3
print("P%d B200_LAT=%s | %s | host single gate %.3f ms %s" % (lam, os.environ.get("B200_LAT", "default"), " | ".join(res), dt,
      hashlib.sha256(np.ascontiguousarray(o).tobytes()).hexdigest()[:8]), flush=True)
