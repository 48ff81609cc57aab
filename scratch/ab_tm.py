"""A/B: blind-rotation kernels v2 / tm on N rotations (timing from the library's CUDA events)."""
import os
import sys
import numpy as np, torch
sys.path.insert(0, ".")
from eoc_tfhe_b200 import CloudKey, PARAMS_80, PARAMS_128, GateOp, set_blind_rotate_kernel
lam = int(sys.argv[1]) if len(sys.argv) > 1 else 80
counts = [int(c) for c in sys.argv[2:]] or [8880, 65536]
P = PARAMS_80 if lam == 80 else PARAMS_128
rng = np.random.default_rng(0)
bk = rng.integers(-2**31, 2**31, size=P.bk_shape, dtype=np.int64).astype(np.int32)
ks = rng.integers(-2**31, 2**31, size=P.ks_shape, dtype=np.int64).astype(np.int32)
ck = CloudKey.from_arrays(P, bk, ks, device=0)
for count in counts:
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    a = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda", generator=g)
    b = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda", generator=g)
    outs = {}
    variants = os.environ.get("AB_VARIANTS", "v2,tm,tmacc").split(",")
    for v in variants:
        set_blind_rotate_kernel(v)
        out = torch.empty_like(a)
        for _ in range(2): ck.gate(GateOp.NAND, a, b, out=out)
        torch.cuda.synchronize(); ck.timing_drain()
        for _ in range(3): ck.gate(GateOp.NAND, a, b, out=out)
        torch.cuda.synchronize()
        tm = ck.timing_drain()
        outs[v] = out.clone()
        print("P%d %s %d items: blind_rotate %.3f ms, keyswitch %.3f ms" % (lam, v, count, tm["blind_rotate"] / tm["calls"], tm["keyswitch"] / tm["calls"]), flush=True)
    print("   equal to %s:" % variants[0], {v: bool(torch.equal(outs[variants[0]], outs[v])) for v in variants[1:]}, flush=True)
