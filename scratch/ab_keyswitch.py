"""A/B of the two key-switch kernels: outputs and non-zero-digit counts compared bit-for-bit, kernel times.
usage: python scratch/ab_keyswitch.py [count] [80|128]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from eoc_tfhe_b200 import CloudKey, PARAMS_80, PARAMS_128, GateOp, set_keyswitch_kernel
count = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
P = PARAMS_128 if (len(sys.argv) > 2 and sys.argv[2] == "128") else PARAMS_80
rng = np.random.default_rng(0)
bk = rng.integers(-2**31, 2**31, size=P.bk_shape, dtype=np.int64).astype(np.int32)
ks = rng.integers(-2**31, 2**31, size=P.ks_shape, dtype=np.int64).astype(np.int32)
ck = CloudKey.from_arrays(P, bk, ks)
# host entry (with non-zero digit counts) on a ragged count
u = rng.integers(-2**31, 2**31, size=(9472 + 37, P.N + 1), dtype=np.int64).astype(np.int32)
res = {}
for v in (1, 2):
    set_keyswitch_kernel(v)
    res[v] = ck.keyswitch(u, with_counts=True)
print("host keyswitch %d samples: outputs identical %s, digit counts identical %s" % (
    len(u), np.array_equal(res[1][0], res[2][0]), np.array_equal(res[1][1], res[2][1])), flush=True)
torch.manual_seed(0)
a = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda")
b = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda")
c = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda")
for op, args in ((GateOp.NAND, (a, b)), (GateOp.MUX, (a[: count // 2], b[: count // 2], c[: count // 2]))):
    outs = {}
    for v in (1, 2):
        set_keyswitch_kernel(v)
        out = torch.empty_like(args[0])
        best = 1e9
        for it in range(2):
            ck.gate(op, *args, out=out)
            torch.cuda.synchronize()
            t = ck.timing_drain()
            best = min(best, t["keyswitch"])
        outs[v] = out.cpu().numpy()
        print("%s x %d, keyswitch kernel %d: %.3f ms (blind rotate %.1f ms)" % (op.name, len(args[0]), v, best, t["blind_rotate"]), flush=True)
    print("   outputs identical:", np.array_equal(outs[1], outs[2]), flush=True)
