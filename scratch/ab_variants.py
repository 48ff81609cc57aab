"""A/B of the blind-rotation kernel variants on one random batch: outputs compared bit-for-bit with v2,
kernel time per variant (CUDA events inside the library).
usage: python scratch/ab_variants.py [count] [80|128] [variants, comma separated]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from eoc_tfhe_b200 import CloudKey, PARAMS_80, PARAMS_128, GateOp, set_blind_rotate_kernel
count = int(sys.argv[1]) if len(sys.argv) > 1 else 4736
P = PARAMS_128 if (len(sys.argv) > 2 and sys.argv[2] == "128") else PARAMS_80
variants = [int(v) for v in sys.argv[3].split(",")] if len(sys.argv) > 3 else [0, 2, 3, 4, 5, 6]
rng = np.random.default_rng(0)
bk = rng.integers(-2**31, 2**31, size=P.bk_shape, dtype=np.int64).astype(np.int32)
ks = rng.integers(-2**31, 2**31, size=P.ks_shape, dtype=np.int64).astype(np.int32)
ck = CloudKey.from_arrays(P, bk, ks)
a = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda")
b = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda")
a[5] = 0; b[5] = 0          # a ciphertext whose bara_i are all zero (every step skipped)
a[6, :100] = 0; b[6, :100] = 0
# extracted samples (before key switch) on a small batch: exact difference between kernels
xs = rng.integers(-2**31, 2**31, size=(148 * 4 + 5, P.n + 1), dtype=np.int64).astype(np.int32)
uref = None
for v in variants:
    set_blind_rotate_kernel(v)
    u = ck.bootstrap_woks(xs)
    if uref is None: uref = u
    d = (u.astype(np.int64) - uref.astype(np.int64) + 2**31) % 2**32 - 2**31
    print("variant %d: woks max |u - u_first| = %d LSB, rms %.2f" % (v, int(np.abs(d).max()), float(np.sqrt((d.astype(np.float64)**2).mean()))), flush=True)
ref = None
for v in variants:
    set_blind_rotate_kernel(v)
    out = torch.empty_like(a)
    best = 1e9
    for it in range(3):
        ck.gate(GateOp.NAND, a, b, out=out)
        torch.cuda.synchronize()
        t = ck.timing_drain()
        best = min(best, t["blind_rotate"])
    o = out.cpu().numpy()
    if ref is None: ref = o
    same = bool(np.array_equal(o, ref))
    print("variant %d: %d items, blind_rotate %.3f ms, keyswitch %.3f ms, identical to first: %s, checksum %d, max |diff| %d" % (v, count, best, t["keyswitch"], same, int(o.astype(np.int64).sum()), int(np.abs(o.astype(np.int64) - ref.astype(np.int64)).max())), flush=True)
    if not same:
        bad = np.nonzero((o != ref).any(axis=1))[0]
        print("   differing rows:", len(bad), bad[:10])
