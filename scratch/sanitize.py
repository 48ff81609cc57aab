"""Small run of every default kernel path for compute-sanitizer (n = 8 fixture: 8 CMux steps per rotation).
usage: python scratch/sanitize.py [sizes]   (compute-sanitizer is closed on the GPU pool; this is the small case to run by hand)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from eoc_tfhe_b200 import CloudKey, GateOp, Params

g = dict(np.load(os.path.join(ROOT, "tests", "golden", "small80.npz")))
n, N, k, l, Bgbit, t, bb = (int(v) for v in g["params"])
P = Params(n=n, N=N, k=k, l=l, Bgbit=Bgbit, ks_t=t, ks_basebit=bb)
ck = CloudKey.from_arrays(P, g["bk"], g["ks"], ks_variance=float(g["ks_variance"]), device=0)

def bits(ct):
    ph = ct[:, n].astype(np.int64) - ct[:, :n].astype(np.int64) @ g["lwe_key"].astype(np.int64)
    return ((ph & 0xFFFFFFFF).astype(np.uint32).astype(np.int32) > 0).astype(np.int32)

a, b, c = g["ct_a"], g["ct_b"], g["ct_c"]
sizes = [int(v) for v in sys.argv[1].split(",")] if len(sys.argv) > 1 else [66, 300, 900, 3000]
for count in sizes:   # latency kernel / four per CTA / six per CTA, balanced / streaming key switch
    rep = (count + len(a) - 1) // len(a)
    A, B, Cc = (np.tile(x, (rep, 1))[:count] for x in (a, b, c))
    out = ck.gate(GateOp.NAND, A, B)
    assert np.array_equal(bits(out), np.tile(1 - (g["bits_a"] & g["bits_b"]), rep)[:count]), count
    out = ck.gate(GateOp.MUX, A, B, Cc)
    assert np.array_equal(bits(out), np.tile(np.where(g["bits_a"] > 0, g["bits_b"], g["bits_c"]), rep)[:count]), count
    print("ok", count, flush=True)
ck.close()
