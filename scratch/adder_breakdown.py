"""Where the 4096-pair 16-bit adder's time goes: host pool path vs device-resident pool vs kernel sums."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from eoc_tfhe_b200 import CloudKey, PARAMS_80, circuits
P = PARAMS_80
rng = np.random.default_rng(0)
bk = rng.integers(-2**31, 2**31, size=P.bk_shape, dtype=np.int64).astype(np.int32)
ks = rng.integers(-2**31, 2**31, size=P.ks_shape, dtype=np.int64).astype(np.int32)
ck = CloudKey.from_arrays(P, bk, ks)
nb, count = 16, 4096
nl = circuits.full_adder(nb)
cx = rng.integers(-2**31, 2**31, size=(count, nb, P.n + 1), dtype=np.int64).astype(np.int32)
cy = rng.integers(-2**31, 2**31, size=(count, nb, P.n + 1), dtype=np.int64).astype(np.int32)
circuits.run_batch(ck, nl, {"x": cx[:64], "y": cy[:64]}, 64)
ck.timing_drain()
for it in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    outs, levels = circuits.run_batch(ck, nl, {"x": cx, "y": cy}, count)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    t = ck.timing_drain()
    print("host pool: %.3f s wall; kernels: blind rotate %.1f ms, key switch %.1f ms, calls %s" % (dt, t["blind_rotate"], t["keyswitch"], t.get("calls")), flush=True)
dx, dy = torch.from_numpy(cx).cuda(), torch.from_numpy(cy).cuda()
for it in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    outs, levels = circuits.run_batch(ck, nl, {"x": dx, "y": dy}, count)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    t = ck.timing_drain()
    print("device pool: %.3f s wall; kernels: blind rotate %.1f ms, key switch %.1f ms" % (dt, t["blind_rotate"], t["keyswitch"]), flush=True)
t0 = time.perf_counter(); g, nw = nl.replicate(count); print("replicate: %.3f s" % (time.perf_counter() - t0))
