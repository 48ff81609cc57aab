"""One key switch of a large batch for ncu.  usage: python scratch/prof_ks.py [count] [80|128]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from eoc_tfhe_b200 import CloudKey, PARAMS_80, PARAMS_128, load
import ctypes as C
count = int(sys.argv[1]) if len(sys.argv) > 1 else 9472
P = PARAMS_128 if (len(sys.argv) > 2 and sys.argv[2] == "128") else PARAMS_80
rng = np.random.default_rng(0)
bk = np.zeros(P.bk_shape, dtype=np.int32)
ks = rng.integers(-2**31, 2**31, size=P.ks_shape, dtype=np.int64).astype(np.int32)
ck = CloudKey.from_arrays(P, bk, ks)
u = torch.randint(-2**31, 2**31 - 1, (count, P.N + 1), dtype=torch.int32, device="cuda")
out = torch.empty((count, P.n + 1), dtype=torch.int32, device="cuda")
lib = load()
for it in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    rc = lib.b200_keyswitch_dev(ck._h, C.c_void_p(out.data_ptr()), C.c_void_p(u.data_ptr()), C.c_int64(count), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    e1.record(); torch.cuda.synchronize()
    print("keyswitch", count, "samples rc", rc, "%.3f ms" % e0.elapsed_time(e1), flush=True)
print("checksum", int(out.to(torch.int64).sum().item()))
