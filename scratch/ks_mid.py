import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, ctypes as C
from eoc_tfhe_b200 import CloudKey, PARAMS_80, load, set_keyswitch_kernel
P = PARAMS_80
rng = np.random.default_rng(0)
ck = CloudKey.from_arrays(P, np.zeros(P.bk_shape, dtype=np.int32), rng.integers(-2**31, 2**31, size=P.ks_shape, dtype=np.int64).astype(np.int32))
lib = load()
for count in (2368, 4096, 8192, 9472, 16384):
    u = torch.randint(-2**31, 2**31 - 1, (count, P.N + 1), dtype=torch.int32, device="cuda")
    res = {}
    for v in (1, 0):
        set_keyswitch_kernel(v)
        out = torch.zeros((count, P.n + 1), dtype=torch.int32, device="cuda")
        best = 1e9
        for it in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            lib.b200_keyswitch_dev(ck._h, C.c_void_p(out.data_ptr()), C.c_void_p(u.data_ptr()), C.c_int64(count), C.c_void_p(torch.cuda.current_stream().cuda_stream))
            e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        res[v] = (best, out.clone())
    print("%6d samples: gather %.3f ms, auto %.3f ms, identical %s" % (count, res[1][0], res[0][0], bool(torch.equal(res[1][1], res[0][1]))), flush=True)
