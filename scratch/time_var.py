import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from eoc_tfhe_b200 import CloudKey, PARAMS_80, PARAMS_128, GateOp
counts = [int(x) for x in sys.argv[1].split(",")]
sets = [(PARAMS_80, "P80")] + ([(PARAMS_128, "P128")] if len(sys.argv) > 2 else [])
for P, name in sets:
    rng = np.random.default_rng(0)
    bk = rng.integers(-2**31, 2**31, size=P.bk_shape, dtype=np.int64).astype(np.int32)
    ks = rng.integers(-2**31, 2**31, size=P.ks_shape, dtype=np.int64).astype(np.int32)
    ck = CloudKey.from_arrays(P, bk, ks)
    for count in counts:
        a = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda")
        b = torch.randint(-2**31, 2**31 - 1, (count, P.n + 1), dtype=torch.int32, device="cuda")
        out = torch.empty_like(a)
        for it in range(2):
            ck.gate(GateOp.NAND, a, b, out=out); torch.cuda.synchronize()
        tm = ck.last_timing_ms()
        print(os.environ.get("B200TFHE_LIB", "default")[-14:], name, "count", count, "br ms %.2f ks ms %.2f" % (tm["blind_rotate"], tm["keyswitch"]), "gates/s %.0f" % (count / (tm["blind_rotate"] + tm["keyswitch"]) * 1e3), flush=True)
    ck.close()
