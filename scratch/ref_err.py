import sys; sys.path.insert(0,'tests')
import numpy as np, reflib
for variant in ("spqlios-fma","nayuki-portable"):
    ref = reflib.RefLib(variant)
    rk = ref.keygen(80, seed=42)
    rng = np.random.default_rng(3)
    acc = rng.integers(-2**31, 2**31, size=(8, 2, rk.N), dtype=np.int64).astype(np.int32)
    mx=[]; 
    for g in range(8):
        f = rk.extern_product_fft(acc[g], 0); e = rk.extern_product_exact(acc[g], 0)
        d = (f.astype(np.int64)-e.astype(np.int64)+2**31)%2**32-2**31
        mx.append((np.abs(d).max(), d.mean(), d.std()))
    print(variant, mx[:4])
