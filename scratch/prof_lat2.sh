# phase clocks of the cluster latency kernel.  Build first (here, no GPU needed):
#   cd eoc_tfhe_b200/csrc && nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-fvisibility=default,-O2 \
#      -DB200_LAT2_PROFILE -x cu -c kernels.cu -o ../../scratch/kernels_prof.o && nvcc -shared -o ../../scratch/libb200tfhe_prof.so \
#      ../../scratch/kernels_prof.o abi.o circuit.o wire.o eoc_abi.o -gencode arch=compute_100a,code=sm_100a -Xlinker -Bsymbolic -lcudart_static -lpthread -ldl -lrt
for m in 1 3; do B200TFHE_LIB=$PWD/scratch/libb200tfhe_prof.so B200_LAT=$m python - <<'PY'
import sys, os
import numpy as np
sys.path.insert(0, ".")
from eoc_tfhe_b200 import CloudKey, PARAMS_80, PARAMS_128, GateOp
for P in (PARAMS_80, PARAMS_128):
    rng = np.random.default_rng(0)
    bk = rng.integers(-2**31, 2**31, size=P.bk_shape, dtype=np.int64).astype(np.int32)
    ks = rng.integers(-2**31, 2**31, size=P.ks_shape, dtype=np.int64).astype(np.int32)
    ck = CloudKey.from_arrays(P, bk, ks, device=0)
    a = rng.integers(-2**31, 2**31, size=(1, P.n + 1), dtype=np.int64).astype(np.int32)
    print("B200_LAT", os.environ["B200_LAT"], "n =", P.n, flush=True)
    for _ in range(2): ck.gate(GateOp.NAND, a, a)
PY
done
