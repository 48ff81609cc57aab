"""eoc_tfhe_b200 -- B200-native TFHE gate bootstrapping behind libtfhe's gate API.

Host-side mirror (Python) of the C ABI in include/b200tfhe.h.  The compute path is the CUDA
library eoc_tfhe_b200/libb200tfhe.so; importing this package never falls back to a CPU
implementation -- `lib.load()` raises if the library is missing.

(The directory is spelled with an underscore because `eoc-tfhe_b200` is not importable.)
"""
from .lib import load, LIB_PATH, B200Error  # noqa: F401
from .engine import (  # noqa: F401
    GateOp, Params, CloudKey, PARAMS_80, PARAMS_128, launch_count, fp64_peak_tflops, int32_add_peak, set_blind_rotate_kernel, set_keyswitch_kernel, experiments_built, fft_roundtrip, poly_mul,
)
from . import circuits, eoc, shard, wire  # noqa: F401
