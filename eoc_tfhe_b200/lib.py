"""ctypes binding of libb200tfhe.so (include/b200tfhe.h). No compute happens in Python."""
import ctypes as C
import os

LIB_PATH = os.environ.get("B200TFHE_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libb200tfhe.so")


class B200Error(RuntimeError):
    pass


class b200_wire_cloudkey_info(C.Structure):
    _fields_ = [("lwe_alpha_min", C.c_double), ("lwe_alpha_max", C.c_double), ("tlwe_alpha_min", C.c_double),
                ("tlwe_alpha_max", C.c_double), ("bk_variance", C.c_double), ("ks_offset", C.c_size_t),
                ("ks_bytes", C.c_size_t), ("bk_offset", C.c_size_t), ("bk_bytes", C.c_size_t), ("total_bytes", C.c_size_t)]


class b200_params(C.Structure):
    _fields_ = [("n", C.c_int32), ("N", C.c_int32), ("k", C.c_int32), ("l", C.c_int32),
                ("Bgbit", C.c_int32), ("ks_t", C.c_int32), ("ks_basebit", C.c_int32),
                ("ks_variance", C.c_double)]


_lib = None

_I32P = C.POINTER(C.c_int32)
_VP = C.c_void_p

# name -> (restype, argtypes); every symbol declared in include/b200tfhe.h
SIGNATURES = {
    "b200_last_error": (C.c_char_p, []),
    "b200_device_count": (C.c_int, []),
    "b200_set_device": (C.c_int, [C.c_int]),
    "b200_cloudkey_upload": (C.c_int, [_VP, C.POINTER(_VP)]),
    "b200_cloudkey_forget": (None, [_VP]),
    "b200_key_create": (C.c_int, [C.POINTER(b200_params), _VP, _VP, C.POINTER(_VP)]),
    "b200_key_create_empty": (C.c_int, [C.POINTER(b200_params), C.POINTER(_VP)]),
    "b200_key_free": (None, [_VP]),
    "b200_key_params": (C.c_int, [_VP, C.POINTER(b200_params)]),
    "b200_key_device_blobs": (C.c_int, [_VP, C.POINTER(_VP), C.POINTER(C.c_size_t), C.POINTER(_VP), C.POINTER(C.c_size_t)]),
    "b200_gate_batch_host": (C.c_int, [_VP, C.c_int, _VP, _VP, _VP, _VP, C.c_int64]),
    "b200_gate_batch_dev": (C.c_int, [_VP, C.c_int, _VP, _VP, _VP, _VP, C.c_int64, _VP]),
    "b200_gate_batch_dev_noalloc": (C.c_int, [_VP, C.c_int, _VP, _VP, _VP, _VP, C.c_int64, _VP]),
    "b200_bootstrap_woks_host": (C.c_int, [_VP, _VP, _VP, C.c_int32, C.c_int64]),
    "b200_keyswitch_host": (C.c_int, [_VP, _VP, _VP, _VP, C.c_int64]),
    "b200_keyswitch_dev": (C.c_int, [_VP, _VP, _VP, C.c_int64, _VP]),
    "b200_cmux_step_host": (C.c_int, [_VP, _VP, C.c_int32, C.c_int32, C.c_int64]),
    "b200_extern_product_host": (C.c_int, [_VP, _VP, C.c_int32, C.c_int64]),
    "b200_circuit_run_host": (C.c_int, [_VP, _VP, C.c_int64, _VP, C.c_int64, _I32P]),
    "b200_circuit_run_dev": (C.c_int, [_VP, _VP, C.c_int64, _VP, C.c_int64, _I32P, _VP]),
    "b200_circuit_run_host_io": (C.c_int, [_VP, _VP, C.c_int64, C.c_int64, _VP, C.c_int64, _VP, _VP, C.c_int64, _VP, _VP]),
    "b200_circuit_levelize": (C.c_int, [_VP, C.c_int64, C.c_int64, _VP, _I32P]),
    "b200_expand_gates": (C.c_int, [_VP, C.c_int64, _VP, C.POINTER(C.c_int64), _VP, C.POINTER(C.c_int64)]),
    "b200_level_run_dev": (C.c_int, [_VP, _VP, _VP, C.c_int64, _VP, C.c_int64, _VP]),
    "b200_last_timing_ms": (C.c_int, [_VP, C.POINTER(C.c_float * 3)]),
    "b200_timing_drain": (C.c_int, [_VP, C.POINTER(C.c_double * 2), C.POINTER(C.c_int32)]),
    "b200_fp64_peak_tflops": (C.c_int, [C.c_int, C.POINTER(C.c_double)]),
    "b200_launch_count": (C.c_int64, []),
    "b200_wire_parse_cloudkey": (C.c_int, [_VP, C.c_size_t, C.POINTER(b200_params), C.POINTER(b200_wire_cloudkey_info)]),
    "b200_key_from_cloudkey_bytes": (C.c_int, [_VP, C.c_size_t, C.POINTER(_VP)]),
    "b200_key_from_cloudkey_file": (C.c_int, [C.c_char_p, C.POINTER(_VP)]),
    "b200_wire_cloudkey_size": (C.c_int64, [C.POINTER(b200_params)]),
    "b200_wire_write_cloudkey": (C.c_int, [_VP, C.c_size_t, C.POINTER(C.c_size_t), C.POINTER(b200_params),
                                           C.POINTER(b200_wire_cloudkey_info), _VP, _VP]),
    "b200_wire_sample_bytes": (C.c_int64, [C.c_int32]),
    "b200_wire_read_samples": (C.c_int, [_VP, C.c_size_t, C.c_int32, C.c_int64, _VP, _VP]),
    "b200_wire_write_samples": (C.c_int, [_VP, C.c_size_t, C.c_int32, C.c_int64, _VP, _VP]),
    "b200_set_blind_rotate_kernel": (C.c_int, [C.c_int]),
    "b200_experiments_built": (C.c_int, []),
    "b200_int32_add_peak": (C.c_int, [C.c_int, C.POINTER(C.c_double)]),
    "b200_multi_create": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_void_p)]),
    "b200_multi_free": (None, [C.c_void_p]),
    "b200_multi_device_count": (C.c_int, [C.c_void_p]),
    "b200_multi_gate_batch_host": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]),
    "b200_fft_roundtrip_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64]),
    "b200_poly_mul_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]),
    "b200_set_keyswitch_kernel": (C.c_int, [C.c_int]),
    "bootsGate_batch": (C.c_int, [C.c_int, _VP, _VP, _VP, _VP, C.c_int32, _VP]),
    # eoc evaluation-side entry points (csrc/eoc_abi.cpp); strings come back as new[] pointers
    "addCiphertexts": (_VP, [C.c_char_p, C.c_char_p, C.c_char_p]),
    "subtractCiphertexts": (_VP, [C.c_char_p, C.c_char_p, C.c_char_p]),
    "addCiphertextBits": (_VP, [C.c_char_p, C.c_char_p, C.c_int32]),
    "subtractCiphertextBits": (_VP, [C.c_char_p, C.c_char_p, C.c_int32]),
    "b200_eoc_free": (None, [_VP]),
    "b200_eoc_set_cloud_key": (C.c_int, [C.c_char_p]),
    "b200_eoc_set_key": (C.c_int, [_VP]),
    "b200_add_bits_host": (C.c_int, [_VP, _VP, _VP, _VP, C.c_int32, C.c_int64, C.c_int]),
}
for _g in ("NAND", "OR", "AND", "XOR", "XNOR", "NOR", "ANDNY", "ANDYN", "ORNY", "ORYN"):
    SIGNATURES["boots" + _g] = (None, [_VP, _VP, _VP, _VP])
SIGNATURES["bootsMUX"] = (None, [_VP, _VP, _VP, _VP, _VP])
SIGNATURES["bootsNOT"] = (None, [_VP, _VP, _VP])
SIGNATURES["bootsCOPY"] = (None, [_VP, _VP, _VP])
SIGNATURES["bootsCONSTANT"] = (None, [_VP, C.c_int32, _VP])
SIGNATURES["tfhe_bootstrap_FFT"] = (None, [_VP, _VP, C.c_int32, _VP])
SIGNATURES["tfhe_bootstrap_woKS_FFT"] = (None, [_VP, _VP, C.c_int32, _VP])
SIGNATURES["lweKeySwitch"] = (None, [_VP, _VP, _VP])


def load():
    """Load libb200tfhe.so (built by __graft_entry__.build()). Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200Error(
            "%s not found: build it with `python __graft_entry__.py` (there is no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_LOCAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise B200Error(load().b200_last_error().decode("utf-8", "replace"))
