"""ctypes binding of libmagic_b200.so (include/magic_b200.h).

The library is the product: there is no Python or CPU fallback.  If it is missing, or if no B200 is visible,
the calls raise -- loudly."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libmagic_b200.so")

MAX_ENC_LAYERS = 8
ACT_TANH, ACT_ELU = 1, 2
METRIC_COSINE, METRIC_SQDIFF = 0, 1
PRECISION_BF16, PRECISION_FP32 = 0, 1
FWD_DISTANCE, FWD_ZMEAN, FWD_RECON = 0, 1, 2
SLOT_M, SLOT_V = 0, 1


class MagicConfig(C.Structure):
    _fields_ = [
        ("D", C.c_int32), ("n_enc", C.c_int32), ("enc", C.c_int32 * MAX_ENC_LAYERS), ("dec", C.c_int32 * 2),
        ("L", C.c_int32), ("act", C.c_int32), ("metric", C.c_int32), ("reciprocal", C.c_int32),
        ("def_weight", C.c_float), ("lr1", C.c_float), ("lr2", C.c_float),
        ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
        ("precision", C.c_int32), ("batch_local", C.c_int32), ("batch_global", C.c_int32),
        ("rank", C.c_int32), ("world", C.c_int32), ("seed", C.c_uint64),
    ]


class MagicStats(C.Structure):
    _fields_ = [
        ("steps", C.c_int64), ("launches_last_step", C.c_int64), ("gemm_flops_per_step", C.c_double),
        ("model_flops_per_pair", C.c_double), ("n_params1", C.c_int64), ("n_params2", C.c_int64),
        ("hbm_bytes", C.c_int64),
    ]


# name -> (restype, argtypes); every symbol include/magic_b200.h declares
_VP, _I32, _I64, _U64 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64
SIGNATURES = {
    "magic_last_error": (C.c_char_p, []),
    "magic_version": (C.c_int, []),
    "magic_create": (C.c_int, [C.POINTER(MagicConfig), _VP, C.POINTER(_VP)]),
    "magic_destroy": (C.c_int, [_VP]),
    "magic_sync": (C.c_int, [_VP]),
    "magic_get_stats": (C.c_int, [_VP, C.POINTER(MagicStats)]),
    "magic_param_count": (C.c_int, [_VP, C.POINTER(_I32)]),
    "magic_param_info": (C.c_int, [_VP, _I32, C.POINTER(C.c_char_p), C.POINTER(_I64 * 2), C.POINTER(_I32)]),
    "magic_set_param": (C.c_int, [_VP, _I32, _VP]),
    "magic_get_param": (C.c_int, [_VP, _I32, _VP]),
    "magic_get_grad": (C.c_int, [_VP, _I32, _I32, _VP]),
    "magic_get_slot": (C.c_int, [_VP, _I32, _I32, _I32, _VP]),
    "magic_set_slot": (C.c_int, [_VP, _I32, _I32, _I32, _VP]),
    "magic_get_step": (C.c_int, [_VP, C.POINTER(_I64)]),
    "magic_set_step": (C.c_int, [_VP, _I64]),
    "magic_load_interleaved": (C.c_int, [_VP, _VP]),
    "magic_load_interleaved_u8": (C.c_int, [_VP, _VP]),
    "magic_partial_fit_host_u8": (C.c_int, [_VP, _VP, _VP, _VP, _VP, _VP]),
    "magic_load_planar_u8": (C.c_int, [_VP, _VP, _VP, _VP]),
    "magic_load_pairs_u8": (C.c_int, [_VP, _VP, _VP, C.c_uint64, C.c_uint64]),
    "magic_load_pairs_bits": (C.c_int, [_VP, _VP, _VP, C.c_int32, C.c_uint64, C.c_uint64]),
    "magic_feed_pairs_bits_gather": (C.c_int, [_VP, _VP, _VP, C.c_int32, C.c_int64, _VP, _VP, C.c_uint64, C.c_uint64, C.c_int32]),
    "magic_feed_pairs_bits_host": (C.c_int, [_VP, _VP, _VP, _VP, C.c_int32, C.c_uint64, C.c_uint64, C.c_int32]),
    "magic_generate_batch": (C.c_int, [_VP, _U64, _U64]),
    "magic_export_batch": (C.c_int, [_VP, _VP, _VP]),
    "magic_export_rotation": (C.c_int, [_VP, _VP]),
    "magic_export_latents": (C.c_int, [_VP, _VP]),
    "magic_cosine_distance": (C.c_int, [_VP, _VP, C.c_int32, C.c_int32, _VP]),
    "magic_cosine_distance_noalloc": (C.c_int, [_VP, _VP, C.c_int32, C.c_int32, _VP, _VP]),
    "magic_swap_batch": (C.c_int, [_VP]),
    "magic_set_option": (C.c_int, [_VP, C.c_char_p, C.c_int32]),
    "magic_input_density": (C.c_int, [_VP, C.POINTER(C.c_double)]),
    "magic_step": (C.c_int, [_VP, _VP, _VP, _VP, _VP]),
    "magic_read_losses": (C.c_int, [_VP, C.POINTER(C.c_float)]),
    "magic_step_grads_only": (C.c_int, [_VP, _VP, _VP, _VP, _VP]),
    "magic_partial_fit_host": (C.c_int, [_VP, _VP, _VP, _VP, _VP, _VP]),
    "magic_forward": (C.c_int, [_VP, _I32, _VP, _VP]),
    "magic_decode": (C.c_int, [_VP, _VP, _VP]),
    "magic_comm_unique_id": (C.c_int, [_VP]),
    "magic_comm_init": (C.c_int, [_VP, _VP]),
    "magic_comm_export_peer": (C.c_int, [_VP, _VP]),
    "magic_comm_import_peers": (C.c_int, [_VP, _VP]),
    "magic_profile_enable": (C.c_int, [_VP, _I32]),
    "magic_profile_report": (C.c_int, [_VP, C.c_char_p, _I64]),
    "magic_test_gemm": (C.c_int, [_I32, _I32, _I32, _I32, _VP, _I64, _VP, _I64, _VP, _I64, _I32, _VP]),
}

_lib = None


class MagicError(RuntimeError):
    pass


def load() -> C.CDLL:
    """dlopen the library and bind every declared symbol (works without a GPU; compute calls do not)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise MagicError(
            "%s is missing: build it with `python -m magic_b200.build` (nvcc, sm_100a). "
            "There is no fallback path." % LIB_PATH)
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        msg = load().magic_last_error()
        raise MagicError(msg.decode() if msg else "magic_b200 error %d" % rc)
