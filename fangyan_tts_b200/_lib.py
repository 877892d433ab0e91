"""ctypes binding of include/fangyan_b200.h.  No CPU fallback: a missing library is a hard error."""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

# FY_LIB points at an alternative build of the same ABI (e.g. the stall-profiling build of tools/gemm_stalls.py)
LIB_PATH = Path(os.environ.get("FY_LIB") or Path(__file__).resolve().parent / "libfangyan_b200.so")


class FyTensor(C.Structure):
    _fields_ = [("name", C.c_char_p), ("h_data", C.c_void_p), ("ndim", C.c_int32), ("shape", C.c_int64 * 4)]


class FyS3Config(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("n_mels", "d_model", "n_head", "n_layer", "d_ff", "fsmn_kernel", "fsq_dim")] + \
               [(n, C.c_float) for n in ("rope_theta", "ln_eps", "fsq_scale", "mlp_ln_eps")]


class FyCamConfig(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("feat_dim", "embed_dim", "fcm_channels", "init_channels", "growth", "bn_size",
                                         "n_blocks")] + \
               [("layers", C.c_int32 * 4), ("kernel", C.c_int32 * 4), ("dilation", C.c_int32 * 4),
                ("seg_len", C.c_int32), ("bn_eps", C.c_float)]


class FyWavInfo(C.Structure):
    _fields_ = [("frames", C.c_int64), ("data_offset", C.c_int64), ("channels", C.c_int32), ("sample_rate", C.c_int32),
                ("format", C.c_int32), ("bits", C.c_int32), ("status", C.c_int32), ("reserved", C.c_int32)]


_P, _I, _Z = C.c_void_p, C.c_int, C.c_size_t
# name -> (restype, argtypes); mirrors include/fangyan_b200.h one to one
PROTOTYPES = {
    "fy_create": (_I, [_I, _I, C.POINTER(_P)]),
    "fy_destroy": (None, [_P]),
    "fy_last_error": (C.c_char_p, []),
    "fy_set_stream": (_I, [_P, _P]),
    "fy_version": (C.c_char_p, []),
    "fy_load_campplus": (_I, [_P, C.POINTER(FyTensor), _I, C.POINTER(FyCamConfig)]),
    "fy_load_s3tok": (_I, [_P, C.POINTER(FyTensor), _I, C.POINTER(FyS3Config)]),
    "fy_wav_scan": (_I, [C.POINTER(C.c_char_p), _I, C.POINTER(FyWavInfo), _I]),
    "fy_wav_read_s16": (_I, [C.POINTER(C.c_char_p), C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(_P), _I, _I,
                            C.POINTER(C.c_int32)]),
    "fy_resample_prepare": (_I, [_P, _I, _I]),
    "fy_resample": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I]),
    "fy_fbank_workspace_bytes": (_Z, [_I, _I]),
    "fy_fbank_cmn": (_I, [_P, _P, _I, _P, _P, _I, _I, _P, _P, _Z]),
    "fy_fbank_cmn_static": (_I, [_P, _P, _I, _P, _P, _I, _I, _P]),
    "fy_logmel_workspace_bytes": (_Z, [_I, _I]),
    "fy_logmel": (_I, [_P, _P, _I, _P, _P, _I, _I, _I, _P, _P, _Z]),
    "fy_matcha_mel": (_I, [_P, _P, _I, _P, _P, _I, _I, _P]),
    "fy_campplus_workspace_bytes": (_Z, [_P, _I, _I]),
    "fy_campplus_forward": (_I, [_P, _P, _P, _P, _I, _P, _P, _Z]),
    "fy_campplus_layout": (_I, [_P, _P, _I, _P, _I]),
    "fy_s3tok_workspace_bytes": (_Z, [_P, _I, _I]),
    "fy_s3tok_encode_fsq": (_I, [_P, _P, _P, _P, _I, _P, _P, _P, _P, _Z]),
    "fy_speaker_sum": (_I, [_P, _P, _P, _I, _I, _P, _P]),
    "fy_speaker_mean_finish": (_I, [_P, _P, _P, _I, _P]),
    "fy_profile_enable": (_I, [_P, _I]),
    "fy_profile_read": (_I, [_P, C.POINTER(C.c_double), _I]),
    "fy_gemm_bf16": (_I, [_P, _P, _P, _P, _I, _I, _I]),
}

_lib = None


class FyError(RuntimeError):
    pass


def load():
    """Load the CUDA library; raises if it has not been built (python -m fangyan_tts_b200.build)."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise FyError(f"{LIB_PATH} is missing: build it with `python -m fangyan_tts_b200.build` "
                          "(there is no CPU fallback)")
        lib = C.CDLL(str(LIB_PATH))
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)  # AttributeError if the symbol is not exported
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def check(status: int):
    if status != 0:
        raise FyError(f"fangyan_b200 error {status}: {load().fy_last_error().decode()}")
