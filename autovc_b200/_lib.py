"""ctypes binding of libautovc_b200.so (include/autovc_b200.h).  Fails loudly when the library is
missing: there is no CPU or PyTorch fallback for the generator path."""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_float, c_int, c_int32, c_int64, c_size_t, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libautovc_b200.so")

AVC_OK = 0
AVC_ERR_INVALID_ARG = -1
AVC_ERR_T_TOO_SHORT = -2
AVC_ERR_CUDA = -3
AVC_ERR_NOT_LOADED = -4
AVC_ERR_WORKSPACE = -5
AVC_ERR_MISSING_WEIGHT = -6
AVC_ERR_UNSUPPORTED = -7
AVC_PREC_BF16 = 0
AVC_PREC_FP32 = 1

EXPORTS = ["avc_abi_version", "avc_create", "avc_load_weights", "avc_workspace_bytes", "avc_forward", "avc_encode", "avc_forward_bn_train",
           "avc_last_launch_count", "avc_plan_builds", "avc_graph_replays", "avc_set_profiling", "avc_get_profile", "avc_debug_packed", "avc_debug_gemm", "avc_destroy", "avc_last_error",
           "avc_se_load_weights", "avc_se_workspace_bytes", "avc_se_forward", "avc_se_forward_static", "avc_se_reduce",
           "avc_mel_create", "avc_mel_num_frames", "avc_mel_filterbank", "avc_mel_forward", "avc_mel_last_error", "avc_mel_destroy",
           "avc_voc_create", "avc_voc_load_weights", "avc_voc_condition_workspace_bytes", "avc_voc_condition", "avc_voc_max_rows",
           "avc_voc_generate_workspace_bytes", "avc_voc_generate", "avc_voc_last_launch_count", "avc_voc_last_error", "avc_voc_destroy"]


class AvcDims(Structure):
    _fields_ = [("dim_neck", c_int32), ("dim_emb", c_int32), ("dim_pre", c_int32), ("freq", c_int32)]


class AvcMelConfig(Structure):
    _fields_ = [("sr", c_int32), ("n_fft", c_int32), ("hop_length", c_int32), ("win_length", c_int32), ("n_mels", c_int32),
                ("fmin", c_float), ("fmax", c_float), ("power", c_int32), ("log_normalize", c_int32), ("min_level_db", c_float),
                ("pad_reflect", c_int32), ("frames_major", c_int32)]


class AvcWeightDesc(Structure):
    _fields_ = [("name", c_char_p), ("data", POINTER(c_float)), ("numel", c_int64)]


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m autovc_b200.build` (nvcc, sm_100a). "
            "autovc_b200 has no CPU/PyTorch fallback.")
    L = ctypes.CDLL(LIB_PATH)
    L.avc_abi_version.restype = c_int
    L.avc_create.argtypes = [POINTER(c_void_p), POINTER(AvcDims), c_int, c_int]
    L.avc_create.restype = c_int
    L.avc_load_weights.argtypes = [c_void_p, POINTER(AvcWeightDesc), c_int]
    L.avc_load_weights.restype = c_int
    L.avc_workspace_bytes.argtypes = [c_void_p, c_int, c_int]
    L.avc_workspace_bytes.restype = c_size_t
    L.avc_forward.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p,
                              c_void_p, c_size_t, c_void_p]
    L.avc_forward.restype = c_int
    L.avc_encode.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_size_t, c_void_p]
    L.avc_encode.restype = c_int
    L.avc_forward_bn_train.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                       c_void_p, c_void_p, c_size_t, c_void_p]
    L.avc_forward_bn_train.restype = c_int
    L.avc_last_launch_count.argtypes = [c_void_p]
    L.avc_last_launch_count.restype = c_int64
    L.avc_plan_builds.argtypes = [c_void_p]
    L.avc_plan_builds.restype = c_int64
    L.avc_graph_replays.argtypes = [c_void_p]
    L.avc_graph_replays.restype = c_int64
    L.avc_set_profiling.argtypes = [c_void_p, c_int]
    L.avc_set_profiling.restype = c_int
    L.avc_get_profile.argtypes = [c_void_p, POINTER(c_float), c_int]
    L.avc_get_profile.restype = c_int
    L.avc_debug_packed.argtypes = [c_void_p, c_char_p, POINTER(c_float), c_int64]
    L.avc_debug_packed.restype = c_int64
    L.avc_debug_gemm.argtypes = [c_void_p, c_char_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_int,
                                 c_int, c_void_p, c_void_p, c_void_p]
    L.avc_debug_gemm.restype = c_int
    L.avc_se_load_weights.argtypes = [c_void_p, POINTER(AvcWeightDesc), c_int]
    L.avc_se_load_weights.restype = c_int
    L.avc_se_workspace_bytes.argtypes = [c_void_p, c_int, c_int]
    L.avc_se_workspace_bytes.restype = c_size_t
    L.avc_se_forward.argtypes = [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_size_t, c_void_p]
    L.avc_se_forward.restype = c_int
    L.avc_se_forward_static.argtypes = [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]
    L.avc_se_forward_static.restype = c_int
    L.avc_se_reduce.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p]
    L.avc_se_reduce.restype = c_int
    L.avc_mel_create.argtypes = [POINTER(c_void_p), POINTER(AvcMelConfig), c_int]
    L.avc_mel_create.restype = c_int
    L.avc_mel_num_frames.argtypes = [c_void_p, c_int64]
    L.avc_mel_num_frames.restype = c_int64
    L.avc_mel_filterbank.argtypes = [c_void_p, POINTER(c_float), c_int64]
    L.avc_mel_filterbank.restype = c_int64
    L.avc_mel_forward.argtypes = [c_void_p, c_void_p, c_int, c_int64, c_int64, c_void_p, c_void_p]
    L.avc_mel_forward.restype = c_int
    L.avc_mel_last_error.argtypes = [c_void_p]
    L.avc_mel_last_error.restype = c_char_p
    L.avc_mel_destroy.argtypes = [c_void_p]
    L.avc_mel_destroy.restype = None
    L.avc_voc_create.argtypes = [POINTER(c_void_p), c_int]
    L.avc_voc_create.restype = c_int
    L.avc_voc_load_weights.argtypes = [c_void_p, POINTER(AvcWeightDesc), c_int]
    L.avc_voc_load_weights.restype = c_int
    L.avc_voc_condition_workspace_bytes.argtypes = [c_void_p, c_int, c_int]
    L.avc_voc_condition_workspace_bytes.restype = c_size_t
    L.avc_voc_condition.argtypes = [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]
    L.avc_voc_condition.restype = c_int
    L.avc_voc_max_rows.argtypes = []
    L.avc_voc_max_rows.restype = c_int
    L.avc_voc_generate_workspace_bytes.argtypes = [c_void_p, c_int, c_int]
    L.avc_voc_generate_workspace_bytes.restype = c_size_t
    L.avc_voc_generate.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_int, c_int, c_void_p,
                                   ctypes.c_uint64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]
    L.avc_voc_generate.restype = c_int
    L.avc_voc_last_launch_count.argtypes = [c_void_p]
    L.avc_voc_last_launch_count.restype = c_int64
    L.avc_voc_last_error.argtypes = [c_void_p]
    L.avc_voc_last_error.restype = c_char_p
    L.avc_voc_destroy.argtypes = [c_void_p]
    L.avc_voc_destroy.restype = None
    L.avc_destroy.argtypes = [c_void_p]
    L.avc_destroy.restype = None
    L.avc_last_error.argtypes = [c_void_p]
    L.avc_last_error.restype = c_char_p
    _lib = L
    return L


def last_error(handle) -> str:
    return lib().avc_last_error(handle).decode("utf-8", "replace")


def raise_for(code: int, handle) -> None:
    """Map avc_status to the exception the reference raises in the same situation."""
    if code == AVC_OK:
        return
    msg = last_error(handle)
    if code == AVC_ERR_T_TOO_SHORT:
        raise ValueError(msg)          # reference: torch.cat() on an empty list -> ValueError
    if code in (AVC_ERR_MISSING_WEIGHT,):
        raise KeyError(msg)
    raise RuntimeError(f"autovc_b200 error {code}: {msg}")
