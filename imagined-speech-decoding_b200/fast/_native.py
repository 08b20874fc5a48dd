"""ctypes binding of libfastb200.so -- the C ABI declared in include/fastb200.h.

There is NO fallback: if the shared library is missing or a call fails, a RuntimeError is raised.  The library is
plain CUDA behind ``extern "C"``; torch is used on this side only for device memory and the current stream.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# FASTB200_LIB: developer override (A/B timing of kernel variants built elsewhere); the default is the in-tree library
LIB_PATH = os.environ.get("FASTB200_LIB") or os.path.join(_PKG, "libfastb200.so")

MAX_ZONES, MAX_CHANS, MAX_LAYERS = 16, 256, 16
MODE_DEFAULT, MODE_TRAIN_HEAD = 0, 1

_i32 = C.c_int32


class ModelDesc(C.Structure):
    """Mirror of ``struct fastb200_model`` (include/fastb200.h)."""
    _fields_ = [
        ("n_chan", _i32), ("window_len", _i32), ("slide_step", _i32), ("n_zones", _i32), ("dim_cnn", _i32),
        ("dim_token", _i32), ("n_heads", _i32), ("n_layers", _i32), ("dim_hidden", _i32), ("n_classes", _i32),
        ("n_pos", _i32), ("n_params", _i32),
        ("zone_size", _i32 * MAX_ZONES), ("zone_off", _i32 * MAX_ZONES), ("zone_chan", _i32 * MAX_CHANS),
        ("off_pos", _i32), ("off_cls", _i32),
        ("off_cnn1_w", _i32 * MAX_ZONES), ("off_cnn1_b", _i32 * MAX_ZONES), ("off_cnn2_w", _i32 * MAX_ZONES),
        ("off_cnn3_w", _i32 * MAX_ZONES), ("off_cnn4_w", _i32 * MAX_ZONES),
        ("off_in_w", _i32), ("off_in_b", _i32),
        ("off_ln1_w", _i32 * MAX_LAYERS), ("off_ln1_b", _i32 * MAX_LAYERS),
        ("off_qkv_w", _i32 * MAX_LAYERS), ("off_qkv_b", _i32 * MAX_LAYERS),
        ("off_out_w", _i32 * MAX_LAYERS), ("off_out_b", _i32 * MAX_LAYERS),
        ("off_ln2_w", _i32 * MAX_LAYERS), ("off_ln2_b", _i32 * MAX_LAYERS),
        ("off_ff1_w", _i32 * MAX_LAYERS), ("off_ff1_b", _i32 * MAX_LAYERS),
        ("off_ff2_w", _i32 * MAX_LAYERS), ("off_ff2_b", _i32 * MAX_LAYERS),
        ("off_last_w", _i32), ("off_last_b", _i32),
    ]


ALT_MAX_STAGES = 6


class AltStage(C.Structure):
    """Mirror of ``struct fastb200_alt_stage``."""
    _fields_ = [("kind", _i32), ("out_ch", _i32), ("k", _i32), ("pad", _i32), ("act", _i32), ("pool_kind", _i32), ("pool", _i32),
                ("w_off", _i32), ("scale_off", _i32), ("shift_off", _i32)]


class AltHeadDesc(C.Structure):
    """Mirror of ``struct fastb200_alt_head`` (eval-mode chain of the alternative zone heads)."""
    _fields_ = [("window_len", _i32), ("n_zones", _i32), ("n_stages", _i32), ("final_kind", _i32),
                ("lin_in", _i32), ("lin_out", _i32), ("lin_w_off", _i32), ("lin_b_off", _i32),
                ("stage", AltStage * ALT_MAX_STAGES),
                ("zone_size", _i32 * MAX_ZONES), ("zone_off", _i32 * MAX_ZONES), ("zone_chan", _i32 * MAX_CHANS),
                ("zone_base", C.c_int64 * MAX_ZONES)]


_P = C.c_void_p
_MP = C.POINTER(ModelDesc)
_AP = C.POINTER(AltHeadDesc)
# name -> (restype, argtypes); every symbol include/fastb200.h declares
SIGNATURES = {
    "fastb200_version": (C.c_int, []),
    "fastb200_sizeof_model": (C.c_int, []),
    "fastb200_last_error": (C.c_char_p, []),
    "fastb200_launch_count": (C.c_int64, []),
    "fastb200_copy_rows_async": (C.c_int, [_P, C.c_int64, _P, C.c_int64, C.c_int64, C.c_int64, _P]),
    "fastb200_window_gather": (C.c_int, [_P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P]),
    "fastb200_head_supported": (C.c_int, [_MP, C.c_int]),
    "fastb200_head_packed_floats": (C.c_int64, [_MP]),
    "fastb200_head_saved_floats": (C.c_int64, [_MP, C.c_int, C.c_int]),
    "fastb200_head_pack": (C.c_int, [_MP, _P, _P, _P]),
    "fastb200_head_fwd": (C.c_int, [_MP, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P, _P]),
    "fastb200_head_bwd": (C.c_int, [_MP, _P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P, _P]),
    "fastb200_head_unpack_grad": (C.c_int, [_MP, _P, _P, _P, _P]),
    "fastb200_head_bwd_tc_supported": (C.c_int, [_MP]),
    "fastb200_head_bwd_tc_work_floats": (C.c_int64, [_MP, C.c_int, C.c_int]),
    "fastb200_head_bwd_tc": (C.c_int, [_MP, _P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P, _P, _P]),
    "fastb200_head_bwd_tc_static_max_floats": (C.c_int64, []),
    "fastb200_head_bwd_tc_static": (C.c_int, [_MP, _P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P]),
    "fastb200_head_tc_supported": (C.c_int, [_MP]),
    "fastb200_head_packed_tc_bytes": (C.c_int64, [_MP, C.c_int]),
    "fastb200_head_pack_tc": (C.c_int, [_MP, _P, _P, C.c_int, _P]),
    "fastb200_head_fwd_tc": (C.c_int, [_MP, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, _P, C.c_int, _P]),
    "fastb200_head_fwd_tc_train": (C.c_int, [_MP, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P, C.c_int, _P]),
    "fastb200_sizeof_alt_head": (C.c_int, []),
    "fastb200_alt_head_fwd": (C.c_int, [_AP, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P]),
    "fastb200_transformer_saved_floats": (C.c_int64, [_MP, C.c_int, C.c_int]),
    "fastb200_transformer_fwd": (C.c_int, [_MP, _P, _P, C.c_int, C.c_int, C.c_int, C.c_float, C.c_uint64, C.c_uint64,
                                           _P, _P, _P, _P]),
    "fastb200_transformer_bwd": (C.c_int, [_MP, _P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_float, C.c_uint64,
                                           C.c_uint64, _P, _P, _P, _P]),
    "fastb200_ce_fwd_bwd": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_float, _P, _P, _P, _P, _P]),
    "fastb200_adamw_flat": (C.c_int, [_P, _P, _P, _P, C.c_int64, _P, C.c_int, C.c_double, C.c_double, C.c_double,
                                      C.c_double, _P, C.c_int, _P]),
    "fastb200_adamw_flat_peers": (C.c_int, [_P, C.POINTER(C.c_void_p), C.c_int, _P, _P, C.c_int64, _P, C.c_int, C.c_double,
                                            C.c_double, C.c_double, C.c_double, _P, C.c_int, _P]),
    "fastb200_trials_from_mat": (C.c_int, [_P, C.c_int, _P, C.c_int, C.c_int, C.c_int, C.c_int, _P]),
    "fastb200_sosfiltfilt": (C.c_int, [_P, C.c_int, _P, C.c_int, C.c_int64, C.c_int, C.POINTER(C.c_double), C.c_int,
                                       C.c_int, _P]),
    "fastb200_resample_poly": (C.c_int, [_P, C.c_int, _P, C.c_int, C.c_int64, C.c_int, C.c_int, _P, C.c_int, C.c_int,
                                         C.c_int, C.c_int, C.POINTER(C.c_double), _P]),
    "fastb200_sosfiltfilt_resample_supported": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.c_int,
                                                          C.c_int, C.c_int, C.c_int]),
    "fastb200_sosfiltfilt_resample": (C.c_int, [_P, C.c_int, _P, C.c_int, C.c_int64, C.c_int, C.POINTER(C.c_double), C.c_int,
                                                C.c_int, C.c_int, C.POINTER(C.c_double), C.c_int, C.c_int, C.c_int, C.c_int,
                                                _P]),
}

_lib = None


def lib():
    """The loaded library (loaded on first use).  Raises if it was not built: there is no CPU fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python imagined-speech-decoding_b200/build.py` "
                "(nvcc, sm_100a).  The FAST hot path has no CPU or eager-PyTorch fallback.")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)          # AttributeError here = header / library mismatch: fail loudly
            fn.restype = res
            fn.argtypes = args
        if handle.fastb200_sizeof_model() != C.sizeof(ModelDesc):
            raise RuntimeError("fastb200_model layout mismatch between include/fastb200.h and fast/_native.py")
        if handle.fastb200_sizeof_alt_head() != C.sizeof(AltHeadDesc):
            raise RuntimeError("fastb200_alt_head layout mismatch between include/fastb200.h and fast/_native.py")
        _lib = handle
    return _lib


_replayed = 0      # launches replayed from captured CUDA graphs (they do not pass through the library)


def note_replayed(n: int):
    global _replayed
    _replayed += int(n)


def launch_count() -> int:
    """Kernels of libfastb200.so launched by this process so far: direct launches (counted inside the library) plus the
    launches of every CUDA-graph replay the Python side reported with ``note_replayed``."""
    return int(lib().fastb200_launch_count()) + _replayed


def check(code: int, what: str):
    if code != 0:
        msg = lib().fastb200_last_error().decode(errors="replace")
        raise RuntimeError(f"{what} failed ({code}): {msg}")


def ptr(t):
    """Device (or host) address of a tensor, or NULL."""
    return None if t is None else t.data_ptr()


def stream_ptr(device=None):
    return torch.cuda.current_stream(device).cuda_stream


def require_cuda(t: torch.Tensor, name: str):
    if not t.is_cuda:
        raise RuntimeError(
            f"{name} is on {t.device}: the FAST hot path runs only as CUDA kernels on a B200 (sm_100a); "
            "move the module and its input to a CUDA device (there is no CPU fallback).")
