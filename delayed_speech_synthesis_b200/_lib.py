"""ctypes binding of libhga_b200.so (include/hga_b200.h).  There is NO CPU fallback: if the library cannot be loaded,
or no CUDA device is visible when a compute entry point is called, an exception is raised."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

from . import _build

HGA_ABI_VERSION = 1
HGA_MAX_SECTIONS = 8
PREC_EXACT, PREC_F64, PREC_F32, PREC_MIXED = 0, 1, 2, 3
PRECISIONS = {"exact": PREC_EXACT, "f64": PREC_F64, "f32": PREC_F32, "mixed": PREC_MIXED}


class HgaError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libhga_b200 error {code}: {msg}")
        self.code = code


class Config(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32), ("device", C.c_int32), ("fs", C.c_int32), ("n_channels", C.c_int32),
        ("frame_len", C.c_int32), ("hop", C.c_int32), ("eps", C.c_double),
        ("n_sections_bp", C.c_int32), ("n_sections_notch", C.c_int32),
        ("sos_bp", C.c_double * (HGA_MAX_SECTIONS * 6)), ("sos_notch", C.c_double * (HGA_MAX_SECTIONS * 6)),
        ("zi_bp", C.c_double * (HGA_MAX_SECTIONS * 2)), ("zi_notch", C.c_double * (HGA_MAX_SECTIONS * 2)),
        ("precision", C.c_int32), ("warmup", C.c_int32), ("warmup_tol", C.c_double),
        ("max_push_rows", C.c_int32), ("reserved", C.c_int32),
    ]


_vp, _i32, _i64, _dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_double
_pi64, _pi32 = C.POINTER(C.c_int64), C.POINTER(C.c_int32)

# name -> (restype, argtypes); every symbol include/hga_b200.h declares
SIGNATURES = {
    "hga_last_error": (C.c_char_p, []),
    "hga_abi_version": (C.c_int, []),
    "hga_device_count": (C.c_int, []),
    "hga_create": (C.c_int, [C.POINTER(Config), C.POINTER(_vp)]),
    "hga_destroy": (C.c_int, [_vp]),
    "hga_reset": (C.c_int, [_vp, _vp]),
    "hga_warmup": (C.c_int, [_vp]),
    "hga_last_plan": (C.c_int, [_vp, _pi64, _pi32, _pi32, _pi32]),
    "hga_num_frames": (_i64, [_vp, _i64]),
    "hga_extract_offline": (C.c_int, [_vp, _vp, _i64, _i32, _pi64, _pi64, _pi64, _vp, _vp, _vp, _i64, _vp]),
    "hga_extract_offline_host": (C.c_int, [_vp, _vp, _i64, _i32, _pi64, _pi64, _pi64, _vp, _vp, _vp, _i64]),
    "hga_stream_frames": (_i64, [_vp, _i64]),
    "hga_stream_push": (C.c_int, [_vp, _vp, _i64, _i64, _vp, _vp, _vp, _i64, _pi64, _vp]),
    "hga_stream_push_host": (C.c_int, [_vp, _vp, _i64, _i64, _vp, _vp, _vp, _i64, _pi64]),
    "hga_get_state": (C.c_int, [_vp, _vp, _vp, _vp, _pi32]),
    "hga_set_state": (C.c_int, [_vp, _vp, _vp, _vp, _pi32]),
    "hga_sosfilt": (C.c_int, [_i32, _vp, _i32, _vp, _i64, _i32, _i64, _vp, _vp, _i64, _vp]),
    "hga_log_power": (C.c_int, [_i32, _vp, _i64, _i64, _i32, _i64, _i64, _i32, _i32, _vp, _vp, _dbl, _vp, _vp, _vp, _i64, _vp]),
    "hga_zscore": (C.c_int, [_i32, _vp, _i64, _i32, _i64, _vp, _vp, _vp, _i64, _vp]),
    "hga_stats_accumulate": (C.c_int, [_i32, _vp, _i64, _i32, _i64, _vp, _vp]),
    "hga_stats_accumulate_static": (C.c_int, [_i32, _vp, _i64, _i32, _i64, _vp, _vp]),
    "hga_stats_merge": (C.c_int, [_i32, _vp, _vp, _i32, _i32, _vp]),
    "hga_stats_finalize": (C.c_int, [_i32, _vp, _i32, _vp, _vp, _vp]),
    "hga_stats_merge_days": (C.c_int, [_i32, _vp, _i32, _i32, _i32, _vp, _vp, _vp]),
    "hga_chanmap_create": (C.c_int, [_i32, _i32, _i32, _pi32, _pi32, _vp, _vp, _i32, _pi32, _pi32, C.POINTER(_vp)]),
    "hga_chanmap_destroy": (C.c_int, [_vp]),
    "hga_chanmap_apply": (C.c_int, [_vp, _vp, _i64, _i64, _vp, _i64, _vp]),
    "hga_extract_offline_stats": (C.c_int, [_vp, _vp, _i64, _i32, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                                            _vp, _vp, _vp, _i64, _vp, _vp, C.POINTER(C.c_int32)]),
    "hga_extract_offline_mapped": (C.c_int, [_vp, _vp, _vp, _i64, _i32, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                                             _vp, _vp, _vp, _i64, _vp, C.POINTER(C.c_int32)]),
    "hga_chanmap_apply_raw": (C.c_int, [_vp, _vp, _i32, _vp, _i64, _i64, _vp, _i64, _vp]),
    "hga_extract_offline_raw_host": (C.c_int, [_vp, _vp, _vp, _i32, _vp, _i64, _i32, _pi64, _pi64, _pi64, _vp, _vp, _vp, _i64]),
    "hga_chanmap_apply_host": (C.c_int, [_vp, _vp, _i64, _i64, _vp, _i64]),
    "hga_packet_header": (C.c_int, [_vp, _i64, _pi32, _pi32]),
    "hga_packet_decode": (C.c_int, [_i32, _vp, _i64, _vp, _i64, _vp]),
    "hga_stream_push_packet": (C.c_int, [_vp, _vp, _i64, _vp, _vp, _vp, _i64, _vp, C.POINTER(_i64)]),
    "hga_launch_count": (_i64, []),
    "hga_set_timing": (C.c_int, [_vp, _i32]),
    "hga_last_kernel_ms": (_dbl, [_vp]),
    "hga_debug_guards": (C.c_int, [C.c_int32]),
    "hga_debug_check_guards": (C.c_int64, []),
    "hga_debug_guarded_buffers": (C.c_int64, []),
    "hga_last_kernel_split_ms": (C.c_int, [_vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
}

_lib = None


def lib_path() -> Path:
    return _build.LIB


def load() -> C.CDLL:
    """Load libhga_b200.so (building it in-tree first if it is missing or stale and nvcc is available)."""
    global _lib
    if _lib is not None:
        return _lib
    if _build.stale() and _build.nvcc_path() is not None:
        _build.build()
    if not _build.LIB.exists():
        raise ImportError(f"{_build.LIB} is missing and cannot be built here (no nvcc): the HGA path has no CPU fallback")
    lib = C.CDLL(str(_build.LIB))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype, fn.argtypes = res, args
    if lib.hga_abi_version() != HGA_ABI_VERSION:
        raise ImportError(f"libhga_b200 ABI {lib.hga_abi_version()} != binding {HGA_ABI_VERSION}")
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        raise HgaError(rc, load().hga_last_error().decode(errors="replace"))


def require_device() -> None:
    if load().hga_device_count() < 1:
        raise HgaError(3, "no CUDA device visible: the HGA path has no CPU fallback")
