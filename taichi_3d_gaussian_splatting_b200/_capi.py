"""ctypes binding of libgs_b200.so, the C ABI declared in include/gs_b200.h.

The library is built in-tree by `build.py` (nvcc, sm_100a).  There is no CPU or
PyTorch fallback: if the library cannot be loaded (and cannot be built), importing
this module raises, and so does every rasteriser call.
"""
from __future__ import annotations

import ctypes as C
import os

from . import build as _build

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = _build.LIB

ABI_VERSION = 113

_p = C.c_void_p
_i = C.c_int
_l = C.c_int64
_f = C.c_float

class gs_view_t(C.Structure):
    """include/gs_b200.h: one view of a gradient batch (device pointers into its forward / backward state)."""
    _fields_ = [("slot_of_point", C.c_void_p), ("point_in_camera", C.c_void_p), ("splat_grads", C.c_void_p),
                ("intrinsics", C.c_void_p), ("q_cam_pc", C.c_void_p), ("t_pc_cam", C.c_void_p),
                ("reserved", C.c_int64 * 2)]


class gs_exchange_t(C.Structure):
    """include/gs_b200.h: what gs_point_bwd_batch_exchange needs to know about the symmetric buffers."""
    _fields_ = [("multicast_base", C.c_void_p), ("local_base", C.c_void_p), ("flags", C.c_void_p * 16),
                ("rank", C.c_int32), ("world", C.c_int32), ("chunks", C.c_int32), ("reducer_ctas", C.c_int32),
                ("epoch", C.c_uint32), ("reserved", C.c_uint32), ("scratch", C.c_void_p)]


# name -> (restype, argtypes); must list every symbol of include/gs_b200.h
SIGNATURES = {
    "gs_abi_version": (_i, []),
    "gs_status_string": (C.c_char_p, [_i]),
    "gs_project_workspace_bytes": (_l, [_l]),
    "gs_sort_workspace_bytes": (_l, [_l]),
    "gs_point_sort_workspace_bytes": (_l, [_l]),
    "gs_emit_workspace_bytes": (_l, [_l]),
    "gs_sort_points_by_depth": (_i, [_p, _l, _p, _i, _p, _p, _p]),
    "gs_emit_tile_keys": (_i, [_p, _p, _l, _l, _i, _p, _p, _p, _p, _i, _p]),
    "gs_sort_pairs32": (_i, [_p, _p, _p, _p, _l, _i, _p, _p, C.POINTER(C.c_int), _p]),
    "gs_tile_ranges32": (_i, [_p, _l, _p, _p, _i, _p]),
    "gs_sorted_keys": (_i, [_p, _p, _p, _l, _p, _p]),
    "gs_bin_workspace_bytes": (_l, [_l, _l, _i, _i]),
    "gs_bin_entries_capacity": (_l, [_l, _l, _i]),
    "gs_bin_tiles": (_i, [_p, _p, _l, _l, _i, _i, _p, _p, _p, _p, _p, _p]),
    "gs_inverse_pose": (_i, [_p, _p, _i, _p, _p, _p]),
    "gs_project_cull_sh": (_i, [_p, _p, _p, _p, _l, _p, _p, _p, _i, _f, _f, _i, _i, _f,
                                _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "gs_point_bwd_batch": (_i, [_p, _p, _p, _l, _l, _p, _i, _i, _f, _f, _f, _f, _f, _p, _p, _p]),
    "gs_point_bwd_all_points": (_i, [_p, _p, _p, _l, C.POINTER(gs_view_t), _i, _f, _f, _f, _f, _f, _p, _p, _p]),
    "gs_exchange_flag_words": (_l, [_i, _i]),
    "gs_exchange_scratch_bytes": (_l, [_i]),
    "gs_point_bwd_batch_exchange": (_i, [_p, _p, _p, _l, _p, _i, _i, _f, _f, _f, _f, _f,
                                         C.POINTER(gs_exchange_t), _p]),
    # no workspace argument: its scratch is a file-scope device array of the library (no allocation)
    "gs_project_cull_sh_static": (_i, [_p, _p, _p, _p, _l, _p, _p, _p, _i, _f, _f, _i, _i, _f,
                                       _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "gs_emit_keys": (_i, [_p, _p, _p, _l, _l, _i, _p, _p, _p]),
    "gs_sort_pairs": (_i, [_p, _p, _p, _p, _l, _i, _i, _p, C.POINTER(C.c_int), _p]),
    "gs_tile_ranges": (_i, [_p, _l, _p, _p, _i, _p]),
    "gs_blend_fwd": (_i, [_i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _p, _l, _p, _p]),
    "gs_blend_bwd_scratch_bytes": (_l, [_i, _i]),
    "gs_blend_bwd": (_i, [_i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _l, _p, _p, _p]),
    "gs_point_bwd": (_i, [_p, _p, _p, _p, _p, _p, _p, _i, _p, _l, _p, _p, _i, _f, _f, _f, _f, _f,
                          _p, _p, _p, _p, _p, _p, _p, _i, _p]),
    "gs_loss_workspace_bytes": (_l, [_i, _i]),
    "gs_loss_maps_bytes": (_l, [_i, _i]),
    "gs_loss_fwd": (_i, [_p, _l, _l, _l, _p, _l, _l, _l, _i, _i, _f, _p, _p, _l, _f, _p, _p, _p, _p]),
    "gs_loss_bwd": (_i, [_p, _l, _l, _l, _p, _l, _l, _l, _i, _i, _f, _p, _p, _p, _p]),
    "gs_scale_reg_bwd": (_i, [_p, _p, _l, _f, _p, _p, _p, _i, _p]),
    "gs_controller_accumulate": (_i, [_p, _l, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "gs_adam_step": (_i, [_p, _p, _p, _p, _l, _f, _f, _f, _f, _l, _p, _i, _p]),
}


class GsError(RuntimeError):
    pass


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH) or os.environ.get("GS_B200_REBUILD") == "1":
        # raises if nvcc is missing or the compile fails: no fallback path exists
        _build.build(force=os.environ.get("GS_B200_REBUILD") == "1")
    lib = C.CDLL(LIB_PATH)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here means header and library disagree
        fn.restype = restype
        fn.argtypes = argtypes
    got = lib.gs_abi_version()
    if got != ABI_VERSION:
        raise GsError(f"libgs_b200.so ABI {got} != expected {ABI_VERSION}; rebuild with build.py")
    return lib


lib = _load()


def check(status: int, what: str) -> None:
    if status != 0:
        msg = lib.gs_status_string(status)
        raise GsError(f"{what} failed: status {status} ({msg.decode() if msg else '?'})")
