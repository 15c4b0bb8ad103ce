"""ctypes bindings of libparticles_cuda.so (include/particles_cuda.h).

The library is the product; this module only declares its symbols.  It is built in-tree by
``build.py`` and loading fails loudly if it is missing — there is no Python or CPU fallback.
"""
from __future__ import annotations

import ctypes
import os

from . import build as _build

PS_OK = 0
PS_ERR_BAD_ARG = -1
PS_ERR_OOM = -2
PS_ERR_CUDA = -3
PS_ERR_NCCL = -4
PS_ERR_CELL_OVERFLOW = -5
PS_ERR_CAPACITY = -6
PS_ERR_NO_DEVICE = -7
PS_ERR_MIGRATION = -8

PS_F64 = 0
PS_F32 = 1


class PsConfig(ctypes.Structure):
    """struct ps_config (include/particles_cuda.h)."""

    _fields_ = [
        ("struct_size", ctypes.c_uint32),
        ("device", ctypes.c_int32),
        ("count", ctypes.c_uint64),
        ("radius", ctypes.c_double),
        ("width", ctypes.c_double),
        ("height", ctypes.c_double),
        ("grid_size", ctypes.c_double),
        ("target_fps", ctypes.c_double),
        ("sub_steps", ctypes.c_uint64),
        ("g", ctypes.c_double),
        ("boundary_damping", ctypes.c_double),
        ("particle_damping", ctypes.c_double),
        ("precision", ctypes.c_int32),
        ("collisions", ctypes.c_int32),
        ("fuse_substeps", ctypes.c_uint32),
        ("contacts_per_particle", ctypes.c_uint32),
        ("slab_capacity_percent", ctypes.c_uint64),
        ("reserved", ctypes.c_uint64 * 3),
    ]


class PsStats(ctypes.Structure):
    _fields_ = [
        ("substeps_done", ctypes.c_uint64),
        ("pairs_last", ctypes.c_uint64),
        ("pairs_total", ctypes.c_uint64),
        ("rounds_last", ctypes.c_uint64),
        ("rounds_total", ctypes.c_uint64),
        ("max_cell_count", ctypes.c_uint32),
        ("reserved", ctypes.c_uint32),
        ("kernel_launches", ctypes.c_uint64),
    ]


# name -> (restype, argtypes); every symbol include/particles_cuda.h declares.
_P = ctypes.c_void_p
_PP = ctypes.POINTER(ctypes.c_void_p)
_U64 = ctypes.c_uint64
_PD = ctypes.POINTER(ctypes.c_double)
_PF = ctypes.POINTER(ctypes.c_float)
_PU64 = ctypes.POINTER(ctypes.c_uint64)
_CFG = ctypes.POINTER(PsConfig)

SYMBOLS = {
    "ps_create": (ctypes.c_int, [_CFG, _PP]),
    "ps_destroy": (None, [_P]),
    "ps_upload": (ctypes.c_int, [_P, _P, _P]),
    "ps_upload_f32": (ctypes.c_int, [_P, _P, _P]),
    "ps_download": (ctypes.c_int, [_P, _P, _P]),
    "ps_download_f32": (ctypes.c_int, [_P, _P, _P]),
    "ps_download_acc": (ctypes.c_int, [_P, _P]),
    "ps_spawn": (ctypes.c_int, [_P, _U64, ctypes.c_double, ctypes.c_double]),
    "ps_export_rects": (ctypes.c_int, [_P, _P]),
    "ps_set_colors": (ctypes.c_int, [_P, _P]),
    "ps_rainbow_colors": (ctypes.c_int, [_P]),
    "ps_get_colors": (ctypes.c_int, [_P, _P]),
    "ps_render": (ctypes.c_int, [_P, ctypes.c_uint32, ctypes.c_uint32, _P]),
    "ps_step": (ctypes.c_int, [_P, _U64]),
    "ps_step_substeps": (ctypes.c_int, [_P, _U64]),
    "ps_frame_host": (ctypes.c_int, [_P, _P, _P, _P, _P, _U64]),
    "ps_frame_host_f32": (ctypes.c_int, [_P, _P, _P, _P, _P, _U64]),
    "ps_toggle_acc": (ctypes.c_int, [_P]),
    "ps_set_acc": (ctypes.c_int, [_P, ctypes.c_int]),
    "ps_get_acc": (ctypes.c_int, [_P]),
    "ps_sync": (ctypes.c_int, [_P]),
    "ps_set_stream": (ctypes.c_int, [_P, _P]),
    "ps_grid_shape": (ctypes.c_int, [_P, _PU64, _PU64]),
    "ps_dt": (ctypes.c_double, [_P]),
    "ps_capture_pairs": (ctypes.c_int, [_P, ctypes.c_int]),
    "ps_contact_pairs": (ctypes.c_int, [_P, _P, _U64, _PU64]),
    "ps_get_stats": (ctypes.c_int, [_P, ctypes.POINTER(PsStats)]),
    "ps_device_state": (ctypes.c_int, [_P, _PP, _PP, _PP, _PP]),
    "ps_last_error": (ctypes.c_char_p, [_P]),
    "ps_status_string": (ctypes.c_char_p, [ctypes.c_int]),
    "ps_abi_version": (ctypes.c_int, []),
    "ps_host_alloc": (ctypes.c_int, [_PP, _U64]),
    "ps_host_free": (ctypes.c_int, [_P]),
    "ps_integrate_bounce_f32": (ctypes.c_int, [_CFG, ctypes.c_int, _P, _P, _P, _P, _U64, _U64, _P]),
    "ps_integrate_bounce_f64": (ctypes.c_int, [_CFG, ctypes.c_int, _P, _P, _P, _P, _U64, _U64, _P]),
    "ps_compute_start_array": (ctypes.c_int, [_P, _P, _U64, _P]),
    "ps_compute_start_array_static": (ctypes.c_int, [_P, _P, _U64, _P]),
    "ps_nccl_unique_id": (ctypes.c_int, [_P]),
    "ps_slab_create": (ctypes.c_int, [_CFG, ctypes.c_int, ctypes.c_int, _P, _PP]),
    "ps_slab_destroy": (None, [_P]),
    "ps_slab_upload_global": (ctypes.c_int, [_P, _P, _P]),
    "ps_slab_spawn": (ctypes.c_int, [_P, _U64, ctypes.c_double, ctypes.c_double]),
    "ps_slab_step": (ctypes.c_int, [_P, _U64]),
    "ps_slab_step_substeps": (ctypes.c_int, [_P, _U64]),
    "ps_slab_set_acc": (ctypes.c_int, [_P, ctypes.c_int]),
    "ps_slab_sync": (ctypes.c_int, [_P]),
    "ps_slab_set_stream": (ctypes.c_int, [_P, _P]),
    "ps_slab_set_timing": (ctypes.c_int, [_P, ctypes.c_int]),
    "ps_slab_get_timing": (ctypes.c_int, [_P, _PD, ctypes.c_uint32, _PU64]),
    "ps_slab_columns": (ctypes.c_int, [_P, _PU64, _PU64]),
    "ps_slab_owned_count": (ctypes.c_int, [_P, _PU64]),
    "ps_slab_download_owned": (ctypes.c_int, [_P, _P, _P, _P, _U64, _PU64]),
    "ps_slab_get_stats": (ctypes.c_int, [_P, ctypes.POINTER(PsStats)]),
    "ps_slab_last_error": (ctypes.c_char_p, [_P]),
    "ps_slabgroup_create": (ctypes.c_int, [_CFG, ctypes.c_int, _PP]),
    "ps_slabgroup_destroy": (None, [_P]),
    "ps_slabgroup_upload_global": (ctypes.c_int, [_P, _P, _P]),
    "ps_slabgroup_set_acc": (ctypes.c_int, [_P, ctypes.c_int]),
    "ps_slabgroup_step_substeps": (ctypes.c_int, [_P, _U64]),
    "ps_slabgroup_owned_count": (ctypes.c_int, [_P, ctypes.c_int, _PU64]),
    "ps_slabgroup_download_owned": (ctypes.c_int, [_P, ctypes.c_int, _P, _P, _P, _U64, _PU64]),
    "ps_slabgroup_get_stats": (ctypes.c_int, [_P, ctypes.c_int, ctypes.POINTER(PsStats)]),
    "ps_slabgroup_last_error": (ctypes.c_char_p, [_P]),
}

# not part of the stable ABI (tuning hooks used by bench.py --sweep)
DEBUG_SYMBOLS = {
    "ps_debug_set_integrate_variant": (ctypes.c_int, [_P, ctypes.c_int]),
    "ps_debug_set_transfer_mode": (ctypes.c_int, [_P, ctypes.c_int]),
}

_lib = None


def lib_path() -> str:
    return _build.LIB_PATH


def load(build_if_missing: bool = True):
    """dlopen libparticles_cuda.so and bind every declared symbol."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        if not build_if_missing:
            raise RuntimeError(f"{path} is missing; run `python __graft_entry__.py` (there is no CPU fallback)")
        _build.build_lib()
    L = ctypes.CDLL(path)
    for table in (SYMBOLS, DEBUG_SYMBOLS):
        for name, (res, args) in table.items():
            fn = getattr(L, name)  # AttributeError here = header and library out of sync
            fn.restype = res
            fn.argtypes = args
    _lib = L
    return L


class PsError(RuntimeError):
    def __init__(self, status: int, text: str):
        super().__init__(f"[{status}] {text}")
        self.status = status
