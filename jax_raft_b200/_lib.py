"""ctypes binding of ``libraftk.so`` — the C ABI declared in ``include/raftk.h``.

There is no fallback: if the shared library is missing or a call returns non-zero this
module raises.  Device memory and streams come from torch (plumbing only); the library
itself knows nothing about torch.
"""

from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_size_t, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libraftk.so")

MAX_LEVELS = 6
STATIC_SCRATCH_BYTES = 16 << 20  # RAFTK_STATIC_SCRATCH_BYTES

# name -> (restype, argtypes); must list every symbol include/raftk.h declares.
_FloatPtrArray = POINTER(c_void_p)
SIGNATURES = {
    "raftk_version": (c_int, []),
    "raftk_error_string": (c_char_p, [c_int]),
    "raftk_corr_pyramid_scratch_bytes": (c_size_t, [c_int, c_int, c_int, c_int, c_int]),
    "raftk_corr_pyramid": (
        c_int,
        [c_void_p, c_void_p, _FloatPtrArray, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p,
         c_size_t, c_void_p],
    ),
    "raftk_corr_pyramid_static": (
        c_int,
        [c_void_p, c_void_p, _FloatPtrArray, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p],
    ),
    "raftk_corr_pyramid_simt": (
        c_int,
        [c_void_p, c_void_p, _FloatPtrArray, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p],
    ),
    "raftk_corr_lookup": (
        c_int,
        [_FloatPtrArray, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p],
    ),
    "raftk_grid_sample": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "raftk_altcorr_pool_fmap2": (c_int, [_FloatPtrArray, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "raftk_altcorr_lookup": (
        c_int,
        [c_void_p, _FloatPtrArray, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
         c_int, c_void_p],
    ),
    "raftk_altcorr_split": (c_int, [c_void_p, c_void_p, ctypes.c_longlong, c_int, c_float, c_void_p]),
    "raftk_altcorr_lookup_tc": (
        c_int,
        [c_void_p, _FloatPtrArray, c_void_p, _FloatPtrArray, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
         c_int, c_void_p],
    ),
    "raftk_gather_rows_split": (
        c_int,
        [_FloatPtrArray, _FloatPtrArray, c_int, ctypes.POINTER(c_int), c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p],
    ),
    "raftk_bias_act_nhwc": (c_int, [c_void_p, c_void_p, ctypes.c_longlong, c_int, c_int, c_void_p]),
    "raftk_instance_norm_scratch_bytes": (c_size_t, [c_int, c_int]),
    "raftk_instance_norm_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_float, c_int, c_void_p, c_size_t, c_void_p]),
    "raftk_instance_norm_residual_nhwc": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, ctypes.c_float, c_int, c_void_p, ctypes.c_size_t, c_void_p]),
    "raftk_add_relu": (c_int, [c_void_p, c_void_p, ctypes.c_longlong, c_void_p]),
    "raftk_gru_gate": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, ctypes.c_longlong, c_int, c_void_p]),
    "raftk_gru_update": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, ctypes.c_longlong, c_int, c_void_p]),
    "raftk_flow_update": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, ctypes.c_longlong, c_void_p]),
    "raftk_convex_upsample_bias": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "raftk_motion_scatter": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, ctypes.c_longlong, c_int, c_void_p]),
    "raftk_lookup_conv_weight_cols": (c_int, [c_int, c_int]),
    "raftk_corr_lookup_conv1x1": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "raftk_set_pdl": (c_int, [c_int]),
    "raftk_set_altcorr_mode": (c_int, [c_int]),
    "raftk_mask_upsample": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "raftk_corr_lookup_update": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                         c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "raftk_pyramid_tiled_slab_floats": (c_size_t, [c_int, c_int, c_int]),
    "raftk_corr_pyramid_tiled": (
        c_int,
        [c_void_p, c_void_p, _FloatPtrArray, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p,
         c_size_t, c_void_p],
    ),
    "raftk_corr_lookup_tiled": (
        c_int,
        [_FloatPtrArray, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p],
    ),
    "raftk_corr_lookup_tiled_update": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                               c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "raftk_pyramid_tile": (c_int, [_FloatPtrArray, _FloatPtrArray, c_int, ctypes.c_longlong, c_int, c_int, c_void_p]),
    "raftk_pyramid_untile": (c_int, [_FloatPtrArray, _FloatPtrArray, c_int, ctypes.c_longlong, c_int, c_int, c_void_p]),
    "raftk_convex_upsample": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "raftk_corr_pyramid_tiled_f16": (
        c_int,
        [c_void_p, c_void_p, _FloatPtrArray, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p,
         c_size_t, c_void_p],
    ),
    "raftk_corr_lookup_tiled_f16": (
        c_int,
        [_FloatPtrArray, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p],
    ),
    "raftk_corr_lookup_tiled_f16_update": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                                   c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "raftk_pyramid_tile_f16": (c_int, [_FloatPtrArray, _FloatPtrArray, c_int, ctypes.c_longlong, c_int, c_int, c_void_p]),
    "raftk_pyramid_untile_f16": (c_int, [_FloatPtrArray, _FloatPtrArray, c_int, ctypes.c_longlong, c_int, c_int, c_void_p]),
}

_lib = None

# Kernel launches issued through this binding (bench.py reports them as ``gpu_launches``).
LAUNCHES = {"calls": 0, "kernels": 0}


def count_launch(kernels: int) -> None:
    LAUNCHES["calls"] += 1
    LAUNCHES["kernels"] += kernels


class RaftkError(RuntimeError):
    """A libraftk call returned a non-zero status."""

    def __init__(self, fn: str, code: int, msg: str):
        super().__init__(f"{fn} failed with status {code}: {msg}")
        self.code = code


def load():
    """Load libraftk.so (once).  Raises ImportError with build instructions if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found. Build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `python -m jax_raft_b200.build`). There is no CPU or eager fallback."
        )
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here == header/library mismatch
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def check(fn_name: str, code: int) -> None:
    if code != 0:
        msg = load().raftk_error_string(code)
        raise RaftkError(fn_name, code, msg.decode() if msg else "?")


def ptr_array(tensors) -> "ctypes.Array":
    """Host array of device pointers for the ``levels`` arguments."""
    arr = (c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = t.data_ptr()
    return arr


def current_stream_ptr(device) -> int:
    import torch

    return torch.cuda.current_stream(device).cuda_stream
