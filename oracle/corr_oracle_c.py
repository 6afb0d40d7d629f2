"""ctypes front end of oracle/corr_oracle.c (OpenMP).  TEST INFRASTRUCTURE — not shipped.

Same functions, layouts and reference citations as oracle/corr_oracle.py; exists so that the timed
CPU baseline uses every host core.  `build()` compiles it with gcc into oracle/_build/ (git-ignored;
it travels to the GPU box with the snapshot, and is rebuilt there on demand).
"""

from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(_HERE, "corr_oracle.c")
LIB = os.path.join(_HERE, "_build", "libcorr_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(SRC):
        return LIB
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    cmd = ["gcc", "-O3", "-fopenmp", "-ffp-contract=off", "-shared", "-fPIC", "-o", LIB, SRC, "-lm"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"gcc failed:\n{res.stderr}")
    return LIB


def _load():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
    return _lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def avg_pool_2x2(x: np.ndarray) -> np.ndarray:
    """(n, h, w, 1) -> (n, h//2, w//2, 1); model.py:444."""
    x = _f32(x)
    n, h, w, c = x.shape
    assert c == 1
    out = np.empty((n, h // 2, w // 2, 1), np.float32)
    _load().oracle_avg_pool2(_ptr(x), _ptr(out), ctypes.c_long(n), h, w)
    return out


def build_pyramid(fmap1: np.ndarray, fmap2: np.ndarray, num_levels: int = 4):
    """model.py:418-446: the volume is a float32 BLAS matmul (NumPy), pooling is the C routine."""
    from . import corr_oracle as O

    assert fmap1.shape == fmap2.shape, "Input feature maps should have the same shapes"
    mfs = O.min_fmap_size(num_levels)
    assert not any(s < mfs for s in fmap1.shape[-3:-1]), "Feature maps are too small to be down-sampled by the correlation pyramid."
    vol = O.corr_volume(fmap1, fmap2)
    b, h, w, _, _, c = vol.shape
    vol = vol.reshape(b * h * w, h, w, c)
    pyr = [vol]
    for _ in range(num_levels - 1):
        pyr.append(avg_pool_2x2(pyr[-1]))
    return pyr


def index_pyramid(pyramid, centroids_coords: np.ndarray, radius: int = 4) -> np.ndarray:
    """model.py:448-470."""
    levels = [_f32(p) for p in pyramid]
    coords = _f32(centroids_coords)
    b, hq, wq, _ = coords.shape
    q = b * hq * wq
    assert levels[0].shape[0] == q
    h, w = levels[0].shape[1:3]
    s = 2 * radius + 1
    out = np.empty((b, hq, wq, len(levels) * s * s), np.float32)
    arr = (ctypes.c_void_p * len(levels))(*[lv.ctypes.data for lv in levels])
    _load().oracle_index_pyramid(arr, len(levels), _ptr(coords), _ptr(out), ctypes.c_long(q), h, w, radius)
    return out


def upsample_flow(flow: np.ndarray, up_mask: np.ndarray | None = None, factor: int = 8) -> np.ndarray:
    """model.py:69-98 (mask branch in C; the bilinear branch defers to the NumPy oracle)."""
    if up_mask is None or factor != 8 or flow.shape[-1] != 2:
        from . import corr_oracle as O

        return O.upsample_flow(flow, up_mask, factor)
    flow, up_mask = _f32(flow), _f32(up_mask)
    b, h, w, _ = flow.shape
    assert up_mask.shape == (b, h, w, 576)
    out = np.empty((b, 8 * h, 8 * w, 2), np.float32)
    _load().oracle_convex_upsample(_ptr(flow), _ptr(up_mask), _ptr(out), b, h, w)
    return out
