"""The C-ABI library: it loads, exports every symbol include/raftk.h declares, and rejects bad
arguments before any launch.  No GPU, no compute calls."""

import ctypes
import os
import re

from conftest import ROOT


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "raftk.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(raftk_\w+)\s*\(", text)))


def test_header_symbols_exported_and_bound(lib):
    from jax_raft_b200 import _lib

    names = _declared_symbols()
    assert len(names) >= 8
    for name in names:
        assert hasattr(lib, name), f"{name} declared in raftk.h but not exported by libraftk.so"
        assert name in _lib.SIGNATURES, f"{name} has no ctypes signature"
    assert set(_lib.SIGNATURES) == set(names)


def test_version_and_error_strings(lib):
    assert lib.raftk_version() >= 1000
    assert lib.raftk_error_string(0) == b"ok"
    assert b"too small" in lib.raftk_error_string(-3)


def test_rejections_happen_before_launch(lib):
    from jax_raft_b200._lib import RaftkError, check
    import pytest

    null = ctypes.c_void_p(0)
    arr = (ctypes.c_void_p * 4)()
    # NULL pointers
    assert lib.raftk_convex_upsample(null, null, null, 1, 8, 8, null) == -1
    assert lib.raftk_corr_lookup(arr, 4, null, null, 16, 16, 16, 4, null) == -1
    # bad sizes
    assert lib.raftk_convex_upsample(ctypes.c_void_p(16), ctypes.c_void_p(16), ctypes.c_void_p(16), 1, 0, 8, null) == -2
    assert lib.raftk_corr_pyramid_simt(ctypes.c_void_p(16), ctypes.c_void_p(16), arr, 9, 1, 4, 16, 16, 8, null) == -2
    # model.py:431-436: h, w >= 2 * 2^(L-1)
    for i in range(4):
        arr[i] = 256
    assert lib.raftk_corr_pyramid_simt(ctypes.c_void_p(16), ctypes.c_void_p(16), arr, 4, 1, 4, 15, 16, 8, null) == -3
    assert lib.raftk_corr_pyramid(ctypes.c_void_p(256), ctypes.c_void_p(256), arr, 4, 1, 4, 15, 16, 8, null, 0, null) == -3
    # misaligned pointer
    assert lib.raftk_convex_upsample(ctypes.c_void_p(4), ctypes.c_void_p(8), ctypes.c_void_p(16), 1, 8, 8, null) == -6
    with pytest.raises(RaftkError, match="too small"):
        check("raftk_corr_pyramid", -3)
