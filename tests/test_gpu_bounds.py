"""Out-of-bounds guard: every kernel family writes exactly its output and nothing around it.
(compute-sanitizer is closed on this pool, so the outputs are carved out of a larger buffer whose
margins hold a sentinel that must survive.)  Shapes are ragged on purpose: odd sizes, row pitches
TMA cannot address, partial tiles, batch > 1."""

import ctypes

import pytest
import torch

import jax_raft_b200 as J
from jax_raft_b200 import _lib

pytestmark = pytest.mark.gpu
SENTINEL = -12345.5
MARGIN = 4096  # floats on each side


def _carve(n, dev, align=64):
    buf = torch.full((n + 2 * MARGIN,), SENTINEL, dtype=torch.float32, device=dev)
    assert MARGIN % align == 0
    return buf, buf[MARGIN:MARGIN + n]


def _intact(buf, n):
    return bool((buf[:MARGIN] == SENTINEL).all()) and bool((buf[MARGIN + n:] == SENTINEL).all())


@pytest.mark.parametrize("shape", [(1, 16, 16, 64), (2, 19, 21, 128), (1, 23, 31, 256), (3, 24, 40, 128), (1, 46, 62, 128)])
@pytest.mark.parametrize("entry", ["raftk_corr_pyramid", "raftk_corr_pyramid_simt"])
def test_pyramid_writes_stay_inside(cuda, lib, shape, entry):
    b, h, w, d = shape
    nq = h * w
    gen = torch.Generator(device=cuda).manual_seed(0)
    f1 = torch.randn((b, h, w, d), device=cuda, generator=gen)
    f2 = torch.randn((b, h, w, d), device=cuda, generator=gen)
    bufs, views = [], []
    for l in range(4):
        n = b * nq * (h >> l) * (w >> l)
        buf, v = _carve(n, cuda)
        bufs.append((buf, n))
        views.append(v)
    stream = ctypes.c_void_p(torch.cuda.current_stream(cuda).cuda_stream)
    with torch.cuda.device(cuda):
        if entry == "raftk_corr_pyramid":
            nbytes = lib.raftk_corr_pyramid_scratch_bytes(b, nq, h, w, d)
            sbuf, scratch = _carve(nbytes // 4 + 64, cuda)
            rc = lib.raftk_corr_pyramid(f1.data_ptr(), f2.data_ptr(), _lib.ptr_array(views), 4, b, nq, h, w, d,
                                        scratch.data_ptr(), nbytes, stream)
        else:
            rc = lib.raftk_corr_pyramid_simt(f1.data_ptr(), f2.data_ptr(), _lib.ptr_array(views), 4, b, nq, h, w, d, stream)
        _lib.check(entry, rc)
        torch.cuda.synchronize(cuda)
    for (buf, n), v in zip(bufs, views):
        assert _intact(buf, n)
        assert bool((v != SENTINEL).all()), "an output cell was never written"
    if entry == "raftk_corr_pyramid":
        assert _intact(sbuf, nbytes // 4 + 64)


@pytest.mark.parametrize("shape,r", [((1, 16, 16), 4), ((2, 19, 21), 3), ((1, 24, 40), 4), ((1, 46, 62), 3)])
def test_lookup_and_altcorr_writes_stay_inside(cuda, lib, shape, r):
    b, h, w = shape
    d = 128
    q = b * h * w
    gen = torch.Generator(device=cuda).manual_seed(1)
    pyr = [torch.randn((q, h >> l, w >> l, 1), device=cuda, generator=gen) for l in range(4)]
    coords = J.make_coords_grid(b, h, w, device=cuda) + (torch.rand((b, h, w, 2), device=cuda, generator=gen) - 0.5) * 30
    c = 4 * (2 * r + 1) ** 2
    stream = ctypes.c_void_p(torch.cuda.current_stream(cuda).cuda_stream)
    buf, out = _carve(q * c, cuda)
    with torch.cuda.device(cuda):
        rc = lib.raftk_corr_lookup(_lib.ptr_array(pyr), 4, coords.data_ptr(), out.data_ptr(), q, h, w, r, stream)
        _lib.check("raftk_corr_lookup", rc)
        torch.cuda.synchronize(cuda)
    assert _intact(buf, q * c) and bool((out != SENTINEL).all())

    f1 = torch.randn((b, h, w, d), device=cuda, generator=gen)
    f2 = torch.randn((b, h, w, d), device=cuda, generator=gen)
    alt = J.AltCorrBlock(4, r)
    _, f2_levels, f1_split, f2_split = alt.build_pyramid(f1, f2)
    buf, out = _carve(q * c, cuda)
    with torch.cuda.device(cuda):
        rc = lib.raftk_altcorr_lookup(f1.data_ptr(), _lib.ptr_array(f2_levels), 4, coords.data_ptr(), out.data_ptr(),
                                      b, h * w, h, w, d, r, stream)
        _lib.check("raftk_altcorr_lookup", rc)
        torch.cuda.synchronize(cuda)
    assert _intact(buf, q * c) and bool((out != SENTINEL).all())
    # the tensor-core entry point (+-15 px per-pixel flow: box and scalar paths both occur)
    buf, out = _carve(q * c, cuda)
    with torch.cuda.device(cuda):
        rc = lib.raftk_altcorr_lookup_tc(f1.data_ptr(), _lib.ptr_array(f2_levels), f1_split.data_ptr(),
                                         _lib.ptr_array(f2_split), 4, coords.data_ptr(), out.data_ptr(), b, h * w, h, w, d, r,
                                         stream)
        _lib.check("raftk_altcorr_lookup_tc", rc)
        torch.cuda.synchronize(cuda)
    assert _intact(buf, q * c) and bool((out != SENTINEL).all())


@pytest.mark.parametrize("shape", [(1, 1, 1), (2, 7, 5), (1, 46, 62), (3, 55, 128)])
def test_upsample_writes_stay_inside(cuda, lib, shape):
    b, h, w = shape
    gen = torch.Generator(device=cuda).manual_seed(2)
    flow = torch.randn((b, h, w, 2), device=cuda, generator=gen)
    mask = torch.randn((b, h, w, 576), device=cuda, generator=gen)
    n = b * 8 * h * 8 * w * 2
    buf, out = _carve(n, cuda)
    with torch.cuda.device(cuda):
        rc = lib.raftk_convex_upsample(flow.data_ptr(), mask.data_ptr(), out.data_ptr(), b, h, w,
                                       ctypes.c_void_p(torch.cuda.current_stream(cuda).cuda_stream))
        _lib.check("raftk_convex_upsample", rc)
        torch.cuda.synchronize(cuda)
    assert _intact(buf, n) and bool((out != SENTINEL).all())
