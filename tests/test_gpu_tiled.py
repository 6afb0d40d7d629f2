"""The tiled pyramid path (include/raftk.h "Tiled pyramid"): raftk_corr_pyramid_tiled, raftk_corr_lookup_tiled(_update),
raftk_pyramid_tile / _untile — the product path since round 2.  Same bars as the row-major entry points: the lookup
bit-exact against the float32 oracle (and therefore equal to the row-major kernel), pooling bounds exact, correlation
values within 1e-3 relative; plus the layout contract itself (cell addresses, zero padding)."""

import ctypes

import numpy as np
import pytest
import torch

import jax_raft_b200 as J
from jax_raft_b200 import _lib
from jax_raft_b200.corr import TiledPyramid
from oracle import corr_oracle as O

pytestmark = pytest.mark.gpu
F32 = np.float32


def _tiled_index(y, x, tiles_x):
    return ((y >> 2) * tiles_x + (x >> 2)) * 16 + (y & 3) * 4 + (x & 3)


def _assert_bit_exact(got, ref):
    assert got.shape == ref.shape
    bad = ~((got == ref) | (np.isnan(got) & np.isnan(ref)))
    assert not bad.any(), f"{bad.sum()} of {bad.size} values differ; max |d| = {np.abs(got - ref)[bad].max()}"


def _lookup_tiled(pyr_np, coords_np, r, dev):
    cb = J.CorrBlock(len(pyr_np), r)
    tp = TiledPyramid.from_dense([torch.from_numpy(p).to(dev) for p in pyr_np])
    out = cb.index_pyramid(tp, torch.from_numpy(coords_np).to(dev))
    torch.cuda.synchronize()
    return out.cpu().numpy()


# ---------------------------------------------------------------------------------------------------------
# layout
# ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("shape", [(3, 16, 16), (2, 19, 21), (1, 46, 62), (2, 55, 128)])
def test_tile_untile_round_trip_and_addresses(cuda, lib, shape):
    q, h, w = shape
    rng = np.random.default_rng(h)
    dense = [rng.standard_normal((q, h >> l, w >> l, 1)).astype(F32) for l in range(4)]
    tp = TiledPyramid.from_dense([torch.from_numpy(d).to(cuda) for d in dense])
    for l, (t, d) in enumerate(zip(tp.levels, dense)):
        lh, lw = h >> l, w >> l
        ty, tx = (lh + 3) // 4, 2 * ((lw + 7) // 8)
        assert tuple(t.shape) == (q, ty, tx, 4, 4)
        assert ty * tx * 16 == lib.raftk_pyramid_tiled_slab_floats(h, w, l)
        flat = t.cpu().numpy().reshape(q, -1)
        ys, xs = np.meshgrid(np.arange(lh), np.arange(lw), indexing="ij")
        # every cell sits at the documented address ...
        assert np.array_equal(flat[:, _tiled_index(ys, xs, tx)], d[..., 0])
        # ... and everything else in the slab is zero
        mask = np.ones(flat.shape[1], bool)
        mask[_tiled_index(ys, xs, tx).ravel()] = False
        assert not flat[:, mask].any()
    for got, want in zip(tp.dense(), dense):
        assert np.array_equal(got.cpu().numpy(), want)


# ---------------------------------------------------------------------------------------------------------
# lookup
# ---------------------------------------------------------------------------------------------------------
def test_tap_order_fixture(cuda, lib, golden):
    g = golden("lookup_tap_order")
    h, w, r = int(g["h"]), int(g["w"]), int(g["radius"])
    pyr = []
    for l in range(4):
        yy, xx = np.meshgrid(np.arange(h >> l), np.arange(w >> l), indexing="ij")
        pyr.append(np.tile((l * 10000 + yy * 100 + xx).astype(F32)[None, :, :, None], (2, 1, 1, 1)))
    _assert_bit_exact(_lookup_tiled(pyr, g["coords"], r, cuda), g["out"])


@pytest.mark.parametrize("r", [3, 4])
def test_border_fixture(cuda, lib, golden, r):
    g = golden(f"lookup_borders_r{r}")
    pyr = [g[f"level{l}"] for l in range(4)]
    _assert_bit_exact(_lookup_tiled(pyr, g["coords"], r, cuda), g["out"])


@pytest.mark.parametrize("shape", [(1, 16, 16), (2, 19, 21), (1, 23, 31), (1, 46, 62), (3, 24, 40)])
@pytest.mark.parametrize("r", [3, 4])
@pytest.mark.parametrize("seed", [0, 1])
def test_random_against_oracle(cuda, lib, shape, r, seed):
    b, h, w = shape
    rng = np.random.default_rng(seed)
    q = b * h * w
    pyr = [rng.standard_normal((q, h >> l, w >> l, 1)).astype(F32) for l in range(4)]
    coords = O.make_coords_grid(b, h, w) + rng.uniform(-8, 8, (b, h, w, 2)).astype(F32)
    far = rng.random((b, h, w)) < 0.02
    coords[far] += rng.choice([-1, 1], size=(far.sum(), 2)).astype(F32) * 500
    _assert_bit_exact(_lookup_tiled(pyr, coords, r, cuda), O.index_pyramid(pyr, coords, r))


def _awkward_coords(rng, b, h, w):
    """As tests/test_gpu_fuzz.py: integers, halves, level borders, values just below an integer (fl(c + d) rounds
    up to an integer: the case the kernel's re-anchored weights exist for), far positions, and non-finite ones."""
    c = O.make_coords_grid(b, h, w) + rng.uniform(-9, 9, (b, h, w, 2)).astype(F32)
    kind = rng.integers(0, 12, (b, h, w))
    c[kind == 0] = np.round(c[kind == 0])
    c[kind == 1] = np.floor(c[kind == 1]) + F32(0.5)
    edge = np.array([-1, 0, w - 1, w, h - 1, h, -0.0], F32)
    m = kind == 2
    c[m] = rng.choice(edge, size=(m.sum(), 2))
    m = kind == 3
    c[m] = np.nextafter(np.round(c[m]), F32(-np.inf)).astype(F32)
    m = kind == 4
    c[m] = (rng.choice([-1, 1], size=(m.sum(), 2)) * 10.0 ** rng.uniform(2, 9, (m.sum(), 2))).astype(F32)
    m = kind == 5
    c[m] = np.nextafter(np.round(c[m] / 8) * 8, F32(-np.inf)).astype(F32)  # just below an integer on every level
    return c


@pytest.mark.parametrize("seed,b,h,w,levels,r", [(s, *c) for s, c in enumerate([
    (1, 8, 8, 1, 0), (1, 8, 9, 2, 1), (2, 9, 8, 2, 2), (1, 16, 16, 4, 3), (1, 16, 20, 4, 4), (3, 17, 19, 4, 4),
    (1, 21, 18, 3, 5), (1, 33, 40, 5, 3), (1, 32, 36, 4, 7), (2, 24, 24, 4, 3), (1, 40, 72, 4, 4), (1, 64, 33, 6, 2),
    (1, 13, 50, 3, 4), (4, 16, 16, 4, 4), (1, 30, 31, 4, 3), (1, 16, 64, 4, 4), (5, 16, 17, 4, 4), (1, 29, 27, 4, 4),
])])
def test_lookup_fuzz(cuda, lib, seed, b, h, w, levels, r):
    rng = np.random.default_rng(1000 + seed)
    q = b * h * w
    pyr = [rng.standard_normal((q, h >> l, w >> l, 1)).astype(F32) for l in range(levels)]
    coords = _awkward_coords(rng, b, h, w)
    if seed % 3 == 0:
        coords[0, 0, 0] = [np.nan, 1.0]
        coords[0, 1, 1] = [2.0, np.inf]
        coords[0, 2, 2] = [-np.inf, -np.inf]
    with np.errstate(invalid="ignore"):
        ref = O.index_pyramid(pyr, coords, r)
    _assert_bit_exact(_lookup_tiled(pyr, coords, r, cuda), ref)


def test_sintel_shape_full_size_and_row_major_agreement(cuda, lib):
    """BASELINE config 2 (440x1024 -> 55x128, N = 7040): oracle on the same inputs, and the row-major kernel."""
    rng = np.random.default_rng(1)
    b, h, w, r = 1, 55, 128, 4
    q = b * h * w
    pyr = [rng.standard_normal((q, h >> l, w >> l, 1)).astype(F32) for l in range(4)]
    coords = O.make_coords_grid(b, h, w) + rng.uniform(-8, 8, (b, h, w, 2)).astype(F32)
    got = _lookup_tiled(pyr, coords, r, cuda)
    _assert_bit_exact(got, O.index_pyramid(pyr, coords, r))
    rm = J.CorrBlock(4, r).index_pyramid([torch.from_numpy(p).to(cuda) for p in pyr], torch.from_numpy(coords).to(cuda))
    _assert_bit_exact(got, rm.cpu().numpy())


@pytest.mark.parametrize("b", [1, 3, 8])
def test_tiled_equals_row_major_at_batch_size(cuda, lib, b):
    """Sizes the oracle is not run on: the two kernels (different layouts, different tap schedules) agree bit for
    bit, a zero pyramid gives zeros, and group tails (Q not a multiple of the queries per warp) are covered."""
    h, w, r = 55, 128, 4
    q = b * h * w
    gen = torch.Generator(device=cuda).manual_seed(b)
    pyr = [torch.randn((q, h >> l, w >> l, 1), device=cuda, generator=gen) for l in range(4)]
    coords = J.make_coords_grid(b, h, w, device=cuda) + (torch.rand((b, h, w, 2), device=cuda, generator=gen) - 0.5) * 16
    cb = J.CorrBlock(4, r)
    tp = TiledPyramid.from_dense(pyr)
    a, c = cb.index_pyramid(tp, coords), cb.index_pyramid(pyr, coords)
    assert bool(((a == c) | (a.isnan() & c.isnan())).all())
    assert not cb.index_pyramid(TiledPyramid.from_dense([torch.zeros_like(t) for t in pyr]), coords).any()


@pytest.mark.parametrize("shape,r", [((1, 16, 24), 4), ((2, 19, 21), 3), ((1, 55, 128), 4)])
def test_lookup_with_fused_coordinate_update(cuda, lib, shape, r):
    """raftk_corr_lookup_tiled_update == raftk_flow_update followed by the lookup, bit for bit, including the
    in-place coordinates, the flow and its pitched copy."""
    b, h, w = shape
    gen = torch.Generator(device=cuda).manual_seed(h + r)
    q = b * h * w
    pyr = [torch.randn((q, h >> l, w >> l, 1), device=cuda, generator=gen) for l in range(4)]
    tp = TiledPyramid.from_dense(pyr)
    c0 = J.make_coords_grid(b, h, w, device=cuda)
    c1 = c0 + (torch.rand((b, h, w, 2), device=cuda, generator=gen) - 0.5) * 10
    c1[0, 1, 1] = 3.0e6  # a far query
    delta = torch.randn((b, h, w, 8), device=cuda, generator=gen)  # padded flow-head output
    bias = torch.randn((8,), device=cuda, generator=gen)
    cb = J.CorrBlock(4, r)
    ref_c1 = c1 + (delta[..., :2] + bias[:2])
    ref_flow = ref_c1 - c0
    ref_out = cb.index_pyramid(pyr, ref_c1)
    flow = torch.full((b, h, w, 2), 9.0, device=cuda)
    flow8 = torch.zeros((b, h, w, 8), device=cuda)
    got_c1 = c1.clone()
    out = cb.index_pyramid_update(tp, c0, got_c1, delta, bias, flow, flow8)
    assert torch.equal(got_c1, ref_c1) and torch.equal(flow, ref_flow)
    assert torch.equal(flow8[..., :2], ref_flow) and not flow8[..., 2:].any()
    assert bool((out == ref_out).all())


def test_lookup_writes_exactly_its_output(cuda, lib):
    """Sentinels around the output of the bulk-copy store path (group tail: Q = 5 queries)."""
    b, h, w, r = 1, 16, 16, 4
    q = 5
    gen = torch.Generator(device=cuda).manual_seed(0)
    pyr = [torch.randn((q, h >> l, w >> l, 1), device=cuda, generator=gen) for l in range(4)]
    tp = TiledPyramid.from_dense(pyr)
    coords = torch.rand((q, 2), device=cuda, generator=gen) * 14
    buf = torch.full((q * 324 + 64,), 7.5, device=cuda)
    out = buf[32:32 + q * 324]
    rc = lib.raftk_corr_lookup_tiled(_lib.ptr_array(tp.levels), 4, coords.data_ptr(), out.data_ptr(), q, h, w, r,
                                     _lib.current_stream_ptr(cuda))
    assert rc == 0
    torch.cuda.synchronize()
    assert bool((buf[:32] == 7.5).all()) and bool((buf[32 + q * 324:] == 7.5).all())
    ref = J.CorrBlock(4, r).index_pyramid(pyr, coords.view(1, 1, q, 2)).view(-1)
    assert bool((out == ref).all())


# ---------------------------------------------------------------------------------------------------------
# volume + pyramid written tiled by the GEMM epilogue
# ---------------------------------------------------------------------------------------------------------
def _rel_err(got, ref64):
    rms = np.sqrt(np.mean(ref64 ** 2))
    return float(np.max(np.abs(got - ref64) / np.maximum(np.abs(ref64), rms)))


def _build_tiled_raw(f1, f2, dev, levels=4, nq_rows=None):
    """raftk_corr_pyramid_tiled through the C ABI into NaN-prefilled slabs (so unwritten cells show)."""
    lib = _lib.load()
    b, h, w, d = f2.shape
    t1 = torch.from_numpy(f1).to(dev)
    t2 = torch.from_numpy(f2).to(dev)
    nq = t1.shape[1] * t1.shape[2]
    tp = TiledPyramid.empty(b * nq, h, w, levels, dev)
    for t in tp.levels:
        t.fill_(float("nan"))
    nbytes = lib.raftk_corr_pyramid_scratch_bytes(b, nq, h, w, d)
    scratch = torch.empty((max(nbytes, 1),), dtype=torch.uint8, device=dev)
    rc = lib.raftk_corr_pyramid_tiled(t1.data_ptr(), t2.data_ptr(), _lib.ptr_array(tp.levels), levels, b, nq, h, w, d,
                                      scratch.data_ptr(), nbytes, _lib.current_stream_ptr(dev))
    assert rc == 0, rc
    torch.cuda.synchronize()
    return tp


@pytest.mark.parametrize("shape", [(1, 16, 16, 128), (2, 17, 23, 64), (1, 46, 62, 128), (1, 24, 40, 256), (1, 33, 65, 64),
                                   (1, 19, 21, 192)])
def test_pyramid_tiled_values_padding_and_row_major_agreement(cuda, lib, shape):
    b, h, w, d = shape
    rng = np.random.default_rng(d + h)
    f1 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    f2 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    tp = _build_tiled_raw(f1, f2, cuda)
    ref64 = O.build_pyramid(f1, f2, dtype=np.float64)
    rm = J.CorrBlock(4, 4, layout="rowmajor").build_pyramid(torch.from_numpy(f1).to(cuda), torch.from_numpy(f2).to(cuda))
    for l, got in enumerate(tp.dense()):
        got = got.cpu().numpy()
        assert got.shape == ref64[l].shape
        assert _rel_err(got, ref64[l]) <= 1e-3, l
        # same MMAs, same pooling order: the row-major kernel's bits
        assert np.array_equal(got, rm[l].cpu().numpy()), l
        # layout contract: tiles that hold a valid cell are fully written, their padding is zero
        lh, lw = h >> l, w >> l
        t = tp.levels[l].cpu().numpy()  # (Q, ty, tx, 4, 4)
        _, ty, tx, _, _ = t.shape
        for tyi in range(ty):
            for txi in range(tx):
                if txi * 4 >= lw:
                    continue  # the all-padding tile that rounds a tile row up to a whole 128-byte line
                tile = t[:, tyi, txi]
                assert not np.isnan(tile).any(), (l, tyi, txi)
                yy, xx = np.meshgrid(tyi * 4 + np.arange(4), txi * 4 + np.arange(4), indexing="ij")
                pad = (yy >= lh) | (xx >= lw)
                assert not tile[:, pad].any(), (l, tyi, txi)


def test_pyramid_tiled_query_row_band_and_small_nq(cuda, lib):
    """Row-sharded mode, including a band of <= 128 queries (single-CTA tiles)."""
    rng = np.random.default_rng(9)
    b, h, w, d = 2, 18, 20, 128
    f1 = rng.standard_normal((b, h, w, d)).astype(F32)
    f2 = rng.standard_normal((b, h, w, d)).astype(F32)
    full = [t.cpu().numpy() for t in _build_tiled_raw(f1, f2, cuda).dense()]
    for y0, y1 in ((5, 12), (3, 6)):
        band = [t.cpu().numpy() for t in _build_tiled_raw(np.ascontiguousarray(f1[:, y0:y1]), f2, cuda).dense()]
        for l in range(4):
            ref = full[l].reshape(b, h, w, h >> l, w >> l, 1)[:, y0:y1].reshape(band[l].shape)
            assert np.array_equal(band[l], ref)


def test_pyramid_tiled_other_level_counts_and_fallbacks(cuda, lib):
    rng = np.random.default_rng(3)
    f1 = rng.standard_normal((1, 16, 24, 64)).astype(F32)
    f2 = rng.standard_normal((1, 16, 24, 64)).astype(F32)
    for levels in (1, 2, 3):
        ref = O.build_pyramid(f1, f2, num_levels=levels, dtype=np.float64)
        for l, got in enumerate(_build_tiled_raw(f1, f2, cuda, levels=levels).dense()):
            assert _rel_err(got.cpu().numpy(), ref[l]) <= 1e-3
    # configurations the tiled epilogue does not cover are refused by the C ABI ...
    arr = (ctypes.c_void_p * 5)(*([256] * 5))
    assert lib.raftk_corr_pyramid_tiled(ctypes.c_void_p(256), ctypes.c_void_p(256), arr, 5, 1, 1024, 32, 32, 64,
                                        ctypes.c_void_p(256), 1 << 30, None) == -4
    assert lib.raftk_corr_pyramid_tiled(ctypes.c_void_p(256), ctypes.c_void_p(256), arr, 4, 1, 256, 16, 16, 96,
                                        ctypes.c_void_p(256), 1 << 30, None) == -4
    # ... and served by CorrBlock through the row-major kernels + raftk_pyramid_tile
    f1 = rng.standard_normal((1, 32, 32, 96)).astype(F32)
    f2 = rng.standard_normal((1, 32, 32, 96)).astype(F32)
    pyr = J.CorrBlock(5, 4).build_pyramid(torch.from_numpy(f1).to(cuda), torch.from_numpy(f2).to(cuda))
    assert isinstance(pyr, TiledPyramid) and len(pyr) == 5
    ref = O.build_pyramid(f1, f2, num_levels=5, dtype=np.float64)
    for l in range(5):
        assert _rel_err(pyr[l].cpu().numpy(), ref[l]) <= 1e-3


def test_sintel_shape_build_then_lookup(cuda, lib):
    """BASELINE config 2 end to end on the tiled path: build_pyramid -> index_pyramid equals the oracle's lookup of
    the GPU's own (untiled) pyramid bit for bit, and the pyramid meets the 1e-3 bar."""
    rng = np.random.default_rng(0)
    b, h, w, d, r = 1, 55, 128, 256, 4
    f1 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    f2 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    cb = J.CorrBlock(4, r)
    pyr = cb.build_pyramid(torch.from_numpy(f1).to(cuda), torch.from_numpy(f2).to(cuda))
    assert isinstance(pyr, TiledPyramid)
    ref = O.build_pyramid(f1, f2)
    dense = [p.cpu().numpy() for p in pyr]
    for l in range(4):
        assert _rel_err(dense[l], ref[l].astype(np.float64)) <= 1e-3
    coords = O.make_coords_grid(b, h, w) + rng.uniform(-8, 8, (b, h, w, 2)).astype(F32)
    got = cb.index_pyramid(pyr, torch.from_numpy(coords).to(cuda)).cpu().numpy()
    _assert_bit_exact(got, O.index_pyramid(dense, coords, r))
