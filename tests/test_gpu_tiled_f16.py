"""The tiled pyramid with float16 storage (include/raftk.h "Tiled pyramid with float16 storage"):
raftk_corr_pyramid_tiled_f16, raftk_corr_lookup_tiled_f16(_update), raftk_pyramid_tile_f16 / _untile_f16.

Bars.  Storage: every stored value is the round-to-nearest-even float16 of the float32 value the float32 tiled path
stores (checked bit for bit against that path), so correlation values stay within 1e-3 relative of the float64
oracle (tolerance below: 1e-3 * max(|ref|, rms(ref)), the bar of tests/test_gpu_pyramid.py); pooling bounds and
padding exact.  Lookup: BIT-EXACT against the float32 oracle evaluated on the stored (rounded) pyramid — window
indexing and tap arithmetic are those of the float32 kernels."""

import ctypes

import numpy as np
import pytest
import torch

import jax_raft_b200 as J
from jax_raft_b200 import _lib
from jax_raft_b200.corr import TiledPyramid
from oracle import corr_oracle as O

from test_gpu_tiled import _assert_bit_exact, _awkward_coords, _rel_err, _tiled_index

pytestmark = pytest.mark.gpu
F32 = np.float32


def _rounded(p):
    """What float16 storage keeps of a float32 array (round to nearest even), back in float32."""
    return p.astype(np.float16).astype(F32)


def _lookup_f16(pyr_np, coords_np, r, dev):
    cb = J.CorrBlock(len(pyr_np), r, storage="fp16")
    tp = TiledPyramid.from_dense([torch.from_numpy(p).to(dev) for p in pyr_np], storage="fp16")
    assert tp.storage == "fp16" and all(t.dtype == torch.float16 for t in tp.levels)
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
    dense = [(rng.standard_normal((q, h >> l, w >> l, 1)) * 3).astype(F32) for l in range(4)]
    tp = TiledPyramid.from_dense([torch.from_numpy(d).to(cuda) for d in dense], storage="fp16")
    for l, (t, d) in enumerate(zip(tp.levels, dense)):
        lh, lw = h >> l, w >> l
        ty, tx = (lh + 3) // 4, 2 * ((lw + 7) // 8)
        assert tuple(t.shape) == (q, ty, tx, 4, 4) and t.dtype == torch.float16
        flat = t.cpu().numpy().reshape(q, -1)
        ys, xs = np.meshgrid(np.arange(lh), np.arange(lw), indexing="ij")
        assert np.array_equal(flat[:, _tiled_index(ys, xs, tx)], d[..., 0].astype(np.float16))  # same element offsets
        mask = np.ones(flat.shape[1], bool)
        mask[_tiled_index(ys, xs, tx).ravel()] = False
        assert not flat[:, mask].any()
    for got, want in zip(tp.dense(), dense):
        assert got.dtype == torch.float32 and np.array_equal(got.cpu().numpy(), _rounded(want))


# ---------------------------------------------------------------------------------------------------------
# lookup
# ---------------------------------------------------------------------------------------------------------
def test_tap_order_fixture(cuda, lib, golden):
    """The tap-order fixture's values (l*10000 + y*100 + x, integers below 2048 per level-local term) survive
    float16 only at level 0; use y*16 + x, exact in float16, and let the oracle recompute the expectation."""
    g = golden("lookup_tap_order")
    h, w, r = int(g["h"]), int(g["w"]), int(g["radius"])
    pyr = []
    for l in range(4):
        yy, xx = np.meshgrid(np.arange(h >> l), np.arange(w >> l), indexing="ij")
        pyr.append(np.tile((l * 512 + yy * 16 + xx).astype(F32)[None, :, :, None], (2, 1, 1, 1)))
        assert np.array_equal(_rounded(pyr[-1]), pyr[-1])
    _assert_bit_exact(_lookup_f16(pyr, g["coords"], r, cuda), O.index_pyramid(pyr, g["coords"], r))


@pytest.mark.parametrize("r", [3, 4])
def test_border_fixture(cuda, lib, golden, r):
    g = golden(f"lookup_borders_r{r}")
    pyr = [_rounded(g[f"level{l}"]) for l in range(4)]
    _assert_bit_exact(_lookup_f16(pyr, g["coords"], r, cuda), O.index_pyramid(pyr, g["coords"], r))


@pytest.mark.parametrize("shape", [(1, 16, 16), (2, 19, 21), (1, 23, 31), (1, 46, 62), (3, 24, 40)])
@pytest.mark.parametrize("r", [3, 4])
def test_random_against_oracle(cuda, lib, shape, r):
    b, h, w = shape
    rng = np.random.default_rng(r)
    q = b * h * w
    pyr = [rng.standard_normal((q, h >> l, w >> l, 1)).astype(F32) for l in range(4)]
    coords = O.make_coords_grid(b, h, w) + rng.uniform(-8, 8, (b, h, w, 2)).astype(F32)
    far = rng.random((b, h, w)) < 0.02
    coords[far] += rng.choice([-1, 1], size=(far.sum(), 2)).astype(F32) * 500
    # the kernel reads float16(pyr): the oracle gets exactly those values
    _assert_bit_exact(_lookup_f16(pyr, coords, r, cuda), O.index_pyramid([_rounded(p) for p in pyr], coords, r))


@pytest.mark.parametrize("seed,b,h,w,levels,r", [(s, *c) for s, c in enumerate([
    (1, 8, 8, 1, 0), (1, 8, 9, 2, 1), (2, 9, 8, 2, 2), (1, 16, 16, 4, 3), (1, 16, 20, 4, 4), (3, 17, 19, 4, 4),
    (1, 21, 18, 3, 5), (1, 33, 40, 5, 3), (1, 32, 36, 4, 7), (2, 24, 24, 4, 3), (1, 40, 72, 4, 4), (1, 64, 33, 6, 2),
    (1, 13, 50, 3, 4), (4, 16, 16, 4, 4), (1, 30, 31, 4, 3), (1, 16, 64, 4, 4), (5, 16, 17, 4, 4), (1, 29, 27, 4, 4),
])])
def test_lookup_fuzz(cuda, lib, seed, b, h, w, levels, r):
    """Fast path (4 levels, r in {3, 4}) and generic path, awkward and non-finite coordinates."""
    rng = np.random.default_rng(2000 + seed)
    q = b * h * w
    pyr = [_rounded(rng.standard_normal((q, h >> l, w >> l, 1)).astype(F32)) for l in range(levels)]
    coords = _awkward_coords(rng, b, h, w)
    if seed % 3 == 0:
        coords[0, 0, 0] = [np.nan, 1.0]
        coords[0, 1, 1] = [2.0, np.inf]
        coords[0, 2, 2] = [-np.inf, -np.inf]
    with np.errstate(invalid="ignore"):
        ref = O.index_pyramid(pyr, coords, r)
    _assert_bit_exact(_lookup_f16(pyr, coords, r, cuda), ref)


@pytest.mark.parametrize("b", [1, 3, 8])
def test_f16_equals_f32_kernels_on_representable_values(cuda, lib, b):
    """Sizes the oracle is not run on (Sintel shape, batch 1-8): on a pyramid whose values are float16-representable
    the float16 kernels return the float32 tiled kernel's bits; a zero pyramid gives zeros."""
    h, w, r = 55, 128, 4
    q = b * h * w
    gen = torch.Generator(device=cuda).manual_seed(b)
    pyr = [torch.randn((q, h >> l, w >> l, 1), device=cuda, generator=gen).half().float() for l in range(4)]
    coords = J.make_coords_grid(b, h, w, device=cuda) + (torch.rand((b, h, w, 2), device=cuda, generator=gen) - 0.5) * 16
    cb = J.CorrBlock(4, r)
    a = cb.index_pyramid(TiledPyramid.from_dense(pyr, storage="fp16"), coords)
    c = cb.index_pyramid(TiledPyramid.from_dense(pyr), coords)
    assert bool(((a == c) | (a.isnan() & c.isnan())).all())
    assert not cb.index_pyramid(TiledPyramid.from_dense([torch.zeros_like(t) for t in pyr], storage="fp16"), coords).any()


@pytest.mark.parametrize("shape,r", [((1, 16, 24), 4), ((2, 19, 21), 3), ((1, 55, 128), 4)])
def test_lookup_with_fused_coordinate_update(cuda, lib, shape, r):
    """raftk_corr_lookup_tiled_f16_update == coordinate update followed by the float16 lookup, bit for bit."""
    b, h, w = shape
    gen = torch.Generator(device=cuda).manual_seed(h + r)
    q = b * h * w
    pyr = [torch.randn((q, h >> l, w >> l, 1), device=cuda, generator=gen) for l in range(4)]
    tp = TiledPyramid.from_dense(pyr, storage="fp16")
    c0 = J.make_coords_grid(b, h, w, device=cuda)
    c1 = c0 + (torch.rand((b, h, w, 2), device=cuda, generator=gen) - 0.5) * 10
    c1[0, 1, 1] = 3.0e6  # a far query
    delta = torch.randn((b, h, w, 8), device=cuda, generator=gen)
    bias = torch.randn((8,), device=cuda, generator=gen)
    cb = J.CorrBlock(4, r, storage="fp16")
    ref_c1 = c1 + (delta[..., :2] + bias[:2])
    ref_flow = ref_c1 - c0
    ref_out = cb.index_pyramid(tp, ref_c1)
    flow = torch.full((b, h, w, 2), 9.0, device=cuda)
    flow8 = torch.zeros((b, h, w, 8), device=cuda)
    got_c1 = c1.clone()
    out = cb.index_pyramid_update(tp, c0, got_c1, delta, bias, flow, flow8)
    assert torch.equal(got_c1, ref_c1) and torch.equal(flow, ref_flow)
    assert torch.equal(flow8[..., :2], ref_flow) and not flow8[..., 2:].any()
    assert bool((out == ref_out).all())


# ---------------------------------------------------------------------------------------------------------
# volume + pyramid written as float16 tiles by the GEMM epilogue
# ---------------------------------------------------------------------------------------------------------
def _build_raw(f1, f2, dev, levels=4, storage="fp16"):
    """The C-ABI entry point into NaN-prefilled slabs (so unwritten cells show)."""
    lib = _lib.load()
    b, h, w, d = f2.shape
    t1 = torch.from_numpy(f1).to(dev)
    t2 = torch.from_numpy(f2).to(dev)
    nq = t1.shape[1] * t1.shape[2]
    tp = TiledPyramid.empty(b * nq, h, w, levels, dev, storage)
    for t in tp.levels:
        t.fill_(float("nan"))
    nbytes = lib.raftk_corr_pyramid_scratch_bytes(b, nq, h, w, d)
    scratch = torch.empty((max(nbytes, 1),), dtype=torch.uint8, device=dev)
    fn = lib.raftk_corr_pyramid_tiled_f16 if storage == "fp16" else lib.raftk_corr_pyramid_tiled
    rc = fn(t1.data_ptr(), t2.data_ptr(), _lib.ptr_array(tp.levels), levels, b, nq, h, w, d, scratch.data_ptr(), nbytes,
            _lib.current_stream_ptr(dev))
    assert rc == 0, rc
    torch.cuda.synchronize()
    return tp


@pytest.mark.parametrize("shape", [(1, 16, 16, 128), (2, 17, 23, 64), (1, 46, 62, 128), (1, 24, 40, 256), (1, 33, 65, 64),
                                   (1, 19, 21, 192)])
def test_pyramid_values_padding_and_float32_agreement(cuda, lib, shape):
    b, h, w, d = shape
    rng = np.random.default_rng(d + h)
    f1 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    f2 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    tp = _build_raw(f1, f2, cuda)
    tp32 = _build_raw(f1, f2, cuda, storage="fp32")
    ref64 = O.build_pyramid(f1, f2, dtype=np.float64)
    for l, got in enumerate(tp.dense()):
        got = got.cpu().numpy()
        assert got.shape == ref64[l].shape
        # the north-star's bar for correlation values: 1e-3 relative (float16 rounding is <= 4.9e-4 of it)
        assert _rel_err(got, ref64[l]) <= 1e-3, l
        lh, lw = h >> l, w >> l
        t = tp.levels[l].cpu().numpy()  # (Q, ty, tx, 4, 4) float16
        t32 = tp32.levels[l].cpu().numpy()
        _, ty, tx, _, _ = t.shape
        for tyi in range(ty):
            for txi in range(tx):
                if txi * 4 >= lw:
                    continue  # the all-padding tile that rounds a tile row up to a whole line
                tile = t[:, tyi, txi]
                assert not np.isnan(tile).any(), (l, tyi, txi)
                # one rounding of the float32 path's value: same accumulators, same float32 pooling of unrounded levels
                assert np.array_equal(tile, t32[:, tyi, txi].astype(np.float16)), (l, tyi, txi)
                yy, xx = np.meshgrid(tyi * 4 + np.arange(4), txi * 4 + np.arange(4), indexing="ij")
                pad = (yy >= lh) | (xx >= lw)
                assert not tile[:, pad].any(), (l, tyi, txi)


def test_pyramid_query_row_band_small_nq_and_level_counts(cuda, lib):
    rng = np.random.default_rng(9)
    b, h, w, d = 2, 18, 20, 128
    f1 = rng.standard_normal((b, h, w, d)).astype(F32)
    f2 = rng.standard_normal((b, h, w, d)).astype(F32)
    full = [t.cpu().numpy() for t in _build_raw(f1, f2, cuda).dense()]
    for y0, y1 in ((5, 12), (3, 6)):  # (3, 6): 60 queries, single-CTA tiles
        band = [t.cpu().numpy() for t in _build_raw(np.ascontiguousarray(f1[:, y0:y1]), f2, cuda).dense()]
        for l in range(4):
            ref = full[l].reshape(b, h, w, h >> l, w >> l, 1)[:, y0:y1].reshape(band[l].shape)
            assert np.array_equal(band[l], ref)
    f1 = rng.standard_normal((1, 16, 24, 64)).astype(F32)
    f2 = rng.standard_normal((1, 16, 24, 64)).astype(F32)
    for levels in (1, 2, 3):
        ref = O.build_pyramid(f1, f2, num_levels=levels, dtype=np.float64)
        for l, got in enumerate(_build_raw(f1, f2, cuda, levels=levels).dense()):
            assert _rel_err(got.cpu().numpy(), ref[l]) <= 1e-3
    # configurations the tiled epilogue does not cover are refused by the C ABI and served by CorrBlock through
    # the row-major kernels + raftk_pyramid_tile_f16
    arr = (ctypes.c_void_p * 5)(*([256] * 5))
    assert lib.raftk_corr_pyramid_tiled_f16(ctypes.c_void_p(256), ctypes.c_void_p(256), arr, 4, 1, 256, 16, 16, 96,
                                            ctypes.c_void_p(256), 1 << 30, None) == -4
    f1 = rng.standard_normal((1, 32, 32, 96)).astype(F32)
    f2 = rng.standard_normal((1, 32, 32, 96)).astype(F32)
    pyr = J.CorrBlock(5, 4, storage="fp16").build_pyramid(torch.from_numpy(f1).to(cuda), torch.from_numpy(f2).to(cuda))
    assert isinstance(pyr, TiledPyramid) and pyr.storage == "fp16" and len(pyr) == 5
    ref = O.build_pyramid(f1, f2, num_levels=5, dtype=np.float64)
    for l in range(5):
        assert _rel_err(pyr[l].cpu().numpy(), ref[l]) <= 1e-3


def test_sintel_shape_build_then_lookup(cuda, lib):
    """BASELINE config 2 on the float16 path: the pyramid meets the 1e-3 bar, and build_pyramid -> index_pyramid
    equals the oracle's lookup of the GPU's own stored pyramid bit for bit."""
    rng = np.random.default_rng(0)
    b, h, w, d, r = 1, 55, 128, 256, 4
    f1 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    f2 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    cb = J.CorrBlock(4, r, storage="fp16")
    pyr = cb.build_pyramid(torch.from_numpy(f1).to(cuda), torch.from_numpy(f2).to(cuda))
    assert isinstance(pyr, TiledPyramid) and pyr.storage == "fp16"
    ref = O.build_pyramid(f1, f2, dtype=np.float64)
    dense = [p.cpu().numpy() for p in pyr]
    for l in range(4):
        assert _rel_err(dense[l], ref[l]) <= 1e-3
    coords = O.make_coords_grid(b, h, w) + rng.uniform(-8, 8, (b, h, w, 2)).astype(F32)
    got = cb.index_pyramid(pyr, torch.from_numpy(coords).to(cuda)).cpu().numpy()
    _assert_bit_exact(got, O.index_pyramid(dense, coords, r))


def test_end_to_end_epe_with_float16_pyramid(cuda, lib):
    """raft_large through the float16 pyramid against the CPU reference (torchvision-init weights): the north-star's
    0.01 px mean-EPE bar, and the distance to the float32-pyramid model."""
    from oracle import raft_cpu

    rng = np.random.default_rng(0)
    net = raft_cpu.build_modules("raft_large", seed=0)
    img1 = rng.uniform(-1, 1, (1, 128, 160, 3)).astype(F32)
    img2 = np.clip(np.roll(img1, (3, -5), axis=(1, 2)) + rng.normal(0, 0.02, img1.shape).astype(F32), -1, 1)
    ref = raft_cpu.forward(net, img1, img2, 6)
    variables = J.variables_from_torch_state_dict(net.state_dict())
    m16, _ = J.raft_large(corr_block=J.CorrBlock(4, 4, storage="fp16"))
    m32, _ = J.raft_large()
    got16 = m16.apply(variables, img1, img2, train=False, num_flow_updates=6).cpu().numpy()
    got32 = m32.apply(variables, img1, img2, train=False, num_flow_updates=6).cpu().numpy()
    epe = float(np.sqrt(((got16[-1] - ref[-1]) ** 2).sum(-1)).mean())
    assert epe <= 0.01, epe
    assert float(np.sqrt(((got16[-1] - got32[-1]) ** 2).sum(-1)).mean()) <= 0.005
