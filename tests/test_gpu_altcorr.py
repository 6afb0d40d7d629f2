"""alt_corr (family 2b): pooled-fmap2 levels and the on-the-fly lookup against the oracle and
against the materialised-volume path.  Index work is shared with the volume lookup (bit-exact
there); values differ by the summation order of the dot products: the float32 FMA kernels
(kernel="simt") within 1e-4 of the correlation scale, the tensor-core kernels (kernel="tc", the
default: bf16 hi/mid split products, as the volume GEMM) within 3e-4 (north-star: 1e-3 relative).
"tc64" / "tc128" force one of the two tensor-core kernels (raftk_set_altcorr_mode: 8 x 8 patches with the
A tile in shared memory, 16 x 8 patches with A in tensor memory); plain "tc" chooses by problem size."""

import numpy as np
import pytest
import torch

import jax_raft_b200 as J
from oracle import corr_oracle as O

pytestmark = pytest.mark.gpu
F32 = np.float32


KERNELS = [("tc", 3e-4), ("tc64", 3e-4), ("tc128", 3e-4), ("simt", 1e-4)]
TC_KERNELS = ["tc", "tc64", "tc128"]


@pytest.fixture(autouse=True)
def _auto_mode_afterwards(lib):
    yield
    lib.raftk_set_altcorr_mode(0)


def _alt_block(lib, kernel, levels, r):
    """AltCorrBlock for a test's kernel name; "tc64" / "tc128" pin the tensor-core kernel for the rest of the test."""
    lib.raftk_set_altcorr_mode({"tc64": 1, "tc128": 2}.get(kernel, 0))
    return J.AltCorrBlock(levels, r, kernel="simt" if kernel == "simt" else "tc")


def _scale_close(got, ref, tol=1e-4):
    rms = float(np.sqrt(np.mean(ref.astype(np.float64) ** 2))) or 1.0
    assert float(np.abs(got - ref).max()) <= tol * max(rms, 1.0), float(np.abs(got - ref).max())


@pytest.mark.parametrize("kernel,tol", KERNELS)
@pytest.mark.parametrize("shape,r", [((1, 16, 16, 128), 4), ((2, 19, 21, 128), 3), ((1, 23, 31, 256), 4)])
def test_against_oracle(cuda, lib, shape, r, kernel, tol):
    b, h, w, d = shape
    rng = np.random.default_rng(h)
    f1 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    f2 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    coords = O.make_coords_grid(b, h, w) + rng.uniform(-8, 8, (b, h, w, 2)).astype(F32)
    alt = _alt_block(lib, kernel, 4, r)
    pyr = alt.build_pyramid(torch.from_numpy(f1).to(cuda), torch.from_numpy(f2).to(cuda))
    assert len(pyr) == (2 if kernel == "simt" else 4)
    ref_levels = O.pool_fmap2(f2)
    for l in range(1, 4):  # pooled fmap2: same ((a+b)+(c+d))/4 order as the oracle -> exact
        np.testing.assert_array_equal(pyr[1][l].cpu().numpy(), ref_levels[l])
    out = alt.index_pyramid(pyr, torch.from_numpy(coords).to(cuda)).cpu().numpy()
    ref = O.index_pyramid(O.build_pyramid(f1, f2), coords, r)
    _scale_close(out, ref, tol)
    # and against the volume path on the GPU
    cb = J.CorrBlock(4, r, kernel="simt")
    vol = cb.index_pyramid(cb.build_pyramid(torch.from_numpy(f1).to(cuda), torch.from_numpy(f2).to(cuda)),
                           torch.from_numpy(coords).to(cuda)).cpu().numpy()
    _scale_close(out, vol, tol)


@pytest.mark.parametrize("kernel", TC_KERNELS + ["simt"])
def test_query_row_band(cuda, lib, kernel):
    rng = np.random.default_rng(8)
    b, h, w, d = 1, 20, 24, 128
    f1 = rng.standard_normal((b, h, w, d)).astype(F32)
    f2 = rng.standard_normal((b, h, w, d)).astype(F32)
    coords = O.make_coords_grid(b, h, w) + rng.uniform(-6, 6, (b, h, w, 2)).astype(F32)
    alt = _alt_block(lib, kernel, 4, 4)
    t1, t2, tc = (torch.from_numpy(x).to(cuda) for x in (f1, f2, coords))
    full = alt.index_pyramid(alt.build_pyramid(t1, t2), tc)
    band = alt.index_pyramid(alt.build_pyramid(t1, t2, query_rows=(7, 15)), tc[:, 7:15].contiguous())
    assert torch.equal(band, full[:, 7:15])
    # already-sliced fmap1 (what a row-sharded rank holds)
    band2 = alt.index_pyramid(alt.build_pyramid(t1[:, 7:15].contiguous(), t2, query_rows=(7, 15)), tc[:, 7:15].contiguous())
    assert torch.equal(band2, band)


def _smooth_flow(b, h, w, rng, amp=6.0):
    """A low-frequency flow field (what RAFT actually sees): neighbouring windows overlap, so the
    tiled kernel takes its shared bounding-box path; random per-pixel flow takes the fallback."""
    yy, xx = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
    fx = amp * np.sin(2 * np.pi * (xx / w + 0.3 * yy / h) + rng.uniform(0, 6)) + rng.uniform(-3, 3)
    fy = amp * np.cos(2 * np.pi * (yy / h - 0.2 * xx / w) + rng.uniform(0, 6)) + rng.uniform(-3, 3)
    return np.tile(np.stack([fx, fy], -1).astype(F32)[None], (b, 1, 1, 1))


@pytest.mark.parametrize("kernel,tol", KERNELS)
@pytest.mark.parametrize("shape,r", [((1, 24, 40, 128), 4), ((2, 19, 21, 64), 3), ((1, 46, 62, 128), 3), ((1, 55, 128, 256), 4)])
def test_smooth_flow_shared_path_against_oracle(cuda, lib, shape, r, kernel, tol):
    b, h, w, d = shape
    rng = np.random.default_rng(h + w)
    f1 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    f2 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    coords = O.make_coords_grid(b, h, w) + _smooth_flow(b, h, w, rng)
    coords[0, 0, 0] = (1e7, 3.0)       # one far-out query: its tile falls back, the rest must not change
    coords[0, h - 1, w - 1] = (-40.0, h + 30.0)
    alt = _alt_block(lib, kernel, 4, r)
    t1, t2, tc = (torch.from_numpy(x).to(cuda) for x in (f1, f2, coords))
    out = alt.index_pyramid(alt.build_pyramid(t1, t2), tc).cpu().numpy()
    ref = O.alt_index(f1, O.pool_fmap2(f2), coords, r)
    _scale_close(out, ref, tol)
    assert not out[0, 0, 0].any()
    # a band of rows gives bit-identical values: each cell's dot product is one fixed fma chain,
    # whatever tile or path computed it
    band = alt.index_pyramid(alt.build_pyramid(t1, t2, query_rows=(3, 17)), tc[:, 3:17].contiguous()).cpu().numpy()
    assert np.array_equal(band, out[:, 3:17])


@pytest.mark.parametrize("kernel", TC_KERNELS)
def test_tc_rough_flow_and_stray_queries(cuda, lib, kernel):
    """The tensor-core kernels' other paths: per-pixel random flow of +-25 px (windows of a patch do not fit one
    box: scalar path; M = 128: first the two half-patch boxes), a patch with one stray far / non-finite query
    (it stays out of the box, its neighbours are unchanged), and levels the windows fall entirely outside of."""
    rng = np.random.default_rng(5)
    b, h, w, d, r = 1, 40, 48, 128, 4
    f1 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    f2 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    t1, t2 = torch.from_numpy(f1).to(cuda), torch.from_numpy(f2).to(cuda)
    alt = _alt_block(lib, kernel, 4, r)
    rough = O.make_coords_grid(b, h, w) + rng.uniform(-25, 25, (b, h, w, 2)).astype(F32)
    out = alt.index_pyramid(alt.build_pyramid(t1, t2), torch.from_numpy(rough).to(cuda)).cpu().numpy()
    _scale_close(out, O.alt_index(f1, O.pool_fmap2(f2), rough, r), 3e-4)
    smooth = O.make_coords_grid(b, h, w) + _smooth_flow(b, h, w, rng)
    base = alt.index_pyramid(alt.build_pyramid(t1, t2), torch.from_numpy(smooth).to(cuda)).cpu().numpy()
    stray = smooth.copy()
    stray[0, 10, 10] = (np.nan, 2.0)
    stray[0, 11, 12] = (5e8, -3.0)
    stray[0, 20, 20] = (-300.0, 900.0)
    with np.errstate(invalid="ignore"):
        ref = O.alt_index(f1, O.pool_fmap2(f2), stray, r)
    got = alt.index_pyramid(alt.build_pyramid(t1, t2), torch.from_numpy(stray).to(cuda)).cpu().numpy()
    keep = np.ones((h, w), bool)
    keep[10, 10] = keep[11, 12] = keep[20, 20] = False
    assert np.array_equal(got[0][keep], base[0][keep])          # the neighbours' bits do not change
    assert np.isnan(got[0, 10, 10]).all() == np.isnan(ref[0, 10, 10]).all()
    assert not got[0, 11, 12].any() and not got[0, 20, 20].any()


def test_tc128_half_patch_boxes_and_agreement_with_tc64(cuda, lib):
    """A flow that stretches along x (0.9 px per px): the windows of a 16 x 8 patch no longer fit one 32-cell-wide box,
    the two 8 x 8 halves do — the M = 128 kernel multiplies each half's box in turn.  Values within the split-product
    error of the oracle, and the two tensor-core kernels within it of each other."""
    rng = np.random.default_rng(11)
    b, h, w, d, r = 1, 32, 96, 128, 4
    f1 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    f2 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    yy, xx = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
    flow = np.stack([0.9 * (xx - w / 2) * 0.5 + 2.0 * np.sin(yy / 5.0), 0.3 * (yy - h / 2)], -1).astype(F32)[None]
    coords = O.make_coords_grid(b, h, w) + flow
    t1, t2, tc = (torch.from_numpy(x).to(cuda) for x in (f1, f2, coords))
    ref = O.alt_index(f1, O.pool_fmap2(f2), coords, r)
    outs = {}
    for kernel in ("tc64", "tc128"):
        alt = _alt_block(lib, kernel, 4, r)
        outs[kernel] = alt.index_pyramid(alt.build_pyramid(t1, t2), tc).cpu().numpy()
        _scale_close(outs[kernel], ref, 3e-4)
    _scale_close(outs["tc128"], outs["tc64"], 1e-4)
