"""Parity of the volume + pyramid kernels (family 1) with the oracle, through the C ABI.

Bars: pooling bounds bit-exact (every level has exactly the reference's shape and
each cell averages exactly its 2^l x 2^l block: checked against float64 block means); correlation
values within 1e-3 relative, with the floor SURVEY.md §7.2 proposes for near-zero entries:
|d| <= 1e-3 * max(|ref|, rms(ref)).  The fp32 FMA kernel is held to 1e-5 on the same measure.
"""

import numpy as np
import pytest
import torch

import jax_raft_b200 as J
from oracle import corr_oracle as O

pytestmark = pytest.mark.gpu
F32 = np.float32


def _build(f1, f2, dev, kernel, levels=4, query_rows=None):
    cb = J.CorrBlock(levels, 4, kernel=kernel)
    pyr = cb.build_pyramid(torch.from_numpy(f1).to(dev), torch.from_numpy(f2).to(dev), query_rows=query_rows)
    torch.cuda.synchronize()
    return [p.cpu().numpy() for p in pyr]


def _rel_err(got, ref64):
    rms = np.sqrt(np.mean(ref64 ** 2))
    return float(np.max(np.abs(got - ref64) / np.maximum(np.abs(ref64), rms)))


TOL = {"simt": 1e-5, "tc": 1e-3}


@pytest.mark.parametrize("kernel", ["simt", "tc"])
@pytest.mark.parametrize("name", ["odd", "small"])
def test_fixture(cuda, lib, golden, kernel, name):
    g = golden(f"pyramid_{name}")
    f1, f2 = g["fmap1"], g["fmap2"]
    _, h, w, _ = f1.shape
    stride = int(g["query_stride"])
    pyr = _build(f1, f2, cuda, kernel)
    ref64 = O.build_pyramid(f1, f2, dtype=np.float64)
    for l in range(4):
        assert pyr[l].shape == (h * w, h >> l, w >> l, 1)
        assert _rel_err(pyr[l][::stride], g[f"level{l}"].astype(np.float64)) <= 2 * TOL[kernel]
        assert _rel_err(pyr[l], ref64[l]) <= TOL[kernel], (kernel, l)


@pytest.mark.parametrize("kernel", ["simt", "tc"])
@pytest.mark.parametrize("shape", [(1, 16, 16, 128), (2, 17, 23, 64), (1, 46, 62, 128), (1, 24, 40, 256)])
def test_random_against_oracle_and_pooling_bounds(cuda, lib, kernel, shape):
    b, h, w, d = shape
    rng = np.random.default_rng(d + h)
    f1 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    f2 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    pyr = _build(f1, f2, cuda, kernel)
    ref64 = O.build_pyramid(f1, f2, dtype=np.float64)
    for l in range(4):
        assert pyr[l].shape == ref64[l].shape
        assert _rel_err(pyr[l], ref64[l]) <= TOL[kernel], (kernel, l)
    # pooling bounds: each level-l cell is the mean of ITS OWN level-0 2^l x 2^l block (VALID floor)
    v0 = pyr[0][..., 0].astype(np.float64)
    for l in range(1, 4):
        s, hl, wl = 2 ** l, h >> l, w >> l
        blk = v0[:, : hl * s, : wl * s].reshape(-1, hl, s, wl, s).mean(axis=(2, 4))
        np.testing.assert_allclose(pyr[l][..., 0], blk, rtol=0, atol=2e-6 * np.abs(v0).max())


@pytest.mark.parametrize("kernel", ["simt", "tc"])
def test_query_row_band(cuda, lib, kernel):
    """Row-sharded mode: the band's pyramid equals the band of the full pyramid."""
    rng = np.random.default_rng(9)
    b, h, w, d = 2, 18, 20, 128
    f1 = rng.standard_normal((b, h, w, d)).astype(F32)
    f2 = rng.standard_normal((b, h, w, d)).astype(F32)
    full = _build(f1, f2, cuda, kernel)
    band = _build(f1, f2, cuda, kernel, query_rows=(5, 12))
    for l in range(4):
        ref = full[l].reshape(b, h, w, h >> l, w >> l, 1)[:, 5:12].reshape(band[l].shape)
        np.testing.assert_allclose(band[l], ref, rtol=0, atol=1e-5)


@pytest.mark.parametrize("kernel", ["simt", "tc"])
def test_other_level_counts(cuda, lib, kernel):
    rng = np.random.default_rng(3)
    f1 = rng.standard_normal((1, 16, 24, 64)).astype(F32)
    f2 = rng.standard_normal((1, 16, 24, 64)).astype(F32)
    for levels in (1, 2, 3):
        pyr = _build(f1, f2, cuda, kernel, levels=levels)
        ref = O.build_pyramid(f1, f2, num_levels=levels, dtype=np.float64)
        assert len(pyr) == levels
        for l in range(levels):
            assert _rel_err(pyr[l], ref[l]) <= TOL[kernel]


def test_sintel_shape_full_size(cuda, lib):
    """BASELINE config 2: 55x128, D = 256 — the full 7040 x 7040 volume against the oracle."""
    rng = np.random.default_rng(0)
    b, h, w, d = 1, 55, 128, 256
    f1 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    f2 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    ref = O.build_pyramid(f1, f2)  # float32 BLAS on the host
    pyr = _build(f1, f2, cuda, "tc")
    for l in range(4):
        assert pyr[l].shape == ref[l].shape == (h * w, h >> l, w >> l, 1)
        assert _rel_err(pyr[l], ref[l].astype(np.float64)) <= 1e-3


def test_checksum_property_at_batch_size(cuda, lib):
    """Size-independent property at B = 4 Sintel-shape pairs (no oracle at this size): the sum of
    volume row q equals f1[q] . sum_j f2[j] / sqrt(D), and each pooled level has the same sum as the
    level-0 cells it covers, scaled by 4^-l."""
    b, h, w, d = 4, 55, 128, 256
    gen = torch.Generator(device=cuda).manual_seed(0)
    f1 = torch.randn((b, h, w, d), device=cuda, generator=gen)
    f2 = torch.randn((b, h, w, d), device=cuda, generator=gen)
    pyr = J.CorrBlock(4, 4).build_pyramid(f1, f2)
    row_sum = pyr[0].double().sum(dim=(1, 2, 3)).view(b, h * w)
    expect = torch.einsum("bqd,bd->bq", f1.view(b, h * w, d).double(), f2.view(b, h * w, d).double().sum(1)) / d ** 0.5
    assert torch.allclose(row_sum, expect, rtol=0, atol=2e-3 * float(expect.abs().mean()) + 1e-2)
    for l in range(1, 4):
        s = 2 ** l
        cover = pyr[0][:, : (h >> l) * s, : (w >> l) * s].double().sum(dim=(1, 2, 3))
        assert torch.allclose(pyr[l].double().sum(dim=(1, 2, 3)) * (4 ** l), cover, rtol=0,
                              atol=1e-3 * float(cover.abs().mean()) + 1e-2)


def test_1080p_shape_properties_and_altcorr_agreement(cuda, lib):
    """Largest named configuration (BASELINE config 4 shape: 1088x1920 -> 136x240, N = 32640,
    5.7 GB pyramid), where the CPU oracle is not run: row-sum checksum of the volume, pooled levels
    consistent with level 0, and the materialised-volume lookup equal to the on-the-fly alt_corr
    lookup (two independent kernels) on a band of query rows."""
    b, h, w, d, r = 1, 136, 240, 256, 4
    gen = torch.Generator(device=cuda).manual_seed(0)
    f1 = torch.randn((b, h, w, d), device=cuda, generator=gen)
    f2 = torch.randn((b, h, w, d), device=cuda, generator=gen)
    cb = J.CorrBlock(4, r)
    pyr = cb.build_pyramid(f1, f2)
    assert [tuple(p.shape) for p in pyr] == [(h * w, h >> l, w >> l, 1) for l in range(4)]
    row_sum = pyr[0].double().sum(dim=(1, 2, 3))
    expect = (f1.view(h * w, d).double() @ f2.view(h * w, d).double().sum(0)) / d ** 0.5
    assert torch.allclose(row_sum, expect, rtol=0, atol=2e-3 * float(expect.abs().mean()) + 1e-2)
    for l in range(1, 4):
        s = 2 ** l
        cover = pyr[0][:, : (h >> l) * s, : (w >> l) * s].double().sum(dim=(1, 2, 3))
        assert torch.allclose(pyr[l].double().sum(dim=(1, 2, 3)) * (4 ** l), cover, rtol=0,
                              atol=1e-3 * float(cover.abs().mean()) + 1e-2)
    coords = J.make_coords_grid(b, h, w, device=cuda) + (torch.rand((b, h, w, 2), device=cuda, generator=gen) - 0.5) * 16
    vol = cb.index_pyramid(pyr, coords)
    alt = J.AltCorrBlock(4, r)
    band = alt.index_pyramid(alt.build_pyramid(f1, f2, query_rows=(40, 57)), coords[:, 40:57].contiguous())
    assert float((band - vol[:, 40:57]).abs().max()) <= 1e-3 * max(1.0, float(vol.abs().mean()))
