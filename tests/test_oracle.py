"""The CPU oracle against the committed golden vectors, torchvision (the upstream the reference
was ported from) and SciPy.  No GPU."""

import numpy as np
import pytest
import torch

from oracle import corr_oracle as O

F32 = np.float32


def test_make_coords_grid_is_xy():
    c = O.make_coords_grid(2, 3, 5)
    assert c.shape == (2, 3, 5, 2) and c.dtype == F32
    assert c[1, 2, 4, 0] == 4 and c[1, 2, 4, 1] == 2  # (x, y): model.py:37-40


def test_tap_order_golden(golden):
    g = golden("lookup_tap_order")
    h, w, r = int(g["h"]), int(g["w"]), int(g["radius"])
    pyr = []
    for l in range(4):
        yy, xx = np.meshgrid(np.arange(h >> l), np.arange(w >> l), indexing="ij")
        pyr.append(np.tile((l * 10000 + yy * 100 + xx).astype(F32)[None, :, :, None], (2, 1, 1, 1)))
    out = O.index_pyramid(pyr, g["coords"], r)
    assert np.array_equal(out, g["out"])
    # channel l*81 + i*9 + j: i is the x offset (model.py:452-455, 463)
    x, y = g["coords"].reshape(-1, 2)[0]
    assert out.reshape(2, -1)[0, 0 * 81 + 8 * 9 + 0] == 0 * 10000 + (y - r) * 100 + (x + r)


@pytest.mark.parametrize("r", [3, 4])
def test_border_golden(golden, r):
    g = golden(f"lookup_borders_r{r}")
    pyr = [g[f"level{l}"] for l in range(4)]
    out = O.index_pyramid(pyr, g["coords"], r)
    assert np.array_equal(out, g["out"])
    # far outside -> exactly zero
    far = np.where(np.abs(g["coords"].reshape(-1, 2)[:, 0]) >= 1e5)[0]
    assert far.size and not out.reshape(-1, out.shape[-1])[far].any()


@pytest.mark.parametrize("name", ["odd", "small"])
def test_pyramid_golden_and_bounds(golden, name):
    g = golden(f"pyramid_{name}")
    f1, f2 = g["fmap1"], g["fmap2"]
    _, h, w, _ = f1.shape
    pyr = O.build_pyramid(f1, f2)
    stride = int(g["query_stride"])
    for l in range(4):
        assert pyr[l].shape == (h * w, h >> l, w >> l, 1)  # VALID pooling floors odd sizes
        np.testing.assert_allclose(pyr[l][::stride], g[f"level{l}"], rtol=2e-5, atol=2e-5)


def test_pyramid_asserts_like_reference():
    f = np.zeros((1, 16, 16, 4), F32)
    with pytest.raises(AssertionError, match="same shapes"):
        O.build_pyramid(f, np.zeros((1, 16, 17, 4), F32))
    with pytest.raises(AssertionError, match="too small"):
        O.build_pyramid(np.zeros((1, 15, 16, 4), F32), np.zeros((1, 15, 16, 4), F32))


def test_oracle_matches_torchvision_corr_block():
    from torchvision.models.optical_flow.raft import CorrBlock

    rng = np.random.default_rng(0)
    b, h, w, d = 2, 19, 22, 32
    f1 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    f2 = (rng.standard_normal((b, h, w, d)) * 1.45).astype(F32)
    for r in (3, 4):
        cb = CorrBlock(num_levels=4, radius=r)
        cb.build_pyramid(torch.from_numpy(f1).permute(0, 3, 1, 2), torch.from_numpy(f2).permute(0, 3, 1, 2))
        pyr = O.build_pyramid(f1, f2)
        for a, t in zip(pyr, cb.corr_pyramid):
            np.testing.assert_allclose(a[..., 0], t[:, 0].numpy(), rtol=1e-5, atol=1e-5)
        coords = O.make_coords_grid(b, h, w) + rng.uniform(-8, 8, (b, h, w, 2)).astype(F32)
        out = O.index_pyramid(pyr, coords, r)
        tv = cb.index_pyramid(torch.from_numpy(coords).permute(0, 3, 1, 2)).permute(0, 2, 3, 1).numpy()
        # torch normalises / denormalises coordinates, so agreement is to ~1e-5 only
        np.testing.assert_allclose(out, tv, rtol=1e-4, atol=1e-4)


def test_grid_sample_matches_scipy_grid_constant():
    from scipy.ndimage import map_coordinates

    rng = np.random.default_rng(5)
    img = rng.standard_normal((3, 7, 9, 1)).astype(F32)
    grid = rng.uniform(-3, 11, (3, 6, 6, 2)).astype(F32)
    out = O.grid_sample(img, grid)
    for n in range(3):
        ref = map_coordinates(img[n, :, :, 0].astype(np.float64),
                              [grid[n, ..., 1].ravel().astype(np.float64), grid[n, ..., 0].ravel().astype(np.float64)],
                              order=1, mode="grid-constant", cval=0.0).reshape(6, 6)
        np.testing.assert_allclose(out[n, ..., 0], ref, rtol=1e-5, atol=1e-6)


def test_upsample_golden_and_torchvision(golden):
    from torchvision.models.optical_flow._utils import upsample_flow as tv_up

    g = golden("upsample")
    flow = g["flow"]
    b, h, w, _ = flow.shape
    for k in range(9):
        m = np.zeros((b, h, w, 576), F32)
        m[..., k * 64:(k + 1) * 64] = 40.0
        np.testing.assert_array_equal(O.upsample_flow(flow, m), g["onehot_out"][k])
    up = O.upsample_flow(flow, g["random_mask"])
    np.testing.assert_array_equal(up, g["random_out"])
    tv = tv_up(torch.from_numpy(flow).permute(0, 3, 1, 2), torch.from_numpy(g["random_mask"]).permute(0, 3, 1, 2))
    np.testing.assert_allclose(up, tv.permute(0, 2, 3, 1).numpy(), rtol=1e-5, atol=1e-5)
    # bilinear branch (raft_small, model.py:79-83)
    tvb = tv_up(torch.from_numpy(flow).permute(0, 3, 1, 2)).permute(0, 2, 3, 1).numpy()
    np.testing.assert_allclose(O.upsample_flow(flow), tvb, rtol=1e-4, atol=1e-4)


def test_alt_index_equals_volume_lookup():
    rng = np.random.default_rng(2)
    b, h, w, d = 1, 17, 20, 16
    f1 = rng.standard_normal((b, h, w, d)).astype(F32)
    f2 = rng.standard_normal((b, h, w, d)).astype(F32)
    coords = O.make_coords_grid(b, h, w) + rng.uniform(-6, 6, (b, h, w, 2)).astype(F32)
    full = O.index_pyramid(O.build_pyramid(f1, f2), coords, 4)
    alt = O.alt_index(f1, O.pool_fmap2(f2), coords, 4)
    np.testing.assert_allclose(alt, full, rtol=1e-4, atol=1e-5)
    rows = O.alt_index(f1, O.pool_fmap2(f2), coords, 4, row_range=(5, 11))
    np.testing.assert_allclose(rows, full[:, 5:11], rtol=1e-4, atol=1e-5)


def test_c_oracle_equals_numpy_oracle():
    """oracle/corr_oracle.c (the multi-threaded CPU baseline) against the NumPy oracle: pooling and
    lookup bit-identical (same float32 operations in the same order), upsampling to expf rounding."""
    from oracle import corr_oracle_c as C

    rng = np.random.default_rng(0)
    b, h, w, d = 2, 19, 22, 32
    f1 = rng.standard_normal((b, h, w, d)).astype(F32)
    f2 = rng.standard_normal((b, h, w, d)).astype(F32)
    p_np, p_c = O.build_pyramid(f1, f2), C.build_pyramid(f1, f2)
    for a, c in zip(p_np, p_c):
        np.testing.assert_array_equal(a, c)
    coords = O.make_coords_grid(b, h, w) + rng.uniform(-8, 8, (b, h, w, 2)).astype(F32)
    coords[0, 0, 0] = (1e7, -3.0)
    coords[0, 1, 1] = (-1e-9, -1e-9)
    for r in (3, 4):
        np.testing.assert_array_equal(O.index_pyramid(p_np, coords, r), C.index_pyramid(p_np, coords, r))
    flow = rng.standard_normal((b, h, w, 2)).astype(F32) * 4
    mask = rng.standard_normal((b, h, w, 576)).astype(F32)
    np.testing.assert_allclose(C.upsample_flow(flow, mask), O.upsample_flow(flow, mask), rtol=0, atol=2e-5)
