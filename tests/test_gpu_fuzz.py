"""Seeded random sweep of shapes, radii, level counts and awkward coordinates through the C ABI against
the oracle: the lookup bit-exact, the pyramid within the 1e-3 relative bar, alt_corr within 1e-3.
Complements the fixed cases of test_gpu_lookup / test_gpu_pyramid / test_gpu_altcorr."""

import numpy as np
import pytest
import torch

import jax_raft_b200 as J
from oracle import corr_oracle as O

pytestmark = pytest.mark.gpu
F32 = np.float32


def _awkward_coords(rng, b, h, w):
    """Grid + noise, with shares of: exact integers, exact halves, the level borders -1 / 0 / size-1 /
    size, values just below an integer, far and very far (|c| up to 1e9) positions, negative zero."""
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
    return c


CASES = [(seed, *shape) for seed, shape in enumerate([
    (1, 8, 8, 1, 0), (1, 8, 9, 2, 1), (2, 9, 8, 2, 2), (1, 16, 16, 4, 3), (1, 16, 20, 4, 4), (3, 17, 19, 4, 4),
    (1, 21, 18, 3, 5), (1, 33, 40, 5, 3), (1, 32, 36, 4, 7), (2, 24, 24, 4, 3), (1, 40, 72, 4, 4), (1, 64, 33, 6, 2),
    (1, 13, 50, 3, 4), (4, 16, 16, 4, 4), (1, 30, 31, 4, 3), (1, 16, 64, 4, 4),
])]


@pytest.mark.parametrize("seed,b,h,w,levels,r", CASES)
def test_lookup_fuzz(cuda, lib, seed, b, h, w, levels, r):
    rng = np.random.default_rng(1000 + seed)
    q = b * h * w
    pyr = [rng.standard_normal((q, h >> l, w >> l, 1)).astype(F32) for l in range(levels)]
    coords = _awkward_coords(rng, b, h, w)
    cb = J.CorrBlock(levels, r)
    got = cb.index_pyramid([torch.from_numpy(p).to(cuda) for p in pyr], torch.from_numpy(coords).to(cuda)).cpu().numpy()
    with np.errstate(invalid="ignore"):
        ref = O.index_pyramid(pyr, coords, r)
    assert got.shape == ref.shape == (b, h, w, levels * (2 * r + 1) ** 2)
    assert np.array_equal(got, ref), f"{(got != ref).sum()} of {got.size} values differ"


@pytest.mark.parametrize("seed,b,h,w,levels,d", [(0, 1, 8, 8, 1, 32), (1, 2, 9, 11, 2, 64), (2, 1, 16, 24, 4, 96), (3, 1, 19, 17, 3, 256),
                                                  (4, 3, 16, 16, 4, 128), (5, 1, 40, 24, 4, 192), (6, 1, 64, 64, 6, 64), (7, 1, 27, 45, 4, 8)])
@pytest.mark.parametrize("kernel", ["tc", "simt"])
def test_pyramid_fuzz(cuda, lib, kernel, seed, b, h, w, levels, d):
    rng = np.random.default_rng(2000 + seed)
    f1 = (rng.standard_normal((b, h, w, d)) * rng.uniform(0.2, 3)).astype(F32)
    f2 = (rng.standard_normal((b, h, w, d)) * rng.uniform(0.2, 3)).astype(F32)
    cb = J.CorrBlock(levels, 4, kernel=kernel)
    pyr = cb.build_pyramid(torch.from_numpy(f1).to(cuda), torch.from_numpy(f2).to(cuda))
    ref = O.build_pyramid(f1, f2, num_levels=levels, dtype=np.float64)
    assert len(pyr) == levels
    for l, (got, want) in enumerate(zip(pyr, ref)):
        got = got.cpu().numpy()
        assert got.shape == want.shape == (b * h * w, h >> l, w >> l, 1)
        rms = max(float(np.sqrt((want ** 2).mean())), 1e-6)
        assert np.abs(got - want).max() / rms <= 1e-3, (kernel, l)


@pytest.mark.parametrize("seed,b,h,w,d,r", [(0, 1, 16, 16, 64, 4), (1, 2, 17, 23, 128, 3), (2, 1, 24, 40, 256, 4), (3, 1, 32, 19, 32, 3)])
def test_altcorr_fuzz(cuda, lib, seed, b, h, w, d, r):
    rng = np.random.default_rng(3000 + seed)
    f1 = rng.standard_normal((b, h, w, d)).astype(F32)
    f2 = rng.standard_normal((b, h, w, d)).astype(F32)
    coords = _awkward_coords(rng, b, h, w)
    cb = J.AltCorrBlock(4, r)
    state = cb.build_pyramid(torch.from_numpy(f1).to(cuda), torch.from_numpy(f2).to(cuda))
    got = cb.index_pyramid(state, torch.from_numpy(coords).to(cuda)).cpu().numpy()
    with np.errstate(invalid="ignore"):
        ref = O.alt_index(f1, O.pool_fmap2(f2, 4), coords, r)
    scale = max(float(np.sqrt((ref ** 2).mean())), 1e-6)
    assert np.abs(got - ref).max() / scale <= 1e-3
