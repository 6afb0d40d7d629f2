"""Parity of the window-lookup kernel (family 2) with the oracle, through the C ABI.
Bar: bit-exact (index work AND float32 arithmetic in the reference's order)."""

import numpy as np
import pytest
import torch

import jax_raft_b200 as J
from oracle import corr_oracle as O

pytestmark = pytest.mark.gpu
F32 = np.float32


def _lookup(pyr_np, coords_np, r, dev):
    cb = J.CorrBlock(len(pyr_np), r)
    pyr = [torch.from_numpy(p).to(dev) for p in pyr_np]
    out = cb.index_pyramid(pyr, torch.from_numpy(coords_np).to(dev))
    torch.cuda.synchronize()
    return out.cpu().numpy()


def _assert_bit_exact(got, ref):
    # -0.0 == +0.0 is accepted (a zero-weight corner outside the staged footprint), nothing else
    assert got.shape == ref.shape
    bad = ~((got == ref) | (np.isnan(got) & np.isnan(ref)))
    assert not bad.any(), f"{bad.sum()} of {bad.size} values differ; max |d| = {np.abs(got - ref)[bad].max()}"


def test_tap_order_fixture(cuda, lib, golden):
    g = golden("lookup_tap_order")
    h, w, r = int(g["h"]), int(g["w"]), int(g["radius"])
    pyr = []
    for l in range(4):
        yy, xx = np.meshgrid(np.arange(h >> l), np.arange(w >> l), indexing="ij")
        pyr.append(np.tile((l * 10000 + yy * 100 + xx).astype(F32)[None, :, :, None], (2, 1, 1, 1)))
    _assert_bit_exact(_lookup(pyr, g["coords"], r, cuda), g["out"])


@pytest.mark.parametrize("r", [3, 4])
def test_border_fixture(cuda, lib, golden, r):
    g = golden(f"lookup_borders_r{r}")
    pyr = [g[f"level{l}"] for l in range(4)]
    _assert_bit_exact(_lookup(pyr, g["coords"], r, cuda), g["out"])


@pytest.mark.parametrize("shape", [(1, 16, 16), (2, 19, 21), (1, 23, 31), (1, 46, 62)])
@pytest.mark.parametrize("r", [3, 4])
@pytest.mark.parametrize("seed", [0, 1])
def test_random_against_oracle(cuda, lib, shape, r, seed):
    b, h, w = shape
    rng = np.random.default_rng(seed)
    q = b * h * w
    pyr = [rng.standard_normal((q, h >> l, w >> l, 1)).astype(F32) for l in range(4)]
    coords = O.make_coords_grid(b, h, w) + rng.uniform(-8, 8, (b, h, w, 2)).astype(F32)
    # a 2 % share forced far out of bounds
    far = rng.random((b, h, w)) < 0.02
    coords[far] += rng.choice([-1, 1], size=(far.sum(), 2)).astype(F32) * 500
    _assert_bit_exact(_lookup(pyr, coords, r, cuda), O.index_pyramid(pyr, coords, r))


def test_other_radii_and_level_counts(cuda, lib):
    rng = np.random.default_rng(4)
    b, h, w = 1, 18, 20
    coords = O.make_coords_grid(b, h, w) + rng.uniform(-4, 4, (b, h, w, 2)).astype(F32)
    for levels, r in ((1, 1), (2, 2), (3, 5), (4, 0)):
        pyr = [rng.standard_normal((b * h * w, h >> l, w >> l, 1)).astype(F32) for l in range(levels)]
        _assert_bit_exact(_lookup(pyr, coords, r, cuda), O.index_pyramid(pyr, coords, r))


def test_empty_batch(cuda, lib):
    cb = J.CorrBlock(4, 4)
    pyr = [torch.zeros((0, 16 >> l, 16 >> l, 1), device=cuda) for l in range(4)]
    out = cb.index_pyramid(pyr, torch.zeros((0, 16, 16, 2), device=cuda))
    assert out.shape == (0, 16, 16, 324)


def test_sintel_shape_full_size(cuda, lib):
    """BASELINE config 2 (440x1024 -> 55x128, N = 7040) against the oracle on the same inputs."""
    rng = np.random.default_rng(1)
    b, h, w, r = 1, 55, 128, 4
    q = b * h * w
    pyr = [rng.standard_normal((q, h >> l, w >> l, 1)).astype(F32) for l in range(4)]
    coords = O.make_coords_grid(b, h, w) + rng.uniform(-8, 8, (b, h, w, 2)).astype(F32)
    _assert_bit_exact(_lookup(pyr, coords, r, cuda), O.index_pyramid(pyr, coords, r))


def test_linearity_at_batch_size(cuda, lib):
    """Size-independent property at a size the oracle is not run on (B = 8 Sintel-shape pairs):
    lookup(a*P1 + P2) == a*lookup(P1) + lookup(P2) for a power-of-two a (exact in float32 up to the
    final additions), and a zero pyramid gives exact zeros."""
    b, h, w, r = 8, 55, 128, 4
    q = b * h * w
    gen = torch.Generator(device=cuda).manual_seed(0)
    p1 = [torch.randn((q, h >> l, w >> l, 1), device=cuda, generator=gen) for l in range(4)]
    coords = J.make_coords_grid(b, h, w, device=cuda) + (torch.rand((b, h, w, 2), device=cuda, generator=gen) - 0.5) * 16
    cb = J.CorrBlock(4, r)
    o1 = cb.index_pyramid(p1, coords)
    o2 = cb.index_pyramid([4.0 * t for t in p1], coords)
    assert torch.equal(o2, 4.0 * o1)
    oz = cb.index_pyramid([torch.zeros_like(t) for t in p1], coords)
    assert not oz.any()


@pytest.mark.parametrize("c", [1, 3, 4])
def test_grid_sample_standalone(cuda, lib, c):
    """raftk_grid_sample (model.py:24-34) against the oracle: bit-exact, including the border cases."""
    rng = np.random.default_rng(c)
    n, h, w = 3, 7, 9
    img = rng.standard_normal((n, h, w, c)).astype(F32)
    grid = rng.uniform(-3, 12, (n, 6, 5, 2)).astype(F32)
    grid[0, 0, :4] = [[-1.0, 0.0], [w - 1.0, h - 1.0], [w - 0.5, 2.0], [1e6, -1e6]]
    got = J.grid_sample(torch.from_numpy(img).to(cuda), torch.from_numpy(grid).to(cuda)).cpu().numpy()
    _assert_bit_exact(got, O.grid_sample(img, grid))


def test_programmatic_dependent_launch_switch(cuda, lib):
    """raftk_set_pdl: the lookup launched behind another kernel with and without programmatic stream
    serialization gives the same bits, eagerly and from a captured graph; the call returns the old setting."""
    b, h, w, r = 1, 24, 40, 4
    gen = torch.Generator(device=cuda).manual_seed(2)
    q = b * h * w
    pyr = [torch.randn((q, h >> l, w >> l, 1), device=cuda, generator=gen) for l in range(4)]
    base = J.make_coords_grid(b, h, w, device=cuda)
    noise = (torch.rand((b, h, w, 2), device=cuda, generator=gen) - 0.5) * 12
    cb = J.CorrBlock(4, r)
    prev = lib.raftk_set_pdl(1)
    try:
        outs = {}
        for on in (0, 1):
            assert lib.raftk_set_pdl(on) in (0, 1)
            coords = base + noise          # produced by the kernel right before the lookup
            eager = cb.index_pyramid(pyr, coords)
            g = torch.cuda.CUDAGraph()
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                with torch.cuda.graph(g, stream=s):
                    c2 = base + noise
                    captured = cb.index_pyramid(pyr, c2)
            g.replay()
            torch.cuda.synchronize()
            assert torch.equal(eager, captured)
            outs[on] = eager
        assert torch.equal(outs[0], outs[1])
        assert lib.raftk_set_pdl(0) == 1 and lib.raftk_set_pdl(1) == 0
    finally:
        lib.raftk_set_pdl(prev)


@pytest.mark.parametrize("shape,r", [((1, 16, 24), 4), ((2, 19, 21), 3), ((1, 55, 128), 4)])
def test_lookup_with_fused_coordinate_update(cuda, lib, shape, r):
    """raftk_corr_lookup_update == raftk_flow_update followed by raftk_corr_lookup, bit for bit, including the
    in-place coordinates, the flow and its pitched copy."""
    b, h, w = shape
    gen = torch.Generator(device=cuda).manual_seed(h + r)
    q = b * h * w
    pyr = [torch.randn((q, h >> l, w >> l, 1), device=cuda, generator=gen) for l in range(4)]
    c0 = J.make_coords_grid(b, h, w, device=cuda)
    c1 = c0 + (torch.rand((b, h, w, 2), device=cuda, generator=gen) - 0.5) * 10
    c1[0, 1, 1] = 3.0e6  # a far query takes the slow path
    delta = torch.randn((b, h, w, 8), device=cuda, generator=gen)  # padded flow-head output
    bias = torch.randn((8,), device=cuda, generator=gen)
    cb = J.CorrBlock(4, r)
    ref_c1 = c1 + (delta[..., :2] + bias[:2])
    ref_flow = ref_c1 - c0
    ref_out = cb.index_pyramid(pyr, ref_c1)
    flow = torch.full((b, h, w, 2), 9.0, device=cuda)
    flow8 = torch.zeros((b, h, w, 8), device=cuda)
    got_c1 = c1.clone()
    out = cb.index_pyramid_update(pyr, c0, got_c1, delta, bias, flow, flow8)
    assert torch.equal(got_c1, ref_c1) and torch.equal(flow, ref_flow)
    assert torch.equal(flow8[..., :2], ref_flow) and not flow8[..., 2:].any()
    assert torch.equal(out, ref_out)
    assert not J.CorrBlock(3, 4).fused_conv_supported  # other configurations: flow_update + lookup
