"""raftk_corr_lookup_conv1x1 (lookup fused with the 1x1 convolution that consumes it).

The fused kernel's lookup half is the bit-exact tap arithmetic of raftk_corr_lookup; its product is
tcgen05 kind::tf32.  Checked (a) tightly against a float64 product of the TF32-truncated operands
(the hardware drops the low 13 mantissa bits of each float32 operand), which pins tap order, K padding
and the swizzled A tile, and (b) against the float32 convolution within the TF32 bar."""

import numpy as np
import pytest
import torch

import jax_raft_b200 as J
from oracle import corr_oracle as O

pytestmark = pytest.mark.gpu
F32 = np.float32


def _tf32_trunc(a):
    return (np.ascontiguousarray(a, F32).view(np.uint32) & np.uint32(0xFFFFE000)).view(F32)


def _case(rng, b, h, w, r, n):
    q = b * h * w
    pyr = [rng.standard_normal((q, h >> l, w >> l, 1)).astype(F32) for l in range(4)]
    coords = O.make_coords_grid(b, h, w) + rng.uniform(-8, 8, (b, h, w, 2)).astype(F32)
    far = rng.random((b, h, w)) < 0.02
    coords[far] += rng.choice([-1, 1], size=(far.sum(), 2)).astype(F32) * 500
    c = 4 * (2 * r + 1) ** 2
    weight = (rng.standard_normal((n, c)) / np.sqrt(c)).astype(F32)
    bias = rng.standard_normal(n).astype(F32)
    return pyr, coords, weight, bias


@pytest.mark.parametrize("b,h,w,r,n", [(1, 16, 16, 4, 256), (2, 19, 21, 3, 96), (1, 23, 31, 4, 64), (1, 46, 62, 3, 96),
                                       (1, 55, 128, 4, 256), (3, 16, 24, 4, 16)])
@pytest.mark.parametrize("relu", [True, False])
def test_fused_lookup_conv_against_oracle(cuda, lib, b, h, w, r, n, relu):
    rng = np.random.default_rng(b * 1000 + h * 10 + r)
    pyr, coords, weight, bias = _case(rng, b, h, w, r, n)
    feats = O.index_pyramid(pyr, coords, r).reshape(-1, weight.shape[1])
    ref32 = feats.astype(np.float64) @ weight.astype(np.float64).T + bias
    ref_tf = _tf32_trunc(feats).astype(np.float64) @ _tf32_trunc(weight).astype(np.float64).T + bias
    if relu:
        ref32, ref_tf = np.maximum(ref32, 0), np.maximum(ref_tf, 0)

    cb = J.CorrBlock(4, r)
    packed = cb.pack_conv_weight(torch.from_numpy(weight).to(cuda))
    got = cb.index_pyramid_conv([torch.from_numpy(p).to(cuda) for p in pyr], torch.from_numpy(coords).to(cuda), packed,
                                torch.from_numpy(bias).to(cuda), relu=relu).cpu().numpy().reshape(-1, n)
    scale = float(np.sqrt((ref32 ** 2).mean())) + 1e-6
    assert np.abs(got - ref_tf).max() / scale <= 2e-5, "differs from the product of the TF32-truncated operands"
    assert np.abs(got - ref32).max() / scale <= 5e-3   # TF32 bar against the float32 convolution


def test_matches_separate_kernels_on_device(cuda, lib):
    """Against raftk_corr_lookup followed by a float32 matmul, at the bench shape."""
    b, h, w, r, n = 1, 55, 128, 4, 256
    gen = torch.Generator(device=cuda).manual_seed(3)
    q = b * h * w
    pyr = [torch.randn((q, h >> l, w >> l, 1), device=cuda, generator=gen) for l in range(4)]
    coords = J.make_coords_grid(b, h, w, device=cuda) + (torch.rand((b, h, w, 2), device=cuda, generator=gen) - 0.5) * 16
    weight = torch.randn((n, 324), device=cuda, generator=gen) / 18
    bias = torch.randn((n,), device=cuda, generator=gen)
    cb = J.CorrBlock(4, r)
    feats = cb.index_pyramid(pyr, coords).reshape(q, 324)
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        ref = torch.relu(feats.double() @ weight.double().T + bias).float()
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    got = cb.index_pyramid_conv(pyr, coords, cb.pack_conv_weight(weight), bias).reshape(q, n)
    assert (got - ref).abs().max() / ref.pow(2).mean().sqrt() <= 5e-3


def test_edge_cases(cuda, lib):
    cb = J.CorrBlock(4, 4)
    w = torch.zeros((32, 324), device=cuda)
    out = cb.index_pyramid_conv([torch.zeros((0, 16 >> l, 16 >> l, 1), device=cuda) for l in range(4)],
                                torch.zeros((0, 16, 16, 2), device=cuda), cb.pack_conv_weight(w), torch.zeros(32, device=cuda))
    assert out.shape == (0, 16, 16, 32)
    # a batch whose last tile is ragged (16 * 17 = 272 = 4 * 64 + 16 queries) and non-finite coordinates
    rng = np.random.default_rng(0)
    pyr, coords, weight, bias = _case(rng, 1, 16, 17, 4, 32)
    coords[0, 3, 3] = [np.inf, 2.0]
    got = cb.index_pyramid_conv([torch.from_numpy(p).to(cuda) for p in pyr], torch.from_numpy(coords).to(cuda),
                                cb.pack_conv_weight(torch.from_numpy(weight).to(cuda)), torch.from_numpy(bias).to(cuda),
                                relu=False).cpu().numpy()
    assert np.isnan(got[0, 3, 3]).all() and np.isfinite(np.delete(got.reshape(-1, 32), 3 * 17 + 3, axis=0)).all()
    assert not J.CorrBlock(3, 4).fused_conv_supported
