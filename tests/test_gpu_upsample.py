"""Parity of the convex upsampling kernel (family 3) with the oracle, through the C ABI.
Tolerance: 2e-6 relative to the flow scale (expf vs numpy exp, summation order) — the north-star's
bar for floating point here is the final 0.01 px EPE; this is four orders tighter."""

import numpy as np
import pytest
import torch

import jax_raft_b200 as J
from oracle import corr_oracle as O

pytestmark = pytest.mark.gpu
F32 = np.float32


def _up(flow, mask, dev):
    out = J.upsample_flow(torch.from_numpy(flow).to(dev), torch.from_numpy(mask).to(dev))
    torch.cuda.synchronize()
    return out.cpu().numpy()


def _close(got, ref):
    scale = max(1.0, float(np.abs(ref).max()))
    np.testing.assert_allclose(got, ref, rtol=0, atol=2e-6 * scale)


def test_onehot_fixture_neighbour_order_and_zero_padding(cuda, lib, golden):
    g = golden("upsample")
    flow = g["flow"]
    b, h, w, _ = flow.shape
    for k in range(9):
        m = np.zeros((b, h, w, 576), F32)
        m[..., k * 64:(k + 1) * 64] = 40.0
        _close(_up(flow, m, cuda), g["onehot_out"][k])
    _close(_up(flow, g["random_mask"], cuda), g["random_out"])


@pytest.mark.parametrize("shape", [(1, 1, 1), (1, 2, 3), (2, 7, 5), (1, 46, 62), (1, 55, 128)])
def test_random_against_oracle(cuda, lib, shape):
    b, h, w = shape
    rng = np.random.default_rng(h * 131 + w)
    flow = (rng.standard_normal((b, h, w, 2)) * 5).astype(F32)
    mask = (rng.standard_normal((b, h, w, 576)) * 2).astype(F32)
    _close(_up(flow, mask, cuda), O.upsample_flow(flow, mask))


def test_convexity_properties_at_batch_size(cuda, lib):
    """Size-independent properties at B = 16 Sintel-shape maps: a constant flow stays constant
    (weights sum to 1) away from the zero-padded border, and output is bounded by 8 * max |flow|."""
    b, h, w = 16, 55, 128
    gen = torch.Generator(device=cuda).manual_seed(0)
    mask = torch.randn((b, h, w, 576), device=cuda, generator=gen) * 3
    const = torch.full((b, h, w, 2), 1.25, device=cuda)
    up = J.upsample_flow(const, mask)
    inner = up[:, 8:-8, 8:-8]
    assert torch.allclose(inner, torch.full_like(inner, 10.0), rtol=0, atol=1e-5)
    flow = torch.randn((b, h, w, 2), device=cuda, generator=gen)
    up = J.upsample_flow(flow, mask)
    assert up.shape == (b, 8 * h, 8 * w, 2)
    assert float(up.abs().max()) <= 8 * float(flow.abs().max()) * (1 + 1e-5)


def test_empty_batch(cuda, lib):
    out = J.upsample_flow(torch.zeros((0, 4, 4, 2), device=cuda), torch.zeros((0, 4, 4, 576), device=cuda))
    assert out.shape == (0, 32, 32, 2)


def test_bias_variant_and_out_buffer(cuda, lib):
    """raftk_convex_upsample_bias: logits mask + bias[channel] give the bits of the plain entry on the
    pre-added mask; ``out=`` writes into the caller's buffer."""
    b, h, w = 2, 9, 14
    gen = torch.Generator(device=cuda).manual_seed(7)
    flow = torch.randn((b, h, w, 2), device=cuda, generator=gen) * 4
    mask = torch.randn((b, h, w, 576), device=cuda, generator=gen) * 2
    bias = torch.randn((576,), device=cuda, generator=gen)
    ref = J.upsample_flow(flow, mask + bias)
    buf = torch.full((3, b, 8 * h, 8 * w, 2), -1.0, device=cuda)
    got = J.upsample_flow(flow, mask, mask_bias=bias, out=buf[1])
    assert got.data_ptr() == buf[1].data_ptr()
    assert torch.equal(buf[1], ref)
    assert (buf[0] == -1).all() and (buf[2] == -1).all()
