"""raftk_mask_upsample (MaskPredictor's 1x1 convolution fused with the convex upsampler).
Reference: the oracle's upsample_flow applied to a float64 product of the TF32-truncated operands (the
tensor core drops the low 13 mantissa bits of each float32 operand) — tight, pins the weight permutation,
the four-pass column order and the TMEM addressing — and to the float32 product within the TF32 bar."""

import numpy as np
import pytest
import torch

import jax_raft_b200 as J
from oracle import corr_oracle as O

pytestmark = pytest.mark.gpu
F32 = np.float32


def _tf32_trunc(a):
    return (np.ascontiguousarray(a, F32).view(np.uint32) & np.uint32(0xFFFFE000)).view(F32)


@pytest.mark.parametrize("b,h,w,k", [(1, 8, 16, 256), (2, 7, 5, 256), (1, 46, 62, 96), (1, 55, 128, 256), (3, 1, 1, 32)])
def test_against_oracle(cuda, lib, b, h, w, k):
    rng = np.random.default_rng(b * 100 + h + k)
    feats = np.maximum(rng.standard_normal((b, h, w, k)), 0).astype(F32)
    weight = (rng.standard_normal((576, k)) * (2.0 / np.sqrt(k))).astype(F32)
    bias = rng.standard_normal(576).astype(F32)
    flow = (rng.standard_normal((b, h, w, 2)) * 4).astype(F32)
    m64 = feats.reshape(-1, k).astype(np.float64)
    mask32 = (m64 @ weight.astype(np.float64).T + bias).reshape(b, h, w, 576)
    mask_tf = (_tf32_trunc(feats).reshape(-1, k).astype(np.float64) @ _tf32_trunc(weight).astype(np.float64).T + bias).reshape(b, h, w, 576)
    ref_tf = O.upsample_flow(flow, mask_tf.astype(F32))
    ref32 = O.upsample_flow(flow, mask32.astype(F32))
    pw, pb = J.pack_mask_predictor(torch.from_numpy(weight).to(cuda), torch.from_numpy(bias).to(cuda))
    got = J.mask_upsample(torch.from_numpy(flow).to(cuda), torch.from_numpy(feats).to(cuda), pw, pb).cpu().numpy()
    scale = max(1.0, float(np.abs(ref32).max()))
    assert got.shape == ref32.shape
    assert np.abs(got - ref_tf).max() <= 2e-5 * scale
    assert np.abs(got - ref32).max() <= 2e-2 * scale  # TF32 logits in front of an exponential


def test_equals_separate_kernels_and_out_buffer(cuda, lib):
    """Against the stand-alone upsampler fed the same TF32-precision logits' float32 counterpart, and the
    out= buffer contract."""
    b, h, w, k = 2, 9, 14, 256
    gen = torch.Generator(device=cuda).manual_seed(4)
    feats = torch.relu(torch.randn((b, h, w, k), device=cuda, generator=gen))
    weight = torch.randn((576, k), device=cuda, generator=gen) / 16
    bias = torch.randn((576,), device=cuda, generator=gen)
    flow = torch.randn((b, h, w, 2), device=cuda, generator=gen) * 3
    mask = (feats.reshape(-1, k).double() @ weight.double().T + bias).float().reshape(b, h, w, 576)
    ref = J.upsample_flow(flow, mask)
    pw, pb = J.pack_mask_predictor(weight, bias)
    buf = torch.full((2, b, 8 * h, 8 * w, 2), -7.0, device=cuda)
    got = J.mask_upsample(flow, feats, pw, pb, out=buf[1])
    assert got.data_ptr() == buf[1].data_ptr() and (buf[0] == -7.0).all()
    assert (got - ref).abs().max() <= 2e-2 * max(1.0, float(ref.abs().max()))
    empty = J.mask_upsample(torch.zeros((0, 4, 4, 2), device=cuda), torch.zeros((0, 4, 4, k), device=cuda), pw, pb)
    assert empty.shape == (0, 32, 32, 2)
    assert lib.raftk_mask_upsample(feats.data_ptr(), pw.data_ptr(), pb.data_ptr(), flow.data_ptr(), buf.data_ptr(), b, h, w, 48, None) == -4


@pytest.mark.parametrize("b,h,w", [(3, 55, 128), (1, 139, 137)])
def test_more_tiles_than_sms(cuda, lib, b, h, w):
    """More 128-pixel tiles than SMs (several waves of CTAs, no pass split): (3, 55, 128) = 165 tiles, (1, 139, 137) =
    149 tiles with a ragged last one.  Checked against the float64 product of the TF32-truncated operands on the first
    and last image rows (tight bar) and everywhere against the stand-alone upsampler on float32 logits (TF32 bar)."""
    k = 256
    gen = torch.Generator(device=cuda).manual_seed(b + h)
    feats = torch.relu(torch.randn((b, h, w, k), device=cuda, generator=gen))
    weight = torch.randn((576, k), device=cuda, generator=gen) * (2.0 / 16)
    bias = torch.randn((576,), device=cuda, generator=gen)
    flow = torch.randn((b, h, w, 2), device=cuda, generator=gen) * 4
    assert (b * h * w + 127) // 128 >= torch.cuda.get_device_properties(cuda).multi_processor_count
    pw, pb = J.pack_mask_predictor(weight, bias)
    got = J.mask_upsample(flow, feats, pw, pb)
    mask32 = (feats.reshape(-1, k).double() @ weight.double().T + bias).float().reshape(b, h, w, 576)
    ref32 = J.upsample_flow(flow, mask32)
    scale = max(1.0, float(ref32.abs().max()))
    assert float((got - ref32).abs().max()) <= 2e-2 * scale
    # tight check on the first and the last image rows of every pair (covers the first tiles and the ragged end)
    f_np, w_np = feats.cpu().numpy(), weight.cpu().numpy()
    mask_tf = (_tf32_trunc(f_np).reshape(-1, k).astype(np.float64) @ _tf32_trunc(w_np).astype(np.float64).T
               + bias.cpu().numpy()).reshape(b, h, w, 576).astype(F32)
    ref_tf = O.upsample_flow(flow.cpu().numpy(), mask_tf)
    g = got.cpu().numpy()
    for rows in (slice(0, 16), slice(8 * h - 16, 8 * h)):
        assert np.abs(g[:, rows] - ref_tf[:, rows]).max() <= 2e-5 * scale
