"""Functional helpers of the reference (``/root/reference/jax_raft/model.py:24-98``), NHWC.

``upsample_flow`` with a mask runs the hand-written sm_100a kernel through the C ABI; the
mask-less bilinear branch (raft_small, model.py:79-83) is out of kernel scope
and stays on the host framework's library path.
"""

from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F

from . import _lib
from .corr import _check_cuda_f32


def make_coords_grid(batch_size: int, h: int, w: int, device="cuda") -> torch.Tensor:
    """``coords[b, y, x] = (x, y)`` float32 (model.py:37-40)."""
    ys, xs = torch.meshgrid(torch.arange(h, device=device), torch.arange(w, device=device), indexing="ij")
    coords = torch.stack([xs, ys], dim=-1).to(torch.float32)
    return coords[None].repeat(batch_size, 1, 1, 1)


def grid_sample(img: torch.Tensor, grid: torch.Tensor) -> torch.Tensor:
    """Bilinear sample of ``img`` (n, h, w, c) at absolute pixel coords ``grid`` (n, gh, gw, 2), given
    as (x, y); zero outside (model.py:24-34)."""
    assert img.shape[-3] > 1  # model.py:25
    assert grid.ndim == 4 and grid.shape[-1] == 2  # model.py:26
    img = _check_cuda_f32("img", img)
    grid = _check_cuda_f32("grid", grid)
    n, h, w, c = img.shape
    assert grid.shape[0] == n
    gh, gw = int(grid.shape[1]), int(grid.shape[2])
    out = torch.empty((n, gh, gw, c), dtype=torch.float32, device=img.device)
    lib = _lib.load()
    with torch.cuda.device(img.device):
        rc = lib.raftk_grid_sample(img.data_ptr(), grid.data_ptr(), out.data_ptr(), n, h, w, c, gh, gw,
                                   _lib.current_stream_ptr(img.device))
    _lib.check("raftk_grid_sample", rc)
    _lib.count_launch(1)
    return out


def resize_with_aligned_corners(image: torch.Tensor, shape, method: str = "bilinear", antialias: bool = True) -> torch.Tensor:
    """NHWC align-corners bilinear resize (model.py:43-66); host-framework code, not a kernel."""
    assert method == "bilinear", "currently only bilinear interpolation is supported"
    x = image.permute(0, 3, 1, 2)
    out = F.interpolate(x, size=(int(shape[1]), int(shape[2])), mode="bilinear", align_corners=True)
    return out.permute(0, 2, 3, 1).contiguous()


def upsample_flow(flow: torch.Tensor, up_mask: Optional[torch.Tensor] = None, factor: int = 8, *,
                  mask_bias: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Upsample flow (B, h, w, C) by ``factor`` (model.py:69-98).

    With ``up_mask`` (B, h, w, 9*factor*factor): convex combination of the 3x3 neighbours with
    softmax weights (model.py:85-98) — the CUDA kernel, which implements factor 8 and C = 2 (the
    only configuration the reference's models use).  Without: ``factor *`` align-corners bilinear
    resize (model.py:79-83).

    Not in the reference's signature: ``mask_bias`` (576,) is added to ``up_mask`` inside the kernel
    (the mask predictor's bias without a separate pass) and ``out`` is a preallocated contiguous
    (B, 8h, 8w, 2) result buffer.
    """
    batch_size, h, w, num_channels = flow.shape
    if up_mask is None:
        x = flow.permute(0, 3, 1, 2)
        up = F.interpolate(x, size=(h * factor, w * factor), mode="bilinear", align_corners=True)
        return (factor * up).permute(0, 2, 3, 1).contiguous()

    assert up_mask.shape == (batch_size, h, w, 9 * factor * factor)
    if factor != 8 or num_channels != 2:
        raise NotImplementedError("the convex upsampling kernel implements factor=8, 2 flow channels")
    flow = _check_cuda_f32("flow", flow)
    up_mask = _check_cuda_f32("up_mask", up_mask)
    if out is None:
        out = torch.empty((batch_size, 8 * h, 8 * w, 2), dtype=torch.float32, device=flow.device)
    else:
        assert out.shape == (batch_size, 8 * h, 8 * w, 2) and out.dtype == torch.float32 and out.is_contiguous()
        assert out.device == flow.device
    lib = _lib.load()
    with torch.cuda.device(flow.device):
        if mask_bias is None:
            rc = lib.raftk_convex_upsample(flow.data_ptr(), up_mask.data_ptr(), out.data_ptr(), batch_size, h, w,
                                           _lib.current_stream_ptr(flow.device))
        else:
            mask_bias = _check_cuda_f32("mask_bias", mask_bias)
            assert mask_bias.shape == (9 * factor * factor,)
            rc = lib.raftk_convex_upsample_bias(flow.data_ptr(), up_mask.data_ptr(), mask_bias.data_ptr(), out.data_ptr(),
                                                batch_size, h, w, _lib.current_stream_ptr(flow.device))
    _lib.check("raftk_convex_upsample", rc)
    _lib.count_launch(1)
    return out


def pack_mask_predictor(weight: torch.Tensor, bias: torch.Tensor):
    """Filters (576, K[, 1, 1]) and bias (576) of MaskPredictor's last convolution (model.py:394-400, any
    constant factor already folded in) -> the pass order ``raftk_mask_upsample`` reads: row
    ``g*144 + k*16 + t`` is mask channel ``k*64 + 16*g + t``."""
    n = weight.shape[0]
    assert n == 576 and bias.shape == (576,)
    w = weight.reshape(9, 4, 16, -1).permute(1, 0, 2, 3).reshape(576, -1)  # (k, g, t) -> (g, k, t)
    b = bias.reshape(9, 4, 16).permute(1, 0, 2).reshape(576)
    return w.to(torch.float32).contiguous(), b.to(torch.float32).contiguous()


def mask_upsample(flow: torch.Tensor, features: torch.Tensor, packed_weight: torch.Tensor, packed_bias: torch.Tensor,
                  out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``upsample_flow(flow, conv1x1(features) + bias)`` in one kernel (TF32 product),
    with ``features`` (B, h, w, K) the mask predictor's hidden activation and the weights from
    ``pack_mask_predictor``.  Not in the reference's API."""
    flow = _check_cuda_f32("flow", flow)
    features = _check_cuda_f32("features", features)
    w = _check_cuda_f32("packed_weight", packed_weight)
    bvec = _check_cuda_f32("packed_bias", packed_bias)
    b, h, wd, c = flow.shape
    assert c == 2 and features.shape[:3] == (b, h, wd)
    k = features.shape[3]
    assert w.shape == (576, k) and bvec.shape == (576,)
    if out is None:
        out = torch.empty((b, 8 * h, 8 * wd, 2), dtype=torch.float32, device=flow.device)
    else:
        assert out.shape == (b, 8 * h, 8 * wd, 2) and out.dtype == torch.float32 and out.is_contiguous()
    lib = _lib.load()
    with torch.cuda.device(flow.device):
        rc = lib.raftk_mask_upsample(features.data_ptr(), w.data_ptr(), bvec.data_ptr(), flow.data_ptr(), out.data_ptr(), b, h, wd,
                                     k, _lib.current_stream_ptr(flow.device))
    _lib.check("raftk_mask_upsample", rc)
    _lib.count_launch(1)
    return out
