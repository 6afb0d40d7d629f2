"""CPU restatement of the reference's whole forward pass.  TEST INFRASTRUCTURE — not shipped.

Follows ``RAFT.__call__`` (``/root/reference/jax_raft/model.py:557-605``) and ``UpdateCell.__call__``
(``:494-510``) step by step.  The hot path — ``build_pyramid``, ``index_pyramid``, ``upsample_flow``
— is the NumPy oracle (``oracle/corr_oracle.py``); the dense-convolution sub-networks (encoders,
update block, mask predictor: out of kernel scope) are torchvision 0.26's RAFT
modules on the CPU in float32, i.e. the implementation the reference was ported from and
validated against (``scripts/validate_sintel.py:257-263``).  Weights are a torchvision
``state_dict`` (shared with the GPU build through the converter's rules).

Used (a) as the end-to-end checker in tests and ``smoke()``, (b) as the timed CPU baseline of
``bench.py`` (``cpu_baseline`` and ``--impl reference``), with per-phase timers so a bounded sample
(fewer iterations) can be extrapolated to the full iteration count.
"""

from __future__ import annotations

import time
from typing import Dict, List, Tuple

import numpy as np
import torch

from . import corr_oracle as O


def build_modules(name: str, state_dict=None, seed: int = 0):
    """torchvision ``raft_large`` / ``raft_small`` sub-modules on the CPU (eval mode)."""
    import torchvision.models.optical_flow as tvof

    torch.manual_seed(seed)
    net = getattr(tvof, name)(weights=None).eval()
    if state_dict is not None:
        net.load_state_dict(state_dict, strict=False)
    return net


def _nchw(x: np.ndarray) -> torch.Tensor:
    return torch.from_numpy(np.ascontiguousarray(x)).permute(0, 3, 1, 2)


def _nhwc(t: torch.Tensor) -> np.ndarray:
    return np.ascontiguousarray(t.permute(0, 2, 3, 1).numpy())


@torch.no_grad()
def forward(net, image1: np.ndarray, image2: np.ndarray, num_flow_updates: int = 12,
            timers: Dict[str, float] | None = None, hot=None) -> np.ndarray:
    """image1/2 (B, H, W, 3) float32 in [-1, 1] -> flows (N, B, H, W, 2).  model.py:557-605.

    ``hot``: the module providing build_pyramid / index_pyramid / upsample_flow — the NumPy oracle
    (default) or its multi-threaded C twin ``oracle.corr_oracle_c`` (the timed CPU baseline)."""
    t = timers if timers is not None else {}
    H_ = hot if hot is not None else O

    def tick(key, t0):
        t[key] = t.get(key, 0.0) + (time.perf_counter() - t0)

    batch_size, h, w, _ = image1.shape
    assert (h, w) == image2.shape[-3:-1], "input images should have the same shape"  # :559
    assert (h % 8 == 0) and (w % 8 == 0), "input image H and W should be divisible by 8"  # :560
    radius = net.corr_block.radius
    num_levels = net.corr_block.num_levels

    t0 = time.perf_counter()
    fmaps = net.feature_encoder(torch.cat([_nchw(image1), _nchw(image2)], dim=0))  # :562
    fmap1, fmap2 = (_nhwc(f) for f in torch.chunk(fmaps, 2, dim=0))  # :563
    tick("encoders", t0)

    t0 = time.perf_counter()
    pyramid = H_.build_pyramid(fmap1, fmap2, num_levels)  # :567
    tick("build_pyramid", t0)

    t0 = time.perf_counter()
    context_out = net.context_encoder(_nchw(image1))  # :569
    hid = net.update_block.hidden_state_size  # :576
    hidden_state = torch.tanh(context_out[:, :hid])  # :582-583
    context = torch.relu(context_out[:, hid:])  # :584
    tick("encoders", t0)

    coords0 = O.make_coords_grid(batch_size, h // 8, w // 8)  # :586
    coords1 = coords0.copy()  # :587
    flows: List[np.ndarray] = []
    for _ in range(num_flow_updates):  # nn.scan(UpdateCell), :589-603
        t0 = time.perf_counter()
        corr_features = H_.index_pyramid(pyramid, coords1, radius)  # :499
        tick("index_pyramid", t0)

        t0 = time.perf_counter()
        flow = coords1 - coords0  # :501
        hidden_state, delta_flow = net.update_block(hidden_state, context, _nchw(corr_features), _nchw(flow))  # :502
        coords1 = coords1 + _nhwc(delta_flow)  # :505
        up_mask = None if net.mask_predictor is None else _nhwc(net.mask_predictor(hidden_state))  # :507
        tick("update_block", t0)

        t0 = time.perf_counter()
        flows.append(H_.upsample_flow(coords1 - coords0, up_mask))  # :508
        tick("upsample_flow", t0)
    return np.stack(flows, axis=0)


def time_pair(net, image1: np.ndarray, image2: np.ndarray, sample_iters: int, full_iters: int,
              hot=None) -> Tuple[float, Dict[str, float]]:
    """Seconds per image pair at ``full_iters`` iterations, measured on a bounded sample of
    ``sample_iters`` iterations: one-off phases (encoders, pyramid) counted once, per-iteration
    phases scaled by ``full_iters / sample_iters``.  Exact when ``sample_iters == full_iters``."""
    timers: Dict[str, float] = {}
    forward(net, image1, image2, sample_iters, timers, hot=hot)
    scale = full_iters / float(sample_iters)
    per_pair = timers["encoders"] + timers["build_pyramid"] + scale * (
        timers["index_pyramid"] + timers["update_block"] + timers["upsample_flow"])
    return per_pair, timers
