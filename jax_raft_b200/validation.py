"""Sintel-style validation harness, mirroring
``/root/reference/scripts/validate_sintel.py``: ``InputPadder`` (``:23-40``), Middlebury ``.flo``
reader (``:42-61``), the EPE / 1-3-5 px metrics and fps accounting of ``validate_sintel_jax``
(``:164-206``).  The dataset itself is not available offline; ``validate_pairs`` takes any iterable of
(image1, image2, flow_gt) arrays, so it runs on synthetic pairs in the tests and on Sintel when the
files exist.
"""

from __future__ import annotations

import time
from typing import Dict, Iterable, Optional, Tuple

import numpy as np


class InputPadder:
    """Pads (..., H, W, C) NHWC images so H and W are divisible by 8 (validate_sintel.py:23-40).

    'sintel' mode splits the padding evenly top/bottom and left/right; any other mode pads the
    bottom only.  Padding replicates the border (torch ``F.pad(mode='replicate')``)."""

    def __init__(self, dims, mode: str = "sintel"):
        self.ht, self.wd = dims[-3], dims[-2]
        pad_ht = (((self.ht // 8) + 1) * 8 - self.ht) % 8
        pad_wd = (((self.wd // 8) + 1) * 8 - self.wd) % 8
        if mode == "sintel":
            self._pad = [pad_wd // 2, pad_wd - pad_wd // 2, pad_ht // 2, pad_ht - pad_ht // 2]
        else:
            self._pad = [pad_wd // 2, pad_wd - pad_wd // 2, 0, pad_ht]

    def pad(self, *inputs):
        l, r, t, b = self._pad
        return [np.pad(x, [(0, 0)] * (x.ndim - 3) + [(t, b), (l, r), (0, 0)], mode="edge") for x in inputs]

    def unpad(self, x):
        ht, wd = x.shape[-3], x.shape[-2]
        l, r, t, b = self._pad
        return x[..., t:ht - b, l:wd - r, :]


def read_flo(path: str) -> Optional[np.ndarray]:
    """Middlebury .flo -> (H, W, 2) float32 (validate_sintel.py:42-61; little-endian)."""
    with open(path, "rb") as f:
        magic = np.fromfile(f, np.float32, count=1)
        if magic.size != 1 or magic[0] != np.float32(202021.25):
            return None
        w = int(np.fromfile(f, np.int32, count=1)[0])
        h = int(np.fromfile(f, np.int32, count=1)[0])
        data = np.fromfile(f, np.float32, count=2 * w * h)
    return data.reshape(h, w, 2)


def write_flo(path: str, flow: np.ndarray) -> None:
    h, w, _ = flow.shape
    with open(path, "wb") as f:
        np.array([202021.25], np.float32).tofile(f)
        np.array([w, h], np.int32).tofile(f)
        np.ascontiguousarray(flow, np.float32).tofile(f)


def epe_metrics(epe_all: np.ndarray) -> Dict[str, float]:
    """Mean end-point error and the 1 / 3 / 5 px inlier rates (validate_sintel.py:196-203)."""
    return {"epe": float(np.mean(epe_all)), "1px": float(np.mean(epe_all < 1)), "3px": float(np.mean(epe_all < 3)),
            "5px": float(np.mean(epe_all < 5))}


def validate_pairs(model, variables, pairs: Iterable[Tuple[np.ndarray, np.ndarray, np.ndarray]], iters: int = 32,
                   mode: str = "sintel") -> Dict[str, float]:
    """The loop of validate_sintel_jax (validate_sintel.py:164-206) over ``pairs`` of uint8-range
    (H, W, 3) frames and (H, W, 2) ground-truth flow: normalise to [-1, 1] (``:177-178``), pad to a
    multiple of 8, time host->device + forward + sync (``:185-186``), drop the first sample from the
    timing (compile / graph capture, ``:187-188``), unpad, accumulate EPE."""
    import torch

    epe_list, times = [], []
    for n, (image1, image2, flow_gt) in enumerate(pairs):
        im1 = (np.asarray(image1, np.float32) / 255.0 * 2.0 - 1.0)[None]
        im2 = (np.asarray(image2, np.float32) / 255.0 * 2.0 - 1.0)[None]
        padder = InputPadder(im1.shape, mode)
        im1, im2 = padder.pad(im1, im2)
        t0 = time.time()
        flows = model.apply(variables, im1, im2, train=False, num_flow_updates=iters)
        torch.cuda.synchronize()
        if n > 0:
            times.append(time.time() - t0)
        flow = padder.unpad(flows[-1].cpu().numpy())[0]
        epe_list.append(np.sqrt(((flow - flow_gt) ** 2).sum(-1)).reshape(-1))
    out = epe_metrics(np.concatenate(epe_list))
    out["fps"] = float(1.0 / np.mean(times)) if times else float("nan")
    return out
