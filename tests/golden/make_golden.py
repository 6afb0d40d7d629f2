"""Generates the committed golden vectors under tests/golden/ (run from the repo root:
``python tests/golden/make_golden.py``).

The reference itself cannot be imported (jax / flax are not installed), so the
vectors are produced by the NumPy oracle and, where an independent implementation of the same
maths exists in this image, cross-checked against it before being written:

* torchvision 0.26 ``CorrBlock`` / ``upsample_flow`` / ``raft_small`` / ``raft_large`` — the
  upstream the reference was ported from and validated against (validate_sintel.py:257-263);
* hand-derived known answers (tap order, border zero-fill, pooling bounds, one-hot masks).

A failed cross-check aborts generation.
"""

from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import corr_oracle as O  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
F32 = np.float32


def encoded_pyramid(q, h, w, levels=4):
    """Level l of every query slab holds l*10000 + y*100 + x, so a sampled value names its source."""
    pyr = []
    for l in range(levels):
        hl, wl = h >> l, w >> l
        yy, xx = np.meshgrid(np.arange(hl), np.arange(wl), indexing="ij")
        slab = (l * 10000 + yy * 100 + xx).astype(F32)
        pyr.append(np.tile(slab[None, :, :, None], (q, 1, 1, 1)))
    return pyr


def gen_tap_order():
    h, w, r = 48, 64, 4
    coords = np.array([[[[24.0, 16.0], [40.0, 32.0]]]], F32)  # (1,1,2,2): integer at every level
    pyr = encoded_pyramid(2, h, w)
    out = O.index_pyramid(pyr, coords, r)
    # known answer: channel l*81 + i*9 + j reads (x/2^l + i - r, y/2^l + j - r) exactly
    for qi, (x, y) in enumerate(coords.reshape(-1, 2)):
        for l in range(4):
            for i in range(9):
                for j in range(9):
                    sx, sy = int(x) // 2 ** l + i - r, int(y) // 2 ** l + j - r
                    inb = 0 <= sx < (w >> l) and 0 <= sy < (h >> l)
                    exp = l * 10000 + sy * 100 + sx if inb else 0.0
                    got = out.reshape(2, -1)[qi, l * 81 + i * 9 + j]
                    assert got == exp, (qi, l, i, j, got, exp)
    np.savez_compressed(os.path.join(OUT, "lookup_tap_order.npz"), h=h, w=w, radius=r, coords=coords, out=out)


def gen_borders():
    h, w = 20, 24
    rng = np.random.default_rng(7)
    pts = []
    for x in (-1.0, -0.5, 0.0, 0.25, w - 1.0, w - 0.5, float(w), w + 7.5, -30.0, 1e5, -1e5):
        for y in (-1.0, -0.5, 0.0, h - 1.0, h - 0.5, float(h), 3.75):
            pts.append((x, y))
    # coordinates whose fl(c + d) rounds up to an integer for some tap (weight-0 corner case)
    pts += [(np.nextafter(F32(0), F32(-1)), 5.0), (-1e-9, -1e-9), (np.nextafter(F32(8), F32(0)), 2.5)]
    coords = np.array(pts, F32).reshape(1, 1, -1, 2)
    q = coords.shape[2]
    for r in (3, 4):
        pyr = [rng.standard_normal((q, h >> l, w >> l, 1)).astype(F32) for l in range(4)]
        out = O.index_pyramid(pyr, coords, r)
        # cross-check against scipy's per-corner zero fill (float64) for a few queries
        from scipy.ndimage import map_coordinates
        d = np.arange(-r, r + 1, dtype=F32)
        for qi in range(0, q, 5):
            c = coords.reshape(-1, 2)[qi]
            for l in range(4):
                cl = (c / F32(2 ** l)).astype(F32)
                X, Y = np.meshgrid((cl[0] + d).astype(F32), (cl[1] + d).astype(F32), indexing="ij")
                ref = map_coordinates(pyr[l][qi, :, :, 0].astype(np.float64),
                                      [Y.ravel().astype(np.float64), X.ravel().astype(np.float64)],
                                      order=1, mode="grid-constant", cval=0.0)
                S = (2 * r + 1) ** 2
                got = out.reshape(q, -1)[qi, l * S:(l + 1) * S]
                assert np.allclose(got, ref, rtol=1e-5, atol=1e-5), (r, qi, l)
        np.savez_compressed(os.path.join(OUT, f"lookup_borders_r{r}.npz"), h=h, w=w, radius=r, coords=coords,
                            out=out, **{f"level{l}": pyr[l] for l in range(4)})


def gen_pyramid():
    from torchvision.models.optical_flow.raft import CorrBlock
    rng = np.random.default_rng(3)
    for name, (h, w, d) in {"odd": (19, 21, 24), "small": (23, 31, 128)}.items():
        f1 = (rng.standard_normal((1, h, w, d)) * 1.45).astype(F32)
        f2 = (rng.standard_normal((1, h, w, d)) * 1.45).astype(F32)
        pyr = O.build_pyramid(f1, f2)
        pyr64 = O.build_pyramid(f1, f2, dtype=np.float64)
        cb = CorrBlock(num_levels=4, radius=4)
        cb.build_pyramid(torch.from_numpy(f1).permute(0, 3, 1, 2), torch.from_numpy(f2).permute(0, 3, 1, 2))
        for l, (a, t) in enumerate(zip(pyr, cb.corr_pyramid)):
            assert a.shape[1:3] == (h >> l, w >> l) == tuple(t.shape[-2:])
            assert np.allclose(a[..., 0], t[:, 0].numpy(), rtol=1e-5, atol=1e-5)
        # pooling bounds: level l cell (i, j) is the mean of the 2^l x 2^l level-0 block
        v0 = pyr64[0][..., 0]
        for l in range(1, 4):
            s = 2 ** l
            hl, wl = h >> l, w >> l
            blk = v0[:, : hl * s, : wl * s].reshape(-1, hl, s, wl, s).mean(axis=(2, 4))
            assert np.allclose(pyr64[l][..., 0], blk, rtol=1e-12, atol=1e-12)
        # keep the fixture small: every 8th query slab only (the inputs are stored in full)
        np.savez_compressed(os.path.join(OUT, f"pyramid_{name}.npz"), fmap1=f1, fmap2=f2, query_stride=8,
                            **{f"level{l}": pyr[l][::8] for l in range(4)})


def gen_upsample():
    from torchvision.models.optical_flow._utils import upsample_flow as tv_up
    rng = np.random.default_rng(11)
    b, h, w = 2, 5, 6
    flow = (rng.standard_normal((b, h, w, 2)) * 4).astype(F32)
    # one-hot masks: plane k gets +40, so softmax picks neighbour k exactly (exp(-40) ~ 4e-18)
    masks, outs = [], []
    for k in range(9):
        m = np.zeros((b, h, w, 576), F32)
        m[..., k * 64:(k + 1) * 64] = 40.0
        up = O.upsample_flow(flow, m)
        dy, dx = k // 3 - 1, k % 3 - 1
        fp = np.zeros((b, h + 2, w + 2, 2), F32)
        fp[:, 1:-1, 1:-1] = flow
        exp = 8 * np.repeat(np.repeat(fp[:, 1 + dy:1 + dy + h, 1 + dx:1 + dx + w], 8, axis=1), 8, axis=2)
        assert np.allclose(up, exp, rtol=1e-6, atol=1e-6), k
        masks.append(m)
        outs.append(up)
    mr = rng.standard_normal((b, h, w, 576)).astype(F32)
    upr = O.upsample_flow(flow, mr)
    tv = tv_up(torch.from_numpy(flow).permute(0, 3, 1, 2), torch.from_numpy(mr).permute(0, 3, 1, 2))
    assert np.allclose(upr, tv.permute(0, 2, 3, 1).numpy(), rtol=1e-5, atol=1e-5)
    np.savez_compressed(os.path.join(OUT, "upsample.npz"), flow=flow, onehot_out=np.stack(outs),
                        random_mask=mr, random_out=upr)


def synthetic_pair(h, w, seed=1):
    """uniform [-1,1] frame; frame 2 = frame 1 rolled by (3, -5) px + N(0, 0.02^2)."""
    rng = np.random.default_rng(seed)
    img1 = rng.uniform(-1, 1, (1, h, w, 3)).astype(F32)
    img2 = np.roll(img1, (3, -5), axis=(1, 2)) + rng.normal(0, 0.02, img1.shape).astype(F32)
    return img1, np.clip(img2, -1, 1).astype(F32)


def gen_e2e():
    import torchvision.models.optical_flow as tvof
    for name, iters in (("raft_small", 12), ("raft_large", 6)):
        torch.manual_seed(0)
        net = getattr(tvof, name)(weights=None).eval()
        img1, img2 = synthetic_pair(128, 160)
        with torch.no_grad():
            flows = net(torch.from_numpy(img1).permute(0, 3, 1, 2), torch.from_numpy(img2).permute(0, 3, 1, 2),
                        num_flow_updates=iters)
        final = flows[-1].permute(0, 2, 3, 1).numpy()
        np.savez_compressed(os.path.join(OUT, f"e2e_{name}.npz"), image1=img1, image2=img2, iters=iters,
                            flow=final.astype(F32))
        print(name, "final flow mean |.|", np.abs(final).mean())


if __name__ == "__main__":
    gen_tap_order()
    gen_borders()
    gen_pyramid()
    gen_upsample()
    gen_e2e()
    for f in sorted(os.listdir(OUT)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(OUT, f)))
