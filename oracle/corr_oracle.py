"""CPU oracle for RAFT's correlation hot path.  TEST INFRASTRUCTURE — not shipped.

A NumPy restatement, in float32, of the functions the reference runs through XLA
(``/root/reference/jax_raft/model.py``).  Every function cites the reference lines it
follows.  Layouts are the reference's (NHWC, coords = (x, y)).

Parity status: **pinned by cross-oracles, unpinned by the reference itself.**  The
reference has no tests, no golden vectors, and cannot be imported here (jax / flax are
not installed), so this oracle is pinned against

* torchvision 0.26 ``CorrBlock`` / ``upsample_flow`` — the upstream implementation the
  reference was ported from and validated against
  (``/root/reference/scripts/validate_sintel.py:257-263``),
* ``scipy.ndimage.map_coordinates(order=1, mode='grid-constant')`` — the per-corner
  zero-fill semantics of ``jax.scipy.ndimage.map_coordinates(mode='constant')``,
* hand-derived known-answer vectors (tap order, borders, odd pooling sizes),

in ``tests/test_oracle.py`` and the committed fixtures under ``tests/golden/``.

Third-party arithmetic restated here (not under /root/reference; ``jax >= 0.5.3`` is
the only pin, ``pyproject.toml:10``): ``jax.scipy.ndimage.map_coordinates`` (order 1:
``lower = floor(c)``, ``upper_weight = c - lower``, ``lower_weight = 1 - upper_weight``;
corner contributions ``(w_y * w_x) * where(valid_y & valid_x, v, 0)`` summed in the order
(y0,x0), (y0,x1), (y1,x0), (y1,x1)); ``flax.linen.avg_pool`` (window sum / window size,
VALID); ``flax.linen.softmax`` (``exp(x - max) / sum``); ``jax.lax.conv_general_dilated_patches``
(3x3 SAME patches, feature order (c, ky, kx)); ``jax.image.scale_and_translate`` (bilinear,
no antialias).
"""

from __future__ import annotations

import numpy as np

F32 = np.float32


# --------------------------------------------------------------------------------------
# model.py:37-40  make_coords_grid
# --------------------------------------------------------------------------------------
def make_coords_grid(batch_size: int, h: int, w: int) -> np.ndarray:
    """``coords[b, y, x] = (x, y)`` float32, shape (B, h, w, 2).  model.py:37-40."""
    ys, xs = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
    coords = np.stack([xs, ys], axis=-1).astype(F32)
    return np.tile(coords[None], (batch_size, 1, 1, 1))


# --------------------------------------------------------------------------------------
# model.py:472-481  CorrBlock._compute_corr_volume
# --------------------------------------------------------------------------------------
def corr_volume(fmap1: np.ndarray, fmap2: np.ndarray, dtype=F32) -> np.ndarray:
    """All-pairs correlation ``corr[b,i,j] = sum_d f1[b,i,d] f2[b,j,d] / sqrt(D)``.

    model.py:474-481.  Returns (B, h, w, h, w, 1).  ``dtype=np.float64`` gives the
    high-precision value the float32 tolerance is measured against.
    """
    assert fmap1.shape == fmap2.shape
    b, h, w, d = fmap1.shape
    f1 = fmap1.reshape(b, h * w, d).astype(dtype)
    f2 = fmap2.reshape(b, h * w, d).astype(dtype)
    corr = np.matmul(f1, np.swapaxes(f2, 1, 2))
    corr = corr.reshape(b, h, w, h, w, 1)
    # true division by np.sqrt(num_channels) (a float64 scalar, weakly typed in jax, so
    # the result stays float32): model.py:481
    return (corr / dtype(np.sqrt(d))).astype(dtype)


def avg_pool_2x2(x: np.ndarray) -> np.ndarray:
    """``nn.avg_pool(x, (2, 2), strides=(2, 2), padding=VALID)`` on (n, h, w, c).

    model.py:444.  Odd trailing rows / columns are dropped (VALID).  flax computes the
    window sum with ``lax.reduce_window(add)`` and divides by the window size; XLA does
    not specify the summation order inside the window, so values (not bounds) are only
    defined to float32 rounding.
    """
    n, h, w, c = x.shape
    ho, wo = h // 2, w // 2
    x = x[:, : 2 * ho, : 2 * wo, :]
    s = (x[:, 0::2, 0::2, :] + x[:, 0::2, 1::2, :]) + (x[:, 1::2, 0::2, :] + x[:, 1::2, 1::2, :])
    return (s / x.dtype.type(4)).astype(x.dtype)


# --------------------------------------------------------------------------------------
# model.py:418-446  CorrBlock.build_pyramid
# --------------------------------------------------------------------------------------
def min_fmap_size(num_levels: int) -> int:
    """model.py:431."""
    return 2 * (2 ** (num_levels - 1))


def build_pyramid(fmap1: np.ndarray, fmap2: np.ndarray, num_levels: int = 4, dtype=F32):
    """List of ``num_levels`` arrays (B*h*w, h_l, w_l, 1), each pooled from the previous.

    model.py:418-446.  Raises AssertionError like the reference for unequal shapes
    (``:426``) and feature maps smaller than ``2 * 2**(L-1)`` (``:431-436``).
    """
    assert fmap1.shape == fmap2.shape, "Input feature maps should have the same shapes"
    mfs = min_fmap_size(num_levels)
    assert not any(s < mfs for s in fmap1.shape[-3:-1]), (
        "Feature maps are too small to be down-sampled by the correlation pyramid. "
        f"H and W of feature maps should be at least {mfs}; got: {fmap1.shape[-3:-1]}."
    )
    vol = corr_volume(fmap1, fmap2, dtype=dtype)
    b, h, w, _, _, c = vol.shape
    vol = vol.reshape(b * h * w, h, w, c)
    pyramid = [vol]
    for _ in range(num_levels - 1):
        vol = avg_pool_2x2(vol)
        pyramid.append(vol)
    return pyramid


# --------------------------------------------------------------------------------------
# model.py:24-34  grid_sample  (jax.scipy.ndimage.map_coordinates order=1 mode='constant')
# --------------------------------------------------------------------------------------
def _linear_nodes(coord: np.ndarray):
    """jax ``_linear_indices_and_weights``: [(index, lower_weight), (index+1, upper_weight)]."""
    coord = coord.astype(F32)
    lower = np.floor(coord)
    upper_w = (coord - lower).astype(F32)
    lower_w = (F32(1) - upper_w).astype(F32)
    idx = lower.astype(np.int64)
    return [(idx, lower_w), (idx + 1, upper_w)]


def grid_sample(img: np.ndarray, grid: np.ndarray) -> np.ndarray:
    """Bilinear sample of ``img`` (n, h, w, c) at absolute pixel coords ``grid`` (n, gh, gw, 2)
    given as (x, y); zero outside.  model.py:24-34.

    Per corner: contribution is ``(w_y * w_x) * v`` with ``v = 0`` unless
    ``0 <= y < h`` and ``0 <= x < w``; the four contributions are added in the order
    (y0,x0), (y0,x1), (y1,x0), (y1,x1), all in float32.
    """
    assert img.shape[-3] > 1  # model.py:25
    assert grid.ndim == 4 and grid.shape[-1] == 2  # model.py:26
    n, h, w, c = img.shape
    gx = grid[..., 0]
    gy = grid[..., 1]
    ynodes = _linear_nodes(gy)
    xnodes = _linear_nodes(gx)
    nidx = np.arange(n).reshape(n, 1, 1)
    out = None
    for yi, wy in ynodes:
        vy = (yi >= 0) & (yi < h)
        for xi, wx in xnodes:
            vx = (xi >= 0) & (xi < w)
            valid = vy & vx
            v = img[nidx, np.clip(yi, 0, h - 1), np.clip(xi, 0, w - 1), :]  # (n, gh, gw, c)
            v = np.where(valid[..., None], v, img.dtype.type(0))
            wgt = (wy * wx).astype(F32)
            contrib = (wgt[..., None].astype(img.dtype) * v).astype(img.dtype)
            out = contrib if out is None else (out + contrib).astype(img.dtype)
    return out


# --------------------------------------------------------------------------------------
# model.py:448-470  CorrBlock.index_pyramid
# --------------------------------------------------------------------------------------
def index_pyramid(pyramid, centroids_coords: np.ndarray, radius: int = 4) -> np.ndarray:
    """Window lookup.  coords (B, h, w, 2) as (x, y) -> (B, h, w, L*(2r+1)^2).

    model.py:448-470.  ``delta = stack(meshgrid(di, dj, indexing='ij'))`` is added to the
    (x, y) centroid, so window element (i, j) samples (x + di[i], y + dj[j]) and lands in
    channel ``l*(2r+1)^2 + i*(2r+1) + j``: the x offset is the slow index.  After each
    level ``centroids_coords = centroids_coords / 2`` (``:466``), exact in float32.
    """
    side = 2 * radius + 1
    di = np.linspace(-radius, radius, side).astype(F32)
    dj = np.linspace(-radius, radius, side).astype(F32)
    delta = np.stack(np.meshgrid(di, dj, indexing="ij"), axis=-1).reshape(1, side, side, 2)
    assert centroids_coords.shape[-1] == 2  # model.py:457
    b, h, w, _ = centroids_coords.shape
    cc = centroids_coords.astype(F32).reshape(b * h * w, 1, 1, 2)
    feats = []
    for vol in pyramid:
        sampling = (cc + delta).astype(F32)
        feats.append(grid_sample(vol, sampling).reshape(b, h, w, -1))
        cc = (cc / F32(2)).astype(F32)
    out = np.concatenate(feats, axis=-1)
    assert out.shape == (b, h, w, len(pyramid) * side * side)  # model.py:469
    return out


# --------------------------------------------------------------------------------------
# model.py:69-98  upsample_flow
# --------------------------------------------------------------------------------------
def softmax(x: np.ndarray, axis: int) -> np.ndarray:
    """flax.linen.softmax: ``exp(x - max) / sum(exp(x - max))`` along ``axis``."""
    m = np.max(x, axis=axis, keepdims=True)
    e = np.exp((x - m).astype(x.dtype)).astype(x.dtype)
    return (e / np.sum(e, axis=axis, keepdims=True, dtype=x.dtype)).astype(x.dtype)


def _patches_3x3_same(x: np.ndarray) -> np.ndarray:
    """``conv_general_dilated_patches(x, (3,3), (1,1), 'SAME')`` for NHWC input.

    Returns (B, h, w, c, 9) with the last axis ``k = (dy+1)*3 + (dx+1)`` — i.e. the
    output feature order (c, ky, kx) of model.py:91-93 after its reshape to
    (B, h, w, c, 9, 1, 1).  Zero padding at the image border.
    """
    b, h, w, c = x.shape
    xp = np.zeros((b, h + 2, w + 2, c), x.dtype)
    xp[:, 1:-1, 1:-1, :] = x
    out = np.empty((b, h, w, c, 9), x.dtype)
    for ky in range(3):
        for kx in range(3):
            out[..., ky * 3 + kx] = xp[:, ky : ky + h, kx : kx + w, :]
    return out


def resize_bilinear_align_corners(x: np.ndarray, new_h: int, new_w: int) -> np.ndarray:
    """model.py:43-66 (``jax.image.scale_and_translate`` with ``scale=(new-1)/(old-1)``,
    ``translation=-(scale/2-0.5)``, bilinear, antialias=False): output pixel i samples
    input position ``i*(old-1)/(new-1)`` with a triangle kernel, i.e. torch's
    ``F.interpolate(mode='bilinear', align_corners=True)`` (``:80``)."""
    b, h, w, c = x.shape

    def axis_weights(old, new):
        pos = np.arange(new, dtype=np.float64) * (old - 1) / (new - 1)
        i0 = np.clip(np.floor(pos).astype(np.int64), 0, old - 1)
        i1 = np.clip(i0 + 1, 0, old - 1)
        w1 = (pos - i0).astype(F32)
        return i0, i1, (F32(1) - w1).astype(F32), w1

    y0, y1, wy0, wy1 = axis_weights(h, new_h)
    x0, x1, wx0, wx1 = axis_weights(w, new_w)
    rows = x[:, y0] * wy0[None, :, None, None] + x[:, y1] * wy1[None, :, None, None]
    out = rows[:, :, x0] * wx0[None, None, :, None] + rows[:, :, x1] * wx1[None, None, :, None]
    return out.astype(x.dtype)


def upsample_flow(flow: np.ndarray, up_mask: np.ndarray | None = None, factor: int = 8) -> np.ndarray:
    """model.py:69-98.  flow (B, h, w, C) -> (B, factor*h, factor*w, C).

    Convex branch (``:85-98``): mask channel ``k*f*f + fy*f + fx``; softmax over k;
    ``out[b, f*y+fy, f*x+fx, c] = sum_k softmax_k * f*flow[b, y+dy, x+dx, c]`` with
    ``k = (dy+1)*3 + (dx+1)`` and zeros outside the image.
    """
    b, h, w, c = flow.shape
    new_h, new_w = h * factor, w * factor
    if up_mask is None:
        return (F32(factor) * resize_bilinear_align_corners(flow, new_h, new_w)).astype(flow.dtype)
    assert up_mask.shape == (b, h, w, 9 * factor * factor)  # model.py:85
    m = up_mask.reshape(b, h, w, 1, 9, factor, factor)
    m = softmax(m, axis=4)
    patches = _patches_3x3_same((flow.dtype.type(factor) * flow).astype(flow.dtype))
    patches = patches.reshape(b, h, w, c, 9, 1, 1)
    up = np.sum(m * patches, axis=4, dtype=flow.dtype)  # (b, h, w, c, f, f)
    assert up.shape == (b, h, w, c, factor, factor)  # model.py:95
    return up.transpose(0, 1, 4, 2, 5, 3).reshape(b, new_h, new_w, c)


# --------------------------------------------------------------------------------------
# "alt_corr": the same lookup computed without materialising the volume.  Not in the
# reference; by linearity of the avg-pool it equals
# index_pyramid(build_pyramid(f1, f2), coords) up to float rounding:
#   pool_l(f1 . f2^T)[q, Y, X] = f1[q] . pool_l(f2)[Y, X]
# --------------------------------------------------------------------------------------
def pool_fmap2(fmap2: np.ndarray, num_levels: int = 4):
    """[fmap2, avgpool2(fmap2), ...] each (B, h_l, w_l, D) — pooled with the same VALID
    2x2 rule as model.py:444, applied to the feature map instead of the volume."""
    levels = [fmap2]
    for _ in range(num_levels - 1):
        levels.append(avg_pool_2x2(levels[-1]))
    return levels


def alt_index(fmap1: np.ndarray, fmap2_levels, centroids_coords: np.ndarray, radius: int = 4,
              row_range=None) -> np.ndarray:
    """On-the-fly lookup for query rows ``row_range = (y_lo, y_hi)`` of fmap1 (default all).

    Builds, per level, only the (rows x h_l*w_l) slice of the volume, then samples it with
    the reference's ``grid_sample`` — so window indexing is the reference's by construction.
    """
    b, h, w, d = fmap1.shape
    y_lo, y_hi = (0, h) if row_range is None else row_range
    f1 = fmap1[:, y_lo:y_hi].reshape(b, (y_hi - y_lo) * w, d)
    cc = centroids_coords[:, y_lo:y_hi]
    pyr = []
    for f2 in fmap2_levels:
        _, hl, wl, _ = f2.shape
        vol = np.matmul(f1, np.swapaxes(f2.reshape(b, hl * wl, d), 1, 2)) / F32(np.sqrt(d))
        pyr.append(vol.astype(F32).reshape(b * f1.shape[1], hl, wl, 1))
    return index_pyramid(pyr, cc, radius)
