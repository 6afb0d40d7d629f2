"""Host-side mirror of the reference's correlation block, backed by libraftk (sm_100a).

``CorrBlock`` keeps the reference's duck-typed contract (``/root/reference/jax_raft/model.py:403-481``,
plug point ``:643-644``, contract ``:530-539``): ``out_channels``,
``build_pyramid(fmap1, fmap2) -> pyramid`` and ``index_pyramid(pyramid, coords)``.  Arrays are
torch CUDA float32 tensors in the reference's NHWC layouts; errors are ``AssertionError`` with the
reference's messages, raised before any launch.
"""

from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib


def _check_cuda_f32(name: str, t: torch.Tensor) -> torch.Tensor:
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a torch.Tensor on a CUDA device, got {type(t)}")
    if not t.is_cuda:
        raise RuntimeError(f"{name} is on {t.device}: the correlation path runs on sm_100a only; there is no CPU fallback")
    if t.dtype != torch.float32:
        raise TypeError(f"{name} must be float32, got {t.dtype}")
    return t.contiguous()


def _check_fmaps(fmap1, fmap2, num_levels, query_rows):
    """The reference's build_pyramid assertions (model.py:426, 431-436), raised before any launch.

    With ``query_rows=(y0, y1)`` fmap1 may be the full map (it is sliced) or already the band
    of rows [y0, y1) (row-sharded callers); fmap2 is always the full map.
    """
    h = fmap2.shape[-3]
    presliced = False
    if query_rows is not None:
        y0, y1 = query_rows
        assert 0 <= y0 <= y1 <= h, f"query_rows {query_rows} outside [0, {h}]"
        rows = fmap1.shape[-3]
        assert rows in (h, y1 - y0), (
            f"fmap1 has {rows} rows: with query_rows={query_rows} it must be the full map ({h} rows, sliced here) "
            f"or exactly the band ({y1 - y0} rows)")
        presliced = rows == y1 - y0 and rows != h  # (0, h) on a full map: slicing is the identity either way
    if presliced:
        assert fmap1.shape[0] == fmap2.shape[0] and fmap1.shape[2:] == fmap2.shape[2:], \
            "Input feature maps should have the same shapes"
    else:
        assert fmap1.shape == fmap2.shape, "Input feature maps should have the same shapes"
    min_fmap_size = 2 * (2 ** (num_levels - 1))
    assert not any(s < min_fmap_size for s in fmap2.shape[-3:-1]), (
        "Feature maps are too small to be down-sampled by the correlation pyramid. "
        f"H and W of feature maps should be at least {min_fmap_size}; got: {tuple(fmap2.shape[-3:-1])}. "
        "Remember that input images to the model are downsampled by 8, so that means their "
        f"dimensions should be at least 8 * {min_fmap_size} = {8 * min_fmap_size}."
    )
    fmap1 = _check_cuda_f32("fmap1", fmap1)
    fmap2 = _check_cuda_f32("fmap2", fmap2)
    if query_rows is not None and not presliced:
        fmap1 = fmap1[:, query_rows[0]:query_rows[1]].contiguous()
    return fmap1, fmap2


class TiledPyramid:
    """The correlation pyramid in the library's tiled layout (``include/raftk.h``: 4 x 4-float tiles, one 64-byte
    L2 fetch granule each).  The reference treats the pyramid as an opaque value handed from ``build_pyramid`` to
    ``index_pyramid`` (model.py:530-539), so this is what :class:`CorrBlock` returns by default.

    ``levels[l]`` is a (Q, tiles_y, tiles_x, 4, 4) tensor, float32 or — ``storage="fp16"`` — float16 (values rounded
    once to nearest even when stored).  Indexing, iterating or ``dense()`` give the reference's row-major float32
    levels (Q, h>>l, w>>l, 1) as copies (``raftk_pyramid_untile[_f16]``), for inspection and tests.
    """

    def __init__(self, levels: List[torch.Tensor], q: int, h: int, w: int):
        self.levels = levels
        self.q, self.h, self.w = q, h, w
        self.storage = "fp16" if levels[0].dtype == torch.float16 else "fp32"
        self._dense = None

    @staticmethod
    def level_shape(q: int, h: int, w: int, l: int):
        lh, lw = h >> l, w >> l
        return (q, (lh + 3) // 4, 2 * ((lw + 7) // 8), 4, 4)

    @classmethod
    def empty(cls, q: int, h: int, w: int, num_levels: int, device, storage: str = "fp32") -> "TiledPyramid":
        assert storage in ("fp32", "fp16")
        lib = _lib.load()
        levels = []
        for l in range(num_levels):
            shape = cls.level_shape(q, h, w, l)
            assert shape[1] * shape[2] * 16 == lib.raftk_pyramid_tiled_slab_floats(h, w, l)
            levels.append(torch.empty(shape, dtype=torch.float16 if storage == "fp16" else torch.float32, device=device))
        return cls(levels, q, h, w)

    @classmethod
    def from_dense(cls, levels: Sequence[torch.Tensor], storage: str = "fp32") -> "TiledPyramid":
        """Row-major float32 levels (Q, h>>l, w>>l[, 1]) -> tiled slabs (``raftk_pyramid_tile[_f16]``)."""
        levels = [_check_cuda_f32(f"pyramid[{i}]", t) for i, t in enumerate(levels)]
        q, h, w = int(levels[0].shape[0]), int(levels[0].shape[1]), int(levels[0].shape[2])
        for l, t in enumerate(levels):
            assert tuple(t.shape[:3]) == (q, h >> l, w >> l), "levels must be (Q, h>>l, w>>l[, 1])"
        out = cls.empty(q, h, w, len(levels), levels[0].device, storage)
        lib = _lib.load()
        fn, name = ((lib.raftk_pyramid_tile_f16, "raftk_pyramid_tile_f16") if storage == "fp16"
                    else (lib.raftk_pyramid_tile, "raftk_pyramid_tile"))
        with torch.cuda.device(levels[0].device):
            rc = fn(_lib.ptr_array(levels), _lib.ptr_array(out.levels), len(levels), q, h, w,
                    _lib.current_stream_ptr(levels[0].device))
        _lib.check(name, rc)
        _lib.count_launch(len(levels))
        return out

    def dense(self) -> List[torch.Tensor]:
        dev = self.levels[0].device
        out = [torch.empty((self.q, self.h >> l, self.w >> l, 1), dtype=torch.float32, device=dev)
               for l in range(len(self.levels))]
        lib = _lib.load()
        fn, name = ((lib.raftk_pyramid_untile_f16, "raftk_pyramid_untile_f16") if self.storage == "fp16"
                    else (lib.raftk_pyramid_untile, "raftk_pyramid_untile"))
        with torch.cuda.device(dev):
            rc = fn(_lib.ptr_array(self.levels), _lib.ptr_array(out), len(out), self.q, self.h, self.w,
                    _lib.current_stream_ptr(dev))
        _lib.check(name, rc)
        _lib.count_launch(len(out))
        return out

    def _cached_dense(self):
        if self._dense is None:
            self._dense = self.dense()
        return self._dense

    def __len__(self):
        return len(self.levels)

    def __getitem__(self, i):
        return self._cached_dense()[i]

    def __iter__(self):
        return iter(self._cached_dense())


class CorrBlock:
    """All-pairs correlation pyramid + radius-r window lookup (model.py:403-481).

    ``layout="tiled"`` (default): ``build_pyramid`` returns a :class:`TiledPyramid` and ``index_pyramid`` reads it
    with the tiled kernels; ``layout="rowmajor"`` keeps the reference's (B*nq, h>>l, w>>l, 1) tensors.  Either way
    ``index_pyramid`` accepts both kinds (a list of row-major levels goes through the row-major kernels).

    ``storage="fp16"`` (tiled layout only): the slabs hold float16 — each correlation value rounded once, relative
    error <= 4.9e-4 (the north-star's bar for correlation values is 1e-3); window indexing, pooling bounds and the
    lookup's float32 arithmetic are unchanged.  Halves the pyramid's memory and the traffic of both kernels."""

    def __init__(self, num_levels: int = 4, radius: int = 4, kernel: str = "tc", layout: str = "tiled",
                 storage: str = "fp32"):
        assert 1 <= num_levels <= _lib.MAX_LEVELS
        assert kernel in ("tc", "simt")
        assert layout in ("tiled", "rowmajor")
        assert storage in ("fp32", "fp16")
        assert storage == "fp32" or layout == "tiled", "float16 storage exists for the tiled layout only"
        self.storage = storage
        self.num_levels = num_levels
        self.radius = radius
        self.out_channels = num_levels * (2 * radius + 1) ** 2  # model.py:416
        self.kernel = kernel
        self.layout = layout

    # -- model.py:418-446 ---------------------------------------------------------------
    def build_pyramid(self, fmap1: torch.Tensor, fmap2: torch.Tensor,
                      query_rows: Optional[Tuple[int, int]] = None) -> List[torch.Tensor]:
        """Returns the pyramid: ``num_levels`` levels of (B*nq, h>>l, w>>l, 1), as a :class:`TiledPyramid` or
        (``layout="rowmajor"``) a list of tensors.

        ``query_rows=(y0, y1)`` builds only the volume rows of fmap1 rows [y0, y1): the
        row-sharded mode of SURVEY.md §8e (fmap2 must be the full, all-gathered map).
        """
        fmap1, fmap2 = _check_fmaps(fmap1, fmap2, self.num_levels, query_rows)
        b, h, w, d = fmap2.shape
        nq = fmap1.shape[1] * w
        lib = _lib.load()
        if self.layout == "tiled":
            direct = self.kernel == "tc" and self.num_levels <= 4 and d % 64 == 0 and d <= 256
            if not direct:  # configurations the tiled epilogue does not cover: row-major kernels, then convert
                return TiledPyramid.from_dense(self._build_rowmajor(fmap1, fmap2), self.storage)
            pyr = TiledPyramid.empty(b * nq, h, w, self.num_levels, fmap1.device, self.storage)
            stream = _lib.current_stream_ptr(fmap1.device)
            fn, name = ((lib.raftk_corr_pyramid_tiled_f16, "raftk_corr_pyramid_tiled_f16") if self.storage == "fp16"
                        else (lib.raftk_corr_pyramid_tiled, "raftk_corr_pyramid_tiled"))
            with torch.cuda.device(fmap1.device):
                nbytes = lib.raftk_corr_pyramid_scratch_bytes(b, nq, h, w, d)
                scratch = torch.empty((max(nbytes, 1),), dtype=torch.uint8, device=fmap1.device)
                rc = fn(fmap1.data_ptr(), fmap2.data_ptr(), _lib.ptr_array(pyr.levels), self.num_levels, b, nq, h, w, d,
                        scratch.data_ptr(), nbytes, stream)
            _lib.check(name, rc)
            _lib.count_launch(2)  # bf16 split + tcgen05 GEMM
            return pyr
        return self._build_rowmajor(fmap1, fmap2)

    def _build_rowmajor(self, fmap1: torch.Tensor, fmap2: torch.Tensor) -> List[torch.Tensor]:
        """fmap1 (B, rows, w, D): the query rows held by this call, already checked and sliced."""
        b, h, w, d = fmap2.shape
        nq = fmap1.shape[1] * w
        lib = _lib.load()
        levels = [
            torch.empty((b * nq, h >> l, w >> l, 1), dtype=torch.float32, device=fmap1.device)
            for l in range(self.num_levels)
        ]
        stream = _lib.current_stream_ptr(fmap1.device)
        with torch.cuda.device(fmap1.device):
            if self.kernel == "simt":
                rc = lib.raftk_corr_pyramid_simt(fmap1.data_ptr(), fmap2.data_ptr(), _lib.ptr_array(levels),
                                                 self.num_levels, b, nq, h, w, d, stream)
                _lib.check("raftk_corr_pyramid_simt", rc)
                _lib.count_launch(self.num_levels)  # volume + (L-1) pooling kernels
            else:
                nbytes = lib.raftk_corr_pyramid_scratch_bytes(b, nq, h, w, d)
                scratch = torch.empty((max(nbytes, 1),), dtype=torch.uint8, device=fmap1.device)
                rc = lib.raftk_corr_pyramid(fmap1.data_ptr(), fmap2.data_ptr(), _lib.ptr_array(levels),
                                            self.num_levels, b, nq, h, w, d, scratch.data_ptr(), nbytes, stream)
                _lib.check("raftk_corr_pyramid", rc)
                _lib.count_launch(2 + max(0, self.num_levels - 4))  # bf16 split + tcgen05 GEMM (+ pools past level 3)
                # scratch was allocated on the launching stream, so the caching allocator
                # cannot hand it to anyone before the kernel has consumed it.
        return levels

    # -- model.py:448-470 ---------------------------------------------------------------
    def index_pyramid(self, corr_pyramid: Sequence[torch.Tensor], centroids_coords: torch.Tensor) -> torch.Tensor:
        assert centroids_coords.shape[-1] == 2
        coords = _check_cuda_f32("centroids_coords", centroids_coords)
        b, hq, wq, _ = coords.shape
        q = b * hq * wq
        assert len(corr_pyramid) == self.num_levels
        out = torch.empty((b, hq, wq, self.out_channels), dtype=torch.float32, device=coords.device)
        lib = _lib.load()
        if isinstance(corr_pyramid, TiledPyramid):
            assert corr_pyramid.q == q, "pyramid and coords disagree on the number of query pixels"
            fn, name = ((lib.raftk_corr_lookup_tiled_f16, "raftk_corr_lookup_tiled_f16") if corr_pyramid.storage == "fp16"
                        else (lib.raftk_corr_lookup_tiled, "raftk_corr_lookup_tiled"))
            with torch.cuda.device(coords.device):
                rc = fn(_lib.ptr_array(corr_pyramid.levels), self.num_levels, coords.data_ptr(), out.data_ptr(), q,
                        corr_pyramid.h, corr_pyramid.w, self.radius, _lib.current_stream_ptr(coords.device))
            _lib.check(name, rc)
            _lib.count_launch(1)
            return out
        lv0 = corr_pyramid[0]
        assert lv0.shape[0] == q, "pyramid and coords disagree on the number of query pixels"
        h, w = int(lv0.shape[1]), int(lv0.shape[2])
        levels = [_check_cuda_f32(f"pyramid[{i}]", t) for i, t in enumerate(corr_pyramid)]
        with torch.cuda.device(coords.device):
            rc = lib.raftk_corr_lookup(_lib.ptr_array(levels), self.num_levels, coords.data_ptr(), out.data_ptr(),
                                       q, h, w, self.radius, _lib.current_stream_ptr(coords.device))
        _lib.check("raftk_corr_lookup", rc)
        _lib.count_launch(1)
        assert out.shape == (b, hq, wq, self.out_channels)
        return out


    def index_pyramid_update(self, corr_pyramid: Sequence[torch.Tensor], coords0: torch.Tensor, coords1: torch.Tensor,
                             delta_pre: torch.Tensor, bias: torch.Tensor, flow: torch.Tensor,
                             flow_pitched: Optional[torch.Tensor] = None) -> torch.Tensor:
        """``coords1 += delta_pre[..., :2] + bias[:2]`` (in place), ``flow = coords1 - coords0`` (also into the
        first two channels of ``flow_pitched``), then ``index_pyramid(corr_pyramid, coords1)`` — UpdateCell's
        tail (model.py:504-506) folded into the next iteration's lookup.  ``delta_pre`` is (B, h, w, C >= 2)
        NHWC memory; 4 levels and radius 3 or 4 only."""
        assert self.fused_conv_supported
        b, hq, wq, _ = coords1.shape
        q = b * hq * wq
        assert coords1.is_contiguous() and coords0.is_contiguous() and flow.is_contiguous() and coords1.dtype == torch.float32

        def pixel_stride(name, t):
            # dense NHWC memory, given either as (B, h, w, C) contiguous or as (B, C, h, w) channels_last; anything
            # else (an NCHW-contiguous tensor, a slice) would hand the kernel wrong strides
            assert t.dtype == torch.float32 and t.dim() == 4 and t.shape[0] == b, f"{name}: 4-D float32, batch {b}"
            if tuple(t.shape[1:3]) == (hq, wq) and t.is_contiguous():
                return int(t.shape[3])
            if tuple(t.shape[2:]) == (hq, wq) and t.is_contiguous(memory_format=torch.channels_last):
                return int(t.shape[1])
            raise AssertionError(f"{name} must be dense NHWC memory: (B, h, w, C) contiguous or (B, C, h, w) channels_last")

        ld_delta = pixel_stride("delta_pre", delta_pre)
        ld_flow = 0 if flow_pitched is None else pixel_stride("flow_pitched", flow_pitched)
        out = torch.empty((b, hq, wq, self.out_channels), dtype=torch.float32, device=coords1.device)
        lib = _lib.load()
        tiled = isinstance(corr_pyramid, TiledPyramid)
        if tiled:
            assert corr_pyramid.q == q
            h, w, levels = corr_pyramid.h, corr_pyramid.w, corr_pyramid.levels
            fn, name = ((lib.raftk_corr_lookup_tiled_f16_update, "raftk_corr_lookup_tiled_f16_update")
                        if corr_pyramid.storage == "fp16"
                        else (lib.raftk_corr_lookup_tiled_update, "raftk_corr_lookup_tiled_update"))
        else:
            lv0 = corr_pyramid[0]
            assert lv0.shape[0] == q
            h, w = int(lv0.shape[1]), int(lv0.shape[2])
            levels = [_check_cuda_f32(f"pyramid[{i}]", t) for i, t in enumerate(corr_pyramid)]
            fn, name = lib.raftk_corr_lookup_update, "raftk_corr_lookup_update"
        with torch.cuda.device(coords1.device):
            rc = fn(_lib.ptr_array(levels), self.num_levels, delta_pre.data_ptr(), ld_delta, bias.data_ptr(), coords0.data_ptr(),
                    coords1.data_ptr(), flow.data_ptr(), None if flow_pitched is None else flow_pitched.data_ptr(),
                    ld_flow, out.data_ptr(), q, h, w, self.radius, _lib.current_stream_ptr(coords1.device))
        _lib.check(name, rc)
        _lib.count_launch(1)
        return out

    # -- lookup fused with the 1x1 convolution that consumes it ----------------
    @property
    def fused_conv_supported(self) -> bool:
        return self.num_levels == 4 and self.radius in (3, 4)

    def pack_conv_weight(self, weight: torch.Tensor) -> torch.Tensor:
        """(N, L*(2r+1)^2[, 1, 1]) filters of the convolution that follows the lookup (convcorr1,
        model.py:275) -> the (N, K') layout ``raftk_corr_lookup_conv1x1`` reads: each level's columns
        padded with zeros to a multiple of 32."""
        assert self.fused_conv_supported
        n = weight.shape[0]
        ss = (2 * self.radius + 1) ** 2
        w = weight.reshape(n, self.num_levels, ss).to(torch.float32)
        kl = _lib.load().raftk_lookup_conv_weight_cols(self.num_levels, self.radius) // self.num_levels
        packed = w.new_zeros((n, self.num_levels, kl))
        packed[:, :, :ss] = w
        return packed.reshape(n, self.num_levels * kl).contiguous()

    def index_pyramid_conv(self, corr_pyramid: Sequence[torch.Tensor], centroids_coords: torch.Tensor,
                           packed_weight: torch.Tensor, bias: torch.Tensor, relu: bool = True) -> torch.Tensor:
        """``act(conv1x1(index_pyramid(corr_pyramid, coords)) + bias)`` in one kernel, (B, h, w, N); TF32
        products.  N a multiple of 16, at most 256."""
        assert self.fused_conv_supported
        assert centroids_coords.shape[-1] == 2
        coords = _check_cuda_f32("centroids_coords", centroids_coords)
        b, hq, wq, _ = coords.shape
        q = b * hq * wq
        assert len(corr_pyramid) == self.num_levels
        assert not isinstance(corr_pyramid, TiledPyramid), \
            "raftk_corr_lookup_conv1x1 reads row-major levels: build the pyramid with CorrBlock(layout='rowmajor')"
        lv0 = corr_pyramid[0]
        assert lv0.shape[0] == q, "pyramid and coords disagree on the number of query pixels"
        h, w = int(lv0.shape[1]), int(lv0.shape[2])
        levels = [_check_cuda_f32(f"pyramid[{i}]", t) for i, t in enumerate(corr_pyramid)]
        weight = _check_cuda_f32("packed_weight", packed_weight)
        bias = _check_cuda_f32("bias", bias)
        n = weight.shape[0]
        lib = _lib.load()
        assert weight.shape[1] == lib.raftk_lookup_conv_weight_cols(self.num_levels, self.radius) and bias.shape == (n,)
        out = torch.empty((b, hq, wq, n), dtype=torch.float32, device=coords.device)
        with torch.cuda.device(coords.device):
            rc = lib.raftk_corr_lookup_conv1x1(_lib.ptr_array(levels), self.num_levels, coords.data_ptr(), weight.data_ptr(),
                                               bias.data_ptr(), out.data_ptr(), q, h, w, self.radius, n, 1 if relu else 0,
                                               _lib.current_stream_ptr(coords.device))
        _lib.check("raftk_corr_lookup_conv1x1", rc)
        _lib.count_launch(1)
        return out


class AltCorrBlock:
    """On-the-fly lookup that never materialises the volume ("alt_corr").

    Same contract as :class:`CorrBlock`; the opaque "pyramid" is ``(fmap1_rows, [fmap2 pooled l times])`` plus,
    for the tensor-core kernel, the bf16 [hi | mid] splits of those maps.  Not in the reference;
    equal to the volume path by linearity of the average pool.

    ``kernel="tc"`` (default): ``raftk_altcorr_lookup_tc`` — every 8 x 8 patch of queries is a tcgen05 GEMM against
    the bounding box of its windows — where it applies (<= 4 levels, radius 3 or 4, D a multiple of 64 up to 256);
    ``kernel="simt"``: the float32 FMA kernels (``raftk_altcorr_lookup``).
    """

    def __init__(self, num_levels: int = 4, radius: int = 4, kernel: str = "tc"):
        assert 1 <= num_levels <= _lib.MAX_LEVELS
        assert kernel in ("tc", "simt")
        self.num_levels = num_levels
        self.radius = radius
        self.kernel = kernel
        self.out_channels = num_levels * (2 * radius + 1) ** 2

    def _tc_supported(self, d: int) -> bool:
        return self.kernel == "tc" and self.num_levels <= 4 and self.radius in (3, 4) and d % 64 == 0 and d <= 256

    def build_pyramid(self, fmap1: torch.Tensor, fmap2: torch.Tensor,
                      query_rows: Optional[Tuple[int, int]] = None, f2_split0: Optional[torch.Tensor] = None):
        """``f2_split0``: the bf16 [hi | mid] split of ``fmap2`` (B, h, w, 2D) when the caller already has it
        (sharding.PeerRowGather writes it while gathering); otherwise it is made here."""
        fmap1, fmap2 = _check_fmaps(fmap1, fmap2, self.num_levels, query_rows)
        b, h, w, d = fmap2.shape
        f2_levels = [fmap2] + [
            torch.empty((b, h >> l, w >> l, d), dtype=torch.float32, device=fmap2.device)
            for l in range(1, self.num_levels)
        ]
        lib = _lib.load()
        stream = _lib.current_stream_ptr(fmap2.device)
        with torch.cuda.device(fmap2.device):
            rc = lib.raftk_altcorr_pool_fmap2(_lib.ptr_array(f2_levels), self.num_levels, b, h, w, d, stream)
            _lib.check("raftk_altcorr_pool_fmap2", rc)
            _lib.count_launch(self.num_levels - 1)
            if not self._tc_supported(d):
                return (fmap1, f2_levels)
            # bf16 [hi | mid] operands of the tensor-core lookup, once per pair (1/sqrt(D) rides on fmap1)
            f1_split = torch.empty(tuple(fmap1.shape[:-1]) + (2 * d,), dtype=torch.bfloat16, device=fmap1.device)
            f2_split = [torch.empty(tuple(t.shape[:-1]) + (2 * d,), dtype=torch.bfloat16, device=t.device) for t in f2_levels]
            todo = [(fmap1, f1_split, float(d) ** -0.5)] + [(s_, d_, 1.0) for s_, d_ in zip(f2_levels, f2_split)]
            if f2_split0 is not None:
                assert f2_split0.dtype == torch.bfloat16 and f2_split0.is_contiguous() and tuple(f2_split0.shape) == (b, h, w, 2 * d)
                f2_split[0] = f2_split0
                del todo[1]
            for src, dst, scale in todo:
                rc = lib.raftk_altcorr_split(src.data_ptr(), dst.data_ptr(), src.numel() // d, d, scale, stream)
                _lib.check("raftk_altcorr_split", rc)
            _lib.count_launch(len(todo))
        return (fmap1, f2_levels, f1_split, f2_split)

    def index_pyramid(self, corr_pyramid, centroids_coords: torch.Tensor) -> torch.Tensor:
        fmap1, f2_levels = corr_pyramid[0], corr_pyramid[1]
        assert centroids_coords.shape[-1] == 2
        coords = _check_cuda_f32("centroids_coords", centroids_coords)
        b, hq, wq, _ = coords.shape
        assert fmap1.shape[:3] == (b, hq, wq), "fmap1 rows and coords disagree"
        _, h, w, d = f2_levels[0].shape
        out = torch.empty((b, hq, wq, self.out_channels), dtype=torch.float32, device=coords.device)
        lib = _lib.load()
        with torch.cuda.device(coords.device):
            if len(corr_pyramid) == 4 and wq == w:
                rc = lib.raftk_altcorr_lookup_tc(fmap1.data_ptr(), _lib.ptr_array(f2_levels), corr_pyramid[2].data_ptr(),
                                                 _lib.ptr_array(corr_pyramid[3]), self.num_levels, coords.data_ptr(),
                                                 out.data_ptr(), b, hq * wq, h, w, d, self.radius,
                                                 _lib.current_stream_ptr(coords.device))
                _lib.check("raftk_altcorr_lookup_tc", rc)
            else:
                rc = lib.raftk_altcorr_lookup(fmap1.data_ptr(), _lib.ptr_array(f2_levels), self.num_levels,
                                              coords.data_ptr(), out.data_ptr(), b, hq * wq, h, w, d, self.radius,
                                              _lib.current_stream_ptr(coords.device))
                _lib.check("raftk_altcorr_lookup", rc)
        _lib.count_launch(1)
        return out
