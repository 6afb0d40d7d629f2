"""Multi-GPU partitioning of the correlation path (one process per GPU, torch.distributed).

Not in the reference (single device); asked for by BASELINE.json:

* **batch sharding** — image pairs are independent, so a batch is split across ranks with no
  data-path collective (:func:`batch_slice`);
* **query-row sharding** (1080p-class inputs) — volume row ``n`` depends on ``fmap1[n]`` and on
  all of ``fmap2``, so rank ``g`` keeps fmap1 rows ``row_range(h, W, g)``, its coords and its
  outputs, and the only exchange is ONE all-gather of ``fmap2`` per pair (:func:`all_gather_rows`,
  NCCL over NVLink on GPUs, gloo in the CPU tests).  Pooled fmap2 levels are recomputed locally.
"""

from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist


def batch_slice(num_pairs: int, world_size: int, rank: int) -> slice:
    """Contiguous, balanced split of ``num_pairs`` (the first ``num_pairs % world_size`` ranks get
    one extra pair)."""
    assert 0 <= rank < world_size
    base, rem = divmod(num_pairs, world_size)
    lo = rank * base + min(rank, rem)
    return slice(lo, lo + base + (1 if rank < rem else 0))


def row_range(h: int, world_size: int, rank: int) -> Tuple[int, int]:
    """Image rows [y0, y1) of the 1/8-resolution grid owned by ``rank`` (balanced, contiguous)."""
    s = batch_slice(h, world_size, rank)
    return s.start, s.stop


def all_gather_rows(local_rows: torch.Tensor, h: int, group: Optional[dist.ProcessGroup] = None) -> torch.Tensor:
    """All-gather row bands (B, rows_g, w, D) of a feature map into the full (B, h, w, D).

    Bands may differ by one row; they are padded to the largest band for the collective (one
    ``all_gather_into_tensor``), then the padding is dropped.
    """
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    b, rows, w, d = local_rows.shape
    y0, y1 = row_range(h, world, rank)
    assert rows == y1 - y0, f"rank {rank} should hold rows [{y0}, {y1}) of {h}, got {rows}"
    max_rows = (h + world - 1) // world
    if b == 1 and h % world == 0:
        # equal bands of a single pair are already the full map's memory order: gather straight into the result
        # (no padding, no concatenation pass; one collective and nothing else)
        full = local_rows.new_empty((1, h, w, d))
        dist.all_gather_into_tensor(full.view(-1), local_rows.contiguous().view(-1), group=group)
        return full
    send = local_rows
    if rows != max_rows:
        send = local_rows.new_zeros((b, max_rows, w, d))
        send[:, :rows] = local_rows
    send = send.contiguous()
    flat = send.new_empty((world * send.numel(),))
    dist.all_gather_into_tensor(flat, send.view(-1), group=group)  # 1-D: same layout on nccl and gloo
    recv = flat.view(world, b, max_rows, w, d)
    bands: List[torch.Tensor] = []
    for g in range(world):
        g0, g1 = row_range(h, world, g)
        bands.append(recv[g, :, : g1 - g0])
    return torch.cat(bands, dim=1).contiguous()


class RowShardedCorr:
    """Correlation lookup for the query rows of one rank (BASELINE.json config 4).

    ``corr_block`` is any block with the reference's contract that accepts ``query_rows=`` in
    ``build_pyramid`` (:class:`jax_raft_b200.AltCorrBlock` for 1080p-class inputs).
    """

    def __init__(self, corr_block, group: Optional[dist.ProcessGroup] = None):
        self.corr_block = corr_block
        self.group = group
        self.out_channels = corr_block.out_channels

    def build_pyramid(self, fmap1_rows: torch.Tensor, fmap2_rows: torch.Tensor, h: int):
        """``fmap1_rows`` / ``fmap2_rows``: this rank's bands (B, rows_g, w, D) of the two maps."""
        fmap2 = all_gather_rows(fmap2_rows, h, self.group)
        y0, y1 = row_range(h, dist.get_world_size(self.group), dist.get_rank(self.group))
        return self.corr_block.build_pyramid(fmap1_rows, fmap2, query_rows=(y0, y1))

    def index_pyramid(self, pyramid, coords_rows: torch.Tensor) -> torch.Tensor:
        """``coords_rows``: (B, rows_g, w, 2) centroids of this rank's query pixels."""
        return self.corr_block.index_pyramid(pyramid, coords_rows)
