"""Multi-GPU partitioning of the correlation path (one process per GPU, torch.distributed).

Not in the reference (single device); asked for by BASELINE.json:

* **batch sharding** — image pairs are independent, so a batch is split across ranks with no
  data-path collective (:func:`batch_slice`);
* **query-row sharding** (1080p-class inputs) — volume row ``n`` depends on ``fmap1[n]`` and on
  all of ``fmap2``, so rank ``g`` keeps fmap1 rows ``row_range(h, W, g)``, its coords and its
  outputs, and the only exchange is ONE all-gather of ``fmap2`` per pair: :func:`all_gather_rows`
  (NCCL over NVLink on GPUs, gloo in the CPU tests), or :class:`PeerRowGather`, one kernel that reads
  the bands from peer memory and writes the gathered map and its operand split in the same pass.
  Pooled fmap2 levels are recomputed locally.
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


class PeerRowGather:
    """The fmap2 exchange over peer memory, fused with the level-0 operand split (GPUs of one NVLink domain).

    Every rank's band lives in a symmetric-memory buffer (``torch.distributed._symmetric_memory``: one
    allocation per rank, each mapped into every peer).  :meth:`gather` runs ONE kernel per rank
    (``raftk_gather_rows_split``, csrc/gather_split.cu) that reads each band where it lives — plain
    16-byte loads, remote bands over NVLink, each element once — and writes the full float32 map and
    its bf16 [hi | mid] split locally.  The ranks synchronise inside that kernel, through flag words
    that sit behind each band in the same buffer (``sync="kernel"``: bands complete before anybody
    reads; the kernel ends when every peer has read this rank's band), or through the handle's
    device-side barrier before and after it (``sync="barrier"``).  No host synchronisation and no
    collective library call on the data path; the rendezvous happens once, at construction
    (collective over ``group``).  All ranks must call :meth:`gather` the same number of times.
    """

    SYNC_BYTES = 256  # RAFTK_GATHER_SYNC_BYTES

    def __init__(self, b: int, h: int, w: int, d: int, device, group: Optional[dist.ProcessGroup] = None,
                 sync: str = "kernel"):
        import torch.distributed._symmetric_memory as symm

        assert sync in ("kernel", "barrier")
        self.sync = sync
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        assert self.world <= 8, "raftk_gather_rows_split takes at most 8 bands"
        self.shape = (b, h, w, d)
        self.row_start = [row_range(h, self.world, g)[0] for g in range(self.world)] + [h]
        max_rows = (h + self.world - 1) // self.world
        band_floats = -(-(b * max_rows * w * d) // 64) * 64  # the flag block behind it stays 256-byte aligned
        self._buf = symm.empty((band_floats + self.SYNC_BYTES // 4,), dtype=torch.float32, device=device)
        self._buf.zero_()
        torch.cuda.synchronize(device)  # flags are zero before any peer can see the buffer
        self._hdl = symm.rendezvous(self._buf, self.group)
        self._ptrs = [int(p) for p in self._hdl.buffer_ptrs]
        self._flag_ptrs = [p + band_floats * 4 for p in self._ptrs]
        self._flags = self._buf[band_floats:].view(torch.int32)
        rows = self.row_start[self.rank + 1] - self.row_start[self.rank]
        #: this rank's band (B, rows_g, w, D): write fmap2's rows here (or pass any tensor to :meth:`gather`)
        self.band = self._buf[: b * rows * w * d].view(b, rows, w, d)
        self._hdl.barrier(channel=0)  # nobody signals a peer that is still zeroing

    def timed_out(self) -> bool:
        """True if a wait inside the kernel ever gave up (a rank missed a call); synchronises the host."""
        return bool(self._flags[18].item())

    def gather(self, fmap2_rows: Optional[torch.Tensor] = None, split: bool = True):
        """Returns ``(fmap2, fmap2_split)``: (B, h, w, D) float32 and (B, h, w, 2D) bf16 (None if not ``split``)."""
        import ctypes

        from . import _lib

        b, h, w, d = self.shape
        if fmap2_rows is not None and fmap2_rows.data_ptr() != self.band.data_ptr():
            assert fmap2_rows.shape == self.band.shape, f"expected a band {tuple(self.band.shape)}, got {tuple(fmap2_rows.shape)}"
            self.band.copy_(fmap2_rows)
        dev = self._buf.device
        full = torch.empty((b, h, w, d), dtype=torch.float32, device=dev)
        halves = torch.empty((b, h, w, 2 * d), dtype=torch.bfloat16, device=dev) if split else None
        ptrs = (ctypes.c_void_p * self.world)(*self._ptrs)
        flags = (ctypes.c_void_p * self.world)(*self._flag_ptrs) if self.sync == "kernel" else None
        starts = (ctypes.c_int * (self.world + 1))(*self.row_start)
        lib = _lib.load()
        with torch.cuda.device(dev):
            if flags is None:
                self._hdl.barrier(channel=0)
            rc = lib.raftk_gather_rows_split(ptrs, flags, self.rank, starts, self.world, full.data_ptr(),
                                             halves.data_ptr() if split else None, b, h, w, d,
                                             _lib.current_stream_ptr(dev))
            _lib.check("raftk_gather_rows_split", rc)
            _lib.count_launch(1)
            if flags is None:
                self._hdl.barrier(channel=1)
        return full, halves


class RowShardedCorr:
    """Correlation lookup for the query rows of one rank (BASELINE.json config 4).

    ``corr_block`` is any block with the reference's contract that accepts ``query_rows=`` in
    ``build_pyramid`` (:class:`jax_raft_b200.AltCorrBlock` for 1080p-class inputs).

    ``exchange="nccl"``: one ``all_gather_into_tensor`` of fmap2 (any backend; the CPU tests run it on gloo).
    ``exchange="peer"``: :class:`PeerRowGather` — the gather and the level-0 split in one kernel over peer memory
    (CUDA ranks of one node only); the buffers are made on the first pair of each shape.
    """

    def __init__(self, corr_block, group: Optional[dist.ProcessGroup] = None, exchange: str = "nccl"):
        assert exchange in ("nccl", "peer")
        self.corr_block = corr_block
        self.group = group
        self.exchange = exchange
        self.out_channels = corr_block.out_channels
        self._peer = {}

    def peer_gather(self, b: int, h: int, w: int, d: int, device) -> PeerRowGather:
        key = (b, h, w, d, str(device))
        if key not in self._peer:
            self._peer[key] = PeerRowGather(b, h, w, d, device, self.group)
        return self._peer[key]

    def build_pyramid(self, fmap1_rows: torch.Tensor, fmap2_rows: torch.Tensor, h: int):
        """``fmap1_rows`` / ``fmap2_rows``: this rank's bands (B, rows_g, w, D) of the two maps."""
        y0, y1 = row_range(h, dist.get_world_size(self.group), dist.get_rank(self.group))
        if self.exchange == "peer":
            b, _, w, d = fmap2_rows.shape
            fmap2, halves = self.peer_gather(b, h, w, d, fmap2_rows.device).gather(fmap2_rows)
            return self.corr_block.build_pyramid(fmap1_rows, fmap2, query_rows=(y0, y1), f2_split0=halves)
        fmap2 = all_gather_rows(fmap2_rows, h, self.group)
        return self.corr_block.build_pyramid(fmap1_rows, fmap2, query_rows=(y0, y1))

    def index_pyramid(self, pyramid, coords_rows: torch.Tensor) -> torch.Tensor:
        """``coords_rows``: (B, rows_g, w, 2) centroids of this rank's query pixels."""
        return self.corr_block.index_pyramid(pyramid, coords_rows)
