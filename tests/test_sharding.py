"""Multi-process host logic on CPU: gloo, world_size 2 (rendezvous on 127.0.0.1)."""

import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from jax_raft_b200 import sharding


def test_partitions_cover_everything():
    for n in (1, 7, 55, 64, 136):
        for world in (1, 2, 3, 4, 8):
            got = []
            for r in range(world):
                s = sharding.batch_slice(n, world, r)
                got += list(range(s.start, s.stop))
                y0, y1 = sharding.row_range(n, world, r)
                assert (y0, y1) == (s.start, s.stop)
            assert got == list(range(n))
    assert sharding.row_range(136, 8, 3) == (51, 68)  # 17 rows per GPU


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, h, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import corr_oracle as O

        rng = np.random.default_rng(0)  # same data on every rank
        b, w, d = 2, 18, 8
        f1 = rng.standard_normal((b, h, w, d)).astype(np.float32)
        f2 = rng.standard_normal((b, h, w, d)).astype(np.float32)
        coords = O.make_coords_grid(b, h, w) + rng.uniform(-5, 5, (b, h, w, 2)).astype(np.float32)
        y0, y1 = sharding.row_range(h, world, rank)
        full = sharding.all_gather_rows(torch.from_numpy(f2[:, y0:y1].copy()), h)
        ok_gather = bool(torch.equal(full, torch.from_numpy(f2)))
        # a single pair with equal bands is gathered straight into the result buffer (no padding / concatenation)
        one = sharding.all_gather_rows(torch.from_numpy(f2[:1, y0:y1].copy()), h)
        ok_gather = ok_gather and bool(torch.equal(one, torch.from_numpy(f2[:1])))
        # the sharded lookup (CPU stand-in for the kernel: the oracle) equals the rows of the full one
        mine = O.alt_index(f1, O.pool_fmap2(full.numpy()), coords, 4, row_range=(y0, y1))
        ref = O.index_pyramid(O.build_pyramid(f1, f2), coords, 4)[:, y0:y1]
        ok_rows = bool(np.allclose(mine, ref, rtol=1e-4, atol=1e-5))
        # batch sharding: concatenating every rank's slice restores the batch, no collective needed
        s = sharding.batch_slice(b, world, rank)
        gathered = [None] * world
        dist.all_gather_object(gathered, (s.start, s.stop))
        ok_batch = sorted(gathered) == [(0, 1), (1, 2)][:world] if world == 2 else True
        ret[rank] = (ok_gather, ok_rows, ok_batch)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("h", [17, 16])
def test_row_sharded_all_gather_world2(h):
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), h, ret), nprocs=world, join=True)
    assert dict(ret) == {0: (True, True, True), 1: (True, True, True)}
