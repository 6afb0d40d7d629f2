"""Two-GPU checks of the sharded paths over NCCL (skipped when fewer than two GPUs are visible):
the fmap2 all-gather, the row-sharded alt_corr lookup against the single-GPU result — with the NCCL
exchange and with the fused gather + split kernel over peer memory — and batch sharding (each rank's
pairs equal the same pairs computed alone)."""

import os
import socket

import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, ret):
    import torch.distributed as dist

    import jax_raft_b200 as J
    from jax_raft_b200 import sharding

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        g = torch.Generator(device=dev).manual_seed(0)
        b, h, w, d = 2, 23, 32, 128
        f1 = torch.randn((b, h, w, d), device=dev, generator=g)
        f2 = torch.randn((b, h, w, d), device=dev, generator=g)
        coords = J.make_coords_grid(b, h, w, device=dev) + (torch.rand((b, h, w, 2), device=dev, generator=g) - 0.5) * 12
        y0, y1 = sharding.row_range(h, world, rank)
        corr = sharding.RowShardedCorr(J.AltCorrBlock(4, 4))
        pyr = corr.build_pyramid(f1[:, y0:y1].contiguous(), f2[:, y0:y1].contiguous(), h)
        band = corr.index_pyramid(pyr, coords[:, y0:y1].contiguous())
        alt = J.AltCorrBlock(4, 4)
        full = alt.index_pyramid(alt.build_pyramid(f1, f2), coords)
        ok_rows = bool(torch.equal(band, full[:, y0:y1]))
        # the same exchange as one kernel over peer memory: gathered map and level-0 split bit-equal to
        # all-gather + raftk_altcorr_split, twice over (the buffers are reused per pair), ragged bands (23 rows)
        peer = sharding.RowShardedCorr(J.AltCorrBlock(4, 4), exchange="peer")
        ok_peer = True
        for rep in range(2):
            f2r = (f2 * (1 + rep))[:, y0:y1].contiguous()
            pyr_p = peer.build_pyramid(f1[:, y0:y1].contiguous(), f2r, h)
            ref_p = alt.build_pyramid(f1[:, y0:y1].contiguous(), f2 * (1 + rep), query_rows=(y0, y1))
            ok_peer &= bool(torch.equal(pyr_p[1][0], f2 * (1 + rep)))
            ok_peer &= all(bool(torch.equal(a, b_)) for a, b_ in zip(pyr_p[3], ref_p[3]))
            ok_peer &= bool(torch.equal(peer.index_pyramid(pyr_p, coords[:, y0:y1].contiguous()),
                                        alt.index_pyramid(ref_p, coords[:, y0:y1].contiguous())))
        # gather only (no split), band written in place
        pg = peer.peer_gather(b, h, w, d, dev)
        pg.band.copy_(f2[:, y0:y1])
        only, none = pg.gather(split=False)
        ok_peer &= none is None and bool(torch.equal(only, f2))
        # batch sharding: this rank's pairs through the volume path == the same pairs computed alone
        sl = sharding.batch_slice(b, world, rank)
        cb = J.CorrBlock(4, 4)
        mine = cb.index_pyramid(cb.build_pyramid(f1[sl].contiguous(), f2[sl].contiguous()), coords[sl].contiguous())
        whole = cb.index_pyramid(cb.build_pyramid(f1, f2), coords)
        ok_batch = bool(torch.equal(mine, whole[sl]))
        ret[rank] = (ok_rows, ok_batch, ok_peer)
    finally:
        dist.destroy_process_group()


def test_row_and_batch_sharding_two_gpus(lib):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, _free_port(), ret), nprocs=2, join=True)
    assert dict(ret) == {0: (True, True, True), 1: (True, True, True)}
