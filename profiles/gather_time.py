"""Config-4 exchange, piece by piece (run under torchrun, >= 2 GPUs):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        profiles/gather_time.py

Times, per call and as the max over ranks (CUDA events around REPS back-to-back calls after a device barrier):
the symmetric-memory barrier alone, raftk_gather_rows_split alone over peer bands (no barriers: the bands do not
change) and over local copies of the bands (its HBM floor), the full PeerRowGather.gather() with the in-kernel flags and with the handle's barriers, the NCCL all-gather
alone, raftk_altcorr_split of level 0 alone, and the two-step sequence."""

import ctypes
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jax_raft_b200 import _lib, sharding  # noqa: E402

REPS = 20


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    B, h, w, D = 1, 136, 240, 256
    g = torch.Generator(device=dev).manual_seed(0)
    f2 = torch.randn((B, h, w, D), device=dev, generator=g)
    y0, y1 = sharding.row_range(h, world, rank)
    rows = f2[:, y0:y1].contiguous()
    pg = sharding.PeerRowGather(B, h, w, D, dev)
    pg.band.copy_(rows)
    pgb = sharding.PeerRowGather(B, h, w, D, dev, sync="barrier")
    pgb.band.copy_(rows)
    full = torch.empty_like(f2)
    halves = torch.empty((B, h, w, 2 * D), dtype=torch.bfloat16, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    starts = (ctypes.c_int * (world + 1))(*pg.row_start)
    peer_ptrs = (ctypes.c_void_p * world)(*pg._ptrs)
    local = [f2[:, pg.row_start[k]:pg.row_start[k + 1]].contiguous() for k in range(world)]
    local_ptrs = (ctypes.c_void_p * world)(*[t.data_ptr() for t in local])

    def kernel(ptrs, split=True):
        rc = lib.raftk_gather_rows_split(ptrs, None, rank, starts, world, full.data_ptr(),
                                         halves.data_ptr() if split else None, B, h, w, D, stream)
        assert rc == 0

    pieces = {
        "symm_barrier": lambda: pg._hdl.barrier(channel=0),
        "kernel_peer_bands": lambda: kernel(peer_ptrs),
        "kernel_peer_bands_gather_only": lambda: kernel(peer_ptrs, False),
        "kernel_local_bands": lambda: kernel(local_ptrs),
        "peer_gather_total": lambda: pg.gather(),
        "peer_gather_total_barrier_sync": lambda: pgb.gather(),
        "nccl_all_gather": lambda: sharding.all_gather_rows(rows, h),
        "split_level0": lambda: lib.raftk_altcorr_split(f2.data_ptr(), halves.data_ptr(), B * h * w, D, 1.0, stream),
        "nccl_two_step": lambda: lib.raftk_altcorr_split(sharding.all_gather_rows(rows, h).data_ptr(), halves.data_ptr(),
                                                         B * h * w, D, 1.0, stream),
    }
    out = {}
    for name, fn in pieces.items():
        for _ in range(3):
            fn()
        best = None
        for _ in range(3):
            dist.barrier()
            torch.cuda.synchronize(dev)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(REPS):
                fn()
            b.record()
            torch.cuda.synchronize(dev)
            t = torch.tensor([a.elapsed_time(b) / REPS], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            best = float(t) if best is None else min(best, float(t))
        out[name + "_us"] = round(best * 1e3, 2)
    full_p, halves_p = pg.gather()
    torch.cuda.synchronize(dev)
    assert torch.equal(full_p, f2) and not pg.timed_out()
    if rank == 0:
        out.update(world=world, shape=[B, h, w, D], reps=REPS, remote_mb=round((f2.numel() - rows.numel()) * 4 / 1e6, 2))
        print(json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
