"""BASELINE config 4 at hot-path level: 1088x1920 (136x240 at 1/8), D=256, alt_corr lookup with the
query rows sharded over the ranks and ONE NCCL all-gather of fmap2 per pair.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 profiles/config4_rowshard.py [--iters 32]

Every rank checks its band against the materialised-volume path on its own GPU (same band, full
fmap2), then rank 0 prints one JSON line: all-gather time, pooled-fmap2 time, per-iteration lookup
time (max over ranks, CUDA events).  The rest of the model at this resolution is not sharded
(BASELINE.json does not define it).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import jax_raft_b200 as J  # noqa: E402
from jax_raft_b200 import sharding  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--iters", type=int, default=32)
ap.add_argument("--h", type=int, default=136)
ap.add_argument("--w", type=int, default=240)
args = ap.parse_args()

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "29511")
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)

B, h, w, D, r = 1, args.h, args.w, 256, 4
g = torch.Generator(device=dev).manual_seed(0)  # same maps on every rank; each keeps only its band
f1 = torch.randn((B, h, w, D), device=dev, generator=g) * 1.45
f2 = torch.randn((B, h, w, D), device=dev, generator=g) * 1.45
# a smooth flow field (what RAFT sees between iterations), +-8 px
yy, xx = torch.meshgrid(torch.arange(h, device=dev), torch.arange(w, device=dev), indexing="ij")
flow = torch.stack([8 * torch.sin(6.28 * (xx / w + 0.3 * yy / h)), 8 * torch.cos(6.28 * (yy / h - 0.2 * xx / w))], -1)[None].float()
y0, y1 = sharding.row_range(h, world, rank)
f1_rows, f2_rows = f1[:, y0:y1].contiguous(), f2[:, y0:y1].contiguous()
coords_rows = (J.make_coords_grid(B, h, w, device=dev) + flow)[:, y0:y1].contiguous()

corr = sharding.RowShardedCorr(J.AltCorrBlock(4, r))


def ev():
    return torch.cuda.Event(enable_timing=True)


def max_ms(x):
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


for _ in range(2):  # warm-up (NCCL channels, module load)
    pyr = corr.build_pyramid(f1_rows, f2_rows, h)
    out = corr.index_pyramid(pyr, coords_rows)
dist.barrier()
torch.cuda.synchronize()

a, b_, c = ev(), ev(), ev()
a.record()
full2 = sharding.all_gather_rows(f2_rows, h)
b_.record()
pyr = corr.corr_block.build_pyramid(f1_rows, full2, query_rows=(y0, y1))
c.record()
lk = []
for _ in range(args.iters):
    s, e = ev(), ev()
    s.record()
    out = corr.index_pyramid(pyr, coords_rows)
    e.record()
    lk.append((s, e))
torch.cuda.synchronize()
gather_ms, pool_ms = max_ms(a.elapsed_time(b_)), max_ms(b_.elapsed_time(c))
lookup_ms = max_ms(sum(s.elapsed_time(e) for s, e in lk) / len(lk))

# parity of this rank's band against the materialised-volume path (tcgen05 GEMM + lookup)
assert torch.equal(full2, f2), "all-gather did not restore fmap2"
vol = J.CorrBlock(4, r)
ref = vol.index_pyramid(vol.build_pyramid(f1, f2, query_rows=(y0, y1)), coords_rows)
scale = float(ref.abs().mean())
err = float((out - ref).abs().max())
ok = torch.tensor([1.0 if err <= 1e-3 * max(scale, 1.0) else 0.0], device=dev)
dist.all_reduce(ok, op=dist.ReduceOp.MIN)

if rank == 0:
    nq = B * (y1 - y0) * w
    print(json.dumps({
        "config": f"config 4 hot path: {8 * h}x{8 * w} ({h}x{w}), D={D}, r={r}, alt_corr, query rows over {world} GPUs",
        "n_gpus": world, "rows_per_rank": y1 - y0, "queries_per_rank": nq,
        "all_gather_fmap2_ms": gather_ms, "all_gather_bytes_total": int(f2.numel() * 4),
        "pool_fmap2_ms": pool_ms, "altcorr_lookup_ms_per_iter": lookup_ms,
        "altcorr_tflops_per_rank": 2.0 * nq * 4 * (2 * r + 2) ** 2 * D / (lookup_ms * 1e-3) / 1e12,
        "band_matches_volume_path": bool(ok.item() > 0), "max_abs_err": err, "corr_scale": scale,
        "timing": "CUDA events, max over ranks"}))
dist.destroy_process_group()
