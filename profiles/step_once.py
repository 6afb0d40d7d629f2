"""One eager raft_large 440x1024 x32-iteration forward between cudaProfilerStart/Stop, for ncu:

    python profiles/step_once.py && \
    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
        --log-file gpurun_out/launches.csv python profiles/step_once.py

``--batch B`` profiles B pairs per step; ``--hot-only`` runs only the three kernel families on
synthetic inputs of the same shapes (for ``ncu --set full -k regex:...`` captures).
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

import jax_raft_b200 as J  # noqa: E402
from bench import H, ITERS, W, synthetic_pairs  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=1)
ap.add_argument("--iters", type=int, default=ITERS)
ap.add_argument("--hot-only", action="store_true")
ap.add_argument("--kernel", default="tc")
ap.add_argument("--storage", default="fp32", help="pyramid storage of the hot-only pass: fp32 | fp16 | both")
ap.add_argument("--extras", action="store_true", help="hot-only: also raftk_mask_upsample and the alt_corr lookup (136x240)")
args = ap.parse_args()

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
if args.hot_only:
    b, h, w, d = args.batch, H // 8, W // 8, 256
    g = torch.Generator(device=dev).manual_seed(0)
    f1 = torch.randn((b, h, w, d), device=dev, generator=g) * 1.45
    f2 = torch.randn((b, h, w, d), device=dev, generator=g) * 1.45
    coords = J.make_coords_grid(b, h, w, device=dev) + (torch.rand((b, h, w, 2), device=dev, generator=g) - 0.5) * 16
    flow = torch.randn((b, h, w, 2), device=dev, generator=g)
    mask = torch.randn((b, h, w, 576), device=dev, generator=g)
    blocks = [J.CorrBlock(4, 4, kernel=args.kernel, storage=s) for s in (("fp32", "fp16") if args.storage == "both" else (args.storage,))]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    if args.extras:
        feats = torch.relu(torch.randn((b, h, w, 256), device=dev, generator=g))
        pw, pb = J.pack_mask_predictor(torch.randn((576, 256, 1, 1), device=dev, generator=g) / 16, torch.randn((576,), device=dev, generator=g))
        up_out = torch.empty((b, 8 * h, 8 * w, 2), device=dev)
        ah, aw = 136, 240
        a1 = torch.randn((1, ah, aw, d), device=dev, generator=g) * 1.45
        a2 = torch.randn((1, ah, aw, d), device=dev, generator=g) * 1.45
        yy, xx = torch.meshgrid(torch.arange(ah, device=dev), torch.arange(aw, device=dev), indexing="ij")
        sm = torch.stack([8 * torch.sin(6.28 * (xx / aw + 0.3 * yy / ah)), 8 * torch.cos(6.28 * (yy / ah - 0.2 * xx / aw))], -1)[None].float()
        ac = (J.make_coords_grid(1, ah, aw, device=dev) + sm).contiguous()
        ab = J.AltCorrBlock(4, 4)
    for i in range(3):
        if i == 2:
            torch.cuda.synchronize()
            torch.cuda.profiler.start()
        for cb in blocks:
            pyr = cb.build_pyramid(f1, f2)
            flush.fill_(0)
            out = cb.index_pyramid(pyr, coords)
            del pyr
        up = J.upsample_flow(flow, mask)
        if args.extras:
            J.mask_upsample(flow, feats, pw, pb, out=up_out)
            ab.index_pyramid(ab.build_pyramid(a1, a2), ac)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
else:
    model, variables = J.raft_large(seed=0)
    img1, img2 = synthetic_pairs(args.batch, seed=1)
    i1, i2 = torch.from_numpy(img1).to(dev), torch.from_numpy(img2).to(dev)
    for i in range(3):
        if i == 2:
            torch.cuda.synchronize()
            torch.cuda.profiler.start()
        flows = model.apply(variables, i1, i2, num_flow_updates=args.iters, use_cuda_graph=False)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    assert torch.isfinite(flows).all()
print("ok")
