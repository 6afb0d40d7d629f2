"""alt_corr lookup at 136x240 (BASELINE config 4 on one GPU), D = 256, r = 4, smooth +-8 px flow: the tensor-core
kernel (raftk_altcorr_lookup_tc) beside the float32 FMA kernel (raftk_altcorr_lookup); graph of 4 launches, best of 5.
    python profiles/altcorr_time.py [h w [rows]]      rows: only the first `rows` query rows (a rank's band of config 4)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import jax_raft_b200 as J  # noqa: E402

dev = torch.device("cuda", 0)
h, w = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (136, 240)
rows = int(sys.argv[3]) if len(sys.argv) > 3 else h
d, r, L = 256, 4, 4
g = torch.Generator(device=dev).manual_seed(0)
f1 = torch.randn((1, h, w, d), device=dev, generator=g) * 1.45
f2 = torch.randn((1, h, w, d), device=dev, generator=g) * 1.45
yy, xx = torch.meshgrid(torch.arange(h, device=dev), torch.arange(w, device=dev), indexing="ij")
flow = torch.stack([8 * torch.sin(6.28 * (xx / w + 0.3 * yy / h)), 8 * torch.cos(6.28 * (yy / h - 0.2 * xx / w))], -1)[None].float()
coords = (J.make_coords_grid(1, h, w, device=dev) + flow).contiguous()
flops = 2.0 * rows * w * L * (2 * r + 2) ** 2 * d
coords = coords[:, :rows].contiguous()
outs = {}
for kernel in ("simt", "tc"):
    ab = J.AltCorrBlock(L, r, kernel=kernel)
    pyr = ab.build_pyramid(f1, f2, query_rows=(0, rows))
    outs[kernel] = ab.index_pyramid(pyr, coords)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s), torch.cuda.graph(graph, stream=s):
        for _ in range(4):
            ab.index_pyramid(pyr, coords)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(5):
        e0.record()
        graph.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / 4)
    e0.record()
    pyr = ab.build_pyramid(f1, f2)
    e1.record()
    torch.cuda.synchronize()
    print(f"{rows}/{h}x{w} {kernel:4s}: lookup {best * 1e3:8.1f} us/iter = {flops / best / 1e9:6.1f} TFLOP/s useful; build_pyramid (pool"
          f"{' + split' if kernel == 'tc' else ''}) {e0.elapsed_time(e1) * 1e3:.1f} us", flush=True)
scale = float(outs["simt"].abs().mean())
print(f"max |tc - simt| = {float((outs['tc'] - outs['simt']).abs().max()):.3e} (mean |value| {scale:.3f})")
