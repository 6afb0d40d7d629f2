"""Event timing of raftk_corr_lookup alone at the Sintel 1/8 shape: 8 coordinate sets in rotation so
consecutive launches do not find their footprints in L2 (8 x 29 MB per pair > 126 MB)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import jax_raft_b200 as J  # noqa: E402

dev = torch.device("cuda", 0)
h, w, r = 55, 128, 4
for b in [int(x) for x in (sys.argv[1:] or ["1", "16"])]:
    g = torch.Generator(device=dev).manual_seed(0)
    q = b * h * w
    pyr = [torch.randn((q, h >> l, w >> l, 1), device=dev, generator=g) for l in range(4)]
    coords = [J.make_coords_grid(b, h, w, device=dev) + (torch.rand((b, h, w, 2), device=dev, generator=g) - 0.5) * 16
              for _ in range(8)]
    cb = J.CorrBlock(4, r)
    for c in coords:
        out = cb.index_pyramid(pyr, c)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for c in coords:
            cb.index_pyramid(pyr, c)
        torch.cuda.synchronize()
        with torch.cuda.graph(graph, stream=s):
            for c in coords:
                out = cb.index_pyramid(pyr, c)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for rep in range(5):
        e0.record()
        for _ in range(4):
            graph.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / 32)
    alg = q * 2904
    print(f"B={b}: {best * 1e3:.2f} us per launch, {alg / best / 1e6:.0f} GB/s algorithmic")
