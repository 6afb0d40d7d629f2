"""Event timing of the lookup alone at the Sintel 1/8 shape, row-major kernel (raftk_corr_lookup) beside the tiled
one (raftk_corr_lookup_tiled): 8 coordinate sets in rotation so consecutive launches do not find their footprints
in L2, (a) back to back in a graph of 8 launches (throughput) and (b) one launch at a time between events
(the inputs are far larger than L2 from batch 4 up; nothing is written to flush it).
    python profiles/lookup_time.py [batch ...]        (fractions on the float32 accounting: 2904 B per query)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import jax_raft_b200 as J  # noqa: E402
from jax_raft_b200.corr import TiledPyramid  # noqa: E402

dev = torch.device("cuda", 0)
h, w, r = 55, 128, 4
PEAK = 6545.9  # MEASURED_PEAKS.json hbm_gbs
for b in [int(x) for x in (sys.argv[1:] or ["1", "16"])]:
    g = torch.Generator(device=dev).manual_seed(0)
    q = b * h * w
    dense = [torch.randn((q, h >> l, w >> l, 1), device=dev, generator=g) for l in range(4)]
    coords = [J.make_coords_grid(b, h, w, device=dev) + (torch.rand((b, h, w, 2), device=dev, generator=g) - 0.5) * 16
              for _ in range(8)]
    cb = J.CorrBlock(4, r)
    tiled = TiledPyramid.from_dense(dense)
    tiled16 = TiledPyramid.from_dense(dense, storage="fp16")
    ref = None
    for name, pyr in (("row-major", dense), ("tiled", tiled), ("tiled f16", tiled16)):
        for c in coords:
            out = cb.index_pyramid(pyr, c)
        torch.cuda.synchronize()
        if ref is None:
            ref = out.clone()
        elif name != "tiled f16":
            assert bool(((out == ref) | (out.isnan() & ref.isnan())).all()), "tiled and row-major lookups differ"
        graph = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
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
        single = []
        for rep in range(24):
            e0.record()
            cb.index_pyramid(pyr, coords[rep % 8])
            e1.record()
            torch.cuda.synchronize()
            single.append(e0.elapsed_time(e1))
        single = sorted(single)[len(single) // 2]
        alg = q * 2904
        print(f"B={b:3d} {name:9s}: back to back {best * 1e3:7.2f} us = {alg / best / 1e6:5.0f} GB/s ({alg / best / 1e6 / PEAK:.3f} of peak);"
              f"  single launch between events {single * 1e3:7.2f} us = {alg / single / 1e6:5.0f} GB/s ({alg / single / 1e6 / PEAK:.3f})",
              flush=True)
    del dense, tiled, tiled16
