"""raftk_corr_lookup_conv1x1 (f1, row-major float32 pyramid) against the lookup + the library's fused 1x1 convolution
(TF32) — with the row-major lookup of round 1 and with the tiled float16 lookup the models use since round 2 — at the
Sintel 1/8 shape: graphs of 8 launches over 8 coordinate sets, best of 5."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import jax_raft_b200 as J  # noqa: E402

dev = torch.device("cuda", 0)
torch.backends.cudnn.benchmark = True
torch.backends.cudnn.allow_tf32 = True
h, w, r, n = 55, 128, 4, 256


def timed(fn, coords):
    for c in coords:
        fn(c)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            for c in coords:
                fn(c)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(5):
        e0.record()
        for _ in range(4):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / 32)
    return best * 1e3


for b in [int(x) for x in (sys.argv[1:] or ["1", "2", "4", "16"])]:
    gen = torch.Generator(device=dev).manual_seed(0)
    q = b * h * w
    pyr = [torch.randn((q, h >> l, w >> l, 1), device=dev, generator=gen) for l in range(4)]
    coords = [J.make_coords_grid(b, h, w, device=dev) + (torch.rand((b, h, w, 2), device=dev, generator=gen) - 0.5) * 16
              for _ in range(8)]
    weight = (torch.randn((n, 324, 1, 1), device=dev, generator=gen) / 18).contiguous(memory_format=torch.channels_last)
    bias = torch.randn((n,), device=dev, generator=gen)
    cb = J.CorrBlock(4, r)
    packed = cb.pack_conv_weight(weight)
    from jax_raft_b200.corr import TiledPyramid  # noqa: E402

    tiled16 = TiledPyramid.from_dense(pyr, storage="fp16")

    def separate_tiled(c):
        f = cb.index_pyramid(tiled16, c).permute(0, 3, 1, 2)
        return torch.cudnn_convolution_relu(f, weight, bias, (1, 1), (0, 0), (1, 1), 1)

    def separate(c):
        f = cb.index_pyramid(pyr, c).permute(0, 3, 1, 2)
        return torch.cudnn_convolution_relu(f, weight, bias, (1, 1), (0, 0), (1, 1), 1)

    def fused(c):
        return cb.index_pyramid_conv(pyr, c, packed, bias)

    print(f"B={b}: row-major lookup + conv1x1 {timed(separate, coords):.1f} us, tiled float16 lookup + conv1x1 "
          f"{timed(separate_tiled, coords):.1f} us, fused (f1) {timed(fused, coords):.1f} us per call", flush=True)
    del tiled16, pyr
