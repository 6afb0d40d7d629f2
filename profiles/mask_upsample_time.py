"""f2: raftk_mask_upsample against the library's 1x1 convolution + raftk_convex_upsample_bias, at the Sintel
1/8 shape; CUDA graphs of 8 calls over 8 input sets (L2 flushed by the working set at large batch), best of 5."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

import jax_raft_b200 as J  # noqa: E402

dev = torch.device("cuda", 0)
torch.backends.cudnn.benchmark = True
torch.backends.cudnn.allow_tf32 = True
h, w, k = 55, 128, 256


def timed(fn, n):
    for i in range(n):
        fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            for i in range(n):
                fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(5):
        e0.record()
        for _ in range(4):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / (4 * n))
    return best * 1e3


for b in [int(x) for x in (sys.argv[1:] or ["1", "4", "16", "64"])]:
    gen = torch.Generator(device=dev).manual_seed(0)
    n = 8 if b <= 16 else 2
    feats = [torch.relu(torch.randn((b, k, h, w), device=dev, generator=gen)).contiguous(memory_format=torch.channels_last)
             for _ in range(n)]
    flow = torch.randn((b, h, w, 2), device=dev, generator=gen)
    weight = (torch.randn((576, k, 1, 1), device=dev, generator=gen) / 16).contiguous(memory_format=torch.channels_last)
    bias = torch.randn((576,), device=dev, generator=gen)
    pw, pb = J.pack_mask_predictor(weight, bias)
    out = torch.empty((b, 8 * h, 8 * w, 2), device=dev)

    def separate(i):
        mask = F.conv2d(feats[i], weight, None).permute(0, 2, 3, 1)
        return J.upsample_flow(flow, mask, mask_bias=bias, out=out)

    def fused(i):
        return J.mask_upsample(flow, feats[i].permute(0, 2, 3, 1), pw, pb, out=out)

    print(f"B={b}: conv1x1 + upsample {timed(separate, n):.1f} us, fused {timed(fused, n):.1f} us per call")
