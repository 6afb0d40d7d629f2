"""Which library kernels each convolution of the update block launches (torch.profiler / CUPTI), with
cuDNN autotune on, at the Sintel 1/8 shape.  Used to find the convolutions that make cuDNN launch
nhwcAddPaddingKernel (channel counts it has to pad)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

dev = torch.device("cuda", 0)
torch.backends.cudnn.benchmark = True
torch.backends.cudnn.allow_tf32 = True
B, h, w = int(sys.argv[1]) if len(sys.argv) > 1 else 1, 55, 128
cases = [  # name, cin, cout, kh, kw, pad, fused relu
    ("convcorr1 1x1 324->256", 324, 256, 1, 1, (0, 0), True), ("convcorr1 1x1 328->256", 328, 256, 1, 1, (0, 0), True),
    ("convcorr2 3x3 256->192", 256, 192, 3, 3, (1, 1), True), ("convcorr2 3x3 256->256", 256, 256, 3, 3, (1, 1), True),
    ("convcorr2 3x3 256->192 nofuse", 256, 192, 3, 3, (1, 1), False), ("convcorr2 3x3 256->128", 256, 128, 3, 3, (1, 1), True),
    ("convflow1 7x7 2->128", 2, 128, 7, 7, (3, 3), True), ("convflow1 7x7 8->128", 8, 128, 7, 7, (3, 3), True),
    ("convflow1 7x7 4->128", 4, 128, 7, 7, (3, 3), True),
    ("convflow2 3x3 128->64", 128, 64, 3, 3, (1, 1), True),
    ("conv 3x3 256->126", 256, 126, 3, 3, (1, 1), False), ("conv 3x3 256->128", 256, 128, 3, 3, (1, 1), False),
    ("gru zr 1x5 384->256", 384, 256, 1, 5, (0, 2), False), ("gru q 1x5 384->128", 384, 128, 1, 5, (0, 2), False),
    ("gru zr 5x1 384->256", 384, 256, 5, 1, (2, 0), False), ("gru q 5x1 384->128", 384, 128, 5, 1, (2, 0), False),
    ("flow head 3x3 128->256", 128, 256, 3, 3, (1, 1), True),
    ("flow head 3x3 256->2", 256, 2, 3, 3, (1, 1), False), ("flow head 3x3 256->8", 256, 8, 3, 3, (1, 1), False),
    ("mask 3x3 128->256", 128, 256, 3, 3, (1, 1), True), ("mask 1x1 256->576", 256, 576, 1, 1, (0, 0), False),
]
for name, cin, cout, kh, kw, pad, relu in cases:
    x = torch.randn((B, cin, h, w), device=dev).contiguous(memory_format=torch.channels_last)
    wt = torch.randn((cout, cin, kh, kw), device=dev).contiguous(memory_format=torch.channels_last) * 0.05
    b = torch.randn((cout,), device=dev)

    def run():
        if relu:
            return torch.cudnn_convolution_relu(x, wt, b, (1, 1), pad, (1, 1), 1)
        return F.conv2d(x, wt, None, padding=pad)

    for _ in range(3):
        run()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(5):
            run()
        torch.cuda.synchronize()
    ks = [(e.key[:60], e.count // 5, e.device_time_total / 5 if hasattr(e, "device_time_total") else e.cuda_time_total / 5)
          for e in prof.key_averages()]
    tot = sum(t for _, _, t in ks)
    print(f"{name:28s} {tot:7.1f} us :: " + " | ".join(f"{k} x{n} {t:.1f}" for k, n, t in ks))
