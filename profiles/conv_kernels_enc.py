"""First encoder convolution (7x7 stride 2, 3 input channels) with the input zero-padded to 4 / 8 channels."""
import sys
import torch
import torch.nn.functional as F
from torch.profiler import ProfilerActivity, profile

dev = torch.device("cuda", 0)
torch.backends.cudnn.benchmark = True
torch.backends.cudnn.allow_tf32 = True
for B in (1, 2):
    for cin in (3, 4, 8):
        x = torch.randn((B, cin, 440, 1024), device=dev).contiguous(memory_format=torch.channels_last)
        wt = torch.randn((64, cin, 7, 7), device=dev).contiguous(memory_format=torch.channels_last) * 0.05
        run = lambda: F.conv2d(x, wt, None, stride=2, padding=3)
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for _ in range(5):
                run()
            torch.cuda.synchronize()
        ks = [(e.key[:50], e.count // 5, e.device_time_total / 5) for e in prof.key_averages() if e.device_time_total > 0]
        print(f"B={B} cin={cin}: {sum(t for _, _, t in ks):7.1f} us :: " + " | ".join(f"{k} x{n} {t:.1f}" for k, n, t in ks))
