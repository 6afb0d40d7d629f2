"""HBM ceilings on this B200 for the roofline notes: pure write, pure read, copy (torch library kernels)."""
import torch
dev = torch.device("cuda", 0)
n = 1 << 30  # 1 GiB
a = torch.empty(n, dtype=torch.uint8, device=dev)
b = torch.empty(n, dtype=torch.uint8, device=dev)
af = a.view(torch.float32)
def timeit(fn, iters=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(iters):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        best = min(best, s.elapsed_time(e))
    return best
t = timeit(lambda: a.zero_());            print(f"memset write   {n / t / 1e6:8.1f} GB/s")
t = timeit(lambda: af.fill_(1.5));        print(f"fill   write   {n / t / 1e6:8.1f} GB/s")
t = timeit(lambda: af.sum());             print(f"sum    read    {n / t / 1e6:8.1f} GB/s")
t = timeit(lambda: b.copy_(a));           print(f"copy   rd+wr   {2 * n / t / 1e6:8.1f} GB/s")
