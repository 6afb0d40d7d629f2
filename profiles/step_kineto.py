"""Per-kernel device times of one eager raft_large 440x1024 x32 forward from torch.profiler (CUPTI activity
records: warm caches, real concurrency between the side streams) — a cross-check of the ncu launch list,
whose times are cold-cache and serialised."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

import jax_raft_b200 as J  # noqa: E402
from bench import ITERS, synthetic_pairs  # noqa: E402

dev = torch.device("cuda", 0)
model, variables = J.raft_large(seed=0)
img1, img2 = synthetic_pairs(1, seed=1)
i1, i2 = torch.from_numpy(img1).to(dev), torch.from_numpy(img2).to(dev)
for _ in range(3):
    model.apply(variables, i1, i2, num_flow_updates=ITERS, use_cuda_graph=False)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    model.apply(variables, i1, i2, num_flow_updates=ITERS, use_cuda_graph=False)
    torch.cuda.synchronize()
rows = [(e.key, e.count, e.device_time_total) for e in prof.key_averages() if e.device_time_total > 0]
tot = sum(t for _, _, t in rows)
print(f"{sum(n for _, n, _ in rows)} kernels/copies, {tot / 1e3:.2f} ms summed device time")
for k, n, t in sorted(rows, key=lambda r: -r[2])[:28]:
    print(f"{100 * t / tot:5.1f} %  x{n:<4d} {t / n:7.1f} us  {k[:90]}")
