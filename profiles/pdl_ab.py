"""Same-process A/B of programmatic dependent launch: two CUDA graphs of the same forward (same cuDNN
engine choices), one captured with raftk_set_pdl(0) and one with raftk_set_pdl(1), replayed alternately."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import jax_raft_b200 as J  # noqa: E402
from jax_raft_b200 import _lib  # noqa: E402
from bench import H, ITERS, W, synthetic_pairs  # noqa: E402

dev = torch.device("cuda", 0)
lib = _lib.load()
model, variables = J.raft_large(seed=0)
img1, img2 = synthetic_pairs(1, seed=1)
i1, i2 = torch.from_numpy(img1).to(dev), torch.from_numpy(img2).to(dev)
P = model.prepare(variables, dev)
with torch.no_grad():
    for _ in range(3):
        model.forward(P, i1, i2, ITERS)
torch.cuda.synchronize()
graphs = {}
for on in (0, 1):
    lib.raftk_set_pdl(on)
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s), torch.no_grad():
        model.forward(P, i1, i2, ITERS)
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=s):
            out = model.forward(P, i1, i2, ITERS)
    graphs[on] = (g, out)
torch.cuda.synchronize()
assert torch.equal(graphs[0][1], graphs[1][1]), "PDL changed the result"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
times = {0: [], 1: []}
for rep in range(12):
    for on in (0, 1):
        flush.fill_(rep)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        graphs[on][0].replay()
        b.record()
        torch.cuda.synchronize()
        if rep >= 2:
            times[on].append(a.elapsed_time(b))
for on in (0, 1):
    t = sorted(times[on])
    print(f"pdl={on}: median {t[len(t) // 2]:.3f} ms, min {t[0]:.3f}, max {t[-1]:.3f}")
