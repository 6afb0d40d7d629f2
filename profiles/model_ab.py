"""Same-process A/B of host-model options: one CUDA graph per variant of the raft_large 440x1024 x32 forward
(the library's convolution engines are chosen once per shape per process, so variants share them wherever
their convolutions coincide), replayed alternately with an L2 flush in between.

    [BATCH=8] python profiles/model_ab.py concurrent=False concurrent=True
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import jax_raft_b200 as J  # noqa: E402
from bench import ITERS, synthetic_pairs  # noqa: E402

dev = torch.device("cuda", 0)
variants = sys.argv[1:] or ["", "split_motion_conv=False"]
img1, img2 = synthetic_pairs(int(os.environ.get("BATCH", "1")), seed=1)
i1, i2 = torch.from_numpy(img1).to(dev), torch.from_numpy(img2).to(dev)
_, variables = J.raft_large(seed=0)
graphs = []
for v in variants:
    kw = {}
    for item in filter(None, v.split(",")):
        k, val = item.split("=")
        kw[k] = {"True": True, "False": False}.get(val, val)
    model, _ = J.raft_large(**kw)
    P = model.prepare(variables, dev)
    with torch.no_grad():
        for _ in range(3):
            model.forward(P, i1, i2, ITERS)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s), torch.no_grad():
        with torch.cuda.graph(g, stream=s):
            out = model.forward(P, i1, i2, ITERS)
    torch.cuda.synchronize()
    graphs.append((v or "default", g, out, model, P))  # the graph reads the model's prepared weights: keep them alive
ref = graphs[0][2][-1]
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
times = {entry[0]: [] for entry in graphs}
for rep in range(12):
    for name, g, *_ in graphs:
        flush.fill_(rep)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        g.replay()
        b.record()
        torch.cuda.synchronize()
        if rep >= 2:
            times[name].append(a.elapsed_time(b))
for name, _, out, *_ in graphs:
    t = sorted(times[name])
    epe = float((out[-1] - ref).pow(2).sum(-1).sqrt().mean())
    print(f"{name:40s} median {t[len(t) // 2]:.3f} ms (min {t[0]:.3f}, max {t[-1]:.3f});  mean EPE vs first variant {epe:.2e} px")
