"""What the window lookup costs INSIDE the captured step, by elimination (no event nodes in the graph, so programmatic
dependent launch and the graph's branch overlap stay as they are in the timed bench step):

    python profiles/lookup_marginal.py

Two CUDA graphs of the same raft_large 440x1024 x32 forward, replayed alternately with an L2 flush in between:
  full   the model as bench.py runs it (float16 slabs, coordinate update fused into the lookup);
  stub   the same with the lookup kernel replaced by the bare coordinate update (raftk_flow_update) and a constant
         feature tensor — every other launch, and the data dependencies between them, unchanged.
(full - stub) / 32 is the marginal in-step cost of one lookup launch over a 2-thread-per-pixel update kernel.
The stub's flow is meaningless; only its time is used.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import jax_raft_b200 as J  # noqa: E402
from bench import ITERS, synthetic_pairs  # noqa: E402
from jax_raft_b200 import _lib  # noqa: E402


class StubLookup:
    """A corr block whose lookups launch nothing but the coordinate update."""

    def __init__(self, inner):
        self.inner = inner
        self.out_channels = inner.out_channels
        self.const = None

    def __getattr__(self, name):
        return getattr(self.inner, name)

    def build_pyramid(self, f1, f2):
        return self.inner.build_pyramid(f1, f2)

    def _features(self, coords):
        if self.const is None:
            self.const = torch.randn(tuple(coords.shape[:3]) + (self.out_channels,), device=coords.device)
        return self.const

    def index_pyramid(self, pyr, coords):
        return self._features(coords)

    def index_pyramid_update(self, pyr, coords0, coords1, delta_pre, bias, flow, flow_pitched=None):
        lib = _lib.load()
        q = coords1.numel() // 2
        ld_delta = delta_pre.shape[1] if delta_pre.shape[2:] == coords1.shape[1:3] else delta_pre.shape[3]
        ld_flow = 0 if flow_pitched is None else (flow_pitched.shape[1] if flow_pitched.shape[2:] == coords1.shape[1:3] else flow_pitched.shape[3])
        rc = lib.raftk_flow_update(delta_pre.data_ptr(), ld_delta, bias.data_ptr(), coords0.data_ptr(), coords1.data_ptr(),
                                   flow.data_ptr(), None if flow_pitched is None else flow_pitched.data_ptr(), ld_flow, q,
                                   _lib.current_stream_ptr(coords1.device))
        _lib.check("raftk_flow_update", rc)
        return self._features(coords1)


def main():
    dev = torch.device("cuda", 0)
    batch = int(os.environ.get("BATCH", "1"))
    img1, img2 = synthetic_pairs(batch, seed=1)
    i1, i2 = torch.from_numpy(img1).to(dev), torch.from_numpy(img2).to(dev)
    _, variables = J.raft_large(seed=0)
    graphs = []
    for name in ("full", "stub"):
        model, _ = J.raft_large()
        if name == "stub":
            model = J.RAFT(model.config, corr_block=StubLookup(model.corr_block), precision=model.precision)
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
        graphs.append((name, g, out, model, P))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    times = {n: [] for n, *_ in graphs}
    for rep in range(22):
        for name, g, *_ in graphs:
            flush.fill_(rep)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            g.replay()
            b.record()
            torch.cuda.synchronize()
            if rep >= 2:
                times[name].append(a.elapsed_time(b))
    med = {n: sorted(t)[len(t) // 2] for n, t in times.items()}
    for n, t in times.items():
        print(f"{n:5s} median {med[n]:.3f} ms (min {min(t):.3f}, max {max(t):.3f})")
    print(f"marginal cost of one lookup launch in the step, batch {batch}: {(med['full'] - med['stub']) / ITERS * 1e3:.2f} us "
          f"(over the bare coordinate update it replaces)")


if __name__ == "__main__":
    main()
