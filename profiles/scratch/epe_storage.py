"""Full-size config 2 (and 1): mean EPE of the final / worst flow vs the CPU oracle for pyramid storage fp32 / fp16."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import jax_raft_b200 as J
from oracle import corr_oracle_c, raft_cpu

def epe(a, b): return float(np.sqrt(((a - b) ** 2).sum(-1)).mean())
for name, h, w, iters in (("raft_small", 368, 496, 12), ("raft_large", 440, 1024, 32)):
    rng = np.random.default_rng(1)
    img1 = rng.uniform(-1, 1, (1, h, w, 3)).astype(np.float32)
    img2 = np.clip(np.roll(img1, (3, -5), axis=(1, 2)) + rng.normal(0, 0.02, img1.shape), -1, 1).astype(np.float32)
    net = raft_cpu.build_modules(name, seed=0)
    ref = raft_cpu.forward(net, img1, img2, iters, hot=corr_oracle_c)
    variables = J.variables_from_torch_state_dict(net.state_dict())
    cfg = J.RAFT_LARGE if name == "raft_large" else J.RAFT_SMALL
    for prec in ("tf32", "fp32"):
        for storage in ("fp32", "fp16"):
            cb = J.CorrBlock(cfg.corr_block_num_levels, cfg.corr_block_radius, storage=storage)
            model, _ = getattr(J, name)(corr_block=cb, precision=prec)
            got = model.apply(variables, img1, img2, train=False, num_flow_updates=iters).cpu().numpy()
            print(name, "conv", prec, "pyramid", storage, "EPE last %.5f worst %.5f" % (epe(got[-1], ref[-1]), max(epe(got[i], ref[i]) for i in range(iters))), flush=True)
