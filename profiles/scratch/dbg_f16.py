import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import jax_raft_b200 as J
dev = torch.device("cuda", 0)
b, h, w, d = 1, 16, 16, 128
rng = np.random.default_rng(0)
f1 = torch.from_numpy(rng.standard_normal((b, h, w, d)).astype(np.float32)).to(dev)
f2 = torch.from_numpy(rng.standard_normal((b, h, w, d)).astype(np.float32)).to(dev)
p32 = J.CorrBlock(4, 4).build_pyramid(f1, f2)
p16 = J.CorrBlock(4, 4, storage="fp16").build_pyramid(f1, f2)
torch.cuda.synchronize()
for l in range(4):
    a = p32.levels[l].half().cpu().numpy(); c = p16.levels[l].cpu().numpy()
    bad = a != c
    lw = w >> l
    for t in range(a.shape[2]):
        if t * 4 >= lw: bad[:, :, t] = False
    print("level", l, a.shape, "bad", bad.sum(), "of", bad.size)
    if bad.any():
        idx = np.argwhere(bad)
        print(" first bad idx (q,ty,tx,iy,ix):", idx[:12].tolist())
        print(" bad per tx:", [int(bad[:, :, t].sum()) for t in range(a.shape[2])], "per iy", [int(bad[:, :, :, i].sum()) for i in range(4)], "per ix", [int(bad[..., i].sum()) for i in range(4)])
        q0 = idx[0][0]
        print(" q", q0, "want", a[q0, 0, :2].ravel()[:32], "\n got", c[q0, 0, :2].ravel()[:32])
