import os, sys, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=dev)
try:
    t = symm.empty((1024,), dtype=torch.float32, device=dev)
    t.fill_(rank + 1)
    hdl = symm.rendezvous(t, dist.group.WORLD)
    print(rank, "rendezvous ok; ptrs", [hex(p) for p in hdl.buffer_ptrs], "attrs", [a for a in dir(hdl) if not a.startswith("_")][:30], flush=True)
    hdl.barrier()
    peer = hdl.get_buffer((rank + 1) % world, (1024,), torch.float32)
    print(rank, "peer value", float(peer[0]), flush=True)
    hdl.barrier()
except Exception as e:
    print(rank, "FAILED", type(e).__name__, e, flush=True)
dist.destroy_process_group()
