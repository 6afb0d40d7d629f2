import torch, os, sys
sys.path.insert(0,'/root/repo')
import jax_raft_b200 as J
dev=torch.device('cuda',0)
b,h,w,d=int(os.environ.get('BB','1')),55,128,256
g=torch.Generator(device=dev).manual_seed(0)
f1=torch.randn((b,h,w,d),device=dev,generator=g); f2=torch.randn((b,h,w,d),device=dev,generator=g)
cb=J.CorrBlock(4,4,layout=os.environ.get('LAYOUT','tiled'),storage=os.environ.get('STORAGE','fp32'))
for _ in range(3): pyr=cb.build_pyramid(f1,f2)
torch.cuda.synchronize()
ts=[]
for _ in range(10):
    s,e=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    s.record(); pyr=cb.build_pyramid(f1,f2); e.record(); torch.cuda.synchronize(); ts.append(s.elapsed_time(e))
print(os.environ.get('LAYOUT','tiled'), os.environ.get('STORAGE','fp32'), 'B=%d'%b, 'build_pyramid (split+gemm) ms: min %.4f med %.4f'%(min(ts), sorted(ts)[5]))
