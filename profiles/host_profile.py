import sys, time, cProfile, pstats, io, torch
sys.path.insert(0,'/root/repo')
import bench
from mm_masking_b200.dICP import ICP
from mm_masking_b200.prefetch import DevicePrefetcher
wl=bench.WORKLOADS['c3']; batch=bench.make_batch(wl,0); dev=torch.device('cuda',0)
host={k: torch.from_numpy(batch[k]).pin_memory() for k in ("source","target","T_init","weight")}
alg=ICP(icp_type="pt2pl",differentiable=True,max_iterations=10,tolerance=1e-5); alg.const_iter=True
call=dict(trim_dist=5.0, loss_fn={"name":"cauchy","metric":1.0}, dim=2)
out_T=torch.empty((32,4,4)).pin_memory(); out_gw=torch.empty((32,2000)).pin_memory()
def step(b):
    w=b["weight"].requires_grad_(True)
    out=alg.icp(b["source"],b["target"],b["T_init"],weight=w,**call)
    out["T"].sum().backward()
    out_T.copy_(out["T"].detach(),non_blocking=True); out_gw.copy_(w.grad,non_blocking=True)
pre=DevicePrefetcher((host for _ in range(300)),dev)
for _ in range(10): step(next(pre))
torch.cuda.synchronize()
t0=time.perf_counter()
for _ in range(100): step(next(pre))
t1=time.perf_counter(); torch.cuda.synchronize(); t2=time.perf_counter()
print('host issue ms/step', (t1-t0)*10, 'total ms/step', (t2-t0)*10)
pr=cProfile.Profile(); pr.enable()
for _ in range(100): step(next(pre))
pr.disable(); torch.cuda.synchronize()
s=io.StringIO(); pstats.Stats(pr,stream=s).sort_stats('cumulative').print_stats(28); print(s.getvalue()[:6000])
