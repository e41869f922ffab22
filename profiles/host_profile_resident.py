"""Host time to ISSUE one resident c3 step (no upload, no sync): forward call, loss + backward; cProfile of 200 steps."""
import sys, time, cProfile, pstats, io, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from mm_masking_b200.dICP import ICP
wl = bench.WORKLOADS['c3']; batch = bench.make_batch(wl, 0); dev = torch.device('cuda', 0)
t = {k: torch.from_numpy(batch[k]).to(dev) for k in ("source", "target", "T_init", "weight")}
w = t["weight"].requires_grad_(True)
alg = ICP(icp_type="pt2pl", differentiable=True, max_iterations=10, tolerance=1e-5); alg.const_iter = True
call = dict(trim_dist=5.0, loss_fn={"name": "cauchy", "metric": 1.0}, dim=2)
def fwd():
    return alg.icp(t["source"], t["target"], t["T_init"], weight=w, **call)
def step():
    w.grad = None
    out = fwd()
    out["T"].sum().backward()
for _ in range(20): step()
torch.cuda.synchronize()
for name, fn in (("forward only (no grad)", lambda: fwd()), ("step (fwd + loss + bwd)", step)):
    best = 1e9
    for rep in range(5):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        if name.startswith("forward"):
            with torch.no_grad():
                for _ in range(50): fn()
        else:
            for _ in range(50): fn()
        t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
        best = min(best, (t1 - t0) / 50 * 1e3)
    print(f"{name}: host issue {best:.3f} ms, with sync {(t2 - t0) / 50 * 1e3:.3f} ms")
pr = cProfile.Profile(); pr.enable()
for _ in range(200): step()
pr.disable(); torch.cuda.synchronize()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats('tottime').print_stats(32); print(s.getvalue()[:7000])
