#!/usr/bin/env python
"""Short driver for ncu: `steps` forward+backward passes of the bench.py workload, nothing else.

    python profiles/prof_step.py --workload c3 --steps 2 [--nn-backend auto|rows|grid|brute]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import bench
    from mm_masking_b200.dICP import ICP
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c3")
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--nn-backend", default="auto")
    args = ap.parse_args()
    wl = bench.WORKLOADS[args.workload]
    batch = bench.make_batch(wl, seed=0)
    dev = torch.device("cuda", 0)
    t = {k: torch.from_numpy(batch[k]).to(dev) for k in ("source", "target", "T_init", "weight")}
    w = t["weight"].requires_grad_(True)
    alg = ICP(icp_type="pt2pl", differentiable=True, max_iterations=wl["iters"], tolerance=1e-5)
    alg.const_iter = True
    alg.nn_backend = bench.NN_BACKENDS[args.nn_backend]
    for _ in range(args.steps):
        w.grad = None
        out = alg.icp(t["source"], t["target"], t["T_init"], weight=w, trim_dist=5.0,
                      loss_fn={"name": "cauchy", "metric": 1.0}, dim=wl["dim"])
        out["T"].sum().backward()
    torch.cuda.synchronize()
    print("ok", float(out["T"].sum()))


if __name__ == "__main__":
    main()
