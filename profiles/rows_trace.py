#!/usr/bin/env python
"""Per-item timeline of the persistent icp_rows_kernel (profiling build: -DMMB_ROWS_TRACE=1, loaded through
MMB_LIBRARY).  For every work item (iteration, pair, query group): when its block asked the queue for it, when the
item number arrived, when the pair's previous iteration was complete (`ready`), and when the block's search ended.

    MMB_LIBRARY=build/variants/libmmb_trace.so python profiles/rows_trace.py --workload c3 --out gpurun_out/trace_c3.npz
"""
import argparse
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import bench
    from mm_masking_b200 import _capi
    from mm_masking_b200.dICP import ICP
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c3")
    ap.add_argument("--out", default="gpurun_out/rows_trace.npz")
    args = ap.parse_args()
    wl = bench.WORKLOADS[args.workload]
    batch = bench.make_batch(wl, seed=0)
    dev = torch.device("cuda", 0)
    t = {k: torch.from_numpy(batch[k]).to(dev) for k in ("source", "target", "T_init", "weight")}
    alg = ICP(icp_type="pt2pl", differentiable=True, max_iterations=wl["iters"], tolerance=1e-5)
    alg.const_iter = True
    for _ in range(3):
        alg.icp(t["source"], t["target"], t["T_init"], weight=t["weight"], trim_dist=5.0,
                loss_fn={"name": "cauchy", "metric": 1.0}, dim=wl["dim"])
    torch.cuda.synchronize()
    N, n = t["source"].shape[:2]
    nblk = (n + 31) // 32
    items = wl["iters"] * N * nblk
    raw = C.CDLL(os.environ["MMB_LIBRARY"])
    buf = np.zeros((items, 4), np.uint64)
    rc = raw.mmb_debug_rows_trace(buf.ctypes.data_as(C.c_void_p), C.c_size_t(buf.nbytes))
    assert rc == 0
    q = np.zeros((items * 32, 4), np.uint32)
    assert raw.mmb_debug_rows_qtrace(q.ctypes.data_as(C.c_void_p), C.c_size_t(q.nbytes)) == 0
    np.savez_compressed(args.out, trace=buf, qtrace=q, N=N, nblk=nblk, K=wl["iters"])
    print("ok", items)


if __name__ == "__main__":
    main()
