#!/usr/bin/env python
"""Headline benchmark: batched weighted point-to-plane ICP, forward + backward.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c3|c5|small]

One "step" = one pass of the hot path over one batch of synthetic scan/map pairs: the call
`ICP.icp(source, target, T_init, weight, trim_dist=5, loss_fn=cauchy(1), dim=2)` that
`mm_masking/icp_weight_policy.py:281` makes, followed by `T.sum().backward()` so that the
per-point weight gradients the UNet needs are produced.  Workload c3 (BASELINE.json
configs[2]): 32 pairs per GPU, 2 000 radar points vs 50 000 lidar map points, 10 iterations
(`const_iter: True` so every pair does all 10).

  value  pairs/s with inputs resident in HBM (CUDA events per step, L2 flushed between steps): the step captured
         once with torch.cuda.graph and replayed (every kernel runs on every replay; checked equal to the eager call;
         falls back to the eager loop if the capture fails, --no-graph forces it); `eager` = the eager loop
  e2e    pairs/s through the same public call with HOST (pinned) inputs: H2D of source, target,
         T_init, weight and D2H of T and d(loss)/d(weight) inside the timed region, every step;
         fed through mm_masking_b200.prefetch.DevicePrefetcher (the upload of step k+1 runs on a
         copy stream under the kernels of step k); one event pair around all K steps
  roofline       the dominant kernel of the default path, the exact row-major cell-grid search fused with the ICP
                 step (icp_rows_kernel, one persistent launch for all iterations): algorithmic bytes per launch /
                 its CUDA-event duration against the measured HBM peak, flagged work-avoiding
  roofline_exhaustive_nn   the exhaustive kernel (nn_search_kernel, --nn-backend brute) on the
                 same data in the same run: 8 flop per (source, target) pair per launch / its
                 CUDA-event duration, against the FP32 FMA-pipe rate measured in the same run by
                 mmb_fp32_probe (MEASURED_PEAKS.json has no CUDA-core FP32 figure)
  cpu_baseline   the unmodified reference (oracle/_ref/dicp_ref, torch CPU) on the box's host cores, or the
                 oracle port when that copy is absent (`kind` says which)

Multi-GPU (torchrun): pairs are independent, so each rank runs its own 32 pairs and there is no
data-path collective; value = all ranks' pairs / max-over-ranks time ("weak").
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: N per GPU, n, m, iterations, dim, three_d
    "c3": dict(N=32, n=2000, m=50000, iters=10, dim=2, three_d=False,
               desc="batched weighted pt2pl ICP, 32 pairs/GPU, 2000 radar pts vs 50000 lidar map pts, 10 it, fwd+bwd"),
    "c5": dict(N=32, n=2000, m=200000, iters=20, dim=3, three_d=True,
               desc="SE(3) pt2pl ICP, 32 pairs/GPU, 2000 pts vs 200000 map pts, 20 it, fwd+bwd"),
    # the map extends to 2-3x the scan footprint: most targets fall outside the cell grid (which is fitted to the
    # scan) into the unsorted outside bucket, scanned in full by every query whose bound reaches the grid border
    "c3wide": dict(N=32, n=2000, m=50000, iters=10, dim=2, three_d=False, scan_range=30.0,
                   desc="as c3, but the scan covers 30 m of a +-80 m map (map extent ~2.7x the scan footprint)"),
    "small": dict(N=4, n=500, m=5000, iters=5, dim=2, three_d=False, desc="smoke-sized"),
}
METRIC = "icp_pairs_per_sec_fwd_bwd"
UNIT = "pairs/s"
NN_BACKENDS = {"auto": 0, "brute": 1, "grid": 2, "rows": 3}
FLOP_PER_PAIR = 8          # 3 sub, 3 mul, 2 add: the distance `torch.cdist` evaluates


def kernel_source_file(kernel):
    """the .cu file under mm_masking_b200/csrc that defines `kernel` (None if none does)"""
    d = os.path.join(ROOT, "mm_masking_b200", "csrc")
    for f in sorted(os.listdir(d)):
        if f.endswith(".cu") and (" " + kernel + "(") in open(os.path.join(d, f)).read():
            return f
    return None


def kernel_source_sha(kernel=None):
    """sha256 (first 16 hex) over the CUDA sources a kernel is built from -- the shared headers and the .cu file that
    defines it (every .cu when `kernel` is None or unknown): the key that ties an ncu capture in
    profiles/traffic.json to the binary that is being timed"""
    import hashlib
    h = hashlib.sha256()
    d = os.path.join(ROOT, "mm_masking_b200", "csrc")
    own = kernel_source_file(kernel) if kernel else None
    for f in sorted(os.listdir(d)):
        if f.endswith(".cuh") or (f.endswith(".cu") and (own is None or f == own)):
            h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:16]


def traffic_lookup(kernel, workload):
    """DRAM bytes per launch of `kernel` at `workload` from the committed ncu captures (profiles/traffic.json,
    written by profiles/ncu_summarise.py).  Returns (bytes or None, info dict)."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    sha = kernel_source_sha(kernel)
    if not os.path.exists(path):
        return None, {"traffic_stale": True, "traffic_source": "profiles/traffic.json missing"}
    best = None
    for e in json.load(open(path)).get("entries", []):
        if e.get("kernel") == kernel and e.get("workload") == workload:
            if best is None or e.get("source_sha") == sha:
                best = e
    if best is None:
        return None, {"traffic_stale": True, "traffic_source": f"no ncu capture of {kernel} at {workload} in profiles/traffic.json"}
    if best.get("source_sha") != sha:
        return None, {"traffic_stale": True, "traffic_source": best.get("capture"), "traffic_capture_sha": best.get("source_sha"),
                      "source_sha": sha, "traffic_of_stale_capture": best.get("dram_bytes_per_launch")}
    return best.get("dram_bytes_per_launch"), {"traffic_stale": False, "traffic_source": best.get("capture"), "source_sha": sha}


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if "-" in part:
            a, b = part.split("-"); cpus.update(range(int(a), int(b) + 1))
        elif part:
            cpus.add(int(part))
    return cpus


def bind_to_gpu_numa_node(local_rank):
    """Run this rank on CPU cores local to its GPU WHILE its pinned host buffers are allocated (first touch places
    them on that NUMA node), so that 8 ranks do not all upload through one socket -- the r01 end-to-end scaling
    loss; at N = 8 the four ranks whose buffers sat on the remote socket uploaded at 21 GB/s against 33 GB/s.
    The PCI topology comes from sysfs; where the container hides it (the GPU-local list covers every allowed core)
    the node is found empirically: the allowed cores are split into four contiguous groups, a pinned 64 MB buffer is
    allocated and first-touched from each, and the group whose buffer uploads fastest wins.  Returns a dict for the
    JSON line; call `restore_affinity` after the allocations."""
    import torch
    allowed = set(os.sched_getaffinity(0))
    info = {"bound": False, "allowed_cpus": len(allowed)}
    try:
        pr = torch.cuda.get_device_properties(local_rank)
        path = f"/sys/bus/pci/devices/{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0/local_cpulist"
        cpus = _parse_cpulist(open(path).read()) & allowed
        if cpus and cpus != allowed:
            os.sched_setaffinity(0, cpus)
            info.update(bound=True, how="sysfs local_cpulist", cpus=len(cpus))
            return info
    except Exception as e:                       # containers often hide sysfs topology
        info["sysfs"] = type(e).__name__
    if len(allowed) < 8:
        info["why"] = "fewer than 8 cores allowed: nothing to choose between"
        return info
    try:
        order = sorted(allowed)
        groups = [set(order[i * len(order) // 4:(i + 1) * len(order) // 4]) for i in range(4)]
        dev = torch.device("cuda", local_rank)
        dst = torch.empty(64 << 20, dtype=torch.uint8, device=dev)
        rates = []
        for g in groups:
            os.sched_setaffinity(0, g)
            buf = torch.empty(64 << 20, dtype=torch.uint8).pin_memory()      # allocated and first-touched from group g
            dst.copy_(buf, non_blocking=True); torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(4):
                dst.copy_(buf, non_blocking=True)
            b.record(); torch.cuda.synchronize()
            rates.append(4 * buf.numel() / (a.elapsed_time(b) * 1e-3) / 1e9)
            del buf
            if hasattr(torch._C, "_host_emptyCache"):
                torch._C._host_emptyCache()          # the probe block must not be recycled for the real buffers
        best = max(range(4), key=lambda i: rates[i])
        info["probe_gbs"] = [round(r, 1) for r in rates]
        if rates[best] > 1.15 * min(rates):
            os.sched_setaffinity(0, groups[best])
            info.update(bound=True, how="empirical: fastest of four core groups for a pinned upload", cpus=len(groups[best]),
                        group=best)
        else:
            restore_affinity(allowed)
            info["why"] = "all core groups upload at the same rate"
        del dst
    except Exception as e:
        restore_affinity(allowed)
        info["probe_error"] = f"{type(e).__name__}: {e}"
    return info


def restore_affinity(allowed):
    """give EVERY thread of the process its full CPU allowance back: worker threads (the torch intra-op pool) that
    were created while the main thread was bound inherit the narrow mask and would keep it -- the CPU baseline of
    this very run would then be timed on a quarter of the cores"""
    try:
        tids = [int(t) for t in os.listdir("/proc/self/task")]
    except Exception:
        tids = [0]
    for tid in tids + [0]:
        try:
            os.sched_setaffinity(tid, allowed)
        except Exception:
            pass


def gather_per_rank(value, device):
    """list of one float per rank (rank order) on every rank"""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return [float(value)]
    t = torch.zeros(dist.get_world_size(), dtype=torch.float64, device=device)
    t[dist.get_rank()] = float(value)
    dist.all_reduce(t)
    return [float(x) for x in t.tolist()]


def make_batch(wl, seed):
    from mm_masking_b200 import synth
    return synth.make_icp_batch(N=wl["N"], n=wl["n"], m=wl["m"], seed=seed, three_d=wl["three_d"],
                                scan_range=wl.get("scan_range", 76.0))


# ---------------------------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits",
                 "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self, t0, t1):
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = [r for t, r in self.rows if t0 <= t <= t1 + 0.15] or [r for _, r in self.rows]
        for r in rows:
            f = [x.strip() for x in r.split(",")]
            try:
                sm.append(float(f[0])); mx = max(mx, float(f[1]))
            except Exception:
                continue
            for name, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ---------------------------------------------------------------------------------------------
# CPU reference arm / cpu_baseline (oracle port; the one place bench.py executes oracle/)
# ---------------------------------------------------------------------------------------------
REF_DIR = os.path.join(ROOT, "oracle", "_ref", "dicp_ref")


def reference_kind():
    """"reference": the UNMODIFIED dICP package copied by oracle/Makefile into the git-ignored oracle/_ref/
    (it travels to the GPU box like the built .so); "port": the oracle restatement, when that copy is absent."""
    return "reference" if os.path.exists(os.path.join(REF_DIR, "dICP", "ICP.py")) else "port"


_REF_ICP = {}


def _reference_icp(iters):
    """the reference's own `dICP.ICP.ICP` (external/dICP/dICP/ICP.py:14-46), constructed as
    `mm_masking/icp_weight_policy.py:54` does; `const_iter` is the reference's own config switch
    (config/dICP_config.yaml:7), set on the instance so that every pair does all iterations."""
    if iters not in _REF_ICP:
        from oracle import ref_loader
        ref_loader._stub_matplotlib()              # the reference imports matplotlib at module top, unused on the path
        if REF_DIR not in sys.path:
            sys.path.insert(0, REF_DIR)
        from dICP.ICP import ICP as RefICP
        alg = RefICP(icp_type="pt2pl", differentiable=True, max_iterations=iters, tolerance=1e-5)
        alg.const_iter = True
        _REF_ICP[iters] = alg
    return _REF_ICP[iters]


def cpu_reference_step(batch, wl, pairs, chunk=8):
    """fwd+bwd of the reference on `pairs` pairs, on the host cores, in chunks of <= 8 pairs (its (N,n,m)
    distance tensor is 0.4 GB per pair at c3).  Returns seconds."""
    import torch
    kind = reference_kind()
    secs = 0.0
    for s0 in range(0, pairs, chunk):
        t = lambda a: torch.tensor(a[s0:min(pairs, s0 + chunk)])
        src, tgt, T0 = t(batch["source"]), t(batch["target"]), t(batch["T_init"])
        w = t(batch["weight"]).requires_grad_(True)
        t0 = time.perf_counter()
        if kind == "reference":
            out = _reference_icp(wl["iters"]).icp(src, tgt, T0, weight=w, trim_dist=5.0,
                                                  loss_fn={"name": "cauchy", "metric": 1.0}, dim=wl["dim"])
        else:
            from oracle import dicp_oracle as O
            out = O.dicp(src, tgt, T0, w, icp_type="pt2pl", max_iterations=wl["iters"], tolerance=1e-5,
                         differentiable=True, trim_dist=5.0, loss_fn={"name": "cauchy", "metric": 1.0},
                         dim=wl["dim"], const_iter=True, nn_mode="cdist")
        out["T"].sum().backward()
        secs += time.perf_counter() - t0
    return secs


def reference_sample_text(kind, pairs, iters):
    what = ("unmodified reference dICP.ICP.icp (oracle/_ref/dicp_ref, torch CPU)" if kind == "reference"
            else "oracle port (torch.cdist+argmin NN)")
    return f"{pairs} pair(s) x {iters} it fwd+bwd per step, fp32, {what}"


def parity_check(alg, call, wl, batch, dev_tensors, pairs=2):
    """The timed call once more with the per-iteration correspondences recorded, against the CPU oracle
    (fp64, exhaustive-search semantics of `external/dICP/dICP/nn.py:32-38` on k-d tree proposals) on the first
    `pairs` pairs of the very batch that was timed.  BASELINE.json tolerances: indices bit-exact, pose 1e-5 m /
    1e-6 rad, d/d weight 1e-4 relative."""
    import numpy as np
    import torch
    from oracle import dicp_oracle as O
    d_src, d_tgt, d_T0, d_w = dev_tensors
    w = d_w.detach().clone().requires_grad_(True)
    alg.record_indices = True
    out = alg.icp(d_src, d_tgt, d_T0, weight=w, **call)
    out["T"].sum().backward()
    idx = alg.last_indices[:, :pairs].cpu().numpy().astype(np.int64)
    alg.record_indices = False
    T = out["T"][:pairs].detach().double().cpu().numpy()
    g = w.grad[:pairs].double().cpu()
    t64 = lambda a: torch.tensor(np.asarray(a[:pairs], dtype=np.float64))
    w_ref = t64(batch["weight"]).requires_grad_(True)
    ref = O.dicp(t64(batch["source"]), t64(batch["target"]), t64(batch["T_init"]), w_ref, icp_type="pt2pl",
                 max_iterations=wl["iters"], tolerance=1e-5, differentiable=True, trim_dist=5.0,
                 loss_fn={"name": "cauchy", "metric": 1.0}, dim=wl["dim"], const_iter=True, nn_mode="kdtree",
                 return_indices=True)
    ref["T"].sum().backward()
    T_ref = ref["T"].detach().numpy()
    dT = float(np.abs(T[:, :3, 3] - T_ref[:, :3, 3]).max())
    dR = np.einsum("bij,bkj->bik", T[:, :3, :3], T_ref[:, :3, :3])
    sk = np.stack([dR[:, 2, 1] - dR[:, 1, 2], dR[:, 0, 2] - dR[:, 2, 0], dR[:, 1, 0] - dR[:, 0, 1]], 1) / 2
    dang = float(np.linalg.norm(sk, axis=1).max())
    scale = max(float(w_ref.grad.abs().max()), 1e-30)
    dgrad = float((g - w_ref.grad).abs().max()) / scale
    mism = int((idx != ref["indices"].numpy()).sum())
    ok = mism == 0 and dT < 1e-5 and dang < 1e-6 and dgrad < 1e-4
    return {"pairs": pairs, "iterations": int(idx.shape[0]), "idx_mismatch": mism, "idx_compared": int(idx.size),
            "dT_max": dT, "dR_max": dang, "dgrad_rel": dgrad, "ok": bool(ok),
            "against": "oracle/dicp_oracle.py fp64, nn_mode=kdtree (exhaustive-search semantics), same batch"}


def measure_also(args, rank, world, local_rank, dev):
    """c5 (BASELINE.json configs[4]: 256 pairs GLOBAL, 200 000-point submaps, SE(3), 20 it; strong scaling), c4
    (configs[3]: training step, global batch 64, NCCL all-reduce of the UNet gradients) and c2 (configs[1]: radar
    preprocessing, batch 8 per GPU) on the ranks of this run.  Returns the dict on rank 0, None elsewhere."""
    import torch
    import torch.distributed as dist
    from mm_masking_b200 import dist as mdist
    from mm_masking_b200.dICP import ICP
    out = {}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- c5, fixed global batch -----------------------------------------------------------------
    try:
        wl = WORKLOADS["c5"]
        G = args.c5_global_pairs
        s, e = mdist.shard_range(G, world, rank)
        sub = dict(wl, N=e - s)
        batch = make_batch(sub, seed=1000 + rank)
        t = {k: torch.from_numpy(batch[k]).to(dev) for k in ("source", "target", "T_init", "weight")}
        w = t["weight"].requires_grad_(True)
        alg = ICP(icp_type="pt2pl", differentiable=True, max_iterations=wl["iters"], tolerance=1e-5)
        alg.const_iter = True
        call = dict(trim_dist=5.0, loss_fn={"name": "cauchy", "metric": 1.0}, dim=wl["dim"])

        def step():
            w.grad = None
            o = alg.icp(t["source"], t["target"], t["T_init"], weight=w, **call)
            o["T"].sum().backward()

        for _ in range(2):
            step()
        steps = 4
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        barrier()
        for a, b in ev:
            a.record(); step(); b.record()
        barrier()
        ranks_ms = gather_per_rank(sum(a.elapsed_time(b) for a, b in ev) / steps, dev)
        par = parity_check(alg, call, wl, batch, (t["source"], t["target"], t["T_init"], w), pairs=1) if rank == 0 else None
        out["c5"] = {"metric": METRIC, "value": G / (max(ranks_ms) * 1e-3), "unit": UNIT, "n_gpus": world, "scaling": "strong",
                     "ms_per_step": max(ranks_ms), "per_rank_ms_per_step": ranks_ms, "steps": steps, "warmup": 2,
                     "config": {"workload": "c5: " + wl["desc"].replace("32 pairs/GPU", f"{G} pairs GLOBAL split over the ranks"),
                                "global_pairs": G, "pairs_per_gpu": e - s, "n": wl["n"], "m": wl["m"],
                                "iterations": wl["iters"], "dim": 3, "l2": "working set 2.4 GB per 256 pairs: larger than L2"},
                     "parity": par}
        del t, w, batch
        torch.cuda.empty_cache()
    except Exception as ex:                                  # an auxiliary measurement must not sink the headline
        out["c5"] = {"error": f"{type(ex).__name__}: {ex}"}
    # ---- c4, training step ----------------------------------------------------------------------
    try:
        import bench_train
        targs = bench_train.parse_args(["--gpus", str(world), "--steps", "6", "--warmup", "3"])
        line = bench_train.run(targs, rank, world, local_rank, own_process_group=False, sample_clocks=False)
        out["c4"] = line
        torch.cuda.empty_cache()
    except Exception as ex:
        out["c4"] = {"error": f"{type(ex).__name__}: {ex}"}
    # ---- c2, radar preprocessing (every rank its own batch of 8) -----------------------------------
    try:
        import bench_radar
        rargs = bench_radar.parse_args(["--batch", "8", "--reps", "10", "--no-cpu-baseline"])
        line = bench_radar.run(rargs, local_rank)
        ms = gather_per_rank(line["ms_per_step"], dev)
        ms_e = gather_per_rank(line["e2e"]["ms_per_step"], dev)
        line.update(value=world * 8 / (max(ms) * 1e-3), n_gpus=world, ms_per_step=max(ms))
        line["e2e"].update(value=world * 8 / (max(ms_e) * 1e-3), ms_per_step=max(ms_e))
        out["c2"] = line
    except Exception as ex:
        out["c2"] = {"error": f"{type(ex).__name__}: {ex}"}
    return out if rank == 0 else None


def run_reference(args, wl, rank, world):
    import torch
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    pairs = max(1, min(wl["N"], args.ref_pairs))
    sub = dict(wl, N=pairs)
    batch = make_batch(sub, seed=0)
    for _ in range(min(args.warmup, 1)):
        cpu_reference_step(batch, wl, 1)
    secs = [cpu_reference_step(batch, wl, pairs) for _ in range(args.steps)]
    total = sum(secs)
    value = pairs * args.steps / total
    kind = reference_kind()
    sample = reference_sample_text(kind, pairs, wl["iters"]) + f", {args.steps} steps"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload + ": " + wl["desc"], "pairs_per_step": pairs, "n": wl["n"], "m": wl["m"],
                   "iterations": wl["iters"], "dim": wl["dim"]},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------
def run_ours(args, wl, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist
    from mm_masking_b200 import _capi, dist as mdist
    from mm_masking_b200.dICP import ICP
    from mm_masking_b200.prefetch import DevicePrefetcher

    _capi.require_cuda()
    lib = _capi.load_library()
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    cpus_allowed = set(os.sched_getaffinity(0))
    numa = bind_to_gpu_numa_node(local_rank)      # before pin_memory()
    if world > 1:
        import datetime
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=180))

    batch = make_batch(wl, seed=rank)
    N, n, m, K = wl["N"], wl["n"], wl["m"], wl["iters"]
    host = {k: torch.from_numpy(batch[k]).pin_memory() for k in ("source", "target", "T_init", "weight")}
    d_src, d_tgt, d_T0 = (host[k].to(dev) for k in ("source", "target", "T_init"))
    d_w = host["weight"].to(dev).requires_grad_(True)
    out_T = torch.empty((N, 4, 4), dtype=torch.float32).pin_memory()
    out_gw = torch.empty((N, n), dtype=torch.float32).pin_memory()
    restore_affinity(cpus_allowed)                # the buffers stay where they were first touched

    alg = ICP(icp_type="pt2pl", differentiable=True, max_iterations=K, tolerance=1e-5)
    alg.const_iter = True
    alg.nn_slices = args.nn_slices
    alg.nn_backend = NN_BACKENDS[args.nn_backend]
    call = dict(trim_dist=5.0, loss_fn={"name": "cauchy", "metric": 1.0}, dim=wl["dim"])
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > 126 MB L2

    def step_resident():
        d_w.grad = None
        out = alg.icp(d_src, d_tgt, d_T0, weight=d_w, **call)
        out["T"].sum().backward()
        return out

    def step_e2e(b):
        w = b["weight"].requires_grad_(True)
        out = alg.icp(b["source"], b["target"], b["T_init"], weight=w, **call)
        out["T"].sum().backward()
        out_T.copy_(out["T"].detach(), non_blocking=True)
        out_gw.copy_(w.grad, non_blocking=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    per_rank = []

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        barrier()
        t0 = time.time()
        for a, b in ev:
            flush.fill_(1)                      # evict the previous step's data from L2 (untimed)
            a.record()
            fn()
            b.record()
        barrier()
        t1 = time.time()
        mine = sum(a.elapsed_time(b) for a, b in ev) / steps
        per_rank.clear(); per_rank.extend(gather_per_rank(mine, dev))
        return max(per_rank), t0, t1

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.3)
    launches0 = lib.mmb_launch_count()
    ms_res, t0, t1 = timed(step_resident, args.steps, args.warmup)
    launches_resident = lib.mmb_launch_count() - launches0          # includes the W warm-up steps

    # The same resident step captured ONCE in a CUDA graph and replayed (const_iter: the call path has no host
    # synchronisation; tests/test_graph_capture_gpu.py checks replay == eager bit for bit).  Reported beside `value`,
    # which stays the eager call: it shows what is left when the host's share of issuing a step is taken out.
    graphed, per_rank_eager = None, list(per_rank)
    g = None
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())

        def step_for_capture():
            out = alg.icp(d_src, d_tgt, d_T0, weight=d_w, **call)
            return out["T"], torch.autograd.grad(out["T"].sum(), d_w)[0]
        with torch.cuda.stream(side):
            for _ in range(3):
                step_for_capture()
        torch.cuda.current_stream().wait_stream(side)
        g = torch.cuda.CUDAGraph()
        # thread_local: a CUDA call from another thread of the process (the NCCL watchdog polling its events, the
        # pinned-memory allocator) must not invalidate the capture
        with torch.cuda.graph(g, capture_error_mode="thread_local"):
            g_T, g_gw = step_for_capture()
    except Exception as exc:                                         # informational: never fails the bench
        graphed, g = {"error": repr(exc)[:200]}, None
    # every rank takes the timed loop (it holds barriers) or none does
    if mdist.max_over_ranks(0.0 if g is not None else 1.0, device=dev) == 0.0:
        ms_graph, _, _ = timed(g.replay, args.steps, args.warmup)
        eager = step_resident()
        torch.cuda.synchronize()
        graphed = {"ms_per_step": ms_graph, "value": world * N / (ms_graph * 1e-3), "unit": UNIT,
                   "same_as_eager": bool(torch.equal(g_T, eager["T"]) and torch.equal(g_gw, d_w.grad)),
                   "per_rank_ms_per_step": list(per_rank),
                   "how": "forward + backward of the resident step captured with torch.cuda.graph, one replay per step"}
    elif graphed is None:
        graphed = {"error": "capture failed on another rank"}
    del g
    per_rank.clear(); per_rank.extend(per_rank_eager)
    # e2e: the same call fed from pinned HOST tensors through the package's DevicePrefetcher (what
    # a user's loop does): every step uploads its own source/target/T_init/weight (no reuse across
    # steps; the upload of step k+1 runs on a copy stream under the kernels of step k) and reads T
    # and d/d weight back.  ONE event pair around all K steps, L2 flushes included, so nothing that
    # happens between steps escapes the clock.
    e2e_warm = max(2, args.warmup // 2)
    pre = DevicePrefetcher((host for _ in range(e2e_warm + args.steps * max(1, args.e2e_repeats) + 1)), dev)
    for _ in range(e2e_warm):
        step_e2e(next(pre))
    barrier()
    up0 = pre.bytes_uploaded
    launches1 = lib.mmb_launch_count()
    # >= 3 repeats of the K-step loop, median reported (20 steps are ~25 ms: too short for one sample of a
    # loop whose uploads overlap the kernels)
    e2e_ms, e2e_host_ms = [], []
    for rep in range(max(1, args.e2e_repeats)):
        ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        h0 = time.perf_counter()
        ea.record()
        for _ in range(args.steps):
            flush.fill_(1)
            step_e2e(next(pre))
        eb.record()
        e2e_host_ms.append(1e3 * (time.perf_counter() - h0) / args.steps)     # host time to ISSUE a step (no sync)
        barrier()
        e2e_rank = gather_per_rank(ea.elapsed_time(eb) / args.steps, dev)
        e2e_ms.append(max(e2e_rank))
    t1 = time.time()
    launches_e2e = lib.mmb_launch_count() - launches1
    h2d_per_step = (pre.bytes_uploaded - up0) // (args.steps * len(e2e_ms))
    ms_e2e = sorted(e2e_ms)[len(e2e_ms) // 2]
    ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # for comparison: the reference's loop shape, upload on the compute stream in front of each call
    ea.record()
    for _ in range(args.steps):
        flush.fill_(1)
        step_e2e({k: v.to(dev, non_blocking=True) for k, v in host.items()})
    eb.record()
    barrier()
    ms_e2e_serial = mdist.max_over_ranks(ea.elapsed_time(eb), device=dev) / args.steps
    clocks = None
    if sampler:
        sampler.stop()
        clocks = sampler.summary(t0, t1)

    per_rank_res = list(per_rank)
    also = None
    if args.also:
        # the other BASELINE.json configs, measured in this process at this N (all ranks take part: c4 holds the
        # only data-path collective of the package, c5 is a fixed GLOBAL batch split over the ranks)
        del flush
        torch.cuda.empty_cache()
        also = measure_also(args, rank, world, local_rank, dev)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    parity = parity_check(alg, call, wl, batch, (d_src, d_tgt, d_T0, d_w), pairs=min(2, N))

    # ---- roofline of the dominant kernel: per-class CUDA-event timing of the same forward -----
    NC = _capi.PROFILE_CLASSES
    f32 = dict(dtype=torch.float32, device=dev)
    w_det = d_w.detach()

    def profile_forward(backend):
        """mean ms per launch of each kernel class (mmb_icp_forward_profiled, events on the launch stream)"""
        p = _capi.IcpParams()
        p.N, p.n, p.m, p.target_cols = N, n, m, 6
        p.dtype, p.icp_type, p.dim, p.max_iterations = _capi.F32, _capi.PT2PL, wl["dim"], K
        p.differentiable, p.const_iter, p.use_trim, p.trim_dist = 1, 1, 1, 5.0
        p.loss_kind, p.loss_metric, p.has_weight, p.save_for_backward = _capi.LOSS_CAUCHY, 1.0, 1, 1
        p.tolerance, p.tanh_steepness, p.match_ratio_thresh = 1e-5, 5.0, 0.1
        p.nn_slices, p.nn_backend = args.nn_slices, backend
        ws_bytes, sv_bytes = lib.mmb_icp_workspace_bytes(p), lib.mmb_icp_saved_bytes(p)
        bufs = dict(ws=torch.empty(ws_bytes, dtype=torch.uint8, device=dev), sv=torch.empty(sv_bytes, dtype=torch.uint8, device=dev),
                    T=torch.empty((N, 4, 4), **f32), pc=torch.empty((N, n, 3), **f32), de=torch.zeros((N, K, 6), **f32),
                    wh=torch.zeros((N, K, n), **f32), co=torch.zeros((N, K), **f32),
                    cv=torch.zeros(N, dtype=torch.uint8, device=dev), it=torch.zeros(N, **f32), mr=torch.zeros(N, **f32),
                    idn=torch.zeros(1, dtype=torch.int32, device=dev))
        ms3, cnt3 = (C.c_float * NC)(), (C.c_int32 * NC)()
        acc_ms, acc_n, reps = [0.0] * NC, [0] * NC, 0
        for rep in range(2 + min(args.steps, 5)):
            flush.fill_(1)
            rc = lib.mmb_icp_forward_profiled(p, _capi.ptr(d_src), _capi.ptr(d_tgt), _capi.ptr(d_T0), _capi.ptr(w_det),
                                              _capi.ptr(bufs["T"]), _capi.ptr(bufs["pc"]), _capi.ptr(bufs["de"]),
                                              _capi.ptr(bufs["wh"]), _capi.ptr(bufs["co"]), _capi.ptr(bufs["cv"]),
                                              _capi.ptr(bufs["it"]), _capi.ptr(bufs["mr"]), _capi.ptr(bufs["idn"]),
                                              _capi.ptr(bufs["sv"]), _capi.ptr(bufs["ws"]), ws_bytes, _capi.stream_ptr(),
                                              ms3, cnt3)
            _capi.check(rc, "mmb_icp_forward_profiled")
            if rep >= 2:
                reps += 1
                for c in range(NC):
                    acc_ms[c] += ms3[c]; acc_n[c] += cnt3[c]
        reps = max(1, reps)
        # the fused row-major kernel runs all K iterations in ONE launch (work queue), the other backends one
        # search launch per iteration
        return dict(nn_ms=acc_ms[0] / max(acc_n[0], 1), nn_launches=max(1, acc_n[0] // reps),
                    step_ms=acc_ms[1] / max(acc_n[1], 1),
                    fwd_ms=sum(acc_ms) / reps, nn_share=acc_ms[0] / max(sum(acc_ms), 1e-9),
                    bound_ms=acc_ms[3] / reps, build_ms=acc_ms[4] / reps, other_ms=acc_ms[2] / reps)

    backend_id = NN_BACKENDS[args.nn_backend]
    mode = "brute"
    if backend_id == 2:
        mode = "grid"
    elif backend_id == 3 or (backend_id == 0 and m >= 8192):
        mode = "rows"
    use_grid = mode != "brute"
    prof = profile_forward(backend_id)
    prof_brute = profile_forward(1) if use_grid else prof       # the exhaustive kernel, same data, same run
    nn_ms, step_ms, fwd_ms = prof["nn_ms"], prof["step_ms"], prof["fwd_ms"]

    # FP32 pipe rate measured here (no CUDA-core FP32 figure in MEASURED_PEAKS.json)
    probe_out = torch.zeros(4, **f32)
    peak = {}
    for packed in (0, 1):
        blocks, iters = 148 * 8, 1 << 15
        best = 1e30
        for _ in range(4):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            _capi.check(lib.mmb_fp32_probe(blocks, iters, packed, _capi.ptr(probe_out), _capi.stream_ptr()), "probe")
            b.record(); torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b))
        peak["ffma2" if packed else "ffma"] = blocks * 256 * iters * 8 * 2 * (2 if packed else 1) / (best * 1e-3) / 1e12
    peak_tf = max(peak.values())
    nn_flop = FLOP_PER_PAIR * N * n * m
    brute_tf = nn_flop / (prof_brute["nn_ms"] * 1e-3) / 1e12

    peaks_file = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm_peak = json.load(open(peaks_file)).get("hbm_gbs") if os.path.exists(peaks_file) else None

    # ---- CPU baseline on this box's host cores (bounded sample) -------------------------------
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        pairs = max(1, min(N, args.ref_pairs))
        cpu_reference_step(batch, wl, 1)                     # warm-up (thread pool, allocator)
        secs = cpu_reference_step(batch, wl, pairs)
        kind = reference_kind()
        cpu = {"value": pairs / secs, "unit": UNIT, "cores": cores, "kind": kind,
               "sample": reference_sample_text(kind, pairs, K) + f", one step of {secs:.1f} s"}

    peak_src = ("mmb_fp32_probe in this run (max of FFMA and FFMA2 chains); "
                "MEASURED_PEAKS.json has no CUDA-core FP32 entry")
    tr_b, tr_b_info = traffic_lookup("nn_search_kernel", args.workload)
    roofline_brute = {"kernel": "nn_search_kernel (--nn-backend brute)", "bound": "fp32", "achieved": brute_tf,
                      "peak": peak_tf, "unit": "TFLOP/s", "frac": brute_tf / peak_tf, "traffic": tr_b, **tr_b_info,
                      "flop_per_launch": nn_flop, "ms_per_launch": prof_brute["nn_ms"], "peak_source": peak_src,
                      "peak_detail": peak, "share_of_forward": prof_brute["nn_share"], "forward_ms": prof_brute["fwd_ms"],
                      "note": "frac counts the 8 flop/pair of the reference's distance; the expansion form executes 4 (dim 2), "
                              "so the FMA pipe itself is ~half as busy as frac says"}
    if use_grid:
        # cell-grid search: HBM/L2-gather bound.  Algorithmic bytes per launch, per pair (DESIGN.md sections 3b, 3c):
        #   grid (search only): n x (12 B source + 4 B previous index + 12 B result) + m x 16 B sorted copy + the start table
        #   rows (search + step fused): n x (12 source + 4 previous index + 4 index + 4 index history + 24 matched row
        #                                    + 4 prior weight + 4 weight history) + m x 16 B + the start table
        #   a launch that runs several iterations (the persistent form of the fused kernel: all K) does that many times
        #   the per-iteration bytes -- every iteration gathers from the sorted copy and the start table again
        per_query = 56 if mode == "rows" else 28
        kname = "icp_rows_kernel" if mode == "rows" else "nn_grid_kernel"
        its_per_launch = max(1, K // prof["nn_launches"])
        grid_bytes = its_per_launch * N * (n * per_query + m * 16 + (65536 + 264) * 4)
        hbm = hbm_peak or 6455.9
        gbs = grid_bytes / (nn_ms * 1e-3) / 1e9
        tr, tr_info = traffic_lookup(kname, args.workload)
        roofline = {"kernel": kname, "bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s",
                    "frac": gbs / hbm, "traffic": tr, **tr_info,
                    "bytes_per_launch": grid_bytes, "ms_per_launch": nn_ms, "iterations_per_launch": its_per_launch,
                    "ms_per_iteration": nn_ms / its_per_launch,
                    "peak_source": "MEASURED_PEAKS.json hbm_gbs" if hbm_peak else "fallback 6455.9 GB/s",
                    "work_avoiding": True,
                    "exhaustive_equivalent_tflops": its_per_launch * nn_flop / (nn_ms * 1e-3) / 1e12,
                    "note": "exact cell-grid search: same indices as the exhaustive kernel with ~1e-3 of its distance "
                            "evaluations; latency/issue-bound gathers from L2, not at the HBM roofline; "
                            "roofline_exhaustive_nn is the compute-bound kernel on the same data in the same run"
                            + ("; this kernel also evaluates the residual rows, the normal equations and the solve "
                               "(no separate step kernel)" if mode == "rows" else ""),
                    "share_of_forward": prof["nn_share"], "forward_ms": fwd_ms,
                    "step_kernel_ms_per_launch": step_ms if mode == "grid" else None, "grid_build_ms": prof["build_ms"],
                    "nn_backend": mode}
    else:
        roofline = dict(roofline_brute, kernel="nn_search_kernel", step_kernel_ms_per_launch=step_ms,
                        bound_pass_ms=prof["bound_ms"], nn_backend=args.nn_backend)

    # ours only: brute = init, pack, sub-pack, bound pass; grid = init, geometry, count, scan, scatter
    h2d = sum(host[k].numel() * 4 for k in host)
    d2h = out_T.numel() * 4 + out_gw.numel() * 4
    # `value`: the resident step -- the public call, the loss and its backward -- replayed from the CUDA graph it was
    # captured into once (every kernel runs again on every replay; same results as the eager call bit for bit,
    # checked above), when the capture succeeded on every rank; else the eager loop.  The eager numbers stay in
    # `eager`: what differs is the host's share of issuing a step, which with 8 ranks on one host makes one rank of
    # the eight fall behind now and then.
    use_graph = bool(not args.no_graph and graphed and graphed.get("same_as_eager") and graphed.get("ms_per_step"))
    ms_head = graphed["ms_per_step"] if use_graph else ms_res
    per_rank_head = graphed["per_rank_ms_per_step"] if use_graph else per_rank_res
    line = {
        "metric": METRIC, "value": world * N / (ms_head * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_head, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload + ": " + wl["desc"], "pairs_per_gpu": N, "n": n, "m": m, "iterations": K,
                   "dim": wl["dim"], "icp_type": "pt2pl", "loss": "cauchy(1.0)", "trim_dist": 5.0, "const_iter": True,
                   "l2": "flushed between steps (256 MiB fill, untimed); per-step CUDA events",
                   "launch": ("resident step captured once with torch.cuda.graph (ICP.icp + loss + backward), one replay "
                              "per step; the eager loop is in `eager`") if use_graph else "eager calls",
                   "parallelism": f"independent pairs, {world} rank(s), no data-path collective"},
        "eager": {"value": world * N / (ms_res * 1e-3), "unit": UNIT, "ms_per_step": ms_res,
                  "per_rank_ms_per_step": per_rank_res},
        "e2e": {"value": world * N / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": h2d_per_step, "d2h_bytes_per_step": d2h,
                "serial_upload_value": world * N / (ms_e2e_serial * 1e-3),
                "repeats_ms_per_step": e2e_ms, "host_issue_ms_per_step": e2e_host_ms,
                "overlap": "H2D of step k+1 on a copy stream under the kernels of step k (DevicePrefetcher); "
                           "K uploads inside the timed region, L2 flush time included"},
        # kernels of this library launched inside the two timed regions behind `value` and `e2e`, COUNTED by the
        # library itself (mmb_launch_count differenced around the regions; warm-up steps excluded pro rata)
        "gpu_launches": int(launches_resident * args.steps // (args.steps + args.warmup) + launches_e2e),
        "gpu_launches_per_step": launches_resident // (args.steps + args.warmup),
        "graphed": graphed,
        "roofline": roofline,
        "roofline_exhaustive_nn": roofline_brute,
        "cpu_baseline": cpu,
        "parity": parity,
        "per_rank_ms_per_step": {"resident": per_rank_head, "e2e_last_repeat": e2e_rank,
                                 "min": min(per_rank_head), "median": sorted(per_rank_head)[len(per_rank_head) // 2],
                                 "max": max(per_rank_head)},
        "host_binding": numa,
        "also": also,
        "clocks": clocks,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    if not parity["ok"]:
        print("bench.py: PARITY FAILED " + json.dumps(parity), file=sys.stderr, flush=True)
        sys.exit(3)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--ref-pairs", type=int, default=0,
                    help="pairs per CPU-reference step (bounded sample); 0 = 8 at c3 (BASELINE.md section 3), 1 at c5")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="`value` from the eager loop even when the CUDA-graph capture succeeded")
    ap.add_argument("--c5-global-pairs", type=int, default=256)
    ap.add_argument("--also", type=int, default=1,
                    help="1: also measure c5 (strong, 256 pairs global), c4 (train steps/s) and c2 (radar scans/s) in this "
                         "process at this N and report them under \"also\"")
    ap.add_argument("--e2e-repeats", type=int, default=3, help="repeats of the K-step end-to-end loop (median reported)")
    ap.add_argument("--nn-backend", default="auto", choices=sorted(NN_BACKENDS),
                    help="nearest-neighbour backend: exact cell grid (auto for m >= 8192) or brute force; same indices")
    ap.add_argument("--nn-slices", type=int, default=0, help="tuning: target slices per query tile (0 = library default)")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    if args.ref_pairs <= 0:
        args.ref_pairs = 1 if wl["m"] > 100000 else min(8, wl["N"])
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, wl, rank, world)
    else:
        run_ours(args, wl, rank, world, local_rank)


if __name__ == "__main__":
    main()
