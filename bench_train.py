#!/usr/bin/env python
"""Training-step benchmark (BASELINE.json configs[3], "train steps/sec"): the full ICP-weight-policy
step of `mm_masking/train_icp_weights.py:31-63` -- UNet mask -> per-point weight sampling -> dICP
(10 it, pt2pl, dim 2) -> pose loss + occupancy (mask_pts) loss -> backward -> Adam step -- on a GLOBAL
batch of 64 pairs sharded over the ranks (strong scaling), gradients averaged with one NCCL all-reduce.

    python bench_train.py [--gpus N] [--steps K] [--warmup W] [--batch 64]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 ... bench_train.py --gpus N

  value   steps/s with the rank's shard resident in HBM (CUDA events per step, max over ranks)
  e2e     steps/s with the shard in pinned host memory: H2D of images, points, map and poses
          inside the timed region, D2H of the loss
The UNet is PyTorch/cuDNN (it is not the product); `breakdown_ms` gives its share next to the CUDA
path of this package (weight sampling, ICP forward/backward, occupancy BEV).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def parse_args(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=64, help="GLOBAL batch (pairs per step over all ranks)")
    ap.add_argument("--n", type=int, default=2000)
    ap.add_argument("--m", type=int, default=50000)
    ap.add_argument("--channels-last", type=int, default=1, help="UNet weights/activations in channels_last (fp32 kept)")
    return ap.parse_args(argv)


def run(args, rank, world, local_rank, own_process_group=True, sample_clocks=True):
    """measure the c4 training step; returns the JSON line (a dict) on rank 0, None elsewhere.  With
    own_process_group=False the caller (bench.py --also) has initialised torch.distributed and keeps it."""
    import torch
    import torch.distributed as dist
    from mm_masking_b200 import _capi, synth, dist as mdist
    from mm_masking_b200.icp_weight_policy import LearnICPWeightPolicy
    from mm_masking_b200 import train_icp_weights as TW
    from bench import ClockSampler
    from mm_masking_b200.prefetch import DevicePrefetcher

    _capi.require_cuda(); _capi.load_library()
    torch.cuda.set_device(local_rank)
    torch.backends.cudnn.benchmark = True            # UNet convolutions (PyTorch/cuDNN, not the product)
    dev = torch.device("cuda", local_rank)
    if world > 1 and own_process_group:
        import datetime
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=90))

    s, e = mdist.shard_range(args.batch, world, rank)
    # every rank generates only its own pairs (seeded by the global pair index range)
    host = synth.make_training_batch(e - s, args.n, args.m, seed=100 + s)
    pin = lambda d: {k: {kk: vv.pin_memory() for kk, vv in v.items()} for k, v in d.items()}
    host = pin(host)
    resident = {k: {kk: vv.to(dev) for kk, vv in v.items()} for k, v in host.items()}

    torch.manual_seed(0)                                    # identical initial weights on every rank
    model = LearnICPWeightPolicy({"device": dev, "max_iter": 10}).to(dev)
    if args.channels_last:
        model = model.to(memory_format=torch.channels_last)     # cuDNN NHWC kernels; arithmetic stays fp32
    model.ICP_alg.const_iter = True
    model.train()
    opt = torch.optim.Adam(model.parameters(), lr=1e-4)
    lw = dict(TW.DEFAULT_LOSS_WEIGHTS)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    loss_host = torch.zeros(1).pin_memory()

    def step_resident():
        loss, _ = TW.train_step(model, resident, opt, lw, device=dev, gt_eye=False)
        return loss

    # e2e: the loop a user writes - train_step over DevicePrefetcher(iterator) - so the upload of the
    # next batch runs on a copy stream under this step's kernels; each step uploads its own batch
    e2e_warm = max(1, args.warmup // 2)
    pre = DevicePrefetcher((host for _ in range(2 * (e2e_warm + args.steps) + 1)), dev)

    def step_e2e():
        loss, _ = TW.train_step(model, next(pre), opt, lw, device=dev, gt_eye=False)
        loss_host.copy_(loss.reshape(1), non_blocking=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, whole=False):
        """per-step events (L2 flush between steps untimed), or with `whole` one event pair around all
        the steps, flushes included (the prefetching loop: nothing between steps escapes the clock)"""
        for _ in range(warmup):
            fn()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        barrier()
        t0 = time.time()
        if whole:
            ev[0][0].record()
        for a, b in ev:
            flush.fill_(1)
            if not whole:
                a.record()
            fn()
            if not whole:
                b.record()
        if whole:
            ev[0][1].record()
        barrier()
        t1 = time.time()
        total = ev[0][0].elapsed_time(ev[0][1]) if whole else sum(a.elapsed_time(b) for a, b in ev)
        return mdist.max_over_ranks(total, device=dev) / steps, t0, t1

    sampler = ClockSampler(local_rank) if rank == 0 and sample_clocks else None
    if sampler:
        sampler.start(); time.sleep(0.3)
    ms_res, t0, t1 = timed(step_resident, args.steps, args.warmup)
    ms_e2e, _, t1 = timed(step_e2e, args.steps, e2e_warm, whole=True)
    clocks = None
    if sampler:
        sampler.stop(); clocks = sampler.summary(t0, t1)

    # ---- breakdown on rank 0 (one untimed extra step, events around the stages) -------------------
    breakdown = None
    if world == 1:                      # the forward holds a collective: never run it on one rank only
        def ev():
            x = torch.cuda.Event(enable_timing=True); x.record(); return x
        bs, bm, bt = resident["loc_data"], resident["map_data"], resident["transforms"]
        opt.zero_grad()
        e0 = ev()
        mask = model(bs, bm, bt["T_ml_init"], mask_only=True)
        e1 = ev()
        from mm_masking_b200.radar_utils import extract_weights, extract_bev_from_pts
        w = extract_weights(mask, bs["raw_pc"])[0]
        e2 = ev()
        T = model.icp(bs["filtered_pc"], bm["pc"], bt["T_ml_init"], w)
        e3 = ev()
        bev = extract_bev_from_pts(bm["pc"])
        e4 = ev()
        loss = (T - torch.eye(4, device=dev)).abs().mean() + torch.nn.functional.binary_cross_entropy(mask, bev)
        loss.backward()
        e5 = ev()
        torch.cuda.synchronize()
        breakdown = {"unet_forward": e0.elapsed_time(e1), "weight_sampling": e1.elapsed_time(e2),
                     "icp_forward": e2.elapsed_time(e3), "occupancy_bev": e3.elapsed_time(e4),
                     "loss_and_backward_all": e4.elapsed_time(e5)}
        opt.zero_grad()

    line = None
    if rank == 0:
        h2d = sum(vv.numel() * vv.element_size() for v in host.values() for vv in v.values())
        n_params = sum(p.numel() for p in model.parameters())
        line = {
            "metric": "train_steps_per_sec", "value": 1e3 / ms_res, "unit": "steps/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_res, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "pairs_per_sec": args.batch * 1e3 / ms_res,
            "config": {"workload": "c4: ICP-weight-policy training step (UNet mask -> weight sampling -> dICP 10 it -> "
                                   "pose + occupancy loss -> backward -> Adam), global batch sharded over the ranks",
                       "global_batch": args.batch, "pairs_per_gpu": e - s, "n": args.n, "m": args.m, "iterations": 10,
                       "image": "640x640", "unet_params": n_params, "unet_memory_format": "channels_last" if args.channels_last else "contiguous",
                       "l2": "flushed between steps (256 MiB fill, untimed); per-step CUDA events",
                       "parallelism": f"dp{world}: pairs sharded, one flat NCCL all-reduce of the UNet gradients, "
                                      "MIN/MAX all-reduce of the input normalisation"},
            "e2e": {"value": 1e3 / ms_e2e, "unit": "steps/s", "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4},
            "breakdown_ms": breakdown, "clocks": clocks,
        }
    if world > 1 and own_process_group:
        dist.destroy_process_group()
    return line


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    line = run(args, rank, world, local_rank)
    if line is not None:
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
