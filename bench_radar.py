#!/usr/bin/env python
"""Radar preprocessing benchmark (BASELINE.json configs[1]): polar 400 x 3360 scans -> 640 x 640
Cartesian BEV + CFAR / K-strongest extraction (+ per-point weight sampling), batch 8 on one B200.

    python bench_radar.py [--batch 8] [--reps 20] [--no-cpu-baseline]

Prints ONE JSON line: scans/s of the whole preprocessing chain (polar->Cartesian of the scan and of
its CFAR image as `icp_weight_dataset.py:351-352` does, plus fused CFAR + point extraction), and
per-stage device times with the HBM roofline fraction of each stage computed from the algorithmic
bytes of SURVEY.md section 8d.  Every stage is timed alone with CUDA events, L2 flushed (256 MiB
fill) before every repetition, median over the repetitions.  H8: at batch 8 the stages
move 20-60 MB, a few microseconds of HBM time, so they are launch/latency bound; `--batch 64` shows
the bandwidth regime.
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

RES = 0.0596
A, R, W = 400, 3360, 640


def algorithmic_bytes(B, n_points_total, n_k_points):
    """bytes per launch; u8 variants read 1 byte per bin"""
    touched = 1810                                   # bins within 107.8 m = 320 * 0.2384 * sqrt(2)
    return {
        "polar_to_cart_f32": B * (A * touched * 4 + A * 4 + W * W * 4),
        "polar_to_cart_u8": B * (A * touched * 1 + A * 4 + W * W * 4),
        "cfar_mask_f32": B * (A * 1308 * 4 + A * R * 4),
        "cfar_mask_u8": B * (A * 1308 * 1 + A * R * 4),
        "cfar_extract_fused_u8": B * (A * 1308 * 1) + n_points_total * 12,
        "cfar_extract_fused_f32": B * (A * 1308 * 4) + n_points_total * 12,
        "extract_pc_from_mask": B * (A * R * 4) + n_points_total * 12,
        "kstrong_u8": B * (A * 1310 * 1) + n_k_points * 12,
        "cfar_extract_op_u8": B * (A * 1308 * 1) + n_points_total * 12,
        "kstrong_op_u8": B * (A * 1310 * 1) + n_k_points * 12,
        "sample_weights_fwd": B * 2000 * 28,
        "sample_weights_bwd": B * 2000 * 28,
    }


def parse_args(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args(argv)


def run(args, local_rank=0):
    """measure the c2 preprocessing chain on cuda:local_rank; returns the JSON line as a dict"""
    import numpy as np
    import torch
    from mm_masking_b200 import _capi, synth, radar_utils as ru

    _capi.require_cuda()
    _capi.load_library()
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    B = args.batch
    base = synth.make_radar_scans(B=min(B, 8), seed=0)
    reps_of = (B + 7) // 8
    raw_h = torch.from_numpy(np.concatenate([base["raw"]] * reps_of)[:B]).pin_memory()
    az_h = torch.from_numpy(np.concatenate([base["azimuths"]] * reps_of)[:B]).pin_memory()
    raw, az = raw_h.to(dev), az_h.to(dev)
    fft = torch.from_numpy(np.concatenate([base["fft"]] * reps_of)[:B]).to(dev)
    mask = ru.cfar_mask(raw, RES, diff=False)
    pts_list = ru.cfar_extract_pc(raw, RES, az)
    n_pts = sum(len(p) for p in pts_list)
    k_list = ru.kstrong_extract_pc(raw, RES, az)
    n_k = sum(len(p) for p in k_list)
    cap_pts = max(len(p) for p in pts_list) + 16
    wmask = torch.rand((B, W, W), device=dev, requires_grad=True)
    spts = torch.zeros((B, 2000, 3), device=dev)
    spts[:, :, :2] = (torch.rand((B, 2000, 2), device=dev) - 0.5) * 150
    gw = torch.rand((B, 2000), device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def sample_bwd():
        w = torch.ops.mmb.sample_weights(wmask, spts, 0.2384)
        return w

    w_keep = torch.ops.mmb.sample_weights(wmask, spts, 0.2384)
    stages = {
        "polar_to_cart_f32": lambda: ru.radar_polar_to_cartesian_diff(fft, az, RES),
        "polar_to_cart_u8": lambda: ru.radar_polar_to_cartesian_diff(raw, az, RES),
        "cfar_mask_f32": lambda: ru.cfar_mask(fft, RES, diff=False),
        "cfar_mask_u8": lambda: ru.cfar_mask(raw, RES, diff=False),
        "cfar_extract_fused_u8": lambda: ru.cfar_extract_pc(raw, RES, az),
        "cfar_extract_fused_f32": lambda: ru.cfar_extract_pc(fft, RES, az),
        "extract_pc_from_mask": lambda: ru.extract_pc(mask, RES, az, None, diff=False),
        "kstrong_u8": lambda: ru.kstrong_extract_pc(raw, RES, az),
        # the same two extractions as fixed-capacity custom ops (padded output + counts, no host read in the call):
        # the device chain alone, without the ragged-list assembly of the reference-shaped wrappers
        "cfar_extract_op_u8": lambda: torch.ops.mmb.cfar_extract(raw, az, None, cap_pts, RES, 101, 2.0, 80.0, 5, 1.0, 0.09,
                                                                 False, 10.0),
        "kstrong_op_u8": lambda: torch.ops.mmb.kstrong_extract(raw, az, RES, 12, 0.0, 0.216, 2.0, 80.0, -0.31),
        "sample_weights_fwd": lambda: torch.ops.mmb.sample_weights(wmask.detach(), spts, 0.2384),
        "sample_weights_bwd": lambda: torch.autograd.grad(w_keep, wmask, gw, retain_graph=True),
    }

    def timed(fn):
        for _ in range(args.warmup):
            fn()
        ts = []
        for _ in range(args.reps):
            flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        ts.sort()
        return ts[len(ts) // 2]

    peaks_file = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm_peak = json.load(open(peaks_file)).get("hbm_gbs", 6650.0) if os.path.exists(peaks_file) else 6650.0
    peak_src = "MEASURED_PEAKS.json hbm_gbs (measured)" if os.path.exists(peaks_file) else "fallback 6650 GB/s"
    bytes_of = algorithmic_bytes(B, n_pts, n_k)
    out_stages = {}
    for name, fn in stages.items():
        ms = timed(fn)
        gbs = bytes_of[name] / (ms * 1e-3) / 1e9
        out_stages[name] = {"ms": ms, "bytes": bytes_of[name], "GB/s": gbs, "frac_hbm": gbs / hbm_peak}

    # the whole chain as the dataset runs it (icp_weight_dataset.py:351-352 + extraction), host -> device -> host
    cart_h = torch.empty((B, W, W), dtype=torch.float32).pin_memory()
    cart_cfar_h = torch.empty((B, W, W), dtype=torch.float32).pin_memory()

    def chain_e2e():
        r = raw_h.to(dev, non_blocking=True)
        a_ = az_h.to(dev, non_blocking=True)
        cart_h.copy_(ru.radar_polar_to_cartesian_diff(r, a_, RES), non_blocking=True)
        m = ru.cfar_mask(r, RES, diff=False)
        cart_cfar_h.copy_(ru.radar_polar_to_cartesian_diff(m, a_, RES), non_blocking=True)
        pcs = ru.cfar_extract_pc(r, RES, a_)
        return [p.cpu() for p in pcs]

    def chain_dev():
        cart = ru.radar_polar_to_cartesian_diff(raw, az, RES)
        m = ru.cfar_mask(raw, RES, diff=False)
        cart_cfar = ru.radar_polar_to_cartesian_diff(m, az, RES)
        return cart, cart_cfar, ru.cfar_extract_pc(raw, RES, az)

    ms_dev = timed(chain_dev)
    ms_e2e = timed(chain_e2e)

    cpu = None
    if not args.no_cpu_baseline:
        from oracle import radar_oracle as RO
        torch.set_num_threads(os.cpu_count() or 1)
        f1, a1, t1 = base["fft"][:1], base["azimuths"][:1], base["timestamps"][:1]
        t0 = time.perf_counter()
        c = RO.polar_to_cartesian(f1, a1, RES)
        m = RO.cfar_mask(f1, RES, diff=False, precision="f32")
        c2 = RO.polar_to_cartesian(m, a1, RES)
        RO.extract_pc(m, RES, a1, t1, diff=False)
        secs = time.perf_counter() - t0
        cpu = {"value": 1.0 / secs, "unit": "scans/s", "cores": os.cpu_count() or 1, "kind": "port",
               "sample": f"1 scan through the numpy/torch oracle chain (cfar window sums as the reference's torch.sum loop), {secs:.2f} s"}

    dom = max(out_stages, key=lambda k: out_stages[k]["ms"])
    line = {
        "metric": "radar_preprocessing_scans_per_sec", "value": B / (ms_dev * 1e-3), "unit": "scans/s", "n_gpus": 1,
        "steps": args.reps, "warmup": args.warmup, "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"c2: {B} Navtech scans 400x3360 (raw uint8 rows) -> 640x640 BEV of scan and of CFAR mask "
                               "+ fused CFAR/peak extraction", "batch": B, "points_per_scan": n_pts / B,
                   "kstrong_points_per_scan": n_k / B, "l2": "flushed before every repetition (256 MiB fill), median"},
        "e2e": {"value": B / (ms_e2e * 1e-3), "unit": "scans/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": raw_h.numel() + az_h.numel() * 4,
                "d2h_bytes_per_step": 2 * B * W * W * 4 + n_pts * 12},
        "roofline": {"kernel": dom, "bound": "hbm", "achieved": out_stages[dom]["GB/s"], "peak": hbm_peak, "unit": "GB/s",
                     "frac": out_stages[dom]["frac_hbm"], "traffic": None, "peak_source": peak_src},
        "stages": out_stages, "cpu_baseline": cpu,
    }
    return line


def main():
    print(json.dumps(run(parse_args())), flush=True)


if __name__ == "__main__":
    main()
