#!/usr/bin/env python
"""Short driver for ncu: every radar preprocessing op once (after one warm-up call) on a batch of raw Navtech scans.

    python profiles/prof_radar.py --batch 64
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import numpy as np
    import torch
    from mm_masking_b200 import synth, radar_utils as ru
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    args = ap.parse_args()
    B = args.batch
    base = synth.make_radar_scans(B=min(B, 8), seed=0)
    reps = (B + 7) // 8
    raw = torch.from_numpy(np.concatenate([base["raw"]] * reps)[:B]).cuda()
    az = torch.from_numpy(np.concatenate([base["azimuths"]] * reps)[:B]).cuda()
    for _ in range(2):
        ru.cfar_mask(raw, 0.0596, diff=False)
        ru.radar_polar_to_cartesian_diff(raw, az, 0.0596)
        ru.cfar_extract_pc(raw, 0.0596, az)
        ru.kstrong_extract_pc(raw, 0.0596, az)
    torch.cuda.synchronize()
    print("ok")


if __name__ == "__main__":
    main()
