"""Seeded synthetic Boreas-shaped data.

No dataset ships with this repo (the reference's loader needs Boreas sequences and
ROS2 bags, `mm_masking/icp_weight_dataset.py:29`), so tests and `bench.py` draw
scan/map pairs and Navtech-style polar scans from a small 2-D world of wall
segments. Shapes, padding conventions and perturbation ranges follow the reference:

* scans are zero-padded to n rows       (`icp_weight_dataset.py:379-381`)
* maps are padded with 1000-valued rows (`icp_weight_dataset.py:395-398`)
* T_init is the ground truth perturbed by x,y ~ U(-pos,pos), yaw ~ U(-rot,rot)
  (`icp_weight_dataset.py:261-267`)

Everything is numpy on the host and deterministic in `seed`.
"""
from __future__ import annotations

import numpy as np

MAP_PAD_VALUE = 1000.0          # `external/dICP/config/dICP_config.yaml:4`
RADAR_RES = 0.0596              # `mm_masking/icp_weight_policy.py:29`
CART_RES = 0.2384               # `mm_masking/radar_utils.py:258`
CART_WIDTH = 640
ENCODER_TICKS = 5600            # `mm_masking/radar_utils.py:22`


def make_world(rng: np.random.Generator, n_walls: int = 60, n_boxes: int = 20,
               half: float = 80.0) -> np.ndarray:
    """Wall segments (S,4) = x0,y0,x1,y1 inside a +-half metre square."""
    segs = []
    for _ in range(n_walls):
        c = rng.uniform(-half, half, 2)
        ang = rng.uniform(0, np.pi)
        length = rng.uniform(5.0, 40.0)
        d = 0.5 * length * np.array([np.cos(ang), np.sin(ang)])
        segs.append(np.concatenate([c - d, c + d]))
    for _ in range(n_boxes):
        c = rng.uniform(-half, half, 2)
        w, h = rng.uniform(2.0, 10.0, 2)
        ang = rng.uniform(0, np.pi)
        R = np.array([[np.cos(ang), -np.sin(ang)], [np.sin(ang), np.cos(ang)]])
        corners = (R @ (0.5 * np.array([[-w, -h], [w, -h], [w, h], [-w, h]]).T)).T + c
        for k in range(4):
            segs.append(np.concatenate([corners[k], corners[(k + 1) % 4]]))
    segs = np.clip(np.asarray(segs), -half, half)
    return segs


def _sample_on_segments(rng, segs, count):
    length = np.hypot(segs[:, 2] - segs[:, 0], segs[:, 3] - segs[:, 1]) + 1e-9
    which = rng.choice(len(segs), size=count, p=length / length.sum())
    t = rng.uniform(0, 1, count)
    p0, p1 = segs[which, 0:2], segs[which, 2:4]
    pts = p0 + t[:, None] * (p1 - p0)
    tangent = (p1 - p0) / length[which, None]
    normal = np.stack([-tangent[:, 1], tangent[:, 0]], axis=1)
    return pts, normal


def sample_map(rng, segs, m: int, m_real: int | None = None, sigma: float = 0.02,
               dup_frac: float = 0.01, three_d: bool = False) -> np.ndarray:
    """Target cloud (m,6) float32: xyz + unit normal; rows >= m_real are pad rows."""
    m_real = m if m_real is None else m_real
    pts, nrm = _sample_on_segments(rng, segs, m_real)
    pts = pts + sigma * rng.standard_normal(pts.shape) * np.abs(nrm)
    out = np.full((m, 6), MAP_PAD_VALUE, dtype=np.float32)
    out[:m_real, 0:2] = pts
    out[:m_real, 2] = rng.uniform(0.0, 3.0, m_real) if three_d else 0.0
    out[:m_real, 3:5] = nrm
    out[:m_real, 5] = 0.0
    if three_d:                       # a ground patch with normal +z
        g = rng.choice(m_real, size=m_real // 5, replace=False)
        out[g, 0:2] = rng.uniform(-60, 60, (len(g), 2))
        out[g, 2] = 0.02 * rng.standard_normal(len(g))
        out[g, 3:6] = np.array([0.0, 0.0, 1.0], dtype=np.float32)
    n_dup = int(dup_frac * m_real)
    if n_dup > 0:                     # exact duplicates exercise the lowest-index tie-break
        src = rng.choice(m_real, size=n_dup, replace=False)
        dst = rng.choice(m_real, size=n_dup, replace=False)
        out[dst] = out[src]
    return out


def se2_matrix(x: float, y: float, yaw: float, z: float = 0.0) -> np.ndarray:
    c, s = np.cos(yaw), np.sin(yaw)
    T = np.eye(4)
    T[0, 0], T[0, 1], T[1, 0], T[1, 1] = c, -s, s, c
    T[0, 3], T[1, 3], T[2, 3] = x, y, z
    return T


def sample_scan(rng, segs, T_ts: np.ndarray, n: int, n_wall: int, n_clutter: int,
                sigma: float = 0.10, max_range: float = 76.0, three_d: bool = False) -> np.ndarray:
    """Source cloud (n,3) float32 in the sensor frame, zero rows after the real points."""
    T_st = np.linalg.inv(T_ts)
    pts, _ = _sample_on_segments(rng, segs, 4 * n_wall)
    sensor = T_ts[0:2, 3]
    keep = np.hypot(pts[:, 0] - sensor[0], pts[:, 1] - sensor[1]) < max_range
    pts = pts[keep][:n_wall]
    pts = pts + sigma * rng.standard_normal(pts.shape)
    clutter = sensor + rng.uniform(-max_range, max_range, (n_clutter, 2))
    xy = np.concatenate([pts, clutter], axis=0)
    z = rng.uniform(0.0, 3.0, len(xy)) if three_d else np.zeros(len(xy))
    p_t = np.concatenate([xy, z[:, None]], axis=1)
    p_s = (T_st[:3, :3] @ p_t.T).T + T_st[:3, 3]
    if not three_d:
        p_s[:, 2] = 0.0
    out = np.zeros((n, 3), dtype=np.float32)
    k = min(n, len(p_s))
    out[:k] = p_s[:k]
    return out


def make_weight_mask(rng, width: int = CART_WIDTH) -> np.ndarray:
    """sigmoid(low-pass noise), max-normalised like `icp_weight_policy.py:192-193`."""
    coarse = rng.standard_normal((width // 16 + 2, width // 16 + 2))
    yy = np.linspace(0, coarse.shape[0] - 1.001, width)
    y0 = np.floor(yy).astype(int)
    fy = yy - y0
    rows = coarse[y0] * (1 - fy)[:, None] + coarse[y0 + 1] * fy[:, None]
    img = rows[:, y0] * (1 - fy)[None, :] + rows[:, y0 + 1] * fy[None, :]
    mask = 1.0 / (1.0 + np.exp(-2.0 * img))
    return (mask / mask.max()).astype(np.float32)


def bilinear_weights(mask: np.ndarray, pts: np.ndarray) -> np.ndarray:
    """Host-side bilinear sample used only to give synthetic scans plausible prior weights."""
    W = mask.shape[0]
    u = -pts[:, 0] / CART_RES + (W - 1) / 2.0      # row
    v = pts[:, 1] / CART_RES + (W - 1) / 2.0       # column
    u0, v0 = np.floor(u).astype(int), np.floor(v).astype(int)
    fu, fv = u - u0, v - v0
    out = np.zeros(len(pts))
    for du, dv, w in ((0, 0, (1 - fu) * (1 - fv)), (0, 1, (1 - fu) * fv),
                      (1, 0, fu * (1 - fv)), (1, 1, fu * fv)):
        r, c = u0 + du, v0 + dv
        ok = (r >= 0) & (r < W) & (c >= 0) & (c < W)
        out[ok] += w[ok] * mask[r[ok], c[ok]]
    out[(pts[:, 0] == 0) & (pts[:, 1] == 0)] = 0.0
    return out.astype(np.float32)


def make_icp_batch(N: int, n: int, m: int, seed: int = 0, *, three_d: bool = False,
                   pos: float = 2.0, rot: float = 0.6, real_frac: float = 0.95,
                   clutter_frac: float = 0.3, weights: str = "mask", scan_range: float = 76.0) -> dict:
    """One batch of scan/map pairs.

    Returns float32 arrays: source (N,n,3), target (N,m,6), T_init (N,4,4), T_gt (N,4,4),
    weight (N,n). `pos`/`rot` are the half-widths of the initial perturbation
    (reference training values 2 m / 0.6 rad: `train_icp_weights.py:364-365`).  `scan_range` is the radius the
    scan covers; the map always covers +-80 m, so a small value gives a map much larger than the scan footprint.
    """
    rng = np.random.default_rng(seed)
    source = np.zeros((N, n, 3), np.float32)
    target = np.zeros((N, m, 6), np.float32)
    T_init = np.zeros((N, 4, 4), np.float32)
    T_gt = np.zeros((N, 4, 4), np.float32)
    weight = np.zeros((N, n), np.float32)
    for b in range(N):
        segs = make_world(rng)
        m_real = m if real_frac >= 1.0 else int(m * rng.uniform(real_frac, 1.0))
        target[b] = sample_map(rng, segs, m, m_real, three_d=three_d)
        gt = se2_matrix(rng.uniform(-3, 3), rng.uniform(-3, 3), rng.uniform(-0.5, 0.5))
        n_real = int(n * rng.uniform(0.9, 1.0))
        n_clutter = int(clutter_frac * n_real)
        source[b] = sample_scan(rng, segs, gt, n, n_real - n_clutter, n_clutter, three_d=three_d, max_range=scan_range)
        pert = se2_matrix(rng.uniform(-pos, pos), rng.uniform(-pos, pos), rng.uniform(-rot, rot))
        T_gt[b] = gt
        T_init[b] = pert @ gt
        if weights == "mask":
            weight[b] = bilinear_weights(make_weight_mask(rng), source[b])
        elif weights == "ones":
            weight[b] = ((source[b, :, 0] != 0) | (source[b, :, 1] != 0)).astype(np.float32)
        else:
            raise ValueError(weights)
    return dict(source=source, target=target, T_init=T_init, T_gt=T_gt, weight=weight)


def make_radar_scans(B: int, seed: int = 0, n_az: int = 400, n_bins: int = 3360) -> dict:
    """Navtech-style scans.

    raw (B,n_az,n_bins+11) uint8 with the byte layout `radar_utils.py:20-27` decodes
    (8 B timestamp, 2 B encoder, 1 B unused, then range bins); plus the decoded
    float32 fft (B,n_az,n_bins), float32 azimuths (B,n_az) in radians and int64 timestamps.
    Encoder ticks are non-uniform (exercises the wobble fix, `radar_utils.py:295-309`).
    """
    rng = np.random.default_rng(seed)
    raw = np.zeros((B, n_az, n_bins + 11), np.uint8)
    step = ENCODER_TICKS / n_az
    for b in range(B):
        ticks = np.round(np.arange(n_az) * step + rng.integers(-1, 2, n_az)).astype(np.int64)
        ticks = np.clip(ticks, 0, ENCODER_TICKS - 1)
        ticks = np.maximum.accumulate(ticks)
        ticks = ticks + np.arange(n_az) * 0            # keep dtype
        for k in range(1, n_az):                       # strictly increasing
            if ticks[k] <= ticks[k - 1]:
                ticks[k] = ticks[k - 1] + 1
        stamps = (1_600_000_000_000_000 + b * 250_000 + np.arange(n_az) * 625).astype(np.int64)
        power = rng.exponential(0.04, (n_az, n_bins))
        n_ret = rng.integers(2, 9, n_az)
        for a in range(n_az):
            starts = rng.integers(40, min(n_bins - 12, 1300), n_ret[a])
            for s in starts:
                wdt = rng.integers(3, 11)
                power[a, s:s + wdt] += rng.uniform(0.3, 1.0)
        img = np.clip(power, 0.0, 1.0)
        raw[b, :, 11:] = np.round(img * 255).astype(np.uint8)
        raw[b, :, 0:8] = stamps.view(np.uint8).reshape(n_az, 8)
        raw[b, :, 8:10] = ticks.astype(np.uint16).view(np.uint8).reshape(n_az, 2)
    fft = raw[:, :, 11:].astype(np.float32) / np.float32(255.0)
    enc = raw[:, :, 8:10].copy().view(np.uint16).reshape(B, n_az)
    azimuths = (enc.astype(np.float64) * (2 * np.pi / ENCODER_TICKS)).astype(np.float32)
    stamps = raw[:, :, 0:8].copy().view(np.int64).reshape(B, n_az) * 1000
    return dict(raw=raw, fft=fft, azimuths=azimuths, timestamps=stamps)


def make_training_batch(N: int, n: int = 2000, m: int = 50000, seed: int = 0, width: int = CART_WIDTH) -> dict:
    """One training batch in the reference's layout (`icp_weight_dataset.py:357-362`):

        {'loc_data': {'fft_data' (N,W,W), 'fft_cfar' (N,W,W), 'raw_pc' (N,n,3), 'filtered_pc' (N,n,3)},
         'map_data': {'pc' (N,m,6)}, 'transforms': {'T_ml_init' (N,4,4), 'T_ml_gt' (N,4,4)}}

    as float32 torch tensors on the host.  Points, map and poses come from `make_icp_batch`; the
    Cartesian radar image is the scan rasterised over an exponential noise floor and the CFAR image
    its thresholded version (the reference stores both pre-resampled, `icp_weight_dataset.py:83`)."""
    import torch
    b = make_icp_batch(N, n, m, seed=seed)
    rng = np.random.default_rng(seed + 1000)
    fft = rng.exponential(0.04, (N, width, width)).astype(np.float32)
    cfar = np.zeros((N, width, width), np.float32)
    for k in range(N):
        p = b["source"][k]
        real = (p[:, 0] != 0) | (p[:, 1] != 0)
        u = np.clip(np.round(-p[real, 0] / CART_RES + (width - 1) / 2).astype(int), 0, width - 1)
        v = np.clip(np.round(p[real, 1] / CART_RES + (width - 1) / 2).astype(int), 0, width - 1)
        fft[k, u, v] += rng.uniform(0.3, 1.0, len(u)).astype(np.float32)
        cfar[k, u, v] = 1.0
    fft = np.clip(fft, 0.0, 1.0)
    t = torch.from_numpy
    return {"loc_data": {"fft_data": t(fft), "fft_cfar": t(cfar), "raw_pc": t(b["source"].copy()),
                         "filtered_pc": t(b["source"].copy())},
            "map_data": {"pc": t(b["target"])},
            "transforms": {"T_ml_init": t(b["T_init"]), "T_ml_gt": t(b["T_gt"])}}


def shard_training_batch(batch: dict, world_size: int, rank: int) -> dict:
    """this rank's contiguous share of every tensor of a `make_training_batch` dict"""
    from .dist import shard_range
    n = batch["transforms"]["T_ml_init"].shape[0]
    s, e = shard_range(n, world_size, rank)
    return {k: {kk: vv[s:e] for kk, vv in v.items()} for k, v in batch.items()}
