"""Generate tests/golden/radar_next.npz from the UNMODIFIED reference `mm_masking/radar_utils.py`
(SURVEY.md 8f rows N2, N3 and the adjoint of K1).

    python oracle/make_golden_radar2.py

TEST INFRASTRUCTURE ONLY.  The reference has no tests for these functions, so the vectors are the
outputs (and autograd gradients) of the reference functions themselves on small seeded inputs:

  bev_*      extract_bev_from_pts (radar_utils.py:142-165), float32 and float64 points
  c2p_*      radar_cartesian_to_polar (:338-372) in float64 (the reference casts its grid to double,
             :370, so it cannot run in float32) + autograd gradient with respect to the image
  p2c_*      radar_polar_to_cartesian_diff (:258-336): autograd gradient with respect to the scan
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_loader  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def make_inputs():
    rng = np.random.default_rng(17)
    # ---- points for the occupancy BEV ---------------------------------------------------------
    N, n = 3, 400
    pts = np.zeros((N, n, 3), np.float64)
    pts[:, :, :2] = rng.uniform(-90, 90, (N, n, 2))               # some outside the 76 m image
    pts[:, -30:, :] = 0.0                                          # batch padding -> centre pixel
    pts[0, 0, :2] = (0.2384 * 10, -0.2384 * 20)                    # on pixel coordinates (floor == ceil)
    pts[0, 1, :2] = (76.288, 76.0496)                              # u = 0 exactly / v = 639
    pts[0, 2, :2] = (-76.05, -76.3)                                # u = 639.0.. / v < 0
    pts[1, 0, :2] = (500.0, 3.0)                                   # one coordinate out of range only
    # ---- Cartesian -> polar ---------------------------------------------------------------------
    B, H, W, A, R = 2, 96, 96, 40, 300
    yy = np.linspace(0, 1, H)
    cart = np.stack([0.5 + 0.5 * np.sin(9 * (k + 1) * yy)[:, None] * np.cos(7 * (k + 2) * yy)[None, :] for k in range(B)])
    cart += 0.05 * rng.standard_normal(cart.shape)
    az = np.sort(rng.uniform(0, 2 * np.pi, (B, A)), axis=1)
    G_c2p = rng.standard_normal((B, A, R))
    # ---- polar -> Cartesian adjoint ---------------------------------------------------------------
    A2, R2, width = 40, 256, 64
    fft = rng.uniform(0, 1, (B, A2, R2))
    az2 = (np.arange(A2) * (2 * np.pi / A2))[None, :] + rng.uniform(-0.02, 0.02, (B, A2))
    az2 = np.sort(np.clip(az2, 0, 2 * np.pi - 1e-3), axis=1)
    G_p2c = rng.standard_normal((B, width, width))
    return dict(pts=pts, cart=cart, az=az, G_c2p=G_c2p, fft=fft, az2=az2, G_p2c=G_p2c)


C2P = dict(radar_resolution=0.0596 * 1.5, cart_resolution=0.2384, polar_pixel_shape=(40, 300))
P2C = dict(radar_resolution=0.05, cart_pixel_width=64)


def main():
    _, _, _, ru = ref_loader.load_reference()
    d = make_inputs()
    out = dict(d)
    for dt, tag in ((torch.float32, "f32"), (torch.float64, "f64")):
        bev = ru.extract_bev_from_pts(torch.tensor(d["pts"], dtype=dt)).numpy()
        out[f"bev_{tag}"] = np.packbits(bev != 0, axis=-1)
        assert set(np.unique(bev)) <= {0.0, 1.0}
    bev = ru.extract_bev_from_pts(torch.tensor(d["pts"], dtype=torch.float32), cart_pixel_width=512).numpy()
    out["bev512_f32"] = np.packbits(bev != 0, axis=-1)

    cart = torch.tensor(d["cart"], dtype=torch.float64, requires_grad=True)
    pol = ru.radar_cartesian_to_polar(cart, torch.tensor(d["az"]), **C2P)
    (pol * torch.tensor(d["G_c2p"])).sum().backward()
    out["c2p_out"] = pol.detach().numpy()
    out["c2p_grad"] = cart.grad.numpy()

    for dt, tag in ((torch.float32, "f32"), (torch.float64, "f64")):
        for cross, wob, name in ((True, True, "full"), (False, False, "plain")):
            fft = torch.tensor(d["fft"], dtype=dt, requires_grad=True)
            c = ru.radar_polar_to_cartesian_diff(fft, torch.tensor(d["az2"], dtype=dt), interpolate_crossover=cross,
                                                 fix_wobble=wob, **P2C)
            (c * torch.tensor(d["G_p2c"], dtype=dt)).sum().backward()
            out[f"p2c_{name}_{tag}_out"] = c.detach().numpy()
            out[f"p2c_{name}_{tag}_grad"] = fft.grad.numpy()
    path = os.path.join(GOLDEN, "radar_next.npz")
    np.savez_compressed(path, **out)
    print("radar_next", os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
