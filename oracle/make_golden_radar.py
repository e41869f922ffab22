"""Generate tests/golden/radar_*.npz from the UNMODIFIED reference `mm_masking/radar_utils.py`.

Run in the build container (where /root/reference is mounted):

    python oracle/make_golden_radar.py

TEST INFRASTRUCTURE ONLY.  The reference has no tests for these functions, so the vectors are
outputs of the reference functions themselves on small seeded synthetic scans
(`mm_masking_b200.synth.make_radar_scans`, reduced geometry so the files stay small):

  radar_cfar_extract.npz   cfar_mask (diff and non-diff, float32 and float64 input) + extract_pc
  radar_weights.npz        extract_weights (values, five statistics, autograd gradient wrt the mask)
  radar_polar_to_cart.npz  radar_polar_to_cartesian_diff (float32 and float64)
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_loader  # noqa: E402
from mm_masking_b200 import synth  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
RES = 0.0596


def main():
    _, _, _, ru = ref_loader.load_reference()
    torch.manual_seed(0)

    # ---- CFAR + extraction: 2 scans, 48 azimuths x 1500 bins (maxcol = 1287 < 1500) -------------
    sc = synth.make_radar_scans(B=2, seed=11, n_az=48, n_bins=1500)
    raw, fft, az, ts = sc["raw"], sc["fft"], sc["azimuths"], sc["timestamps"]
    # the reference's own decoder on scan 0 (radar_utils.py:20-27)
    f0, a0, t0 = ru.load_radar(raw[0])
    out = dict(raw=raw, load_fft0=f0, load_az0=a0, load_ts0=t0)
    x32 = torch.from_numpy(fft)
    for diff in (False, True):
        for dt, name in ((torch.float32, "f32"), (torch.float64, "f64")):
            m = ru.cfar_mask(x32.to(dt), RES, diff=diff).numpy()
            key = f"mask_{'diff' if diff else 'hard'}_{name}"
            out[key] = np.packbits(m != 0, axis=-1) if not diff else m.astype(np.float32)
    # extraction on the fp64-input masks (see radar_oracle.py precision policy)
    for diff in (False, True):
        m = ru.cfar_mask(x32.double(), RES, diff=diff).float()
        pcs = ru.extract_pc(m, RES, torch.from_numpy(az), torch.from_numpy(ts), diff=diff)
        tag = "diff" if diff else "hard"
        out[f"pc_{tag}_counts"] = np.array([len(p) for p in pcs], np.int64)
        out[f"pc_{tag}"] = np.concatenate([p.numpy() for p in pcs], axis=0).astype(np.float32)
    # with a per-scan transform
    T_ab = torch.eye(4).repeat(2, 1, 1)
    T_ab[0, :3, :3] = torch.tensor([[0.0, -1.0, 0.0], [1.0, 0.0, 0.0], [0.0, 0.0, 1.0]])
    T_ab[:, :3, 3] = torch.tensor([[1.5, -2.0, 0.25], [-3.0, 0.5, 0.0]])
    m = ru.cfar_mask(x32.double(), RES, diff=False).float()
    pcs = ru.extract_pc(m, RES, torch.from_numpy(az), torch.from_numpy(ts), T_ab=T_ab, diff=False)
    out["T_ab"] = T_ab.numpy()
    out["pc_hard_T"] = np.concatenate([p.numpy() for p in pcs], axis=0).astype(np.float32)
    np.savez_compressed(os.path.join(GOLDEN, "radar_cfar_extract.npz"), **out)

    # ---- extract_weights: 3 masks 640x640 would be 4.9 MB; a smooth low-rank mask keeps it small -----
    rng = np.random.default_rng(5)
    N, n = 3, 300
    yy = np.linspace(0, 1, 640, dtype=np.float32)
    masks = np.stack([(0.5 + 0.5 * np.sin(7 * (k + 1) * yy)[:, None] * np.cos(5 * (k + 2) * yy)[None, :]) for k in range(N)])
    masks = masks.astype(np.float32)
    pts = np.zeros((N, n, 3), np.float32)
    pts[:, :, :2] = rng.uniform(-85, 85, (N, n, 2))            # some fall outside the 76 m image
    pts[:, -20:, :] = 0.0                                       # batch padding
    pts[0, 0, :2] = (76.1688, -76.1688)                         # pixel-centre corners
    pts[0, 1, :2] = (-76.1688, 76.1688)
    pts[0, 2, :2] = (0.1192, 0.1192)
    mt = torch.from_numpy(masks).requires_grad_(True)
    w, dnz, nz, mean_w, max_w, min_w = ru.extract_weights(mt, torch.from_numpy(pts))
    G = torch.from_numpy(rng.standard_normal((N, n)).astype(np.float32))
    (w * G).sum().backward()
    np.savez_compressed(os.path.join(GOLDEN, "radar_weights.npz"), pts=pts, G=G.numpy(), w=w.detach().numpy(),
                        stats=np.array([float(dnz), float(nz), float(mean_w), float(max_w), float(min_w)]),
                        grad_mask_nz_idx=np.stack(np.nonzero(mt.grad.numpy()), axis=1).astype(np.int32),
                        grad_mask_nz_val=mt.grad.numpy()[np.nonzero(mt.grad.numpy())])

    # ---- polar -> Cartesian: 2 scans 400 x 3360 uint8 in, 160 x 160 out --------------------------
    sc = synth.make_radar_scans(B=2, seed=13)
    fftp = torch.from_numpy(sc["fft"]); azp = torch.from_numpy(sc["azimuths"])
    out = dict(raw=sc["raw"])
    for width in (160, 161):
        c32 = ru.radar_polar_to_cartesian_diff(fftp, azp, RES, cart_pixel_width=width).numpy()
        c64 = ru.radar_polar_to_cartesian_diff(fftp.double(), azp.double(), RES, cart_pixel_width=width).numpy()
        out[f"cart{width}_f32"] = c32.astype(np.float32)
        out[f"cart{width}_f64"] = c64.astype(np.float64)
    c32 = ru.radar_polar_to_cartesian_diff(fftp, azp, RES, cart_pixel_width=96, interpolate_crossover=False,
                                           fix_wobble=False).numpy()
    out["cart96_plain_f32"] = c32
    np.savez_compressed(os.path.join(GOLDEN, "radar_polar_to_cart.npz"), **out)
    for f in ("radar_cfar_extract", "radar_weights", "radar_polar_to_cart"):
        print(f, os.path.getsize(os.path.join(GOLDEN, f + ".npz")) // 1024, "KiB")


if __name__ == "__main__":
    main()
