"""Second, independent golden for K1: the reference's numpy + `cv2.remap` twin
`mm_masking/radar_utils.py:197-256` (`radar_polar_to_cartesian`, taken from pyboreas
`external/pyboreas/pyboreas/utils/radar.py:59-122`), run UNMODIFIED in the build container (cv2 4.13).
TEST INFRASTRUCTURE ONLY.

    python oracle/make_golden_radar_cv2.py   ->  tests/golden/radar_cv2.npz

`cv2.remap(INTER_LINEAR)` quantises the fractional sample position to 1/32 pixel (INTER_BITS = 5) and treats
samples outside the scan as 0 without the differentiable variant's azimuth wrap-padding, so this is a
cross-check to a stated tolerance, not a bit-level oracle: see tests/test_radar_cv2_crosscheck*.py."""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_loader  # noqa: E402
from mm_masking_b200 import synth  # noqa: E402

RES = 0.0596


def main():
    _, _, _, ru = ref_loader.load_reference()
    sc = synth.make_radar_scans(B=2, seed=41, n_az=400, n_bins=1024)
    out = dict(raw=sc["raw"])
    for b in range(2):
        for width in (160, 96):      # even widths only: for odd widths the two reference functions use different pixel centres (radar_utils.py:218 `//` vs :408 `/`)
            cart = ru.radar_polar_to_cartesian(sc["fft"][b], sc["azimuths"][b].reshape(-1, 1), RES, cart_resolution=0.2384,
                                               cart_pixel_width=width, interpolate_crossover=False, fix_wobble=True)
            out[f"cart_{b}_{width}"] = np.asarray(cart, np.float32)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "radar_cv2.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
