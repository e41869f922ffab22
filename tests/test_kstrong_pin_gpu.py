"""GPU: `kstrong_rows_kernel` (through the C ABI) against the reference's literal C++ K-strongest
(`detector.inl:39-85` compiled unmodified, goldens tests/golden/kstrong_ref.npz): candidate set, threshold
(>= at a threshold between two adjacent floats, mean divided by the double N), bin range, rho, count and the
saturated-first order; the one thing that stays a choice is the order the reference leaves undefined."""
import numpy as np
import pytest
import torch

import kstrong_pin as KP

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["full", "topk", "scan"])
def test_kernel_selection_is_consistent_with_the_literal_reference(name):
    from mm_masking_b200 import radar_utils as ru
    g = KP.load()
    scan = torch.from_numpy(g[name + "_scan"][None]).cuda()
    az = torch.from_numpy(g[name + "_azimuths"][None].astype(np.float32)).cuda()
    pts, rows, bins = ru.kstrong_extract_pc(scan, KP.RES, az, return_bins=True, **KP.params(g, name))
    ours = {}
    for r, j in zip(rows[0].tolist(), bins[0].tolist()):
        ours.setdefault(r, []).append(j)
    KP.check(g, name, ours)
    if name == "full":          # rho of every point equals a reference rho of the same row (bit-exact float)
        ref = KP.ref_bins_per_row(g, name)
        p = pts[0].cpu().numpy()
        for i, (r, j) in enumerate(zip(rows[0].tolist(), bins[0].tolist())):
            assert j in ref[r]
            rho = np.float32(float(np.float32(j) * np.float32(KP.RES)) + (-0.31))
            phi = np.float32(g[name + "_azimuths"][r])
            np.testing.assert_allclose(p[i, :2], [rho * np.cos(phi), rho * np.sin(phi)], rtol=0, atol=1.6e-5)
