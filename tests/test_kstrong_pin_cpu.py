"""CPU: the K-strongest restatement (`oracle/radar_oracle.py:kstrongest`) against the reference's literal C++
(`detector.inl:39-85` compiled unmodified, `oracle/Makefile` -> tests/golden/kstrong_ref.npz): pinned except
for the order the reference itself leaves undefined."""
import numpy as np
import pytest

from oracle import radar_oracle as RO
import kstrong_pin as KP


def _oracle_rows(g, name):
    scan = g[name + "_scan"][None]
    az = g[name + "_azimuths"][None].astype(np.float32)
    pts, rows, bins = RO.kstrongest(scan, az, KP.RES, **KP.params(g, name))[0]
    out = {}
    for r, j in zip(rows.tolist(), bins.tolist()):
        out.setdefault(r, []).append(j)
    return out, pts, rows, bins


@pytest.mark.parametrize("name", ["full", "topk", "scan"])
def test_oracle_selection_is_consistent_with_the_literal_reference(name):
    g = KP.load()
    ours, _, _, _ = _oracle_rows(g, name)
    KP.check(g, name, ours)


def test_oracle_topk_is_the_strongest_subset_of_the_reference_candidate_set():
    """`full` pins the candidate set; with k below the candidate count the restatement keeps the k strongest
    of exactly that set (descending value, ties to the lower bin)."""
    g = KP.load()
    scan = g["full_scan"]
    ref = KP.ref_bins_per_row(g, "full")
    prm = dict(KP.params(g, "full"), kstrong=4)
    pts, rows, bins = RO.kstrongest(scan[None], g["full_azimuths"][None].astype(np.float32), KP.RES, **prm)[0]
    for row, cand in ref.items():
        mine = bins[rows == row].tolist()
        want = sorted(cand, key=lambda j: (-float(scan[row, j]), j))[:4]
        assert mine == want


def test_oracle_rho_and_cartesian_match_the_reference_points():
    """rho bit-exact (float(j)*res + double offset -> float); x = rho cos(phi), y = rho sin(phi) in float as
    `navtech_extraction_module.cpp:33-39` (std::cos/std::sin of a float are the float overloads)."""
    g = KP.load()
    scan, az = g["full_scan"], g["full_azimuths"]
    ours, pts, rows, bins = _oracle_rows(g, "full")
    ref_rho = {}
    for r, p, t in zip(g["full_rho"], g["full_phi"], g["full_t"]):
        row = int(np.nonzero(g["full_times"] == t)[0][0])
        ref_rho.setdefault(row, {})[float(r)] = float(p)
    for i, (row, j) in enumerate(zip(rows.tolist(), bins.tolist())):
        rho = np.float32(float(np.float32(j) * np.float32(KP.RES)) + (-0.31))
        assert float(rho) in ref_rho[row]
        phi = np.float32(ref_rho[row][float(rho)])
        assert phi == np.float32(az[row])
        np.testing.assert_allclose(pts[i, :2], [rho * np.cos(phi), rho * np.sin(phi)], rtol=0, atol=1.6e-5)
