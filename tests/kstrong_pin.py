"""Shared checker: what the reference's literal K-strongest (`detector.inl:39-85`, compiled unmodified by
oracle/Makefile, outputs in tests/golden/kstrong_ref.npz) DEFINES, applied to any implementation's selection.

An implementation hands over, per scan row, the list of selected bins in its output order.  Checked:
  * rho of every golden point is float(j)*res + range_offset for an integer bin j (recovers the bins);
  * `full` (k = candidates per row): the selected set EQUALS the reference's set, row by row -- pins the bin
    range, the double-N mean, the >= comparison at a threshold that sits between two adjacent floats;
  * `topk`/`scan` (more candidates than k): same count as the reference (k), every selected bin is a
    candidate of the reference threshold (>= the weakest value the reference returned... see below), and
    saturated bins come first in both;
  * what stays a choice: the order among bins whose intensity truncates to the same int (undefined in the
    reference) -- the implementations order by descending intensity, ties to the lower bin."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "kstrong_ref.npz")
RES = 0.0596


def load():
    z = np.load(GOLDEN)
    return {k: z[k] for k in z.files}


def ref_bins_per_row(g, name):
    """{row: [bins in the reference's output order]} recovered from the golden rho / timestamp."""
    rho, t, times = g[name + "_rho"], g[name + "_t"], g[name + "_times"]
    off = g[name + "_params"][5]
    cols = g[name + "_scan"].shape[1]
    table = (np.arange(cols, dtype=np.float32) * np.float32(RES)).astype(np.float64) + off
    table = table.astype(np.float32)                       # p.rho is a float
    out = {}
    for r, tt in zip(rho, t):
        row = int(np.nonzero(times == tt)[0][0])
        j = np.nonzero(table == r)[0]
        assert len(j) == 1, "golden rho is not float(j)*res + range_offset for a unique bin"
        out.setdefault(row, []).append(int(j[0]))
    return out


def params(g, name):
    k, t2, t3, minr, maxr, off = g[name + "_params"]
    return dict(kstrong=int(k), threshold2=float(t2), threshold3=float(t3), minr=float(minr), maxr=float(maxr),
                range_offset=float(off))


def check(g, name, ours):
    """ours: {row: [bins]} from the implementation under test."""
    scan = g[name + "_scan"]
    ref = ref_bins_per_row(g, name)
    prm = params(g, name)
    k = prm["kstrong"]
    checked = 0
    for row in range(scan.shape[0]):
        mine = ours.get(row, [])
        theirs = ref.get(row, [])
        vals = scan[row]
        if name == "full":
            assert sorted(mine) == sorted(theirs), f"row {row}: candidate set differs from the reference"
            assert len(set(theirs)) == k
            checked += 1
            continue
        # topk / scan: threshold2 == 0, so thres == threshold3 exactly and the candidate set is known
        assert prm["threshold2"] == 0.0
        resf = float(np.float32(RES))
        js = np.arange(int(prm["minr"] / resf), scan.shape[1])
        js = js[js < prm["maxr"] / resf]
        cand = [int(j) for j in js if float(vals[j]) >= prm["threshold3"]]
        if len(cand) < k:
            continue                      # the literal reference reads past the end here: undefined
        assert len(theirs) == k and len(set(theirs)) == k and set(theirs) <= set(cand), f"row {row}: reference"
        assert len(mine) == k and set(mine) <= set(cand), f"row {row}: not k candidates of the reference threshold"
        n_sat = min(k, sum(1 for j in cand if vals[j] == 1.0))
        assert all(vals[j] == 1.0 for j in theirs[:n_sat])          # defined part of the reference order
        assert all(vals[j] == 1.0 for j in mine[:n_sat])
        if n_sat < k:
            assert sorted(j for j in mine if vals[j] == 1.0) == sorted(j for j in theirs if vals[j] == 1.0)
        # the choice the reference leaves open: descending intensity, ties to the lower bin
        assert mine == sorted(cand, key=lambda j: (-float(vals[j]), j))[:k], f"row {row}: not the k strongest"
        checked += 1
    assert checked >= (scan.shape[0] if name == "full" else 6), (name, checked)
