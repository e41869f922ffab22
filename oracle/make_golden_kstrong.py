"""Golden vectors for K-strongest from the UNMODIFIED reference C++.  TEST INFRASTRUCTURE ONLY.

`oracle/_ref/kstrong_ref` (built by oracle/Makefile) is
`external/vtr3/main/src/vtr_radar/include/vtr_radar/detector/detector.inl:39-85` compiled as it lies in
/root/reference against the shim headers of oracle/ref_shim/.  This script feeds it seeded scans and stores
scans + returned polar points in tests/golden/kstrong_ref.npz.  Run in the build container:

    make -C oracle && python oracle/make_golden_kstrong.py

What the literal C++ defines and what it does not (note on R4):
  defined    the bin range (int(minr/res) <= j < maxr/res), the mean (double accumulation, divided by the
             DOUBLE N = maxr/res - minr/res), thres = mean*threshold2 + threshold3, the candidate set
             {j : scan[j] >= thres}, rho = float(j)*res + range_offset, the number of points min(k, |set|)
             when |set| >= k, and that saturated bins (value 1.0 -> int 1) come before all others;
  undefined  the order among bins whose intensity truncates to the same int (the comparator takes
             pair<int,float> but the vector holds pair<float,int>, `detector.inl:28-31,55,68`), and the
             points of a row with fewer than k candidates (`intens[j]` is read past the end, `:72-73`).
The cases below stay inside the defined part:
  full   every row has exactly c candidates and k = c, so the output IS the candidate set; one bin sits
         exactly on the threshold (the smallest float >= thres), one just below; strong bins just outside
         the range; threshold2 != 0 so that dividing by the bin count instead of the double N would move
         the threshold across those two bins.
  topk   k = 12 with 14..60 candidates per row, 0..8 of them saturated.
  scan   a Navtech-style synthetic scan with the deployed parameters (k 12, threshold2 0, threshold3
         0.216, `gen_data/config/radar_lidar_config.yaml:91-94`); rows with fewer than k candidates are
         recorded but flagged (their literal output is undefined behaviour).
"""
from __future__ import annotations

import os
import struct
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
BIN = os.path.join(ROOT, "oracle", "_ref", "kstrong_ref")
RES = 0.0596
COLS = 3360


def run_reference(scan, azimuths, times, res, k, t2, t3, minr, maxr, range_offset):
    rows, cols = scan.shape
    with tempfile.TemporaryDirectory() as tmp:
        fin, fout = os.path.join(tmp, "in.bin"), os.path.join(tmp, "out.bin")
        with open(fin, "wb") as f:
            f.write(struct.pack("<iiif", 0, rows, cols, np.float32(res)))
            f.write(np.asarray([k, t2, t3, minr, maxr, range_offset, 0, 0], np.float64).tobytes())
            f.write(np.asarray(azimuths, np.float64).tobytes())
            f.write(np.asarray(times, np.int64).tobytes())
            f.write(np.ascontiguousarray(scan, np.float32).tobytes())
        subprocess.check_call([BIN, fin, fout])
        raw = open(fout, "rb").read()
    count = struct.unpack("<i", raw[:4])[0]
    rec = np.frombuffer(raw[4:], dtype=np.dtype([("rho", "<f4"), ("phi", "<f4"), ("t", "<i8")]), count=count)
    return rec["rho"].copy(), rec["phi"].copy(), rec["t"].copy()


def bin_range(res, minr, maxr, cols):
    resf = float(np.float32(res))
    mincol = minr / resf
    if mincol > cols or mincol < 0:
        mincol = 0.0
    maxcol = maxr / resf
    if maxcol > cols or maxcol < 0:
        maxcol = float(cols)
    js = np.arange(int(mincol), cols)
    return js[js < maxcol], maxcol - mincol


def seq_mean(row, js, N):
    acc = 0.0
    for v in row[js]:
        acc += float(v)                      # `mean += raw_scan.at<float>(i, j)` in double
    return acc / N


def case_full(rng, rows=10, c=9, t2=1.5, t3=0.2, minr=2.0, maxr=80.0):
    js, N = bin_range(RES, minr, maxr, COLS)
    scan = (rng.integers(0, 12, (rows, COLS)).astype(np.float32) / np.float32(255.0))
    for i in range(rows):
        inner = rng.choice(js[2:-2], size=c + 1, replace=False)
        strong, edge_hi, edge_lo = inner[:c - 1], inner[c - 1], inner[c]
        scan[i, strong] = rng.integers(140, 256, c - 1).astype(np.float32) / np.float32(255.0)
        if i % 2 == 0:                                   # first / last bin of the range carry a strong return
            scan[i, js[0]] = scan[i, strong[0]]; scan[i, strong[0]] = 0.0
            scan[i, js[-1]] = scan[i, strong[1]]; scan[i, strong[1]] = 0.0
        scan[i, js[0] - 1] = 1.0                         # just outside the range: must never be selected
        scan[i, js[-1] + 1] = 1.0
        scan[i, edge_hi] = scan[i, edge_lo] = np.float32(t3)
        for _ in range(4):                               # bracket the threshold with adjacent floats
            thres = seq_mean(scan[i], js, N) * t2 + t3
            hi = np.float32(thres)
            if float(hi) < thres:
                hi = np.nextafter(hi, np.float32(2.0))
            scan[i, edge_hi] = hi
            scan[i, edge_lo] = np.nextafter(hi, np.float32(0.0))
        thres = seq_mean(scan[i], js, N) * t2 + t3
        assert float(scan[i, edge_hi]) >= thres > float(scan[i, edge_lo])
        assert int((scan[i, js].astype(np.float64) >= thres).sum()) == c
        wrong = scan[i, js].astype(np.float64).sum() / len(js) * t2 + t3   # mean over the bin count instead of N
        assert not (float(scan[i, edge_hi]) >= wrong > float(scan[i, edge_lo]))
    return scan, dict(k=c, t2=t2, t3=t3, minr=minr, maxr=maxr)


def case_topk(rng, rows=12, k=12, t3=0.216):
    js, _ = bin_range(RES, 2.0, 80.0, COLS)
    scan = (rng.integers(0, 40, (rows, COLS)).astype(np.float32) / np.float32(255.0))
    for i in range(rows):
        c = int(rng.integers(14, 61))
        cand = rng.choice(js, size=c, replace=False)
        scan[i, cand] = rng.integers(60, 255, c).astype(np.float32) / np.float32(255.0)
        scan[i, cand[: int(rng.integers(0, 9))]] = 1.0
    return scan, dict(k=k, t2=0.0, t3=t3, minr=2.0, maxr=80.0)


def case_scan(rows=24):
    from mm_masking_b200 import synth
    d = synth.make_radar_scans(1, seed=77)
    return d["fft"][0, :rows].copy(), dict(k=12, t2=0.0, t3=0.216, minr=2.0, maxr=80.0)


def main():
    rng = np.random.default_rng(2024)
    out = {}
    for name, (scan, prm) in (("full", case_full(rng)), ("topk", case_topk(rng)), ("scan", case_scan())):
        rows = scan.shape[0]
        az = (np.arange(rows) * (2 * np.pi / 400) + 0.01).astype(np.float64)
        times = (1_600_000_000_000_000 + np.arange(rows) * 625).astype(np.int64)     # unique per row: labels the row
        rho, phi, t = run_reference(scan, az, times, RES, prm["k"], prm["t2"], prm["t3"], prm["minr"], prm["maxr"], -0.31)
        out[name + "_scan"] = scan
        out[name + "_azimuths"] = az
        out[name + "_times"] = times
        out[name + "_params"] = np.asarray([prm["k"], prm["t2"], prm["t3"], prm["minr"], prm["maxr"], -0.31], np.float64)
        out[name + "_rho"], out[name + "_phi"], out[name + "_t"] = rho, phi, t
        print(name, "rows", rows, "points", len(rho))
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "kstrong_ref.npz"), **out)


if __name__ == "__main__":
    main()
