"""CPU oracle for the radar preprocessing rows of the hot path.  TEST INFRASTRUCTURE ONLY.

numpy restatements of the reference's `mm_masking/radar_utils.py` functions that the CUDA
kernels replace, each citing the lines it follows, plus the K-strongest detector whose only
specification is C++ (`external/vtr3/main/src/vtr_radar/include/vtr_radar/detector/detector.inl:39-85`).
Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline legs may import this
file; the product (`mm_masking_b200/`) never does.

Pinning: the reference has NO test for any of these functions, so they are
pinned against outputs of the unmodified reference functions run in the build container
(`oracle/make_golden_radar.py` -> `tests/golden/radar_*.npz`, checked by
`tests/test_radar_oracle_golden.py`).  K-strongest: **parity unpinned** -- the reference C++
cannot be compiled here (OpenCV/PCL headers absent) and has no test; the restatement follows the
intended semantics recorded in SURVEY.md section 8a note (top-k by intensity, ties to the lower
range bin) and is pinned only by hand-worked rows in the tests.

Precision policy: integer structure (mask bits, run boundaries, selected
bins, point order) is compared bit-exactly; interpolated / trigonometric floats to a few ulp.
Window sums are evaluated in float64 (`precision="f64"`, exact for uint8-derived scans whatever
the summation order) -- the reference's float32 `torch.sum` order depends on the host's SIMD
width, so its own mask is not bit-stable across machines; `precision="f32"` reproduces the
reference arithmetic literally for the golden comparison.
"""
from __future__ import annotations

import numpy as np

CART_RES = 0.2384      # radar_utils.py:258 default cart_resolution
ENCODER_TO_RAD = 2 * np.pi / 5600


def load_radar(raw: np.ndarray):
    """`radar_utils.py:20-27`: Navtech row = 8 B timestamp, 2 B encoder, 1 B unused, range bins."""
    raw = np.ascontiguousarray(raw)
    timestamps = np.frombuffer(raw[:, :8].tobytes(), dtype=np.int64) * 1000
    azimuths = np.frombuffer(raw[:, 8:10].tobytes(), dtype=np.uint16) * ENCODER_TO_RAD
    fft = np.divide(raw[:, 11:], 255.0, dtype=np.float32)
    return fft, azimuths, timestamps


def cfar_columns(n_bins: int, res: float, width=101, minr=2.0, maxr=80.0, guard=5):
    """`radar_utils.py:34-40`: half-width and the [mincol, maxcol) range that gets a threshold."""
    width = width + 1 if width % 2 == 0 else width
    w2 = width // 2
    mincol = max(0, int(minr / res + w2 + guard + 1))
    maxcol = min(n_bins, int(maxr / res - w2 - guard))
    return w2, mincol, maxcol


def cfar_threshold(raw: np.ndarray, res: float, width=101, minr=2.0, maxr=80.0, guard=5,
                   a_thresh=1.0, b_thresh=0.09, precision="f64") -> np.ndarray:
    """GO-CFAR threshold map (`radar_utils.py:42-60`): 1000 outside [mincol, maxcol)."""
    assert raw.ndim == 3
    w2, mincol, maxcol = cfar_columns(raw.shape[2], res, width, minr, maxr, guard)
    dt = np.float64 if precision == "f64" else np.float32
    x = raw.astype(dt)
    thres = np.full(raw.shape, 1000.0, dtype=dt)
    if precision == "f64":
        c = np.concatenate([np.zeros(raw.shape[:2] + (1,), dt), np.cumsum(x, axis=2)], axis=2)
        j = np.arange(mincol, maxcol)
        left = c[:, :, j - guard] - c[:, :, j - w2 - guard]                    # sum x[j-w2-guard : j-guard]
        right = c[:, :, j + w2 + guard + 1] - c[:, :, j + guard + 1]           # sum x[j+guard+1 : j+w2+guard+1]
    else:
        import torch
        xt = torch.from_numpy(np.ascontiguousarray(x))
        left = np.stack([torch.sum(xt[:, :, j - w2 - guard:j - guard], axis=2).numpy()
                         for j in range(mincol, maxcol)], axis=2)
        right = np.stack([torch.sum(xt[:, :, j + guard + 1:j + w2 + guard + 1], axis=2).numpy()
                          for j in range(mincol, maxcol)], axis=2)
    stat = np.maximum(left, right) / dt(w2)
    thres[:, :, mincol:maxcol] = dt(a_thresh) * stat + dt(b_thresh)
    return thres


def cfar_mask(raw: np.ndarray, res: float, width=101, minr=2.0, maxr=80.0, guard=5, a_thresh=1.0,
              b_thresh=0.09, diff=True, steep_fact=10.0, precision="f64") -> np.ndarray:
    """`radar_utils.py:29-69`.  Returns float32 like the reference."""
    thres = cfar_threshold(raw, res, width, minr, maxr, guard, a_thresh, b_thresh, precision)
    x = raw.astype(thres.dtype)
    if diff:
        soft = 0.5 * np.tanh(thres.dtype.type(steep_fact) * (x - thres) + 2.5) + 0.5        # :64
        return np.where(np.abs(soft) > 0.99, soft, 0.0).astype(np.float32)                  # hardshrink :65
    return (x > thres).astype(np.float32)                                                   # :67


def peak_entries(mask: np.ndarray, res: float, diff=True, steep_fact=10.0) -> np.ndarray:
    """`radar_utils.py:74,167-185`: the (B,A,R) array whose non-zero entries are run starts/ends."""
    mask = mask.astype(np.float32)
    arr = (np.float32(res) * np.arange(mask.shape[2], dtype=np.float32)) * mask
    if diff:
        zero = (1 - np.tanh(np.float32(steep_fact) * arr)).astype(np.float32)
    else:
        zero = (arr == 0).astype(np.float32)
    out = np.zeros_like(arr)
    out[:, :, :-1] = arr[:, :, :-1] * zero[:, :, 1:] + arr[:, :, 1:] * zero[:, :, :-1]
    return out


def extract_pc(mask: np.ndarray, res: float, azimuths: np.ndarray, times: np.ndarray, T_ab=None,
               diff=True, steep_fact=10.0, return_structure=False):
    """`radar_utils.py:71-106`: consecutive non-zero peak entries of the flattened scan are
    averaged in pairs into (rho, phi, t) and mapped to (rho cos phi, rho sin phi, 0).

    Returns a list of (P_b,3) float32 arrays; with return_structure also, per scan, the int
    arrays (row, bin) of every non-zero entry in flattened order (the bit-exact part).
    """
    peaks = peak_entries(mask, res, diff, steep_fact)
    B, A, R = peaks.shape
    out, structure = [], []
    for b in range(B):
        flat = peaks[b].reshape(-1)
        nz = np.nonzero(flat)[0]
        if len(nz) % 2:
            raise ValueError("odd number of peak entries (a run touches the last range bin)")
        rows, bins = nz // R, nz % R
        val = flat[nz]
        phi = azimuths[b].astype(np.float32)[rows]
        rho = (val[1::2] + val[0::2]) / np.float32(2.0)
        ang = (phi[1::2] + phi[0::2]) / np.float32(2.0)
        pc = np.stack([rho * np.cos(ang), rho * np.sin(ang), np.zeros_like(rho)], axis=1).astype(np.float32)
        if T_ab is not None:
            T = np.asarray(T_ab[b], dtype=np.float32)
            pc = (T[:3, :3] @ pc.T).T + T[:3, 3]
        out.append(pc.astype(np.float32))
        structure.append((rows.astype(np.int32), bins.astype(np.int32)))
    return (out, structure) if return_structure else out
