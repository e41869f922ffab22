"""The radar oracle (oracle/radar_oracle.py) against vectors produced by the UNMODIFIED reference
`mm_masking/radar_utils.py` (oracle/make_golden_radar.py).  CPU only."""
import hashlib

import numpy as np
import pytest

from conftest import load_golden
from mm_masking_b200 import synth
from oracle import radar_oracle as RO

RES = 0.0596


@pytest.fixture(scope="module")
def cfar_golden():
    d, _ = load_golden("radar_cfar_extract")
    fft, az, ts = [], [], []
    for b in range(d["raw"].shape[0]):
        f, a, t = RO.load_radar(d["raw"][b])
        fft.append(f); az.append(a); ts.append(t)
    return d, np.stack(fft), np.stack(az).astype(np.float32), np.stack(ts)


def test_load_radar_matches_reference(cfar_golden):
    d, fft, az, ts = cfar_golden
    f0, a0, t0 = RO.load_radar(d["raw"][0])
    assert np.array_equal(f0, d["load_fft0"]) and np.array_equal(a0, d["load_az0"]) and np.array_equal(t0, d["load_ts0"])


def test_cfar_columns_defaults():
    assert RO.cfar_columns(3360, RES) == (50, 89, 1287)


def test_cfar_hard_mask(cfar_golden):
    d, fft, az, ts = cfar_golden
    ref64 = np.unpackbits(d["mask_hard_f64"], axis=-1)[..., :fft.shape[2]].astype(bool)
    ref32 = np.unpackbits(d["mask_hard_f32"], axis=-1)[..., :fft.shape[2]].astype(bool)
    ours = RO.cfar_mask(fft, RES, diff=False, precision="f64") != 0
    assert np.array_equal(ours, ref64)                           # bit-exact against the reference run in fp64
    # the float32 reference differs from its own float64 run only where x is within rounding of thres
    assert (ref32 != ref64).sum() <= 2e-5 * ref64.size
    lit = RO.cfar_mask(fft, RES, diff=False, precision="f32") != 0
    assert (lit != ref32).sum() <= 2e-5 * ref32.size             # literal fp32 restatement (sum order is host dependent)
    assert ours.sum() > 100


def test_cfar_diff_mask(cfar_golden):
    d, fft, az, ts = cfar_golden
    ours = RO.cfar_mask(fft, RES, diff=True, precision="f64")
    ref = d["mask_diff_f64"]
    assert np.array_equal(ours != 0, ref != 0)
    np.testing.assert_allclose(ours, ref, atol=2e-7)
    assert (np.abs(d["mask_diff_f32"] - ref) > 1e-5).sum() <= 2e-5 * ref.size


@pytest.mark.parametrize("diff", [False, True])
def test_extract_pc(cfar_golden, diff):
    d, fft, az, ts = cfar_golden
    tag = "diff" if diff else "hard"
    mask = RO.cfar_mask(fft, RES, diff=diff, precision="f64")
    pcs = RO.extract_pc(mask, RES, az, ts, diff=diff)
    assert [len(p) for p in pcs] == d[f"pc_{tag}_counts"].tolist()
    np.testing.assert_allclose(np.concatenate(pcs), d[f"pc_{tag}"], atol=2e-5)


def test_extract_pc_with_transform(cfar_golden):
    d, fft, az, ts = cfar_golden
    mask = RO.cfar_mask(fft, RES, diff=False, precision="f64")
    pcs = RO.extract_pc(mask, RES, az, ts, T_ab=d["T_ab"], diff=False)
    np.testing.assert_allclose(np.concatenate(pcs), d["pc_hard_T"], atol=3e-5)


def _weight_inputs():
    N = 3
    yy = np.linspace(0, 1, 640, dtype=np.float32)
    masks = np.stack([(0.5 + 0.5 * np.sin(7 * (k + 1) * yy)[:, None] * np.cos(5 * (k + 2) * yy)[None, :]) for k in range(N)])
    return masks.astype(np.float32)


def test_extract_weights():
    d, _ = load_golden("radar_weights")
    masks = _weight_inputs()
    w, stats = RO.extract_weights(masks, d["pts"])
    np.testing.assert_allclose(w, d["w"], atol=2e-6)
    assert np.all(w[:, -20:] == 0)                               # batch padding
    ref = d["stats"]      # diff_mean_num_non0, mean_num_non0, mean_w, max_w, min_w
    np.testing.assert_allclose([stats["diff_mean_num_non0"], stats["mean_num_non0"], stats["mean_w"], stats["max_w"],
                                stats["min_w"]], ref, rtol=1e-5, atol=1e-6)
    g = RO.extract_weights_backward(d["G"], masks.shape, d["pts"])
    ref_g = np.zeros(masks.shape, np.float64)
    idx = d["grad_mask_nz_idx"]
    ref_g[idx[:, 0], idx[:, 1], idx[:, 2]] = d["grad_mask_nz_val"]
    np.testing.assert_allclose(g, ref_g, atol=1e-6)


def test_polar_to_cartesian():
    d, _ = load_golden("radar_polar_to_cart")
    sc = synth.make_radar_scans(B=2, seed=13)
    assert np.array_equal(np.frombuffer(hashlib.sha1(sc["raw"].tobytes()).digest(), dtype=np.uint8), d["raw_sha1"])
    for width in (160, 161):
        c32 = RO.polar_to_cartesian(sc["fft"], sc["azimuths"], RES, cart_pixel_width=width)
        np.testing.assert_allclose(c32, d[f"cart{width}_f32"], atol=3e-6)
        c64 = RO.polar_to_cartesian(sc["fft"].astype(np.float64), sc["azimuths"].astype(np.float64), RES, cart_pixel_width=width)
        np.testing.assert_allclose(c64, d[f"cart{width}_f64"], atol=1e-12)
    c = RO.polar_to_cartesian(sc["fft"], sc["azimuths"], RES, cart_pixel_width=96, interpolate_crossover=False,
                              fix_wobble=False)
    np.testing.assert_allclose(c, d["cart96_plain_f32"], atol=3e-6)


def test_kstrongest_hand_worked_rows():
    """No reference test and no compilable reference: pin the restatement on rows worked by hand."""
    R = 1500
    fft = np.zeros((1, 3, R), np.float32)
    az = np.array([[0.0, np.pi / 2, np.pi]], np.float32)
    # row 0: three bins above 0.216, one equal pair -> strongest first, tie to the lower bin
    fft[0, 0, [100, 200, 300, 400]] = [0.5, 0.9, 0.5, 0.2]
    # row 1: nothing passes; row 2: a bin below minr (j < 33) and one beyond maxr (j > 1342) are ignored
    fft[0, 1, 500] = 0.21
    fft[0, 2, [10, 1400, 700]] = [1.0, 1.0, 0.3]
    (pts, rows, bins), = RO.kstrongest(fft, az, RES, kstrong=12)
    assert rows.tolist() == [0, 0, 0, 2] and bins.tolist() == [200, 100, 300, 700]
    rho = np.float32(200) * np.float32(RES) - 0.31
    np.testing.assert_allclose(pts[0], [rho, 0.0, 0.0], atol=1e-5)
    np.testing.assert_allclose(pts[3], [-(np.float32(700) * np.float32(RES) - 0.31), 0.0, 0.0], atol=1e-4)
    (pts2, rows2, bins2), = RO.kstrongest(fft, az, RES, kstrong=2)
    assert bins2.tolist() == [200, 100, 700]
    # the mean term: threshold2 = 1 adds mean = sum / N with N = 80/res - 2/res (a double, not the bin count)
    resf = float(np.float32(RES))
    N = 80.0 / resf - 2.0 / resf
    t = (0.5 + 0.9 + 0.5 + 0.2) / N + 0.216
    assert t < 0.5
    (_, _, bins3), = RO.kstrongest(fft, az, RES, kstrong=12, threshold2=1.0)
    assert bins3.tolist()[:3] == [200, 100, 300]
