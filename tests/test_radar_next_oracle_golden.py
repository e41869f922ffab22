"""Oracle restatements of SURVEY.md 8f rows N2 (`extract_bev_from_pts`), N3 (`radar_cartesian_to_polar`)
and of the adjoints of both resamplers, against vectors from the UNMODIFIED reference
(oracle/make_golden_radar2.py -> tests/golden/radar_next.npz).  CPU only."""
import numpy as np
import pytest

from conftest import load_golden
from oracle import radar_oracle as RO

C2P = dict(radar_resolution=0.0596 * 1.5, cart_resolution=0.2384, polar_pixel_shape=(40, 300))


@pytest.fixture(scope="module")
def g():
    return load_golden("radar_next")[0]


@pytest.mark.parametrize("tag,dt", [("f32", np.float32), ("f64", np.float64)])
def test_bev_occupancy_bit_exact(g, tag, dt):
    ref = np.unpackbits(g[f"bev_{tag}"], axis=-1)[..., :640].astype(bool)
    ours = RO.extract_bev_from_pts(g["pts"].astype(dt))
    assert ours.dtype == dt and np.array_equal(ours != 0, ref)
    assert not ours[:, 320, 320].any()


def test_bev_other_width_keeps_default_centre(g):
    ref = np.unpackbits(g["bev512_f32"], axis=-1)[..., :512].astype(bool)
    assert np.array_equal(RO.extract_bev_from_pts(g["pts"].astype(np.float32), cart_pixel_width=512) != 0, ref)


def test_cartesian_to_polar_and_adjoint(g):
    out = RO.cartesian_to_polar(g["cart"], g["az"], **C2P)
    np.testing.assert_allclose(out, g["c2p_out"], atol=1e-13, rtol=0)
    grad = RO.cartesian_to_polar_backward(g["G_c2p"], g["az"], C2P["radar_resolution"], g["cart"].shape[1:])
    np.testing.assert_allclose(grad, g["c2p_grad"], atol=1e-12, rtol=0)


@pytest.mark.parametrize("name,cross,wob", [("full", True, True), ("plain", False, False)])
def test_polar_to_cartesian_adjoint(g, name, cross, wob):
    fft, az = g["fft"], g["az2"]
    out = RO.polar_to_cartesian(fft, az, 0.05, cart_pixel_width=64, interpolate_crossover=cross, fix_wobble=wob)
    np.testing.assert_allclose(out, g[f"p2c_{name}_f64_out"], atol=1e-12, rtol=0)
    grad = RO.polar_to_cartesian_backward(g["G_p2c"], az, fft.shape[2], 0.05, interpolate_crossover=cross, fix_wobble=wob)
    np.testing.assert_allclose(grad, g[f"p2c_{name}_f64_grad"], atol=1e-12, rtol=0)
    g32 = RO.polar_to_cartesian_backward(g["G_p2c"].astype(np.float32), az.astype(np.float32), fft.shape[2], 0.05,
                                         interpolate_crossover=cross, fix_wobble=wob)
    ref32 = g[f"p2c_{name}_f32_grad"]
    assert np.abs(g32 - ref32).max() <= 2e-4 * np.abs(ref32).max()      # fp32 tap weights, cancellation near bin edges
