"""GPU parity tests of the radar kernels (through the C ABI via mm_masking_b200.radar_utils) against
the reference's golden vectors (tests/golden/radar_*.npz) and the CPU oracle on seeded scans.
Integer structure (mask bits, point counts/order, selected bins) is compared bit-exactly; sampled
and trigonometric floats to a few ulp (tolerances stated at each assert)."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from mm_masking_b200 import synth
from oracle import radar_oracle as RO

pytestmark = pytest.mark.gpu
RES = 0.0596


def _ru():
    from mm_masking_b200 import radar_utils
    return radar_utils


@pytest.fixture(scope="module")
def cfar_golden():
    d, _ = load_golden("radar_cfar_extract")
    dec = [RO.load_radar(r) for r in d["raw"]]
    return d, np.stack([x[0] for x in dec]), np.stack([x[1] for x in dec]).astype(np.float32), np.stack([x[2] for x in dec])


@pytest.mark.parametrize("kind", ["f32", "f64", "u8"])
def test_cfar_hard_mask_bit_exact(cfar_golden, kind):
    d, fft, az, ts = cfar_golden
    ref = np.unpackbits(d["mask_hard_f64"], axis=-1)[..., :fft.shape[2]].astype(bool)
    x = {"f32": torch.from_numpy(fft), "f64": torch.from_numpy(fft).double(), "u8": torch.from_numpy(d["raw"])}[kind]
    m = _ru().cfar_mask(x.cuda(), RES, diff=False)
    assert m.dtype == torch.float32 and m.shape == fft.shape
    assert np.array_equal(m.cpu().numpy() != 0, ref)
    assert set(np.unique(m.cpu().numpy()).tolist()) <= {0.0, 1.0}


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_cfar_diff_mask(cfar_golden, dtype):
    d, fft, az, ts = cfar_golden
    m = _ru().cfar_mask(torch.from_numpy(fft).to(dtype).cuda(), RES, diff=True)
    assert m.dtype == dtype
    ref = d["mask_diff_f64"]
    got = m.cpu().numpy()
    tol = 1e-6 if dtype == torch.float32 else 1e-7                   # golden stored as float32
    near_cut = np.abs(np.where(ref != 0, ref, got) - 0.99) < 1e-5   # hardshrink cut at 0.99: support may flip within rounding
    assert np.array_equal((got != 0) | near_cut, (ref != 0) | near_cut)
    np.testing.assert_allclose(np.where(near_cut, 0, got), np.where(near_cut, 0, ref), atol=tol)


def test_cfar_full_size_matches_oracle():
    sc = synth.make_radar_scans(B=2, seed=3)
    m = _ru().cfar_mask(torch.from_numpy(sc["raw"]).cuda(), RES, diff=False).cpu().numpy()
    assert np.array_equal(m, RO.cfar_mask(sc["fft"], RES, diff=False))
    m2 = _ru().cfar_mask(torch.from_numpy(sc["fft"]).cuda(), RES, width=60, guard=2, minr=5.0, maxr=150.0, a_thresh=1.2,
                         b_thresh=0.05, diff=False).cpu().numpy()
    assert np.array_equal(m2, RO.cfar_mask(sc["fft"], RES, width=60, guard=2, minr=5.0, maxr=150.0, a_thresh=1.2,
                                           b_thresh=0.05, diff=False))


@pytest.mark.parametrize("diff", [False, True])
def test_extract_pc_golden(cfar_golden, diff):
    d, fft, az, ts = cfar_golden
    ru = _ru()
    tag = "diff" if diff else "hard"
    mask = ru.cfar_mask(torch.from_numpy(fft).double().cuda(), RES, diff=diff).float()
    pcs = ru.extract_pc(mask, RES, torch.from_numpy(az).cuda(), torch.from_numpy(ts).cuda(), diff=diff)
    assert [len(p) for p in pcs] == d[f"pc_{tag}_counts"].tolist()               # bit-exact structure
    got = torch.cat(pcs).cpu().numpy()
    assert got.dtype == np.float32 and np.all(got[:, 2] == 0)
    np.testing.assert_allclose(got, d[f"pc_{tag}"], atol=1.6e-5)                 # 2 ulp at 80 m
    # fused form: identical points, the mask never materialised
    fused = ru.cfar_extract_pc(torch.from_numpy(fft).double().cuda(), RES, torch.from_numpy(az).double().cuda(), diff=diff)
    assert [len(p) for p in fused] == [len(p) for p in pcs]
    np.testing.assert_allclose(torch.cat(fused).cpu().numpy(), got, atol=1.6e-5)


def test_extract_pc_transform_and_fused_raw(cfar_golden):
    d, fft, az, ts = cfar_golden
    ru = _ru()
    mask = ru.cfar_mask(torch.from_numpy(d["raw"]).cuda(), RES, diff=False)
    pcs = ru.extract_pc(mask, RES, torch.from_numpy(az).cuda(), torch.from_numpy(ts).cuda(), T_ab=torch.from_numpy(d["T_ab"]),
                        diff=False)
    np.testing.assert_allclose(torch.cat(pcs).cpu().numpy(), d["pc_hard_T"], atol=3e-5)
    fused = ru.cfar_extract_pc(torch.from_numpy(d["raw"]).cuda(), RES, torch.from_numpy(az).cuda(),
                               T_ab=torch.from_numpy(d["T_ab"]).cuda())
    for a, b in zip(fused, pcs):
        assert torch.equal(a, b)                                                  # same kernels, same bits


def test_extract_pc_structure_full_size():
    sc = synth.make_radar_scans(B=3, seed=5)
    ru = _ru()
    mask = RO.cfar_mask(sc["fft"], RES, diff=False)
    ref, structure = RO.extract_pc(mask, RES, sc["azimuths"], sc["timestamps"], diff=False, return_structure=True)
    pcs = ru.cfar_extract_pc(torch.from_numpy(sc["raw"]).cuda(), RES, torch.from_numpy(sc["azimuths"]).cuda())
    assert [len(p) for p in pcs] == [len(p) for p in ref]
    for got, want in zip(pcs, ref):
        np.testing.assert_allclose(got.cpu().numpy(), want, atol=1.6e-5)
    assert sum(len(p) for p in ref) > 3000


def test_extract_pc_edge_cases():
    ru = _ru()
    az = torch.linspace(0, 6.2, 8).repeat(2, 1).cuda()
    empty = torch.zeros((2, 8, 300), device="cuda")
    pcs = ru.extract_pc(empty, RES, az, None, diff=False)
    assert [tuple(p.shape) for p in pcs] == [(0, 3), (0, 3)]
    one = empty.clone()
    one[1, 3, 10:13] = 1.0                  # a single run: one point at the mean of bins 10 and 12
    one[1, 3, 20] = 1.0                     # a one-bin run
    pcs = ru.extract_pc(one, RES, az, None, diff=False)
    assert len(pcs[0]) == 0 and len(pcs[1]) == 2
    rho = pcs[1].norm(dim=1).cpu().numpy()
    np.testing.assert_allclose(rho, [np.float32(RES) * 11, np.float32(RES) * 20], rtol=1e-6)
    bad = empty.clone()
    bad[0, 2, 299] = 1.0                    # a run touching the last bin: the reference's pairing breaks (odd count)
    with pytest.raises(ValueError):
        ru.extract_pc(bad, RES, az, None, diff=False)


def _weight_masks():
    yy = np.linspace(0, 1, 640, dtype=np.float32)
    return np.stack([(0.5 + 0.5 * np.sin(7 * (k + 1) * yy)[:, None] * np.cos(5 * (k + 2) * yy)[None, :])
                     for k in range(3)]).astype(np.float32)


def test_extract_weights_forward_backward():
    d, _ = load_golden("radar_weights")
    ru = _ru()
    mask = torch.from_numpy(_weight_masks()).cuda().requires_grad_(True)
    w, dnz, nz, mean_w, max_w, min_w = ru.extract_weights(mask, torch.from_numpy(d["pts"]).cuda())
    np.testing.assert_allclose(w.detach().cpu().numpy(), d["w"], atol=2e-6)
    assert torch.all(w[:, -20:] == 0)
    np.testing.assert_allclose([float(dnz), float(nz), float(mean_w), float(max_w), float(min_w)], d["stats"], rtol=1e-5,
                               atol=1e-6)
    assert dnz.requires_grad and not nz.requires_grad
    (w * torch.from_numpy(d["G"]).cuda()).sum().backward()
    ref = np.zeros(mask.shape, np.float32)
    idx = d["grad_mask_nz_idx"]
    ref[idx[:, 0], idx[:, 1], idx[:, 2]] = d["grad_mask_nz_val"]
    np.testing.assert_allclose(mask.grad.cpu().numpy(), ref, atol=1e-6)           # <= 1e-4 relative of the O(1) taps
    # float64 against the oracle
    w64, _ = RO.extract_weights(_weight_masks().astype(np.float64), d["pts"].astype(np.float64))
    got = ru.extract_weights(torch.from_numpy(_weight_masks()).double().cuda(), torch.from_numpy(d["pts"]).double().cuda())[0]
    np.testing.assert_allclose(got.cpu().numpy(), w64, atol=1e-13)


def test_polar_to_cartesian_golden():
    d, _ = load_golden("radar_polar_to_cart")
    sc = synth.make_radar_scans(B=2, seed=13)
    ru = _ru()
    fft, az = torch.from_numpy(sc["fft"]).cuda(), torch.from_numpy(sc["azimuths"]).cuda()
    for width in (160, 161):
        c32 = ru.radar_polar_to_cartesian_diff(fft, az, RES, cart_pixel_width=width)
        np.testing.assert_allclose(c32.cpu().numpy(), d[f"cart{width}_f32"], atol=3e-6)
        c64 = ru.radar_polar_to_cartesian_diff(fft.double(), az.double(), RES, cart_pixel_width=width)
        np.testing.assert_allclose(c64.cpu().numpy(), d[f"cart{width}_f64"], atol=1e-12)
        raw = ru.radar_polar_to_cartesian_diff(torch.from_numpy(sc["raw"]).cuda(), az, RES, cart_pixel_width=width)
        assert torch.equal(raw, c32)                                              # on-the-fly byte/255 decode
    c = ru.radar_polar_to_cartesian_diff(fft, az, RES, cart_pixel_width=96, interpolate_crossover=False, fix_wobble=False)
    np.testing.assert_allclose(c.cpu().numpy(), d["cart96_plain_f32"], atol=3e-6)


def test_polar_to_cartesian_full_size_matches_oracle():
    sc = synth.make_radar_scans(B=1, seed=17)
    c = _ru().radar_polar_to_cartesian_diff(torch.from_numpy(sc["raw"]).cuda(), torch.from_numpy(sc["azimuths"]).cuda(), RES)
    assert c.shape == (1, 640, 640)
    np.testing.assert_allclose(c.cpu().numpy(), RO.polar_to_cartesian(sc["fft"], sc["azimuths"], RES), atol=3e-6)


def test_kstrongest_matches_oracle():
    sc = synth.make_radar_scans(B=2, seed=23)
    ru = _ru()
    for kw in (dict(), dict(kstrong=3, threshold2=1.5, threshold3=0.1), dict(kstrong=40, minr=0.0, maxr=500.0),
               dict(kstrong=20, threshold3=0.01)):          # most bins pass: more candidates than the shared list holds
        ref = RO.kstrongest(sc["fft"], sc["azimuths"], RES, **kw)
        pts, rows, bins = ru.kstrong_extract_pc(torch.from_numpy(sc["raw"]).cuda(), RES, torch.from_numpy(sc["azimuths"]).cuda(),
                                                return_bins=True, **kw)
        pts2 = ru.kstrong_extract_pc(torch.from_numpy(sc["fft"]).cuda(), RES, torch.from_numpy(sc["azimuths"]).cuda(), **kw)
        for b in range(2):
            assert np.array_equal(rows[b].cpu().numpy(), ref[b][1]) and np.array_equal(bins[b].cpu().numpy(), ref[b][2])
            np.testing.assert_allclose(pts[b].cpu().numpy(), ref[b][0], atol=1.6e-5)
            assert torch.equal(pts2[b], pts[b])
        assert sum(len(r[1]) for r in ref) > 100


def test_kstrongest_hand_worked_rows():
    R = 1500
    fft = np.zeros((1, 3, R), np.float32)
    az = np.array([[0.0, np.pi / 2, np.pi]], np.float32)
    fft[0, 0, [100, 200, 300, 400]] = [0.5, 0.9, 0.5, 0.2]
    fft[0, 1, 500] = 0.21
    fft[0, 2, [10, 1400, 700]] = [1.0, 1.0, 0.3]
    pts, rows, bins = _ru().kstrong_extract_pc(torch.from_numpy(fft).cuda(), RES, torch.from_numpy(az).cuda(), return_bins=True)
    assert rows[0].tolist() == [0, 0, 0, 2] and bins[0].tolist() == [200, 100, 300, 700]
    rho = np.float32(200) * np.float32(RES) - 0.31
    np.testing.assert_allclose(pts[0][0].cpu().numpy(), [rho, 0.0, 0.0], atol=1e-5)


def test_mask_gradient_is_deterministic_and_matches_the_atomic_form():
    """the adjoint of the weight sampler sums colliding taps in tap order (sort by pixel, no
    floating-point atomics): two runs give identical bits, and the values equal the atomic kernel's up to
    summation order.  Half of the points are exact duplicates, so many pixels receive several taps."""
    import ctypes as C
    from mm_masking_b200 import _capi
    g = torch.Generator().manual_seed(8)
    N, n, W = 3, 4000, 640
    pts = torch.zeros((N, n, 3))
    pts[:, :, :2] = (torch.rand((N, n, 2), generator=g) - 0.5) * 40          # dense: ~170 x 170 pixels for 4000 points
    pts[:, n // 2:] = pts[:, :n // 2]
    pts[:, :5] = 0.0                                                          # padding rows
    pts = pts.cuda()
    gw = torch.randn((N, n), generator=g).cuda()
    a = torch.ops.mmb.sample_weights_bwd(gw, pts, W, W, 0.2384)
    b = torch.ops.mmb.sample_weights_bwd(gw, pts, W, W, 0.2384)
    assert torch.equal(a, b)
    lib = _capi.load_library()
    ref = torch.zeros((N, W, W), device="cuda")
    _capi.check(lib.mmb_sample_weights_bwd(N, n, 3, W, W, _capi.F32, 0.2384, _capi.ptr(gw), _capi.ptr(pts), _capi.ptr(ref),
                                           _capi.stream_ptr()), "atomic form")
    assert (a != 0).sum() > 1000
    np.testing.assert_allclose(a.cpu().numpy(), ref.cpu().numpy(), rtol=0, atol=2e-5)
    # through autograd, against the oracle's adjoint
    mask = torch.rand((N, W, W), generator=g).cuda().requires_grad_(True)
    w = torch.ops.mmb.sample_weights(mask, pts, 0.2384)
    (w * gw).sum().backward()
    assert torch.equal(mask.grad, a)
