"""CPU: size-independent properties of the oracle restatements (oracle/), the same properties the GPU
parity tests lean on at sizes where no golden vector exists: rigid-motion equivariance of the ICP
solve, batch == loop (the reference's own `tests/test_ICP_inputs.py` property), target-permutation
invariance of the nearest neighbour, <A x, y> = <x, A^T y> for every resampler/adjoint pair,
idempotence of the occupancy BEV, and the K-strongest ordering contract."""
import numpy as np
import pytest
import torch

from oracle import dicp_oracle as O
from oracle import radar_oracle as RO


def _scene(N=2, n=120, m=700, seed=0):
    """a 30 x 20 m room with an inner wall: map points + wall normals, scan = map subset seen from a moved pose"""
    g = torch.Generator().manual_seed(seed)
    walls = torch.tensor([[-15, -10, 15, -10], [15, -10, 15, 10], [15, 10, -15, 10], [-15, 10, -15, -10], [-3, -10, -3, 2]],
                         dtype=torch.float64)
    a, b = walls[:, :2], walls[:, 2:]
    d = b - a
    nrm_w = torch.stack((d[:, 1], -d[:, 0]), dim=1) / d.norm(dim=1, keepdim=True)
    wi = torch.randint(0, walls.shape[0], (N, m), generator=g)
    s = torch.rand((N, m, 1), generator=g, dtype=torch.float64)
    xy = a[wi] + s * d[wi] + 0.01 * torch.randn((N, m, 2), generator=g, dtype=torch.float64)
    z = torch.zeros((N, m, 1), dtype=torch.float64)
    target = torch.cat((xy, z, nrm_w[wi], z), dim=2)
    pick = torch.randint(0, m, (N, n), generator=g)
    pts = torch.gather(target[:, :, :3], 1, pick[:, :, None].expand(N, n, 3)).clone()
    pts[:, :, :2] += 0.03 * torch.randn((N, n, 2), generator=g, dtype=torch.float64)
    xi = torch.tensor([[0.4, -0.3, 0, 0, 0, 0.05], [-0.2, 0.5, 0, 0, 0, -0.04]], dtype=torch.float64)[:N]
    T_true = torch.stack([O.se3_exp(x) for x in xi])
    src = torch.einsum('bij,bnj->bni', torch.linalg.inv(T_true)[:, :3, :3], pts - T_true[:, None, :3, 3])
    w = 0.2 + 0.8 * torch.rand((N, n), generator=g, dtype=torch.float64)
    return src, target, torch.eye(4, dtype=torch.float64).repeat(N, 1, 1), w


KW = dict(icp_type="pt2pl", max_iterations=6, tolerance=1e-12, differentiable=True, trim_dist=5.0,
          loss_fn={"name": "cauchy", "metric": 1.0}, dim=2, const_iter=True)


def test_icp_is_equivariant_under_a_rigid_motion_of_the_map_frame():
    src, tgt, T0, w = _scene()
    G = O.se3_exp(torch.tensor([3.0, -7.0, 0, 0, 0, 0.8], dtype=torch.float64))
    R, t = G[:3, :3], G[:3, 3]
    tgt2 = torch.cat((tgt[:, :, :3] @ R.T + t, tgt[:, :, 3:] @ R.T), dim=2)
    a = O.dicp(src, tgt, T0, w, return_indices=True, **KW)
    b = O.dicp(src, tgt2, G @ T0, w, return_indices=True, **KW)
    assert torch.equal(a["indices"], b["indices"])
    np.testing.assert_allclose((G @ a["T"]).numpy(), b["T"].numpy(), atol=1e-9)
    np.testing.assert_allclose(a["weights"].numpy(), b["weights"].numpy(), atol=1e-9)


def test_icp_batch_equals_loop():
    src, tgt, T0, w = _scene()
    kw = dict(KW, const_iter=False, tolerance=1e-6, max_iterations=30)
    both = O.dicp(src, tgt, T0, w, **kw)
    assert bool(both["stats"]["converged"].all())
    for b in range(src.shape[0]):
        one = O.dicp(src[b:b + 1], tgt[b:b + 1], T0[b:b + 1], w[b:b + 1], **kw)
        np.testing.assert_allclose(one["T"][0].numpy(), both["T"][b].numpy(), atol=1e-8)   # test_ICP_inputs.py:88-90


def test_zero_weights_leave_the_initial_pose():
    src, tgt, T0, w = _scene(N=1)
    T0 = O.se3_exp(torch.tensor([0.3, 0.1, 0, 0, 0, 0.02], dtype=torch.float64))[None]
    out = O.dicp(src, tgt, T0, torch.zeros_like(w), **KW)
    np.testing.assert_allclose(out["T"].numpy(), T0.numpy(), atol=1e-6)


def test_nearest_neighbour_follows_a_permutation_of_the_targets():
    g = torch.Generator().manual_seed(3)
    q = torch.randn((2, 50, 3), generator=g, dtype=torch.float64)
    t = torch.randn((2, 400, 3), generator=g, dtype=torch.float64)
    perm = torch.randperm(400, generator=g)
    i0 = O.nearest_index(q, t)
    i1 = O.nearest_index(q, t[:, perm])
    assert torch.equal(perm[i1], i0)                      # no ties in random doubles
    # ties go to the lowest index: duplicate every target behind the originals
    assert torch.equal(O.nearest_index(q, torch.cat((t, t), dim=1)), i0)


def _az(B, A, seed):
    rng = np.random.default_rng(seed)
    ticks = np.round(np.arange(A) * (5600.0 / A) + rng.integers(-1, 2, size=(B, A)))
    return np.sort(ticks * (2 * np.pi / 5600.0), axis=1)


@pytest.mark.parametrize("cross,wob", [(True, True), (False, True), (True, False)])
def test_polar_to_cartesian_adjoint_is_the_transpose(cross, wob):
    rng = np.random.default_rng(0)
    B, A, R, W = 1, 40, 96, 32
    az = _az(B, A, 1)
    x = rng.random((B, A, R))
    y = rng.standard_normal((B, W, W))
    Ax = RO.polar_to_cartesian(x, az, 0.5, cart_pixel_width=W, interpolate_crossover=cross, fix_wobble=wob)
    ATy = RO.polar_to_cartesian_backward(y, az, R, 0.5, interpolate_crossover=cross, fix_wobble=wob)
    np.testing.assert_allclose(np.sum(Ax * y), np.sum(x * ATy), rtol=1e-10)


def test_cartesian_to_polar_adjoint_is_the_transpose():
    rng = np.random.default_rng(2)
    B, A, R, W = 1, 24, 60, 48
    az = _az(B, A, 3)
    x = rng.random((B, W, W))
    y = rng.standard_normal((B, A, R))
    Ax = RO.cartesian_to_polar(x, az, 0.5, polar_pixel_shape=(A, R))
    ATy = RO.cartesian_to_polar_backward(y, az, 0.5, (W, W))
    np.testing.assert_allclose(np.sum(Ax * y), np.sum(x * ATy), rtol=1e-10)


def test_weight_sampling_adjoint_is_the_transpose():
    rng = np.random.default_rng(4)
    mask = rng.random((2, 640, 640)).astype(np.float32)    # the adjoint restatement works on fp32 taps
    pts = np.zeros((2, 30, 3), dtype=np.float32)
    pts[:, :, :2] = rng.uniform(-70, 70, size=(2, 30, 2)).astype(np.float32)
    pts[0, 5] = 0.0                                        # padding row: samples nothing
    gw = rng.standard_normal((2, 30))
    w = RO.extract_weights(mask, pts)
    w = w[0] if isinstance(w, tuple) else w
    gm = RO.extract_weights_backward(gw, mask.shape, pts)
    np.testing.assert_allclose(np.sum(w.astype(np.float64) * gw), np.sum(mask.astype(np.float64) * gm), rtol=2e-6)


def test_occupancy_bev_is_idempotent_and_ignores_padding():
    rng = np.random.default_rng(5)
    pc = np.zeros((2, 200, 6), dtype=np.float32)
    pc[:, :150, :2] = rng.uniform(-90, 90, size=(2, 150, 2)).astype(np.float32)   # some fall outside +-76 m
    bev = RO.extract_bev_from_pts(pc)
    assert bev.shape == (2, 640, 640) and set(np.unique(bev)) <= {0.0, 1.0}
    assert np.array_equal(RO.extract_bev_from_pts(np.concatenate((pc, pc), axis=1)), bev)
    assert np.array_equal(RO.extract_bev_from_pts(pc[:, ::-1].copy()), bev)
    assert not bev[:, 320, 320].any()                     # the pixel that collects padding / out-of-range points


def test_kstrongest_keeps_at_most_k_per_azimuth_strongest_first():
    rng = np.random.default_rng(6)
    A, R, res, k = 6, 1500, 0.0596, 5
    fft = (rng.random((1, A, R)) * 0.2).astype(np.float32)
    fft[0, 0, 100:103] = (0.9, 0.7, 0.8)                  # 3 bins over the threshold
    fft[0, 1, 200:210] = np.linspace(0.5, 0.95, 10)       # 10 bins: only the 5 strongest stay
    fft[0, 2, 300] = fft[0, 2, 400] = 0.6                 # a tie: lower bin first
    az = np.linspace(0, 2 * np.pi, A, endpoint=False)[None].astype(np.float32)
    out = RO.kstrongest(fft, az, res, kstrong=k)
    pts = out[0] if isinstance(out, (list, tuple)) else out
    pts = np.asarray(pts[0] if isinstance(pts, (list, tuple)) else pts)
    rho = np.hypot(pts[:, 0], pts[:, 1])
    bins = np.rint((rho + 0.31) / res).astype(int)
    assert bins.tolist() == [100, 102, 101, 209, 208, 207, 206, 205, 300, 400]


@pytest.mark.parametrize("dim", [2, 3])
def test_kdtree_nn_mode_equals_exact_mode(dim):
    """`nn_mode="kdtree"` (the checker used at c3 / c5 map sizes) returns the indices of the exhaustive
    `nn_mode="exact"`: synthetic map with 1 % duplicated points (exact ties -> lowest index), pad rows, and
    queries that sit exactly on targets."""
    from mm_masking_b200 import synth
    batch = synth.make_icp_batch(N=2, n=400, m=6000, seed=5, three_d=dim == 3)
    tgt = torch.tensor(batch["target"][:, :, :3], dtype=torch.float64)
    if dim == 2:
        tgt = tgt * torch.tensor([1.0, 1.0, 0.0], dtype=torch.float64)
    g = torch.Generator().manual_seed(3)
    q = tgt[:, torch.randint(0, 6000, (400,), generator=g)] + 0.3 * torch.randn((2, 400, 3), generator=g, dtype=torch.float64)
    q[:, :40] = tgt[:, torch.randint(0, 6000, (40,), generator=g)]            # distance exactly 0, some duplicated
    q[:, 40:80] += 30.0 * torch.randn((2, 40, 3), generator=g, dtype=torch.float64)   # far from everything
    if dim == 2:
        q[:, :, 2] = 0.0
    tgt[1, 100:110] = tgt[1, 99]                                              # an 11-fold tie: the full-scan fallback
    q[1, 0] = tgt[1, 99]
    a = O.nearest_index(q, tgt, "exact")
    b = O.nearest_index(q, tgt, "kdtree")
    assert torch.equal(a, b)
    assert int(b[1, 0]) <= 99


def test_kdtree_mode_gives_the_same_icp_run():
    src, tgt, T0, w = _scene()
    a = O.dicp(src, tgt, T0, w, return_indices=True, **KW)
    b = O.dicp(src, tgt, T0, w, return_indices=True, nn_mode="kdtree", **KW)
    assert torch.equal(a["indices"], b["indices"]) and torch.equal(a["T"], b["T"])
