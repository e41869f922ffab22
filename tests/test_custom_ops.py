"""`torch.ops.mmb.*` (SURVEY.md section 8b "PyTorch custom ops"): registration, fake-tensor tracing (CPU, no
kernel runs) and -- on the GPU -- `torch.library.opcheck` plus equality with the reference-shaped wrappers."""
import numpy as np
import pytest
import torch

from mm_masking_b200 import _capi, ops  # noqa: F401

OPS = ["icp_forward", "icp_backward", "nn_search", "sample_weights", "sample_weights_bwd", "polar_to_cart",
       "polar_to_cart_bwd", "cart_to_polar", "cart_to_polar_bwd", "cfar_mask", "cfar_extract", "kstrong_extract",
       "bev_from_points"]
ICP_ARGS = (_capi.PT2PL, 2, 5, 1e-5, True, True, True, 5.0, _capi.LOSS_CAUCHY, 1.0, 5.0, 0.1, 0, _capi.NN_AUTO)


def test_ops_are_registered_with_schemas():
    for name in OPS:
        op = getattr(torch.ops.mmb, name)
        assert "mmb::" + name in str(op.default._schema)
    assert "Tensor? weight" in str(torch.ops.mmb.icp_forward.default._schema)


def test_fake_tensor_shapes_trace_without_a_device():
    """the fake implementations give the output shapes / dtypes with no kernel launch (what torch.compile and
    torch.export trace through)"""
    from torch._subclasses.fake_tensor import FakeTensorMode
    with FakeTensorMode():
        src, tgt = torch.empty((3, 100, 3), device="cuda"), torch.empty((3, 900, 6), device="cuda")
        T0, w = torch.empty((3, 4, 4), device="cuda"), torch.empty((3, 100), device="cuda")
        T, pc, block, saved = torch.ops.mmb.icp_forward(src, tgt, T0, w, *ICP_ARGS, True)
        assert T.shape == (3, 4, 4) and pc.shape == (3, 100, 3) and saved.numel() > 5 * 3 * 100 * 4
        hist = ops._carve(block, ops.icp_history_specs(3, 100, 5, _capi.PT2PL, torch.float32))
        assert [tuple(h.shape) for h in hist] == [(3, 5, 6), (3, 5, 100), (3, 5), (3,), (3,), (3,), (1,)]
        g = torch.ops.mmb.icp_backward(src, tgt, w, saved, T, None, False, False, True, True, *ICP_ARGS)
        assert [tuple(t.shape) for t in g] == [(0,), (0,), (3, 4, 4), (3, 100)]
        assert torch.ops.mmb.nn_search(src, tgt, 3, 0, 0).dtype == torch.int32
        mask = torch.empty((3, 64, 64), device="cuda")
        assert torch.ops.mmb.sample_weights(mask, src, 0.2384).shape == (3, 100)
        assert torch.ops.mmb.sample_weights_bwd(w, src, 64, 64, 0.2384).shape == (3, 64, 64)
        raw = torch.empty((2, 40, 500 + 11), dtype=torch.uint8, device="cuda")
        az, tab = torch.empty((2, 40), device="cuda"), torch.empty((96, 96), device="cuda")
        assert torch.ops.mmb.polar_to_cart(raw, az, tab, tab, 96, 0.0596, True, True).shape == (2, 96, 96)
        assert torch.ops.mmb.cfar_mask(raw, 0.0596, 101, 2.0, 80.0, 5, 1.0, 0.09, False, 10.0).shape == (2, 40, 500)
        pts, cnt = torch.ops.mmb.cfar_extract(raw, az, None, 300, 0.0596, 101, 2.0, 80.0, 5, 1.0, 0.09, False, 10.0)
        assert pts.shape == (2, 300, 3) and cnt.dtype == torch.int32
        pts, rows, bins, cnt = torch.ops.mmb.kstrong_extract(raw, az, 0.0596, 12, 0.0, 0.216, 2.0, 80.0, -0.31)
        assert pts.shape == (2, 480, 3) and rows.shape == bins.shape == (2, 480)
        assert torch.ops.mmb.bev_from_points(src, 64).shape == (3, 64, 64)


@pytest.mark.gpu
def test_opcheck_and_wrappers_on_the_gpu():
    from mm_masking_b200 import radar_utils as ru, synth
    from mm_masking_b200.dICP import ICP
    batch = synth.make_icp_batch(N=2, n=200, m=3000, seed=7, pos=0.5, rot=0.1)
    t = {k: torch.tensor(batch[k], device="cuda") for k in ("source", "target", "T_init", "weight")}
    w = t["weight"].clone().requires_grad_(True)
    # the op called directly == the reference-shaped wrapper
    T, pc, block, saved = torch.ops.mmb.icp_forward(t["source"], t["target"], t["T_init"], w, *ICP_ARGS, True)
    T.sum().backward()
    alg = ICP(icp_type="pt2pl", differentiable=True, max_iterations=5, tolerance=1e-5)
    alg.const_iter = True
    w2 = t["weight"].clone().requires_grad_(True)
    out = alg.icp(t["source"], t["target"], t["T_init"], weight=w2, trim_dist=5.0,
                  loss_fn={"name": "cauchy", "metric": 1.0}, dim=2)
    out["T"].sum().backward()
    assert torch.equal(T, out["T"]) and torch.equal(w.grad, w2.grad)
    # opcheck: schema, fake implementation, autograd registration
    torch.library.opcheck(torch.ops.mmb.icp_forward.default,
                          (t["source"], t["target"], t["T_init"], t["weight"].clone().requires_grad_(True), *ICP_ARGS, True),
                          test_utils=("test_schema", "test_faketensor", "test_autograd_registration"))
    torch.library.opcheck(torch.ops.mmb.nn_search.default, (t["source"], t["target"], 2, 0, 0))
    mask = torch.rand((2, 640, 640), device="cuda", requires_grad=True)
    torch.library.opcheck(torch.ops.mmb.sample_weights.default, (mask, t["source"], 0.2384))
    sc = synth.make_radar_scans(B=2, seed=3, n_az=40, n_bins=600)
    raw, az = torch.from_numpy(sc["raw"]).cuda(), torch.from_numpy(sc["azimuths"]).cuda()
    rng, ang = ru.form_cart_range_angle_grid(cart_pixel_width=96, dtype=torch.float32, device="cuda")
    fft = torch.from_numpy(sc["fft"]).cuda().requires_grad_(True)
    torch.library.opcheck(torch.ops.mmb.polar_to_cart.default, (fft, az, rng.contiguous(), ang.contiguous(), 96, 0.0596, True, True))
    torch.library.opcheck(torch.ops.mmb.kstrong_extract.default, (raw, az, 0.0596, 12, 0.0, 0.216, 2.0, 80.0, -0.31))
    # fixed-capacity extraction ops == the ragged wrappers
    lists = ru.cfar_extract_pc(raw, 0.0596, az)
    pts, ent = torch.ops.mmb.cfar_extract(raw, az, None, max(len(p) for p in lists) + 5, 0.0596, 101, 2.0, 80.0, 5, 1.0, 0.09,
                                          False, 10.0)
    for b, p in enumerate(lists):
        assert int(ent[b]) == 2 * len(p) and torch.equal(pts[b, :len(p)], p)
    klists = ru.kstrong_extract_pc(raw, 0.0596, az)
    kp, rows, bins, cnt = torch.ops.mmb.kstrong_extract(raw, az, 0.0596, 12, 0.0, 0.216, 2.0, 80.0, -0.31)
    for b, p in enumerate(klists):
        assert int(cnt[b]) == len(p) and torch.equal(kp[b, :len(p)], p)
