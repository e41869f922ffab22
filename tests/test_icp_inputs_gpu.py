"""GPU: the reference's input-handling tests (`external/dICP/tests/test_ICP_inputs.py`) run against
this package, plus the goldens the reference produced for the same ragged inputs."""
import numpy as np
import pytest
import torch

from oracle import dicp_oracle as O

pytestmark = pytest.mark.gpu
TOL = 1e-8
LOSS = {"name": "huber", "metric": 1.0}


def _alg(**kw):
    from mm_masking_b200.dICP import ICP
    base = dict(icp_type="pt2pl", differentiable=True, max_iterations=25, tolerance=TOL)
    base.update(kw)
    return ICP(**base)


def _t(a, grad=False):
    return torch.tensor(np.asarray(a), dtype=torch.float64, requires_grad=grad)


def test_input_types(golden):
    """list-of-ragged == per-sample loop == reference (`test_ICP_inputs.py:36-110`)"""
    d, _ = golden("ragged_input_types")
    src = [_t(d[f"in_source_{i}"]) for i in range(3)]
    tgt = [_t(d[f"in_target_{i}"], True) for i in range(3)]
    T0 = torch.eye(4, dtype=torch.float64).repeat(3, 1, 1)
    alg = _alg()
    batch = alg.icp(src, tgt, T0, trim_dist=5.0, loss_fn=LOSS, dim=2)
    loop_T, loop_mr = [], []
    for i in range(3):
        r = alg.icp(src[i], tgt[i], T0[i], trim_dist=5.0, loss_fn=LOSS, dim=2)
        loop_T.append(r["T"][0].detach()); loop_mr.append(r["stats"]["matched_ratio"][0].item())
    loop_T = torch.stack(loop_T)
    err = torch.stack([O.se3_log(loop_T[i] @ torch.linalg.inv(batch["T"][i].detach())) for i in range(3)])
    assert float(torch.linalg.norm(err)) < TOL
    assert np.linalg.norm(np.array(loop_mr) - batch["stats"]["matched_ratio"].numpy()) < TOL
    np.testing.assert_allclose(batch["T"].detach().numpy(), d["ref_T"], atol=1e-9)
    np.testing.assert_allclose(batch["stats"]["matched_ratio"].numpy(), d["ref_matched_ratio"], atol=1e-12)
    np.testing.assert_array_equal(batch["stats"]["iterations"].numpy(), d["ref_iterations"])
    np.testing.assert_array_equal(batch["stats"]["converged"].numpy(), d["ref_converged"])
    assert batch["deltas"].shape == d["ref_deltas"].shape
    np.testing.assert_allclose(batch["deltas"].numpy(), d["ref_deltas"], atol=1e-9)
    np.testing.assert_allclose(batch["weights"].numpy(), d["ref_weights"], atol=1e-9)
    np.testing.assert_allclose(batch["costs"].numpy(), d["ref_costs"], atol=1e-9, rtol=1e-9)
    np.testing.assert_allclose(batch["pc"].detach().numpy(), d["ref_pc"], atol=1e-8)
    np.testing.assert_allclose(loop_T.numpy(), d["ref_loop_T"], atol=1e-9)


def test_zero_inputs(golden):
    """empty source and/or target returns the initial guess (`test_ICP_inputs.py:113-155`)"""
    d, _ = golden("fixture_pt2pl_diff")
    source = _t(d["in_source"][0], True); target = _t(d["in_target"][0], True)
    src = [source, [], []]
    tgt = [[], target, []]
    T0 = torch.eye(4, dtype=torch.float64).repeat(3, 1, 1)
    alg = _alg()
    for i in range(3):
        r = alg.icp(src[i], tgt[i], T0[i], trim_dist=5.0, loss_fn=None, dim=2)
        assert np.linalg.norm(r["T"][0].detach().numpy() - np.eye(4)) < TOL
    r = alg.icp(src, tgt, T0, trim_dist=5.0, loss_fn=None, dim=2)
    assert np.linalg.norm(r["T"].detach().numpy() - T0.numpy()) < TOL
    assert r["T"].shape == (3, 4, 4)


def test_weight_inputs(golden):
    """None / ones / zero-weighted extras give one pose (`test_ICP_inputs.py:157-211`)"""
    d, _ = golden("ragged_weight_inputs")
    src = [_t(d[f"in_source_{i}"], True) for i in range(3)]
    tgt = [_t(d["in_target"], True) for _ in range(3)]
    wts = [None, _t(d["in_weight_1"], True), _t(d["in_weight_2"], True)]
    T0 = torch.eye(4, dtype=torch.float64).repeat(3, 1, 1)
    alg = _alg()
    loop = torch.stack([alg.icp(src[i], tgt[i], T0[i], weight=wts[i], trim_dist=5.0, loss_fn=LOSS, dim=2)["T"][0].detach()
                        for i in range(3)])
    batch = alg.icp(src, tgt, T0, weight=wts, trim_dist=5.0, loss_fn=LOSS, dim=2)["T"].detach()
    assert np.linalg.norm((batch - loop).numpy()) < TOL
    assert np.linalg.norm((loop[0] - loop[1]).numpy()) < TOL and np.linalg.norm((loop[0] - loop[2]).numpy()) < TOL
    np.testing.assert_allclose(batch.numpy(), d["ref_T"], atol=1e-9)


def test_diff_vs_nondiff(golden):
    """`test_ICP_inputs.py:213-252`"""
    d, _ = golden("fixture_pt2pl_diff")
    source = _t(d["in_source"][0][:50]); target = _t(d["in_target"][0][:55])
    T0 = torch.eye(4, dtype=torch.float64)
    for loss_fn in ({"name": "huber", "metric": 1.0}, {"name": "cauchy", "metric": 0.5}):
        a = _alg(differentiable=True).icp(source, target, T0, trim_dist=5.0, loss_fn=loss_fn, dim=2)["T"][0]
        b = _alg(differentiable=False).icp(source, target, T0, trim_dist=5.0, loss_fn=loss_fn, dim=2)["T"][0]
        assert float(torch.linalg.norm(O.se3_log(a @ torch.linalg.inv(b)))) < TOL


def test_reference_exceptions():
    alg = _alg()
    s = torch.zeros((10, 3), dtype=torch.float64); t = torch.zeros((12, 6), dtype=torch.float64)
    T0 = torch.eye(4, dtype=torch.float64)
    with pytest.raises(ValueError):
        alg.icp(torch.zeros((10, 4), dtype=torch.float64), t, T0)
    with pytest.raises(ValueError):
        alg.icp(s, torch.zeros((12, 5), dtype=torch.float64), T0)
    with pytest.raises(ValueError):
        alg.icp(s, t, torch.eye(3, dtype=torch.float64))
    with pytest.raises(AssertionError):
        alg.icp(s, t, T0, weight=torch.ones(9, dtype=torch.float64).unsqueeze(0).repeat(2, 1))
    with pytest.raises(AssertionError):
        alg.icp([s], [t], [T0], weight=[torch.ones(10), torch.ones(10)])
    with pytest.raises(AssertionError):
        alg.icp(s, t[:, :3], T0)                       # pt2pl needs normals (ICP.py:101)
    with pytest.raises(AssertionError):
        alg.icp(s.float(), t, T0)                      # dtype mismatch (ICP.py:94)
    with pytest.raises(ValueError):
        alg.icp(s, t, T0, loss_fn={"name": "tukey", "metric": 1.0})
