"""SURVEY.md 8f row N1 on CPU: the policy mirror (UNet) and the training / validation losses against
outputs of the UNMODIFIED reference (oracle/make_golden_train.py -> tests/golden/train_step.npz).
The CUDA ops of the step (weight sampling, ICP, occupancy BEV) are covered by the gpu tests; here
`extract_bev_from_pts` is the oracle's restatement."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import make_golden_train as MG
from oracle import radar_oracle as RO


@pytest.fixture(scope="module")
def g():
    return load_golden("train_step")[0]


def test_policy_mirror_reproduces_reference_unet(g):
    from mm_masking_b200.icp_weight_policy import LearnICPWeightPolicy
    d = MG.make_inputs()
    torch.manual_seed(7)
    model = LearnICPWeightPolicy(MG.PARAMS)
    model.eval()
    sd = model.state_dict()
    assert list(sd.keys()) == [str(k) for k in g["sd_keys"]]                 # a reference checkpoint loads
    sums = np.array([[float(v.double().sum()), float(v.double().abs().sum())] for v in sd.values()])
    if not np.allclose(sums, g["sd_sums"], rtol=1e-12, atol=0):
        pytest.skip("torch RNG stream differs from the build container: parameters cannot be reproduced from the seed")
    scan = {"fft_data": torch.from_numpy(d["fft"]).clone(), "fft_cfar": torch.from_numpy(d["cfar"]).clone(),
            "raw_pc": torch.zeros(2, 4, 3), "filtered_pc": torch.zeros(2, 4, 3)}
    with torch.no_grad():
        mask = model(scan, {"pc": torch.zeros(2, 4, 6)}, torch.eye(4).repeat(2, 1, 1), mask_only=True).numpy()
    np.testing.assert_allclose(mask, g["unet_mask"], atol=2e-6, rtol=0)


@pytest.mark.parametrize("i", [0, 1])
@pytest.mark.parametrize("gt_eye", [True, False])
def test_training_and_validation_loss(g, i, gt_eye, monkeypatch):
    from mm_masking_b200 import train_icp_weights as TW
    monkeypatch.setattr(TW, "extract_bev_from_pts", lambda pc: torch.from_numpy(RO.extract_bev_from_pts(pc.numpy())))
    d = MG.make_inputs()
    scan = {"fft_data": torch.from_numpy(d["fft640"]), "fft_cfar": torch.from_numpy(d["cfar640"])}
    T = torch.from_numpy(d["T_pred"]).clone().requires_grad_(True)
    m = torch.from_numpy(d["mask"]).clone().requires_grad_(True)
    loss, comp = TW.eval_training_loss(T, m, torch.tensor(100.25), torch.from_numpy(d["T_gt"]), scan,
                                       {"pc": torch.from_numpy(d["pts"])}, MG._Model, loss_weights=MG.WEIGHTS[i], gt_eye=gt_eye)
    loss.sum().backward()
    tag = f"w{i}_{'eye' if gt_eye else 'gt'}"
    np.testing.assert_allclose(loss.detach().numpy().reshape(-1), g[tag + "_loss"], rtol=1e-6)
    np.testing.assert_allclose([float(comp[k]) for k in ("rot", "trans", "fft", "mask_pts", "cfar", "num_pts")],
                               g[tag + "_comp"], rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(T.grad.numpy(), g[tag + "_gT"], rtol=1e-5, atol=1e-8)
    gm = m.grad.numpy()
    np.testing.assert_allclose([gm.sum(), np.abs(gm).sum(), (gm * d["mask"]).sum()], g[tag + "_gmask_sum"], rtol=1e-4, atol=1e-7)
    val = TW.eval_validation_loss(torch.from_numpy(d["T_pred"]), torch.from_numpy(d["T_gt"]), gt_eye=gt_eye).numpy()
    np.testing.assert_allclose(val, g[tag + "_val"], rtol=1e-6)


def test_epoch_loops_stage_batches_only_for_cuda_devices():
    from mm_masking_b200 import train_icp_weights as TW
    it = [{"a": torch.zeros(2)}]
    assert TW._staged(it, "cpu", True) is it
    assert TW._staged(it, "cuda", False) is it
