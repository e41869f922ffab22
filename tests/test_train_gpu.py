"""GPU: one full ICP-weight-policy training step through the CUDA path --
UNet mask -> weight sampling (K6) -> dICP (K3-K7) -> pose + occupancy (N2) loss -> backward -> Adam."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_training_step_updates_the_unet_and_matches_the_oracle_icp():
    from mm_masking_b200 import synth, train_icp_weights as TW
    from mm_masking_b200.icp_weight_policy import LearnICPWeightPolicy
    from oracle import dicp_oracle as O
    batch = synth.make_training_batch(2, n=300, m=9000, seed=4)
    torch.manual_seed(1)
    model = LearnICPWeightPolicy({"device": "cuda", "max_iter": 4}).cuda().train()
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    before = [p.detach().clone() for p in model.parameters()]

    # the pose the policy returns equals the oracle's dICP on the weights the mask gives
    dev = {k: {kk: vv.cuda() for kk, vv in v.items()} for k, v in batch.items()}
    T_pred, mask, num_non0 = model(dev["loc_data"], dev["map_data"], dev["transforms"]["T_ml_init"])
    assert mask.shape == (2, 640, 640) and float(mask.max()) <= 1.0 + 1e-6
    from mm_masking_b200.radar_utils import extract_weights
    w = extract_weights(mask.detach(), dev["loc_data"]["raw_pc"])[0]
    t64 = lambda a: a.detach().double().cpu()
    ref = O.dicp(t64(dev["loc_data"]["filtered_pc"]), t64(dev["map_data"]["pc"]), t64(dev["transforms"]["T_ml_init"]),
                 t64(w), icp_type="pt2pl", max_iterations=4, tolerance=1e-5, differentiable=True, trim_dist=5.0,
                 loss_fn={"name": "cauchy", "metric": 1.0}, dim=2)
    assert (t64(T_pred) - ref["T"]).abs().max() < 1e-5

    loss, comp = TW.train_step(model, batch, opt, device="cuda", gt_eye=False)
    assert torch.isfinite(loss).all() and set(comp) == {"rot", "trans", "fft", "mask_pts", "cfar", "num_pts"}
    assert float(comp["mask_pts"]) > 0 and float(comp["trans"]) > 0
    grads = [p.grad for p in model.parameters()]
    assert all(g is not None and torch.isfinite(g).all() for g in grads)
    assert any(float(g.abs().max()) > 0 for g in grads)
    assert any(not torch.equal(a, b.detach()) for a, b in zip(before, model.parameters()))

    # the epoch loop: batches come through the DevicePrefetcher (host batches in, same losses out)
    from mm_masking_b200.prefetch import DevicePrefetcher, pin_batch
    cpu_batch = {k: {kk: (vv.cpu() if torch.is_tensor(vv) else vv) for kk, vv in v.items()} for k, v in batch.items()}
    staged = TW._staged([cpu_batch], "cuda", True)
    assert isinstance(staged, DevicePrefetcher) and TW._staged(staged, "cuda", True) is staged
    mean_loss, mean_comp = TW.train_policy(model, [pin_batch(cpu_batch), cpu_batch], opt, device="cuda", gt_eye=False)
    assert torch.isfinite(mean_loss).all() and set(mean_comp) == set(comp)

    model.eval()
    val = TW.validate_policy(model, [batch], gt_eye=False, device="cuda")
    assert val[0].shape == (1, 3) and torch.isfinite(val[0]).all()
    val_host = TW.validate_policy(model, [cpu_batch], gt_eye=False, device="cuda")
    assert torch.allclose(val_host[0], val[0], atol=1e-6)
