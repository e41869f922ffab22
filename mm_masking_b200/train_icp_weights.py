"""Mirror of the training / validation step of `mm_masking/train_icp_weights.py`:
`train_policy` (:22-69), `validate_policy` (:71-176, without the Neptune figures),
`eval_training_loss` (:179-253) and `eval_validation_loss` (:255-274), same signatures and
return values, over the CUDA path of this package.

Data parallelism: one process per GPU, each rank iterates over ITS shard of every
batch; the pairs never mix on the data path, losses are batch means, so equal shards + the mean of
the gradients over ranks (`dist.allreduce_mean_`, one flat NCCL all-reduce of the ~1.8 M UNet
parameters) reproduce the single-process step.
"""
from __future__ import annotations

import torch

from . import dist as mdist
from .prefetch import DevicePrefetcher
from .radar_utils import extract_bev_from_pts

DEFAULT_LOSS_WEIGHTS = {"icp_rot": 1.0, "icp_trans": 1.0, "fft": 0.0, "mask_pts": 1.0, "cfar": 0.0, "num_pts": 0.0}


def eval_training_loss(T_pred, mask, num_non0, batch_T_gt, batch_scan, batch_map, model, loss_weights=DEFAULT_LOSS_WEIGHTS,
                       icp_loss_only_iter=0, gt_eye=True, epoch=0):
    """`train_icp_weights.py:179-253`"""
    mask_criterion = torch.nn.BCELoss()
    z = lambda: torch.zeros(1, dtype=T_pred.dtype, device=T_pred.device)
    loss_rot, loss_trans, loss_fft, loss_mask_pts, loss_cfar, loss_num_pts = z(), z(), z(), z(), z(), z()

    if loss_weights['icp_rot'] > 0.0 or loss_weights['icp_trans'] > 0.0:
        eye = torch.eye(4, dtype=T_pred.dtype, device=T_pred.device)
        if gt_eye:
            xi_wedge = T_pred - eye
        else:
            xi_wedge = torch.matmul(T_pred, torch.inverse(batch_T_gt)) - eye
        xi_r = xi_wedge[:, 0:2, 3]
        xi_theta = xi_wedge[:, 1, 0].unsqueeze(-1)
        loss_rot = torch.norm(xi_theta, dim=1).mean()
        loss_trans = torch.norm(xi_r, dim=1).mean()
    if icp_loss_only_iter <= 0 or (icp_loss_only_iter > 0 and epoch < icp_loss_only_iter) or \
            (loss_weights['icp_rot'] <= 0 and loss_weights['icp_trans'] <= 0):
        if loss_weights['fft'] > 0.0:
            fft_data = batch_scan['fft_data'].to(mask.device)
            mean_azimuth = torch.mean(fft_data, dim=(1, 2), keepdim=True)
            fft_mask = torch.where(fft_data > 3.0 * mean_azimuth, torch.ones_like(fft_data), torch.zeros_like(fft_data))
            loss_fft = mask_criterion(mask, fft_mask)
        if loss_weights['cfar'] > 0.0:
            loss_cfar = mask_criterion(mask, batch_scan['fft_cfar'].to(mask.device))
        if loss_weights['mask_pts'] > 0.0:
            map_pts_mask = extract_bev_from_pts(batch_map['pc'].to(mask.device))       # N2 kernel
            loss_mask_pts = mask_criterion(mask, map_pts_mask)
        if loss_weights['num_pts'] > 0.0:
            loss_num_pts = model.mean_all_pts - num_non0

    loss = loss_weights['icp_rot'] * loss_rot + loss_weights['icp_trans'] * loss_trans \
        + loss_weights['fft'] * loss_fft + loss_weights['mask_pts'] * loss_mask_pts \
        + loss_weights['cfar'] * loss_cfar + loss_weights['num_pts'] * loss_num_pts
    comp = {"rot": (loss_weights['icp_rot'] * loss_rot).detach(), "trans": (loss_weights['icp_trans'] * loss_trans).detach(),
            "fft": (loss_weights['fft'] * loss_fft).detach(), "mask_pts": (loss_weights['mask_pts'] * loss_mask_pts).detach(),
            "cfar": (loss_weights['cfar'] * loss_cfar).detach(), "num_pts": (loss_weights['num_pts'] * loss_num_pts).detach()}
    return loss, comp


def eval_validation_loss(T_pred, batch_T_gt, gt_eye=True):
    """`train_icp_weights.py:255-274`: (norm, rotation, translation) error means."""
    eye = torch.eye(4, dtype=T_pred.dtype, device=T_pred.device)
    xi_wedge = T_pred - eye if gt_eye else torch.matmul(T_pred, torch.inverse(batch_T_gt)) - eye
    xi_r = xi_wedge[:, 0:2, 3]
    xi_theta = xi_wedge[:, 1, 0].unsqueeze(-1)
    xi_stack = torch.cat((xi_theta, xi_r), dim=1)
    return torch.hstack((torch.norm(xi_stack, dim=1).mean(), torch.norm(xi_theta, dim=1).mean(),
                         torch.norm(xi_r, dim=1).mean()))


def train_step(model, batch, opt, loss_weights=DEFAULT_LOSS_WEIGHTS, device='cuda', epoch=None, icp_loss_only_iter=0,
               gt_eye=True):
    """One optimiser step on one (already sharded) batch: the body of the loop of `train_policy`
    (`train_icp_weights.py:31-63`) plus the gradient mean over ranks.  Returns (loss, components).

    Under `torch.distributed` the step equals the single-process step on the whole batch also when the
    shards are unequal (batch % world != 0): the reference's losses are batch means
    (`train_icp_weights.py:199-200,226`), so the local loss is weighted by n_local * world / n_global before
    backward and the plain mean of the gradients over ranks is then the gradient of the global mean.  The
    UNet gradients are all-reduced bucket by bucket from gradient-ready hooks while backward is still running
    (`dist.BucketedGradReducer`; every parameter takes part on every rank, missing gradients zero-filled, so
    the collectives line up), and the returned loss is the mean over the WHOLE batch."""
    batch_scan, batch_map, batch_T = batch['loc_data'], batch['map_data'], batch['transforms']
    batch_T_init = batch_T['T_ml_init'].to(device)
    opt.zero_grad()
    T_pred, mask, num_non0 = model(batch_scan, batch_map, batch_T_init)
    batch_T_gt = batch_T['T_ml_gt'].to(device)
    loss, loss_comp = eval_training_loss(T_pred, mask, num_non0, batch_T_gt, batch_scan, batch_map, model,
                                         loss_weights=loss_weights, icp_loss_only_iter=icp_loss_only_iter, gt_eye=gt_eye,
                                         epoch=epoch if epoch is not None else 0)
    n_local = int(batch_T_init.shape[0])
    reducer = None
    if mdist.world()[1] > 1:
        reducer = getattr(model, "_mmb_grad_reducer", None)
        if reducer is None:
            reducer = model._mmb_grad_reducer = mdist.BucketedGradReducer(list(model.parameters()))
        (loss * mdist.shard_loss_scale(n_local, device=loss.device)).backward()
        reducer.finish()
    else:
        loss.backward()
    opt.step()
    return mdist.global_mean(loss.detach(), n_local), loss_comp


def _staged(iterator, device, prefetch):
    """with a CUDA device the batches come through `DevicePrefetcher`: batch k+1 uploads on a copy stream
    under the kernels of batch k, and the `.to(device)` calls of the loop body become no-ops"""
    if prefetch and torch.device(device).type == 'cuda' and not isinstance(iterator, DevicePrefetcher):
        return DevicePrefetcher(iterator, device)
    return iterator


def train_policy(model, iterator, opt, loss_weights=DEFAULT_LOSS_WEIGHTS, device='cpu', epoch=None, icp_loss_only_iter=0,
                 gt_eye=True, prefetch=True):
    """`train_icp_weights.py:22-69`"""
    model.train()
    loss_hist = 0.0
    loss_comp_hist = []
    count = 0
    for batch in _staged(iterator, device, prefetch):
        loss, loss_comp = train_step(model, batch, opt, loss_weights, device, epoch, icp_loss_only_iter, gt_eye)
        loss_hist += loss
        loss_comp_hist.append(loss_comp)
        count += 1
    mean_loss = loss_hist / count
    mean_loss_comp = {k: sum(d[k] for d in loss_comp_hist) / len(loss_comp_hist) for k in loss_comp_hist[0].keys()}
    return mean_loss, mean_loss_comp


def validate_policy(model, iterator, gt_eye=True, device='cpu', binary=False, neptune_run=None, epoch=None,
                    prefetch=True):
    """`train_icp_weights.py:71-176` without the Neptune figures."""
    model.eval()
    val_acc = torch.zeros((1, 3), device=device)
    mean_num_pc, max_w, min_w, mean_w, count = 0.0, 0.0, 1000.0, 0.0, 0
    with torch.no_grad():
        for batch in _staged(iterator, device, prefetch):
            batch_T = batch['transforms']
            T_pred, mask, _ = model(batch['loc_data'], batch['map_data'], batch_T['T_ml_init'].to(device), binary=binary)
            mean_num_pc += model.mean_num_pts
            max_w = model.max_w if model.max_w > max_w else max_w
            min_w = model.min_w if model.min_w < min_w else min_w
            mean_w += model.mean_w
            val_acc += eval_validation_loss(T_pred, batch_T['T_ml_gt'].to(device), gt_eye=gt_eye)
            count += 1
    return val_acc / count, mean_num_pc / count, mean_w / count, max_w, min_w
