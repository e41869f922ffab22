"""Mirror of the reference caller `mm_masking/icp_weight_policy.py`.

`LearnICPWeightPolicy` keeps the reference's constructor parameters, module layout (so a reference
`state_dict` loads), forward signature and return values (`icp_weight_policy.py:24-288`).  The UNet
is plain PyTorch (it is a dense conv net, not the product); what changes underneath is

  * `extract_weights`  -> `mmb_sample_weights_fwd/bwd`      (radar_utils, K6)
  * `ICP.icp`          -> `mmb_icp_forward/backward`        (dICP, K3-K7)
  * data parallelism: each rank holds a contiguous shard of the batch; the one place the reference
    mixes samples -- the per-channel min/max (or mean/std) of the network input over the WHOLE
    batch, `:150-159` -- becomes a MIN/MAX (SUM) all-reduce of a few floats, so a sharded step
    computes what the single-process step computes.

Left out: the Neptune/matplotlib plotting of `:216-258` (logging, not the path).
"""
from __future__ import annotations

import os.path as osp

import torch
import torch.distributed as dist
import torch.nn as nn
from torch.nn import ModuleList

from .dICP.ICP import ICP
from .radar_utils import extract_weights, form_cart_range_angle_grid, form_polar_range_grid


def weights_init(m):
    """`icp_weight_policy.py:15-22`"""
    classname = m.__class__.__name__
    if classname.find('Conv') != -1:
        try:
            nn.init.xavier_uniform_(m.weight.data)
            nn.init.zeros_(m.bias)
        except AttributeError:
            print("Skipping initialization of ", classname)


DEFAULT_PARAMS = {
    # `train_icp_weights.py:357-420` defaults that matter on the path
    "icp_type": "pt2pl", "fft_input": True, "cfar_input": True, "range_input": True,
    "network_input_type": "cartesian", "network_output_type": "cartesian", "leaky": False, "dropout": 0.0,
    "batch_norm": False, "float_type": torch.float32, "device": "cuda", "init_weights": True,
    "normalize": "minmax", "log_transform": False, "a_thresh": 1.0, "b_thresh": 0.09, "gt_eye": True,
    "max_iter": 10, "loss_icp_rot_weight": 1.0, "loss_icp_trans_weight": 1.0, "norm_weights": True,
}


def _dist_on() -> bool:
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


class LearnICPWeightPolicy(nn.Module):
    def __init__(self, params):
        super().__init__()
        params = dict(DEFAULT_PARAMS, **params)
        self.res = 0.0596
        network_inputs = {"fft": params["fft_input"], "cfar": params["cfar_input"], "range": params["range_input"]}
        device = params["device"]
        self.use_ICP_4_train = params["loss_icp_rot_weight"] > 0.0 and params["loss_icp_trans_weight"] > 0.0

        config_path = osp.join(osp.dirname(osp.abspath(__file__)), 'config', 'dICP_config.yaml')
        self.ICP_alg = ICP(icp_type=params["icp_type"], config_path=config_path, differentiable=True,
                           max_iterations=params["max_iter"], tolerance=1e-5)
        self.ICP_alg_inference = ICP(icp_type=params["icp_type"], config_path=config_path, differentiable=False,
                                     max_iterations=50, tolerance=1e-5)
        self.float_type = params["float_type"]
        self.device = device
        self.network_inputs = network_inputs
        if params["network_input_type"] == 'cartesian':
            self.range_mask, _ = form_cart_range_angle_grid(device=device)
        elif params["network_input_type"] == 'polar':
            self.range_mask = form_polar_range_grid(polar_resolution=self.res, device=device)
        self.network_input_type = params["network_input_type"]
        self.network_output_type = params["network_output_type"]
        self.leaky = params["leaky"]
        self.dropout = params["dropout"]
        self.batch_norm = params["batch_norm"]
        self.normalize_type = params["normalize"]
        self.log_transform = params["log_transform"]
        self.a_thres = params["a_thresh"]
        self.b_thres = params["b_thresh"]
        self.gt_eye = params["gt_eye"]
        self.norm_weights = params["norm_weights"]

        self.mean_num_pts = 0.0
        self.max_w = 0.0
        self.min_w = 0.0
        self.mean_w = 0.0
        self.mean_all_pts = 0.0

        init_c_num = network_inputs['fft'] + network_inputs['cfar'] + network_inputs['range']
        enc_channels = [init_c_num, 8, 16, 32, 64, 128, 256]
        dec_channels = [256, 128, 64, 32, 16, 8]
        self.encoder = ModuleList([self.conv_block(enc_channels[i], enc_channels[i + 1], i)
                                   for i in range(len(enc_channels) - 1)])
        self.decoder = ModuleList([self.conv_block(dec_channels[i], dec_channels[i + 1])
                                   for i in range(len(dec_channels) - 1)])
        self.final_layer = nn.Sequential(nn.Conv2d(dec_channels[-1], 1, kernel_size=1), nn.Sigmoid())
        if params["init_weights"]:
            self.apply(weights_init)

    def conv_block(self, in_channels, out_channels, i=0):
        """`icp_weight_policy.py:104-125`"""
        relu_layer = nn.LeakyReLU(0.1) if self.leaky else nn.ReLU()
        modules = [nn.Conv2d(in_channels, out_channels, kernel_size=3, padding=1), relu_layer]
        if self.batch_norm:
            modules.append(nn.BatchNorm2d(out_channels))
        modules += [nn.Conv2d(out_channels, out_channels, kernel_size=3, padding=1), relu_layer]
        if self.batch_norm:
            modules.append(nn.BatchNorm2d(out_channels))
        if self.dropout > 0.0:
            modules.append(nn.Dropout(p=self.dropout))
        if i > 0:
            modules.append(nn.MaxPool2d(kernel_size=2, stride=2))
        return nn.Sequential(*modules)

    # -- whole-batch statistics of the network input (`:150-159`), exact under sharding ----------
    @staticmethod
    def _global_minmax(x):
        c_max, c_min = torch.max(x), torch.min(x)
        if _dist_on():
            dist.all_reduce(c_max, op=dist.ReduceOp.MAX)
            dist.all_reduce(c_min, op=dist.ReduceOp.MIN)
        return c_min, c_max

    @staticmethod
    def _global_mean_std(x):
        if not _dist_on():
            return torch.mean(x), torch.std(x)
        s = torch.stack([x.sum(dtype=torch.float64), (x.double() ** 2).sum(), torch.tensor(float(x.numel()), device=x.device, dtype=torch.float64)])
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
        mean = s[0] / s[2]
        var = (s[1] - s[2] * mean * mean) / (s[2] - 1)           # unbiased, like torch.std
        return mean.to(x.dtype), var.clamp_min(0).sqrt().to(x.dtype)

    def unet(self, input_data):
        """Encoder / decoder with the reference's skip wiring (`:161-186`)."""
        enc_layers = []
        for layer in self.encoder:
            enc_layers.append(input_data)
            input_data = layer(input_data)
        enc_layers.reverse()
        for i, decoder_layer in enumerate(self.decoder):
            skip_con = enc_layers[i]
            input_data = nn.functional.interpolate(input_data, size=(skip_con.shape[2], skip_con.shape[3]),
                                                   mode='bilinear', align_corners=True)      # UpsamplingBilinear2d
            input_data = decoder_layer(input_data)
            input_data = torch.cat([enc_layers[i], input_data], dim=1)
            input_data = decoder_layer(input_data)
        return self.final_layer(input_data).squeeze(1)

    def forward(self, batch_scan, batch_map, T_init, binary=False, override_mask=None, neptune_run=None, epoch=0,
                batch_idx=0, mask_only=False):
        fft_data = batch_scan['fft_data'].to(self.device)
        fft_cfar = batch_scan['fft_cfar'].to(self.device)
        scan_pc_raw = batch_scan['raw_pc'].to(self.device)
        map_pc = batch_map['pc'].to(self.device)

        if override_mask is None:
            input_data = None
            if self.network_inputs['fft']:
                input_data = fft_data.unsqueeze(1)
            if self.network_inputs['cfar']:
                input_data = torch.cat([input_data, fft_cfar.unsqueeze(1)], dim=1)
            if self.network_inputs['range']:
                range_stack = self.range_mask.to(input_data.dtype).expand(input_data.shape[0], 1, -1, -1)
                input_data = torch.cat([input_data, range_stack], dim=1)
            if self.log_transform:
                input_data = torch.log(input_data + 1e-6)
            chans = []
            for c in range(input_data.shape[1]):
                x = input_data[:, c, :, :]
                if "minmax" in self.normalize_type:
                    c_min, c_max = self._global_minmax(x)
                    x = (x - c_min) / (c_max - c_min)
                elif "standardize" in self.normalize_type:
                    c_mean, c_std = self._global_mean_std(x)
                    x = (x - c_mean) / c_std
                chans.append(x)
            weight_mask = self.unet(torch.stack(chans, dim=1))
        else:
            weight_mask = override_mask

        if self.norm_weights:
            weight_mask = weight_mask / torch.amax(weight_mask, dim=(1, 2), keepdim=True)
        if binary:
            weight_mask = torch.where(weight_mask > 0.5, 1.0, 0.0)
        if mask_only:
            return weight_mask

        weights, diff_mean_num_non0, mean_num_non0, mean_w, max_w, min_w = extract_weights(weight_mask, scan_pc_raw)
        self.mean_num_pts = mean_num_non0
        self.max_w = max_w
        self.min_w = min_w
        self.mean_w = mean_w
        non0_pts = (scan_pc_raw[:, :, 0] != 0.0) * (scan_pc_raw[:, :, 1] != 0.0)
        self.mean_all_pts = torch.sum(non0_pts) / scan_pc_raw.shape[0]

        scan_pc_filt = batch_scan['filtered_pc'].to(self.device)
        if self.training and not self.use_ICP_4_train:
            return T_init, weight_mask, diff_mean_num_non0
        T_est = self.icp(scan_pc_filt, map_pc, T_init, weights)
        return T_est, weight_mask, diff_mean_num_non0

    def icp(self, scan_pc, map_pc, T_init, weights):
        """`icp_weight_policy.py:277-288`"""
        loss_fn = {"name": "cauchy", "metric": 1.0}
        trim_dist = 5.0
        alg = self.ICP_alg if self.training else self.ICP_alg_inference
        icp_result = alg.icp(scan_pc, map_pc, T_init=T_init, weight=weights, trim_dist=trim_dist, loss_fn=loss_fn, dim=2)
        return icp_result['T']
