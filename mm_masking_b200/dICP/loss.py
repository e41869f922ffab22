"""Robust-weight functions with the reference's interface (`external/dICP/dICP/loss.py:4-56`).

Inside `ICP.icp` these weights are evaluated in the fused CUDA step kernel
(`csrc/icp_kernels.cu: row_weight`); this class exists so code that builds a `loss` object
directly keeps working.  It is elementwise tensor arithmetic on whatever device `err` lives on.
"""
import torch


class loss:
    def __init__(self, name="huber", metric=None, differentiable=True, tanh_steepness=10.0):
        self.name = name
        self.metric = metric
        self.differentiable = differentiable
        self.tanh_steepness = tanh_steepness

    def get_weight(self, err):
        if self.name == "huber":
            if self.differentiable:
                return self.metric ** 2 / (self.metric ** 2 + err ** 2)
            return torch.where(err.abs() > self.metric, self.metric / err.abs(), torch.ones_like(err))
        if self.name == "cauchy":
            return 1.0 / (1.0 + (err / self.metric) ** 2)
        if self.name == "trim":
            axis = 1 if err.dim() == 2 else 2
            rho = torch.linalg.norm(err, dim=axis)
            if self.differentiable:
                return 0.5 * torch.tanh(self.tanh_steepness * (self.metric - rho.squeeze()) - 3.0) + 0.5
            shape = (err.shape[0], 1) if err.dim() == 2 else (err.shape[0], err.shape[1])
            return torch.where(rho > self.metric, 0.0, 1.0).to(err.dtype).reshape(shape)
        raise ValueError("Invalid loss name: {}".format(self.name))
