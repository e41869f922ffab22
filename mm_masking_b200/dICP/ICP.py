"""Batched weighted differentiable ICP with the reference's module surface.

Mirrors `external/dICP/dICP/ICP.py` of utiasASRL/mm_masking: same constructor, same
`icp(source, target, T_init, weight, trim_dist, loss_fn, dim)` call, same result dict
(`ICP.py:284-291`), same exceptions for malformed input.  The arithmetic runs in the CUDA
library (`csrc/icp_kernels.cu`) through the C ABI of `include/mm_masking_b200.h`:

  icp()  ->  batch_size_handling()  (host: ragged lists -> dense padded tensors, ICP.py:295-497)
         ->  torch.ops.mmb.icp_forward   (custom op, mm_masking_b200/ops.py)
               forward : mmb_icp_forward  -- all iterations enqueued without host sync
               autograd: torch.ops.mmb.icp_backward -> mmb_icp_backward -- adjoint sweep over the saved iterations

CPU tensors are accepted for drop-in use (the reference's tests pass CPU tensors): they are
moved to the current CUDA device, and results are returned on the inputs' device.  Without a
CUDA device or the built library the call raises; there is no CPU implementation here.
"""
from __future__ import annotations

import os.path as osp

import torch
import yaml

from .. import _capi
from .. import ops as _ops  # noqa: F401  (registers torch.ops.mmb.*)
from .nn import nn

_LOSS_KIND = {"huber": _capi.LOSS_HUBER, "cauchy": _capi.LOSS_CAUCHY}


class ICP:
    def __init__(self, config_path=None, icp_type='pt2pl', max_iterations=100, tolerance=1e-12, differentiable=True):
        if config_path is None:
            config_path = osp.join(osp.dirname(osp.abspath(__file__)), '..', 'config', 'dICP_config.yaml')
        with open(config_path, 'r') as f:
            self.config = yaml.safe_load(f)

        self.icp_type = icp_type
        self.max_iterations = max_iterations
        self.tolerance = tolerance
        self.const_iter = self.config['dICP']['parameters']['const_iter']
        self.verbose = self.config['dICP']['logging']['verbose']
        self.target_pad_val = self.config['dICP']['parameters']['target_pad_val']
        self.match_ratio_thresh = self.config['dICP']['logging']['matched_ratio_thresh']
        self.diff = differentiable
        self.nn_slices = 0          # 0: the library picks the target split of the NN search
        self.nn_backend = _capi.NN_AUTO   # exact cell grid for large maps, brute force otherwise (same indices)
        self.record_indices = False       # tests: keep the per-iteration correspondences in `last_indices` (needs a grad input)
        self.last_indices = None

        use_gumbel = self.config['dICP']['functionality']['gumbel']
        if use_gumbel:
            raise NotImplementedError("config functionality.gumbel=True: the soft nearest neighbour "
                                      "(reference nn.py:43-70, 'not yet integrated') is outside the CUDA path")
        self.nn = nn(self.diff, use_gumbel=use_gumbel,
                     eps=self.config['dICP']['functionality']['gumbel_eps'],
                     tau=self.config['dICP']['functionality']['gumbel_tau'])

    def icp(self, source, target, T_init, weight=None, trim_dist=None, loss_fn=None, dim=3):
        return self.dICP(source, target, T_init, weight, trim_dist, loss_fn, dim)

    def dICP(self, source, target, T_init, weight=None, trim_dist=None, loss_fn=None, dim=3):
        """Same contract as the reference's `ICP.dICP` (`ICP.py:48-293`).

        Returns {"pc" (N,n,3), "T" (N,4,4), "costs" (N,it,1), "deltas" (N,it,6,1),
        "weights" (N,it,R,1), "stats": {"converged", "iterations", "matched_ratio"}} with
        R = n for pt2pl and 3n for pt2pt, `it` the number of iterations executed.
        """
        _capi.require_cuda()
        source, target, T_init, w_init, has_weight = self.batch_size_handling(source, target, T_init, weight)
        assert source.dtype == target.dtype == T_init.dtype
        if source.dtype not in (torch.float32, torch.float64):
            raise TypeError("source/target/T_init must be float32 or float64")
        if self.icp_type == 'pt2pl':
            assert target.shape[2] == 6
        if dim not in (2, 3):
            raise ValueError("dim must be 2 or 3")
        loss_kind, loss_metric = _capi.LOSS_NONE, 1.0
        if loss_fn is not None:
            if loss_fn['name'] not in _LOSS_KIND:
                raise ValueError("Invalid loss name: {}".format(loss_fn['name']))
            loss_kind, loss_metric = _LOSS_KIND[loss_fn['name']], loss_fn['metric']
        if source.shape[0] != target.shape[0] or source.shape[0] != T_init.shape[0]:
            raise ValueError("source, target and T_init must have the same batch size")

        out_device = source.device
        dev = out_device if out_device.type == 'cuda' else torch.device('cuda', torch.cuda.current_device())
        weight_arg = None
        if has_weight:
            weight_arg = w_init.to(device=dev, dtype=source.dtype)
        src_d, tgt_d, T0_d = source.to(dev), target.to(dev), T_init.to(dev)
        needs_grad = torch.is_grad_enabled() and any(
            t is not None and t.requires_grad for t in (src_d, tgt_d, T0_d, weight_arg))
        use_trim = trim_dist is not None and trim_dist >= 0.0
        K = int(self.max_iterations)
        icp_code = _capi.PT2PT if self.icp_type == 'pt2pt' else _capi.PT2PL
        T, pc, block, saved = torch.ops.mmb.icp_forward(
            src_d, tgt_d, T0_d, weight_arg,
            icp_code, int(dim), K, float(self.tolerance),
            bool(self.diff), bool(self.const_iter), bool(use_trim), float(trim_dist) if use_trim else 0.0,
            int(loss_kind), float(loss_metric), float(self.config['dICP']['parameters']['tanh_steepness']),
            float(self.match_ratio_thresh), int(self.nn_slices), int(self.nn_backend), bool(needs_grad))
        deltas, weights, costs, converged, iterations, matched, iters_done = _ops._carve(
            block.detach(), _ops.icp_history_specs(src_d.shape[0], src_d.shape[1], K, icp_code, src_d.dtype))

        # test hook: the correspondences of every iteration, (K,N,n) int32 -- the first block of `saved`
        self.last_indices = None
        if self.record_indices and saved.numel():
            N_, n_ = src_d.shape[0], src_d.shape[1]
            self.last_indices = saved.detach()[:K * N_ * n_ * 4].view(torch.int32).view(K, N_, n_)
        if self.const_iter:
            it = int(self.max_iterations)    # every pair runs all iterations (dICP_config.yaml:7): no host sync at all
        else:
            it = int(iters_done.item())      # the one host sync of the call: history lengths (ICP.py:273-275)
        if self.verbose:
            print("ICP converged in {} iterations".format(it))
        back = (lambda t: t) if out_device == dev else (lambda t: t.to(out_device))
        return {
            "pc": back(pc),
            "T": back(T),
            "costs": back(costs[:, :it].unsqueeze(-1)),
            "deltas": back(deltas[:, :it].unsqueeze(-1)),
            "weights": back(weights[:, :it].unsqueeze(-1)),
            "stats": {
                "converged": back(converged.to(torch.bool)),
                "iterations": back(iterations),
                "matched_ratio": back(matched),
            },
        }

    def batch_size_handling(self, source, target, T_init=None, weight=None):
        """Normalise tensor / batched tensor / ragged list inputs to dense padded tensors.

        Follows the rules of the reference (`ICP.py:295-497`): sources are zero-padded and their
        padding gets weight 0; targets are padded with max(source) * target_pad_val so padding is
        never a nearest neighbour; an empty source or target becomes a one-point cloud with zero
        weight.  Returns (source (N,n,3), target (N,m,3|6), T_init (N,4,4), weight (N,n), has_weight)
        where `has_weight` is False when every prior weight is implicitly 1.  Unlike the reference
        the per-point weights are not repeated for pt2pt (the kernels index rows per point).
        """
        if weight is not None:
            if isinstance(source, list):
                assert len(source) == len(weight), "weight must be list of same length as source"
            else:
                assert source.shape[0] == weight.shape[0], "weight must have same number of rows as source"

        if source is None or target is None or len(source) == 0 or len(target) == 0:
            z = torch.zeros
            return (z((1, 1, 3)), z((1, 1, 6)), torch.eye(4).unsqueeze(0), z((1, 1)), True)

        ref = None
        for t_i in (target if isinstance(target, list) else [target]):
            if t_i is not None and len(t_i) > 0:
                ref = t_i
                break
        pt_device = ref.device if ref is not None else torch.device('cpu')
        pt_dtype = ref.dtype if ref is not None else torch.float32
        target_dim = ref.shape[-1] if ref is not None else 6
        has_weight = weight is not None

        # ---- source and prior weights ------------------------------------------------------------
        if isinstance(source, list):
            has_weight = True                   # ragged rows need the zero weights of the padding
            clouds, priors = [], []
            for i, s_i in enumerate(source):
                empty = len(s_i) == 0
                if empty:
                    s_i = torch.zeros((1, 3), dtype=pt_dtype, device=pt_device)
                elif i > 0 and (s_i.dim() != 2 or s_i.shape[1] not in (3, 6)):
                    raise ValueError("source list must contain (n x 3/6) tensors")
                w_i = torch.ones(s_i.shape[0], dtype=pt_dtype, device=pt_device)
                if weight is not None and not empty and weight[i] is not None:
                    assert len(weight[i]) == s_i.shape[0], "weight must have same number of rows as source"
                    w_i = weight[i] * w_i
                if empty:
                    w_i = 0.0 * w_i
                clouds.append(s_i[:, :3])
                priors.append(w_i)
            n_max = max(c.shape[0] for c in clouds)
            source_batch = torch.stack([torch.cat((c, c.new_zeros((n_max - c.shape[0], 3)))) for c in clouds])
            w = torch.stack([torch.cat((p_i, p_i.new_zeros(n_max - p_i.shape[0]))) for p_i in priors])
        elif source.dim() == 2 and source.shape[1] in (3, 6):
            source_batch = source[:, :3].unsqueeze(0)
            w = weight.unsqueeze(0) if weight is not None else None
        elif source.dim() == 3 and source.shape[2] in (3, 6):
            source_batch = source[:, :, :3]
            w = weight
        else:
            raise ValueError("source must be (n x 3/6) or (N x n x 3/6) or list len(N) (n_N x 3/6)")

        # ---- target ------------------------------------------------------------------------------
        if isinstance(target, list):
            pad_value = torch.max(source_batch) * self.target_pad_val
            rows = []
            for i, t_i in enumerate(target):
                if len(t_i) == 0:
                    t_i = torch.zeros((1, target_dim), dtype=pt_dtype, device=pt_device)
                    if w is None:
                        w = torch.ones(source_batch.shape[:2], dtype=pt_dtype, device=pt_device)
                        has_weight = True
                    w = w.clone()
                    w[i, :] = 0.0
                elif i > 0 and (t_i.dim() != 2 or t_i.shape[1] != target_dim):
                    raise ValueError("target list must contain (m x 3/6) tensors. All tensors must have same number of columns")
                rows.append(t_i)
            m_max = max(t.shape[0] for t in rows)
            target_batch = torch.stack([
                torch.cat((t, pad_value.to(t.dtype) * t.new_ones((m_max - t.shape[0], t.shape[1])))) for t in rows])
        elif target.dim() == 2 and target.shape[1] in (3, 6):
            target_batch = target.unsqueeze(0)
        elif target.dim() == 3 and target.shape[2] in (3, 6):
            target_batch = target
        else:
            raise ValueError("target must be (m x 3/6) or (N x m x 3/6) or list len(N) (m_N x 3/6)")

        # ---- initial transform -------------------------------------------------------------------
        if T_init is None:
            T_init_batch = None
        elif isinstance(T_init, list):
            T_init_batch = torch.stack(T_init, dim=0)
        elif T_init.shape == (4, 4):
            T_init_batch = T_init.unsqueeze(0)
        elif T_init.dim() == 3 and T_init.shape[1:] == (4, 4):
            T_init_batch = T_init
        else:
            raise ValueError("T_init must be (4 x 4) or (N x 4 x 4) or list len(N) (4 x 4)")

        if w is None:
            w = torch.ones(source_batch.shape[:2], dtype=pt_dtype, device=pt_device)
        return source_batch, target_batch, T_init_batch, w, has_weight
