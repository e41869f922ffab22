"""Batched weighted differentiable ICP with the reference's module surface.

Mirrors `external/dICP/dICP/ICP.py` of utiasASRL/mm_masking: same constructor, same
`icp(source, target, T_init, weight, trim_dist, loss_fn, dim)` call, same result dict
(`ICP.py:284-291`), same exceptions for malformed input.  The arithmetic runs in the CUDA
library (`csrc/icp_kernels.cu`) through the C ABI of `include/mm_masking_b200.h`:

  icp()  ->  batch_size_handling()  (host: ragged lists -> dense padded tensors, ICP.py:295-497)
         ->  _IcpFunction.apply     (torch.autograd.Function)
               forward : mmb_icp_forward  -- all iterations enqueued without host sync
               backward: mmb_icp_backward -- adjoint sweep over the saved iterations

CPU tensors are accepted for drop-in use (the reference's tests pass CPU tensors): they are
moved to the current CUDA device, and results are returned on the inputs' device.  Without a
CUDA device or the built library the call raises; there is no CPU implementation here.
"""
from __future__ import annotations

import os.path as osp

import torch
import yaml

from .. import _capi
from .nn import nn

_LOSS_KIND = {"huber": _capi.LOSS_HUBER, "cauchy": _capi.LOSS_CAUCHY}


def _zeros_block(device, specs):
    """zero tensors of the given (shape, dtype) list as 256-byte aligned views of one zero-filled buffer"""
    spans, total = [], 0
    for shape, dt in specs:
        count = 1
        for d in shape:
            count *= int(d)
        nbytes = count * torch.empty((), dtype=dt).element_size()
        spans.append((total, nbytes))
        total += (nbytes + 255) // 256 * 256
    flat = torch.zeros(total, dtype=torch.uint8, device=device)
    return [flat[o:o + nb].view(dt).view(shape) for (o, nb), (shape, dt) in zip(spans, specs)]


class _IcpFunction(torch.autograd.Function):
    """source (N,n,3), target (N,m,3|6), T_init (N,4,4), weight (N,n) or None -> T, pc (+ histories)."""

    @staticmethod
    def forward(ctx, source, target, T_init, weight, cfg):
        lib = _capi.load_library()
        dev = source.device
        dtype = source.dtype
        N, n, _ = source.shape
        m, cols = target.shape[1], target.shape[2]
        K = cfg["max_iterations"]
        rows = 3 * n if cfg["icp_type"] == "pt2pt" else n
        needs_grad = any(t is not None and t.requires_grad for t in (source, target, T_init, weight))

        p = _capi.IcpParams()
        p.N, p.n, p.m, p.target_cols = N, n, m, cols
        p.dtype = _capi.F32 if dtype == torch.float32 else _capi.F64
        p.icp_type = _capi.PT2PT if cfg["icp_type"] == "pt2pt" else _capi.PT2PL
        p.dim = cfg["dim"]
        p.max_iterations = K
        p.differentiable = int(cfg["differentiable"])
        p.const_iter = int(cfg["const_iter"])
        p.use_trim = int(cfg["trim_dist"] is not None and cfg["trim_dist"] >= 0.0)
        p.trim_dist = float(cfg["trim_dist"]) if p.use_trim else 0.0
        p.loss_kind = cfg["loss_kind"]
        p.loss_metric = float(cfg["loss_metric"])
        p.has_weight = int(weight is not None)
        p.save_for_backward = int(needs_grad)
        p.nn_slices = int(cfg.get("nn_slices", 0))
        p.nn_backend = int(cfg.get("nn_backend", _capi.NN_AUTO))
        p.tolerance = float(cfg["tolerance"])
        p.tanh_steepness = float(cfg["tanh_steepness"])
        p.match_ratio_thresh = float(cfg["match_ratio_thresh"])

        src = source.detach().contiguous()
        tgt = target.detach().contiguous()
        T0 = T_init.detach().contiguous()
        w = None if weight is None else weight.detach().contiguous()

        with torch.cuda.device(dev):
            ws_bytes = lib.mmb_icp_workspace_bytes(p)
            sv_bytes = lib.mmb_icp_saved_bytes(p)
            if ws_bytes == 0:
                raise _capi.MmbError("mmb_icp_workspace_bytes rejected the parameters")
            workspace = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
            saved = torch.empty(sv_bytes, dtype=torch.uint8, device=dev) if sv_bytes else None
            T_out = torch.empty((N, 4, 4), dtype=dtype, device=dev)
            pc = torch.empty((N, n, 3), dtype=dtype, device=dev)
            # the zero-initialised outputs (rows of iterations that are not executed stay zero) are carved
            # from ONE zero-filled block: one fill launch instead of seven
            deltas, weights, costs, converged, iterations, matched, iters_done = _zeros_block(dev, (
                ((N, K, 6), dtype), ((N, K, rows), dtype), ((N, K), dtype), ((N,), torch.uint8), ((N,), dtype),
                ((N,), dtype), ((1,), torch.int32)))
            rc = lib.mmb_icp_forward(p, _capi.ptr(src), _capi.ptr(tgt), _capi.ptr(T0), _capi.ptr(w),
                                     _capi.ptr(T_out), _capi.ptr(pc), _capi.ptr(deltas), _capi.ptr(weights),
                                     _capi.ptr(costs), _capi.ptr(converged), _capi.ptr(iterations),
                                     _capi.ptr(matched), _capi.ptr(iters_done), _capi.ptr(saved),
                                     _capi.ptr(workspace), ws_bytes, _capi.stream_ptr())
            _capi.check(rc, "mmb_icp_forward")

        if cfg.get("record_indices") and saved is not None:
            # test hook: the correspondences of every executed iteration, (K,N,n) int32 -- the first block of `saved`
            cfg["indices_out"] = saved[:K * N * n * 4].view(torch.int32).view(K, N, n)
        ctx.params = p
        ctx.has_weight = weight is not None
        ctx.save_for_backward(src, tgt, w if w is not None else src.new_empty(0), saved if saved is not None else src.new_empty(0))
        ctx.mark_non_differentiable(deltas, weights, costs, converged, iterations, matched, iters_done)
        return T_out, pc, deltas, weights, costs, converged, iterations, matched, iters_done

    @staticmethod
    def backward(ctx, g_T, g_pc, *unused):
        lib = _capi.load_library()
        src, tgt, w, saved = ctx.saved_tensors
        p = ctx.params
        dev = src.device
        need_src, need_tgt, need_T, need_w = ctx.needs_input_grad[:4]
        need_w = need_w and ctx.has_weight
        g_T = None if g_T is None else g_T.contiguous()
        g_pc = None if g_pc is None else g_pc.contiguous()
        with torch.cuda.device(dev):
            g_src = torch.empty_like(src) if need_src else None
            g_tgt = torch.zeros_like(tgt) if need_tgt else None
            g_T0 = torch.empty((p.N, 4, 4), dtype=src.dtype, device=dev) if need_T else None
            g_w = torch.empty((p.N, p.n), dtype=src.dtype, device=dev) if need_w else None
            rc = lib.mmb_icp_backward(p, _capi.ptr(src), _capi.ptr(tgt), _capi.ptr(w if ctx.has_weight else None),
                                      _capi.ptr(saved), _capi.ptr(g_T), _capi.ptr(g_pc), _capi.ptr(g_src),
                                      _capi.ptr(g_tgt), _capi.ptr(g_T0), _capi.ptr(g_w), _capi.ptr(None), 0,
                                      _capi.stream_ptr())
            _capi.check(rc, "mmb_icp_backward")
        return g_src, g_tgt, g_T0, g_w, None


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
        cfg = dict(icp_type=self.icp_type, max_iterations=int(self.max_iterations), tolerance=self.tolerance,
                   differentiable=bool(self.diff), const_iter=bool(self.const_iter), trim_dist=trim_dist,
                   loss_kind=loss_kind, loss_metric=loss_metric, dim=int(dim),
                   tanh_steepness=self.config['dICP']['parameters']['tanh_steepness'],
                   match_ratio_thresh=self.match_ratio_thresh, nn_slices=self.nn_slices,
                   nn_backend=self.nn_backend, record_indices=self.record_indices)
        T, pc, deltas, weights, costs, converged, iterations, matched, iters_done = _IcpFunction.apply(
            source.to(dev), target.to(dev), T_init.to(dev), weight_arg, cfg)

        self.last_indices = cfg.get("indices_out")
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
