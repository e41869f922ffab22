"""PyTorch custom operators (`torch.ops.mmb.*`) over the C ABI of include/mm_masking_b200.h.

SURVEY.md section 8(b) "PyTorch custom ops": every kernel entry point of the library is registered with
`torch.library` (schema inferred from the annotations, CUDA kernel only), has a fake (meta) implementation so
that it traces under FakeTensor / `torch.compile` / `torch.export`, and -- where the reference differentiates
through the call -- an autograd formula that is itself a registered op:

    mmb::icp_forward      external/dICP/dICP/ICP.py:48-293 (all iterations, no host sync)   grad: mmb::icp_backward
    mmb::icp_backward     autograd of the unrolled loop ICP.py:130-250 (hand-derived adjoint)
    mmb::nn_search        external/dICP/dICP/nn.py:32-38 (index only; the gather stays a torch op)
    mmb::sample_weights   mm_masking/radar_utils.py:108-126                                grad: mmb::sample_weights_bwd
    mmb::polar_to_cart    radar_utils.py:258-336                                            grad: mmb::polar_to_cart_bwd
    mmb::cart_to_polar    radar_utils.py:338-372                                            grad: mmb::cart_to_polar_bwd
    mmb::cfar_mask        radar_utils.py:29-69
    mmb::cfar_extract     radar_utils.py:29-106 in one pass; fixed-capacity output (points, counts): no host sync
    mmb::kstrong_extract  vtr_radar detector.inl:39-85; fixed-capacity output (points, rows, bins, counts)
    mmb::bev_from_points  radar_utils.py:142-165

The module-level wrappers of this package (`dICP.ICP`, `dICP.nn`, `radar_utils`) call these ops, so the reference-
shaped API and `torch.ops.mmb.*` are the same code path.  There is no CPU kernel: the ops are registered for
CUDA only and raise on other devices.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
from torch import Tensor

from . import _capi

_DT = {torch.float32: _capi.F32, torch.float64: _capi.F64}

# The ops are registered through `torch.library.Library` (define + CUDA impl + fake + autograd) rather than the
# `torch.library.custom_op` decorator: same dispatcher entries, schemas (inferred from the annotations), fake
# tensors and opcheck results, but a call goes dispatcher -> kernel function without the decorator's extra Python
# layers (torch._dynamo.disable wrapper, default filling), which cost ~0.1 ms per call -- a tenth of the c3 step.
_LIB = torch.library.Library("mmb", "DEF")


class _Op:
    """the slice of the `CustomOpDef` interface this module uses"""

    def __init__(self, qualname, fn):
        self.qualname, self.name, self.fn = qualname, qualname.split("::")[1], fn
        self.schema = torch.library.infer_schema(fn, mutates_args=())
        _LIB.define(self.name + self.schema)
        _LIB.impl(self.name, fn, "CUDA")
        self.__doc__ = fn.__doc__

    def register_fake(self, fake):
        torch.library.register_fake(self.qualname, fake, lib=_LIB)
        return fake

    def register_autograd(self, backward, setup_context=None):
        """What `torch.library.register_autograd` installs -- an Autograd-key kernel that runs the op below autograd
        inside an `autograd.Function` -- without its `fill_defaults` pass over the schema (0.1 ms a call for the 19
        arguments of mmb::icp_forward; no schema of this module has default values, so there is nothing to fill)."""
        op = getattr(torch.ops.mmb, self.name).default
        below, after = torch._C._AutoDispatchBelowAutograd, torch._C._after_autograd_keyset

        class Generated(torch.autograd.Function):
            @staticmethod
            def forward(ctx, keyset, *args):
                with below():
                    out = op.redispatch(keyset & after, *args)
                if setup_context is not None:
                    setup_context(ctx, args, out)
                return out

            @staticmethod
            def backward(ctx, *grads):
                needs = ctx.needs_input_grad
                ctx.needs_input_grad = needs[1:]                 # the formula sees the op's own arguments
                try:
                    res = backward(ctx, *grads)
                finally:
                    ctx.needs_input_grad = needs
                return (None, *res) if isinstance(res, tuple) else (None, res)

        Generated.__name__ = Generated.__qualname__ = "GeneratedBackwardFor_mmb_" + self.name

        def autograd_kernel(keyset, *args):
            if torch.is_grad_enabled() and torch._C._any_requires_grad(*args):
                return Generated.apply(keyset, *args)
            with below():
                return op.redispatch(keyset & after, *args)

        _LIB.impl(self.name, autograd_kernel, "Autograd", with_keyset=True)

    def __call__(self, *args, **kwargs):
        return getattr(torch.ops.mmb, self.name)(*args, **kwargs)


def _custom_op(qualname):
    return lambda fn: _Op(qualname, fn)


def _opt(t: Optional[Tensor]):
    return None if t is None or t.numel() == 0 else t


# ---------------------------------------------------------------------------------------------
# weighted differentiable ICP
# ---------------------------------------------------------------------------------------------
_PARAMS = {}        # argument tuple -> (IcpParams, workspace bytes, saved bytes): the struct is read-only after it is built


def _icp_params(source, target, has_weight, save, icp_type, dim, max_iterations, tolerance, differentiable, const_iter,
                use_trim, trim_dist, loss_kind, loss_metric, tanh_steepness, match_ratio_thresh, nn_slices, nn_backend):
    return _icp_plan(source, target, has_weight, save, icp_type, dim, max_iterations, tolerance, differentiable, const_iter,
                     use_trim, trim_dist, loss_kind, loss_metric, tanh_steepness, match_ratio_thresh, nn_slices,
                     nn_backend)[0]


def _icp_plan(source, target, has_weight, save, *rest):
    """(IcpParams, workspace bytes, saved bytes) of a call, cached by its shapes, dtype and options"""
    key = (source.shape[0], source.shape[1], target.shape[1], target.shape[2], source.dtype, bool(has_weight), bool(save),
           *rest)
    hit = _PARAMS.get(key)
    if hit is None:
        if len(_PARAMS) > 256:
            _PARAMS.clear()
        p = _build_icp_params(source, target, has_weight, save, *rest)
        lib = _capi.load_library()
        hit = _PARAMS[key] = (p, lib.mmb_icp_workspace_bytes(p), lib.mmb_icp_saved_bytes(p))
    return hit


def _build_icp_params(source, target, has_weight, save, icp_type, dim, max_iterations, tolerance, differentiable, const_iter,
                      use_trim, trim_dist, loss_kind, loss_metric, tanh_steepness, match_ratio_thresh, nn_slices,
                      nn_backend):
    p = _capi.IcpParams()
    p.N, p.n, p.m, p.target_cols = source.shape[0], source.shape[1], target.shape[1], target.shape[2]
    p.dtype = _DT[source.dtype]
    p.icp_type, p.dim, p.max_iterations = int(icp_type), int(dim), int(max_iterations)
    p.differentiable, p.const_iter, p.use_trim = int(differentiable), int(const_iter), int(use_trim)
    p.trim_dist = float(trim_dist) if use_trim else 0.0
    p.loss_kind, p.loss_metric = int(loss_kind), float(loss_metric)
    p.has_weight, p.save_for_backward = int(has_weight), int(save)
    p.nn_slices, p.nn_backend = int(nn_slices), int(nn_backend)
    p.tolerance, p.tanh_steepness, p.match_ratio_thresh = float(tolerance), float(tanh_steepness), float(match_ratio_thresh)
    return p


_LAYOUTS = {}


def _layout(specs):
    """256-byte aligned (offset, bytes) of every (shape, dtype) in one flat byte buffer, and its total size"""
    try:
        hit = _LAYOUTS.get(specs)
    except TypeError:                                   # a list of specs: not cached
        return _layout_of(specs)
    if hit is None:
        if len(_LAYOUTS) > 256:
            _LAYOUTS.clear()
        hit = _LAYOUTS[specs] = _layout_of(specs)
    return hit


def _layout_of(specs):
    spans, total = [], 0
    for shape, dt in specs:
        count = 1
        for d in shape:
            count *= int(d)
        nbytes = count * dt.itemsize
        spans.append((total, nbytes))
        total += (nbytes + 255) // 256 * 256
    return spans, total


def _carve(flat, specs):
    """views of the given (shape, dtype) list into a flat byte buffer laid out by `_layout`"""
    spans, _ = _layout(specs)
    typed, out = {}, []                                  # one reinterpretation per dtype, one as_strided per tensor
    for (o, _nb), (shape, dt) in zip(spans, specs):
        tv = typed.get(dt)
        if tv is None:
            tv = typed[dt] = flat.view(dt)
        strides, acc = [], 1
        for d in reversed(shape):
            strides.append(acc)
            acc *= int(d)
        out.append(tv.as_strided(shape, strides[::-1], tv.storage_offset() + o // dt.itemsize))
    return out


def _zeros_block(device, specs):
    """zero tensors of the given (shape, dtype) list as views of ONE zero-filled buffer (one fill launch)"""
    return _carve(torch.zeros(_layout(specs)[1], dtype=torch.uint8, device=device), specs)


def icp_history_specs(N, n, K, icp_type, dtype):
    """the zero-initialised outputs of mmb::icp_forward inside its `block` output: deltas (N,K,6), weights
    (N,K,R), costs (N,K), converged (N) u8, iterations (N), matched_ratio (N), iters_done (1) i32"""
    rows = 3 * n if icp_type == _capi.PT2PT else n
    return (((N, K, 6), dtype), ((N, K, rows), dtype), ((N, K), dtype), ((N,), torch.uint8), ((N,), dtype),
            ((N,), dtype), ((1,), torch.int32))


@_custom_op("mmb::icp_forward")
def icp_forward(source: Tensor, target: Tensor, T_init: Tensor, weight: Optional[Tensor], icp_type: int, dim: int,
                max_iterations: int, tolerance: float, differentiable: bool, const_iter: bool, use_trim: bool,
                trim_dist: float, loss_kind: int, loss_metric: float, tanh_steepness: float, match_ratio_thresh: float,
                nn_slices: int, nn_backend: int, save: bool) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """source (N,n,3), target (N,m,3|6), T_init (N,4,4), weight (N,n) or None ->
    (T (N,4,4), pc (N,n,3), block, saved).  `block` is one zero-initialised byte buffer holding the histories and
    statistics (`icp_history_specs` / `_carve`: deltas, weights, costs, converged, iterations, matched_ratio,
    iters_done) -- one buffer so that one fill launch clears them and no two outputs of the op share storage;
    `saved` is the opaque state for mmb::icp_backward (empty unless `save`)."""
    lib = _capi.load_library()
    dev, dtype = source.device, source.dtype
    N, n, _ = source.shape
    K = int(max_iterations)
    p, ws_bytes, sv_bytes = _icp_plan(source, target, weight is not None, save, icp_type, dim, max_iterations, tolerance,
                                      differentiable, const_iter, use_trim, trim_dist, loss_kind, loss_metric,
                                      tanh_steepness, match_ratio_thresh, nn_slices, nn_backend)
    src, tgt, T0 = source.contiguous(), target.contiguous(), T_init.contiguous()
    w = None if weight is None else weight.contiguous()
    with _capi.on_device(dev):
        if ws_bytes == 0:
            raise _capi.MmbError("mmb_icp_workspace_bytes rejected the parameters")
        workspace = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        saved = torch.empty(sv_bytes, dtype=torch.uint8, device=dev)
        T_out = torch.empty((N, 4, 4), dtype=dtype, device=dev)
        pc = torch.empty((N, n, 3), dtype=dtype, device=dev)
        # rows of iterations that are not executed stay zero
        specs = icp_history_specs(N, n, K, icp_type, dtype)
        spans, block_bytes = _layout(specs)
        block = torch.zeros(block_bytes, dtype=torch.uint8, device=dev)
        base = block.data_ptr()                          # deltas, weights, costs, converged, iterations, matched, iters_done
        hist = [_capi.C.c_void_p(base + o) for o, _ in spans]
        rc = lib.mmb_icp_forward(p, _capi.ptr(src), _capi.ptr(tgt), _capi.ptr(T0), _capi.ptr(w),
                                 _capi.ptr(T_out), _capi.ptr(pc), *hist, _capi.ptr(saved if sv_bytes else None),
                                 _capi.ptr(workspace), ws_bytes, _capi.stream_ptr())
        _capi.check(rc, "mmb_icp_forward")
    return T_out, pc, block, saved


@icp_forward.register_fake
def _(source, target, T_init, weight, icp_type, dim, max_iterations, tolerance, differentiable, const_iter, use_trim,
      trim_dist, loss_kind, loss_metric, tanh_steepness, match_ratio_thresh, nn_slices, nn_backend, save):
    N, n, _ = source.shape
    e = source.new_empty
    p = _icp_params(source, target, weight is not None, save, icp_type, dim, max_iterations, tolerance, differentiable,
                    const_iter, use_trim, trim_dist, loss_kind, loss_metric, tanh_steepness, match_ratio_thresh,
                    nn_slices, nn_backend)
    sv_bytes = _capi.load_library().mmb_icp_saved_bytes(p)          # a host-side size computation, no device work
    block_bytes = _layout(icp_history_specs(N, n, max_iterations, icp_type, source.dtype))[1]
    return (e((N, 4, 4)), e((N, n, 3)), e((block_bytes,), dtype=torch.uint8), e((sv_bytes,), dtype=torch.uint8))


@_custom_op("mmb::icp_backward")
def icp_backward(source: Tensor, target: Tensor, weight: Optional[Tensor], saved: Tensor, grad_T: Optional[Tensor],
                 grad_pc: Optional[Tensor], need_source: bool, need_target: bool, need_T_init: bool, need_weight: bool,
                 icp_type: int, dim: int, max_iterations: int, tolerance: float, differentiable: bool, const_iter: bool,
                 use_trim: bool, trim_dist: float, loss_kind: int, loss_metric: float, tanh_steepness: float,
                 match_ratio_thresh: float, nn_slices: int, nn_backend: int) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """Adjoint sweep over the iterations saved by mmb::icp_forward -> (g_source, g_target, g_T_init, g_weight);
    a gradient that was not asked for comes back as an empty tensor."""
    lib = _capi.load_library()
    dev, dtype = source.device, source.dtype
    p = _icp_params(source, target, weight is not None, True, icp_type, dim, max_iterations, tolerance, differentiable,
                    const_iter, use_trim, trim_dist, loss_kind, loss_metric, tanh_steepness, match_ratio_thresh,
                    nn_slices, nn_backend)
    g_T = None if grad_T is None else grad_T.contiguous()
    g_pc = None if grad_pc is None else grad_pc.contiguous()
    need_weight = need_weight and weight is not None
    with _capi.on_device(dev):
        g_src = torch.empty_like(source) if need_source else source.new_empty(0)
        g_tgt = torch.zeros_like(target) if need_target else source.new_empty(0)
        g_T0 = torch.empty((p.N, 4, 4), dtype=dtype, device=dev) if need_T_init else source.new_empty(0)
        g_w = torch.empty((p.N, p.n), dtype=dtype, device=dev) if need_weight else source.new_empty(0)
        rc = lib.mmb_icp_backward(p, _capi.ptr(source), _capi.ptr(target), _capi.ptr(weight), _capi.ptr(saved),
                                  _capi.ptr(g_T), _capi.ptr(g_pc), _capi.ptr(_opt(g_src)), _capi.ptr(_opt(g_tgt)),
                                  _capi.ptr(_opt(g_T0)), _capi.ptr(_opt(g_w)), _capi.ptr(None), 0, _capi.stream_ptr())
        _capi.check(rc, "mmb_icp_backward")
    return g_src, g_tgt, g_T0, g_w


@icp_backward.register_fake
def _(source, target, weight, saved, grad_T, grad_pc, need_source, need_target, need_T_init, need_weight, *params):
    e = source.new_empty
    N, n, _ = source.shape
    return (torch.empty_like(source) if need_source else e(0), torch.empty_like(target) if need_target else e(0),
            e((N, 4, 4)) if need_T_init else e(0), e((N, n)) if need_weight and weight is not None else e(0))


def _icp_setup(ctx, inputs, output):
    source, target, T_init, weight = inputs[:4]
    ctx.params = tuple(inputs[4:18])
    ctx.has_weight = weight is not None
    saved = output[3]
    ctx.save_for_backward(source, target, weight if weight is not None else source.new_empty(0), saved)
    ctx.set_materialize_grads(False)


def _icp_backward(ctx, g_T, g_pc, *unused):
    source, target, weight, saved = ctx.saved_tensors
    if saved.numel() == 0:
        raise _capi.MmbError("mmb::icp_forward was called with save=False: no state for backward")
    need = list(ctx.needs_input_grad[:4])
    g_src, g_tgt, g_T0, g_w = torch.ops.mmb.icp_backward(
        source, target, weight if ctx.has_weight else None, saved, g_T, g_pc, need[0], need[1], need[2],
        need[3] and ctx.has_weight, *ctx.params)
    out = [g_src if need[0] else None, g_tgt if need[1] else None, g_T0 if need[2] else None,
           g_w if need[3] and ctx.has_weight else None]
    return (*out, *([None] * 15))


icp_forward.register_autograd(_icp_backward, setup_context=_icp_setup)


# ---------------------------------------------------------------------------------------------
# nearest neighbour
# ---------------------------------------------------------------------------------------------
@_custom_op("mmb::nn_search")
def nn_search(query: Tensor, target: Tensor, dim: int, nn_slices: int, nn_backend: int) -> Tensor:
    """(N,n) int32 index of the nearest row of target (N,m,3|6) for every row of query (N,n,3); exact squared
    differences, lowest index on ties (`nn.py:32-35`)."""
    lib = _capi.load_library()
    xq, yt = query.contiguous(), target.contiguous()
    N, n, _ = xq.shape
    m, cols = yt.shape[1], yt.shape[2]
    with torch.cuda.device(xq.device):
        dt = _DT[xq.dtype]
        ws_bytes = lib.mmb_nn_workspace_bytes(N, n, m, dt, nn_slices, nn_backend)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=xq.device)
        idx = torch.empty((N, n), dtype=torch.int32, device=xq.device)
        rc = lib.mmb_nn_search(N, n, m, cols, dt, dim, nn_slices, nn_backend, _capi.ptr(xq), _capi.ptr(yt), _capi.ptr(idx),
                               _capi.ptr(ws), ws_bytes, _capi.stream_ptr())
        _capi.check(rc, "mmb_nn_search")
    return idx


@nn_search.register_fake
def _(query, target, dim, nn_slices, nn_backend):
    return query.new_empty(query.shape[:2], dtype=torch.int32)


# ---------------------------------------------------------------------------------------------
# per-point weight sampling (K6)
# ---------------------------------------------------------------------------------------------
@_custom_op("mmb::sample_weights")
def sample_weights(mask: Tensor, pts: Tensor, cart_res: float) -> Tensor:
    """mask (N,H,W), pts (N,n,>=2) -> w (N,n): bilinear sample, padded points (x = y = 0) get 0"""
    lib = _capi.load_library()
    N, H, W = mask.shape
    n, cols = pts.shape[1], pts.shape[2]
    m, q = mask.contiguous(), pts.contiguous()
    with torch.cuda.device(m.device):
        w = torch.empty((N, n), dtype=m.dtype, device=m.device)
        _capi.check(lib.mmb_sample_weights_fwd(N, n, cols, H, W, _DT[m.dtype], float(cart_res), _capi.ptr(m), _capi.ptr(q),
                                               _capi.ptr(w), _capi.stream_ptr()), "mmb_sample_weights_fwd")
    return w


@sample_weights.register_fake
def _(mask, pts, cart_res):
    return mask.new_empty((mask.shape[0], pts.shape[1]))


@_custom_op("mmb::sample_weights_bwd")
def sample_weights_bwd(grad_w: Tensor, pts: Tensor, H: int, W: int, cart_res: float) -> Tensor:
    """adjoint of mmb::sample_weights with respect to the mask: (N,n) -> (N,H,W).  Deterministic: the taps are
    sorted by pixel and summed in tap order (mmb_sample_weights_bwd_det), no floating-point atomics"""
    lib = _capi.load_library()
    g, q = grad_w.contiguous(), pts.contiguous()
    N = g.shape[0]
    with torch.cuda.device(g.device):
        gm = torch.zeros((N, H, W), dtype=g.dtype, device=g.device)
        ws_bytes = lib.mmb_sample_weights_bwd_det_workspace_bytes(N, q.shape[1], H, W, _DT[g.dtype])
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=g.device)
        _capi.check(lib.mmb_sample_weights_bwd_det(N, q.shape[1], q.shape[2], H, W, _DT[g.dtype], float(cart_res),
                                                   _capi.ptr(g), _capi.ptr(q), _capi.ptr(gm), _capi.ptr(ws), ws_bytes,
                                                   _capi.stream_ptr()), "mmb_sample_weights_bwd_det")
    return gm


@sample_weights_bwd.register_fake
def _(grad_w, pts, H, W, cart_res):
    return grad_w.new_empty((grad_w.shape[0], H, W))


def _sw_setup(ctx, inputs, output):
    mask, pts, cart_res = inputs
    ctx.save_for_backward(pts)
    ctx.hw, ctx.cart_res = (mask.shape[1], mask.shape[2]), cart_res


def _sw_backward(ctx, g):
    (pts,) = ctx.saved_tensors
    gm = torch.ops.mmb.sample_weights_bwd(g, pts, ctx.hw[0], ctx.hw[1], ctx.cart_res) if ctx.needs_input_grad[0] else None
    return gm, None, None


sample_weights.register_autograd(_sw_backward, setup_context=_sw_setup)


# ---------------------------------------------------------------------------------------------
# polar <-> Cartesian resampling (K1, N3)
# ---------------------------------------------------------------------------------------------
@_custom_op("mmb::polar_to_cart")
def polar_to_cart(scan: Tensor, azimuths: Tensor, rng: Tensor, ang: Tensor, width: int, radar_resolution: float,
                  interpolate_crossover: bool, fix_wobble: bool) -> Tensor:
    """scan (B,A,R) float or (B,A,R+11) uint8 raw Navtech rows -> (B,width,width) in the dtype of `azimuths`;
    rng/ang: the pixel-centre range / angle tables of `form_cart_range_angle_grid`"""
    lib = _capi.load_library()
    x = scan.contiguous()
    u8 = x.dtype == torch.uint8
    B, A = x.shape[0], x.shape[1]
    R = x.shape[2] - 11 if u8 else x.shape[2]
    dt = _DT[azimuths.dtype]
    with torch.cuda.device(x.device):
        out = torch.empty((B, width, width), dtype=azimuths.dtype, device=x.device)
        _capi.check(lib.mmb_polar_to_cart(B, A, R, width, dt, int(u8), _capi.ptr(x), _capi.ptr(azimuths.contiguous()),
                                          _capi.ptr(rng.contiguous()), _capi.ptr(ang.contiguous()), float(radar_resolution),
                                          int(interpolate_crossover), int(fix_wobble), _capi.ptr(out), _capi.stream_ptr()),
                    "mmb_polar_to_cart")
    return out


@polar_to_cart.register_fake
def _(scan, azimuths, rng, ang, width, radar_resolution, interpolate_crossover, fix_wobble):
    return azimuths.new_empty((scan.shape[0], width, width))


@_custom_op("mmb::polar_to_cart_bwd")
def polar_to_cart_bwd(grad_out: Tensor, azimuths: Tensor, rng: Tensor, ang: Tensor, R: int, radar_resolution: float,
                      interpolate_crossover: bool, fix_wobble: bool) -> Tensor:
    """adjoint of mmb::polar_to_cart with respect to a float scan: (B,width,width) -> (B,A,R)"""
    lib = _capi.load_library()
    g = grad_out.contiguous()
    B, width = g.shape[0], g.shape[1]
    A = azimuths.shape[1]
    with torch.cuda.device(g.device):
        gs = torch.zeros((B, A, R), dtype=g.dtype, device=g.device)
        _capi.check(lib.mmb_polar_to_cart_bwd(B, A, R, width, _DT[g.dtype], _capi.ptr(g), _capi.ptr(azimuths.contiguous()),
                                              _capi.ptr(rng.contiguous()), _capi.ptr(ang.contiguous()),
                                              float(radar_resolution), int(interpolate_crossover), int(fix_wobble),
                                              _capi.ptr(gs), _capi.stream_ptr()), "mmb_polar_to_cart_bwd")
    return gs


@polar_to_cart_bwd.register_fake
def _(grad_out, azimuths, rng, ang, R, radar_resolution, interpolate_crossover, fix_wobble):
    return grad_out.new_empty((grad_out.shape[0], azimuths.shape[1], R))


def _p2c_setup(ctx, inputs, output):
    scan, az, rng, ang, width, res, cross, wobble = inputs
    ctx.save_for_backward(az, rng, ang)
    ctx.meta = (scan.dtype == torch.uint8, scan.shape[2], res, cross, wobble)


def _p2c_backward(ctx, g):
    az, rng, ang = ctx.saved_tensors
    u8, R, res, cross, wobble = ctx.meta
    gs = None
    if ctx.needs_input_grad[0] and not u8:
        gs = torch.ops.mmb.polar_to_cart_bwd(g, az, rng, ang, R, res, cross, wobble)
    return gs, None, None, None, None, None, None, None


polar_to_cart.register_autograd(_p2c_backward, setup_context=_p2c_setup)


@_custom_op("mmb::cart_to_polar")
def cart_to_polar(cart: Tensor, azimuths: Tensor, rng: Tensor, cart_res: float) -> Tensor:
    """cart (B,H,W), azimuths (B,A), rng (R) range of every polar bin -> (B,A,R)"""
    lib = _capi.load_library()
    c = cart.contiguous()
    B, H, W = c.shape
    A, R = azimuths.shape[1], rng.shape[0]
    with torch.cuda.device(c.device):
        out = torch.empty((B, A, R), dtype=c.dtype, device=c.device)
        _capi.check(lib.mmb_cart_to_polar(B, A, R, H, W, _DT[c.dtype], float(cart_res), _capi.ptr(c),
                                          _capi.ptr(azimuths.contiguous()), _capi.ptr(rng.contiguous()), _capi.ptr(out),
                                          _capi.stream_ptr()), "mmb_cart_to_polar")
    return out


@cart_to_polar.register_fake
def _(cart, azimuths, rng, cart_res):
    return cart.new_empty((cart.shape[0], azimuths.shape[1], rng.shape[0]))


@_custom_op("mmb::cart_to_polar_bwd")
def cart_to_polar_bwd(grad_out: Tensor, azimuths: Tensor, rng: Tensor, H: int, W: int, cart_res: float) -> Tensor:
    lib = _capi.load_library()
    g = grad_out.contiguous()
    B, A, R = g.shape
    with torch.cuda.device(g.device):
        gc = torch.zeros((B, H, W), dtype=g.dtype, device=g.device)
        _capi.check(lib.mmb_cart_to_polar_bwd(B, A, R, H, W, _DT[g.dtype], float(cart_res), _capi.ptr(g),
                                              _capi.ptr(azimuths.contiguous()), _capi.ptr(rng.contiguous()), _capi.ptr(gc),
                                              _capi.stream_ptr()), "mmb_cart_to_polar_bwd")
    return gc


@cart_to_polar_bwd.register_fake
def _(grad_out, azimuths, rng, H, W, cart_res):
    return grad_out.new_empty((grad_out.shape[0], H, W))


def _c2p_setup(ctx, inputs, output):
    cart, az, rng, cart_res = inputs
    ctx.save_for_backward(az, rng)
    ctx.meta = (cart.shape[1], cart.shape[2], cart_res)


def _c2p_backward(ctx, g):
    az, rng = ctx.saved_tensors
    H, W, cart_res = ctx.meta
    gc = torch.ops.mmb.cart_to_polar_bwd(g, az, rng, H, W, cart_res) if ctx.needs_input_grad[0] else None
    return gc, None, None, None


cart_to_polar.register_autograd(_c2p_backward, setup_context=_c2p_setup)


# ---------------------------------------------------------------------------------------------
# detection: GO-CFAR mask, fused CFAR + peak extraction, K-strongest, occupancy BEV
# ---------------------------------------------------------------------------------------------
def _cfar_struct(B, A, R, dtype, scan_u8, res, width, minr, maxr, guard, a_thresh, b_thresh, diff, steep_fact):
    p = _capi.CfarParams()
    p.B, p.A, p.R, p.dtype, p.scan_u8 = B, A, R, dtype, int(scan_u8)
    p.width, p.guard, p.diff = int(width), int(guard), int(bool(diff))
    p.res, p.minr, p.maxr = float(res), float(minr), float(maxr)
    p.a_thresh, p.b_thresh, p.steep_fact = float(a_thresh), float(b_thresh), float(steep_fact)
    return p


def _scan_shape(x):
    u8 = x.dtype == torch.uint8
    return u8, x.shape[0], x.shape[1], (x.shape[2] - 11 if u8 else x.shape[2]), (_capi.F32 if u8 else _DT[x.dtype])


@_custom_op("mmb::cfar_mask")
def cfar_mask(scan: Tensor, res: float, width: int, minr: float, maxr: float, guard: int, a_thresh: float,
              b_thresh: float, diff: bool, steep_fact: float) -> Tensor:
    lib = _capi.load_library()
    x = scan.contiguous()
    u8, B, A, R, dt = _scan_shape(x)
    p = _cfar_struct(B, A, R, dt, u8, res, width, minr, maxr, guard, a_thresh, b_thresh, diff, steep_fact)
    out_dtype = (torch.float32 if u8 else x.dtype) if diff else torch.float32
    with torch.cuda.device(x.device):
        mask = torch.empty((B, A, R), dtype=out_dtype, device=x.device)
        _capi.check(lib.mmb_cfar_mask(p, _capi.ptr(x), _capi.ptr(mask), _capi.stream_ptr()), "mmb_cfar_mask")
    return mask


@cfar_mask.register_fake
def _(scan, res, width, minr, maxr, guard, a_thresh, b_thresh, diff, steep_fact):
    u8 = scan.dtype == torch.uint8
    R = scan.shape[2] - 11 if u8 else scan.shape[2]
    out_dtype = (torch.float32 if u8 else scan.dtype) if diff else torch.float32
    return scan.new_empty((scan.shape[0], scan.shape[1], R), dtype=out_dtype)


@_custom_op("mmb::cfar_extract")
def cfar_extract(scan: Tensor, azimuths: Tensor, T_ab: Optional[Tensor], capacity: int, res: float, width: int,
                 minr: float, maxr: float, guard: int, a_thresh: float, b_thresh: float, diff: bool,
                 steep_fact: float) -> Tuple[Tensor, Tensor]:
    """GO-CFAR + peak extraction in one pass, fixed capacity, no host synchronisation:
    -> (points (B,capacity,3), entries (B) int32 = 2 x the number of points of every scan).  Scan b holds
    entries[b] / 2 points (azimuth-major, range ascending); rows beyond that -- and points beyond `capacity` --
    are not written."""
    lib = _capi.load_library()
    x = scan.contiguous()
    u8, B, A, R, dt = _scan_shape(x)
    p = _capi.ExtractParams()
    p.B, p.A, p.R = B, A, R
    p.dtype, p.diff, p.fused = dt, int(bool(diff)), 1
    p.res, p.steep_fact = float(res), float(steep_fact)
    p.cfar = _cfar_struct(B, A, R, dt, u8, res, width, minr, maxr, guard, a_thresh, b_thresh, diff, steep_fact)
    wdt = torch.float32 if dt == _capi.F32 else torch.float64
    az = azimuths.to(wdt).contiguous()
    Tm = None if T_ab is None else T_ab.to(wdt).contiguous()
    cap = max(int(capacity), 1)
    with torch.cuda.device(x.device):
        ws_bytes = lib.mmb_extract_workspace_bytes(p)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
        totals = torch.empty(B, dtype=torch.int32, device=x.device)
        _capi.check(lib.mmb_extract_count(p, _capi.ptr(x), _capi.ptr(ws), ws_bytes, _capi.ptr(totals), _capi.stream_ptr()),
                    "mmb_extract_count")
        ent_val = torch.empty((B, 2 * cap), dtype=wdt, device=x.device)
        ent_row = torch.empty((B, 2 * cap), dtype=torch.int32, device=x.device)
        pts = torch.zeros((B, cap, 3), dtype=wdt, device=x.device)
        _capi.check(lib.mmb_extract_write(p, _capi.ptr(az), _capi.ptr(Tm), _capi.ptr(ws), ws_bytes, 2 * cap,
                                          _capi.ptr(ent_val), _capi.ptr(ent_row), cap, _capi.ptr(pts), _capi.stream_ptr()),
                    "mmb_extract_write")
    return pts, totals


@cfar_extract.register_fake
def _(scan, azimuths, T_ab, capacity, res, width, minr, maxr, guard, a_thresh, b_thresh, diff, steep_fact):
    wdt = torch.float32 if scan.dtype in (torch.uint8, torch.float32) else torch.float64
    return (scan.new_empty((scan.shape[0], max(capacity, 1), 3), dtype=wdt),
            scan.new_empty((scan.shape[0],), dtype=torch.int32))


@_custom_op("mmb::kstrong_extract")
def kstrong_extract(scan: Tensor, azimuths: Tensor, res: float, k: int, threshold2: float, threshold3: float, minr: float,
                    maxr: float, range_offset: float) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """K-strongest per azimuth, fixed capacity A*k, no host synchronisation:
    -> (points (B,A*k,3), rows (B,A*k) i32, bins (B,A*k) i32, counts (B) i32)"""
    lib = _capi.load_library()
    x = scan.contiguous()
    u8, B, A, R, _ = _scan_shape(x)
    if not u8 and x.dtype != torch.float32:
        x = x.float()
    wdt = azimuths.dtype if azimuths.dtype in _DT else torch.float32
    p = _capi.KStrongParams()
    p.B, p.A, p.R, p.dtype, p.scan_u8, p.k = B, A, R, _DT[wdt], int(u8), int(k)
    p.res, p.minr, p.maxr = float(res), float(minr), float(maxr)
    p.threshold2, p.threshold3, p.range_offset = float(threshold2), float(threshold3), float(range_offset)
    az = azimuths.to(wdt).contiguous()
    cap = max(A * int(k), 1)
    with torch.cuda.device(x.device):
        ws_bytes = lib.mmb_kstrong_workspace_bytes(p)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
        totals = torch.empty(B, dtype=torch.int32, device=x.device)
        _capi.check(lib.mmb_kstrong_select(p, _capi.ptr(x), _capi.ptr(ws), ws_bytes, _capi.ptr(totals), _capi.stream_ptr()),
                    "mmb_kstrong_select")
        pts = torch.zeros((B, cap, 3), dtype=wdt, device=x.device)
        rows = torch.zeros((B, cap), dtype=torch.int32, device=x.device)
        bins = torch.zeros((B, cap), dtype=torch.int32, device=x.device)
        _capi.check(lib.mmb_kstrong_write(p, _capi.ptr(az), _capi.ptr(ws), ws_bytes, cap, _capi.ptr(pts), _capi.ptr(rows),
                                          _capi.ptr(bins), _capi.stream_ptr()), "mmb_kstrong_write")
    return pts, rows, bins, totals


@kstrong_extract.register_fake
def _(scan, azimuths, res, k, threshold2, threshold3, minr, maxr, range_offset):
    B, A = scan.shape[0], scan.shape[1]
    wdt = azimuths.dtype if azimuths.dtype in _DT else torch.float32
    cap = max(A * k, 1)
    i32 = dict(dtype=torch.int32)
    return (scan.new_empty((B, cap, 3), dtype=wdt), scan.new_empty((B, cap), **i32), scan.new_empty((B, cap), **i32),
            scan.new_empty((B,), **i32))


@_custom_op("mmb::bev_from_points")
def bev_from_points(pc: Tensor, width: int) -> Tensor:
    lib = _capi.load_library()
    pts = pc.contiguous()
    N, n, cols = pts.shape
    with torch.cuda.device(pts.device):
        bev = torch.empty((N, width, width), dtype=pts.dtype, device=pts.device)
        _capi.check(lib.mmb_bev_from_points(N, n, cols, int(width), _DT[pts.dtype], 0.2384, 320.0, _capi.ptr(pts),
                                            _capi.ptr(bev), _capi.stream_ptr()), "mmb_bev_from_points")
    return bev


@bev_from_points.register_fake
def _(pc, width):
    return pc.new_empty((pc.shape[0], width, width))
