"""Radar utilities with the reference's function surface (`mm_masking/radar_utils.py`).

Same names, arguments, defaults and return values as the reference functions that sit on the hot
path; the arithmetic runs in `csrc/radar_kernels.cu` through the C ABI of
`include/mm_masking_b200.h`:

  cfar_mask                       -> mmb_cfar_mask                (radar_utils.py:29-69)
  extract_pc                      -> mmb_extract_count / _write   (radar_utils.py:71-106)
  cfar_extract_pc (new, fused)    -> same kernels, the mask is never materialised
  extract_weights                 -> mmb_sample_weights_fwd / _bwd (radar_utils.py:108-140)
  radar_polar_to_cartesian_diff   -> mmb_polar_to_cart            (radar_utils.py:258-336)
  kstrong_extract_pc (new)        -> mmb_kstrong_select / _write  (vtr_radar detector.inl:39-85)

`load_radar`, `point_to_cart_idx`, `form_cart_range_angle_grid` and `pol_2_cart` are cheap host /
elementwise helpers kept for call compatibility.  CPU tensors are accepted (moved to the current
CUDA device, results returned on the input's device); without a CUDA device or the built library
every function raises -- there is no CPU implementation here.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _capi
from . import ops as _ops  # noqa: F401  (registers torch.ops.mmb.*)

_DT = {torch.float32: _capi.F32, torch.float64: _capi.F64}


def _device_of(t: torch.Tensor) -> torch.device:
    _capi.require_cuda()
    return t.device if t.is_cuda else torch.device("cuda", torch.cuda.current_device())


def _check_dtype(t: torch.Tensor, what: str):
    if t.dtype not in _DT:
        raise TypeError(f"{what} must be float32 or float64")


# ---------------------------------------------------------------------------------------------
# host helpers (same results as the reference functions of the same name)
# ---------------------------------------------------------------------------------------------
def load_radar(raw_img):
    """`radar_utils.py:20-27`: decode a Navtech polar PNG into fft (A,R) f32, azimuths, timestamps."""
    raw = np.ascontiguousarray(np.asarray(raw_img))
    timestamps = np.frombuffer(raw[:, :8].tobytes(), dtype=np.int64) * 1000
    azimuths = np.frombuffer(raw[:, 8:10].tobytes(), dtype=np.uint16) * (2 * np.pi / 5600)
    fft_data = np.divide(raw[:, 11:], 255.0, dtype=np.float32)
    return fft_data, azimuths, timestamps


def point_to_cart_idx(pc, cart_resolution=0.2384, cart_pixel_width=640, min_to_plus_1=False):
    """`radar_utils.py:374-397`: (x, y) -> pixel coordinates (u up, v right)."""
    u = -pc[:, :, 0] / cart_resolution
    v = pc[:, :, 1] / cart_resolution
    if min_to_plus_1:
        return torch.stack((v, u), dim=2) / (cart_pixel_width - 1) * 2
    return torch.stack((u, v), dim=2) + cart_pixel_width / 2


_GRID_CACHE: dict = {}


def form_cart_range_angle_grid(cart_resolution=0.2384, cart_pixel_width=640, dtype=None, device='cpu'):
    """`radar_utils.py:399-419`: range (m) and angle (rad, 0..2pi) of every Cartesian pixel centre.

    Batch-invariant, so it is evaluated once per (resolution, width, dtype) on the host -- with the
    same torch calls as the reference, which keeps the tables bit-identical to it -- and cached on
    the device that asks for it.
    """
    key = (float(cart_resolution), int(cart_pixel_width), dtype, str(device))
    hit = _GRID_CACHE.get(key)
    if hit is not None:
        return hit
    if (cart_pixel_width % 2) == 0:
        cart_min_range = (cart_pixel_width / 2 - 0.5) * cart_resolution
    else:
        cart_min_range = cart_pixel_width / 2 * cart_resolution
    kw = {} if dtype is None else {"dtype": dtype}
    coords = torch.linspace(-cart_min_range, cart_min_range, cart_pixel_width, **kw)
    Y, X = torch.meshgrid(coords, -1 * coords, indexing='xy')
    sample_range = torch.sqrt(Y * Y + X * X)
    sample_angle = torch.arctan2(Y, X)
    sample_angle = sample_angle + torch.where(sample_angle < 0, 2. * torch.pi, 0.0)
    out = (sample_range.to(device).contiguous(), sample_angle.to(sample_range.dtype).to(device).contiguous())
    _GRID_CACHE[key] = out
    return out


def pol_2_cart(pointcloud):
    """`radar_utils.py:187-195`."""
    rho, phi = pointcloud[:, 0], pointcloud[:, 1]
    return torch.stack((rho * torch.cos(phi), rho * torch.sin(phi), torch.zeros_like(rho)), dim=1)


# ---------------------------------------------------------------------------------------------
# K2a: GO-CFAR mask and peak extraction
# ---------------------------------------------------------------------------------------------
def _warn_no_grad(t, what):
    """The soft (diff=True) mask and peak extraction exist in the reference to be differentiated through; the CUDA
    path implements their forward values only (the reference has no live diff=True call site).  A
    caller who would back-propagate through them is told so instead of silently getting no gradient."""
    if isinstance(t, torch.Tensor) and t.requires_grad and torch.is_grad_enabled():
        import warnings
        warnings.warn(f"mm_masking_b200.radar_utils.{what}: the input requires grad but this op is forward-only here; "
                      "its output carries no autograd graph", RuntimeWarning, stacklevel=3)


def _cfar_params(B, A, R, dtype, scan_u8, res, width, minr, maxr, guard, a_thresh, b_thresh, diff, steep_fact):
    p = _capi.CfarParams()
    p.B, p.A, p.R, p.dtype, p.scan_u8 = B, A, R, dtype, int(scan_u8)
    p.width, p.guard, p.diff = int(width), int(guard), int(bool(diff))
    p.res, p.minr, p.maxr = float(res), float(minr), float(maxr)
    p.a_thresh, p.b_thresh, p.steep_fact = float(a_thresh), float(b_thresh), float(steep_fact)
    return p


def _scan_args(raw_scans: torch.Tensor):
    """(tensor on the CUDA device, dtype code, is_u8, B, A, R) for a (B,A,R) float scan or a
    (B,A,R+11) uint8 batch of raw Navtech rows."""
    assert raw_scans.ndim == 3, "raw_scans must be 3D"
    dev = _device_of(raw_scans)
    x = raw_scans.detach().to(dev).contiguous()
    if x.dtype == torch.uint8:
        return x, _capi.F32, True, x.shape[0], x.shape[1], x.shape[2] - 11
    _check_dtype(x, "raw_scans")
    return x, _DT[x.dtype], False, x.shape[0], x.shape[1], x.shape[2]


def cfar_mask(raw_scans, res, width=101, minr=2.0, maxr=80.0, guard=5, a_thresh=1.0, b_thresh=0.09, diff=True,
              steep_fact=10.0):
    """GO-CFAR detection mask, same contract as the reference (`radar_utils.py:29-69`).

    raw_scans (B,A,R) float32/float64 -- or (B,A,R+11) uint8 raw Navtech rows, decoded on the fly.
    Returns (B,A,R): float32 ones/zeros when `diff` is False, the soft mask in the scan dtype
    otherwise.  Window sums are exact (fp64 prefix sums), see DESIGN.md section 6.
    """
    lib = _capi.load_library()
    _warn_no_grad(raw_scans, "cfar_mask")
    x, dt, u8, B, A, R = _scan_args(raw_scans)
    p = _cfar_params(B, A, R, dt, u8, res, width, minr, maxr, guard, a_thresh, b_thresh, diff, steep_fact)
    out_dtype = (torch.float32 if u8 else x.dtype) if diff else torch.float32
    with torch.cuda.device(x.device):
        mask = torch.empty((B, A, R), dtype=out_dtype, device=x.device)
        _capi.check(lib.mmb_cfar_mask(p, _capi.ptr(x), _capi.ptr(mask), _capi.stream_ptr()), "mmb_cfar_mask")
    return mask if raw_scans.is_cuda else mask.to(raw_scans.device)


_EXTRACT_CAPACITY = 32768        # points per scan written before the totals are known (radar scans give 2k-15k)


def _extract(src, p, azimuth_angles, T_ab, out_device):
    """count -> allocate -> write; returns the list of (P_b,3) tensors."""
    lib = _capi.load_library()
    dev = src.device
    wdt = torch.float32 if p.dtype == _capi.F32 else torch.float64
    az = azimuth_angles.detach().to(device=dev, dtype=wdt).contiguous()
    Tm = None if T_ab is None else torch.as_tensor(T_ab).detach().to(device=dev, dtype=wdt).contiguous()
    B, A = p.B, p.A
    with torch.cuda.device(dev):
        ws_bytes = lib.mmb_extract_workspace_bytes(p)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        totals = torch.empty(B, dtype=torch.int32, device=dev)
        _capi.check(lib.mmb_extract_count(p, _capi.ptr(src), _capi.ptr(ws), ws_bytes, _capi.ptr(totals),
                                          _capi.stream_ptr()), "mmb_extract_count")

        def write(cap_entries):
            ent_val = torch.empty((B, cap_entries), dtype=wdt, device=dev)
            ent_row = torch.empty((B, cap_entries), dtype=torch.int32, device=dev)
            out_pts = torch.empty((B, cap_entries // 2, 3), dtype=wdt, device=dev)
            _capi.check(lib.mmb_extract_write(p, _capi.ptr(az), _capi.ptr(Tm), _capi.ptr(ws), ws_bytes, cap_entries,
                                              _capi.ptr(ent_val), _capi.ptr(ent_row), cap_entries // 2, _capi.ptr(out_pts),
                                              _capi.stream_ptr()), "mmb_extract_write")
            return out_pts

        # the points are written with a generous fixed capacity BEFORE the per-scan totals are read, so the one
        # host synchronisation of the call (the output is a list of ragged clouds) sits at the very end instead of
        # between the two launches; a scan with more points than the capacity is rewritten (not seen on radar data)
        cap = 2 * _EXTRACT_CAPACITY
        pts = write(cap)
        tot = totals.cpu().tolist()
        if any(t % 2 for t in tot):
            raise ValueError("odd number of peak entries in a scan (a detection run touches the last range bin)")
        if max(tot) > cap:
            pts = write(max(tot))
    out = [pts[b, :tot[b] // 2] for b in range(B)]
    return out if out_device == dev else [o.to(out_device) for o in out]


def extract_pc(thres_mask, res, azimuth_angles, azimuth_times, T_ab=None, diff=True, steep_fact=10.0):
    """Detection mask -> list of (P_b,3) point clouds, same contract as `radar_utils.py:71-106`:
    every run of non-zero mask bins in an azimuth row gives one point at the mean of its first and
    last range, azimuth-major, range ascending.  `azimuth_times` is accepted for call
    compatibility (the reference carries it through the averaging and then drops it)."""
    assert thres_mask.ndim == 3
    _warn_no_grad(thres_mask, "extract_pc")
    dev = _device_of(thres_mask)
    wdt = torch.promote_types(thres_mask.dtype, azimuth_angles.dtype)
    if wdt not in _DT:
        raise TypeError("thres_mask / azimuth_angles must be float32 or float64")
    m = thres_mask.detach().to(device=dev, dtype=wdt).contiguous()
    p = _capi.ExtractParams()
    p.B, p.A, p.R = m.shape
    p.dtype, p.diff, p.fused = _DT[wdt], int(bool(diff)), 0
    p.res, p.steep_fact = float(res), float(steep_fact)
    return _extract(m, p, azimuth_angles, T_ab, thres_mask.device)


def cfar_extract_pc(raw_scans, res, azimuth_angles, azimuth_times=None, T_ab=None, width=101, minr=2.0, maxr=80.0,
                    guard=5, a_thresh=1.0, b_thresh=0.09, diff=False, steep_fact=10.0):
    """`extract_pc(cfar_mask(raw_scans, ...), ...)` in one pass over the scan: the mask is evaluated
    on the fly in shared memory and never written to HBM.  Same points as the two-call form."""
    x, dt, u8, B, A, R = _scan_args(raw_scans)
    p = _capi.ExtractParams()
    p.B, p.A, p.R = B, A, R
    p.dtype, p.diff, p.fused = dt, int(bool(diff)), 1
    p.res, p.steep_fact = float(res), float(steep_fact)
    p.cfar = _cfar_params(B, A, R, dt, u8, res, width, minr, maxr, guard, a_thresh, b_thresh, diff, steep_fact)
    return _extract(x, p, azimuth_angles, T_ab, raw_scans.device)


# ---------------------------------------------------------------------------------------------
# K2b: K-strongest
# ---------------------------------------------------------------------------------------------
def kstrong_extract_pc(raw_scans, res, azimuth_angles, kstrong=12, threshold2=0.0, threshold3=0.216, minr=2.0,
                       maxr=80.0, range_offset=-0.31, return_bins=False):
    """K-strongest detector of vtr_radar (`detector.inl:39-85`, parameters of
    `gen_data/config/radar_lidar_config.yaml:83-94`): per azimuth the up-to-k bins at or above
    mean*threshold2 + threshold3, strongest first (ties to the nearer bin), as (x, y, 0) points.
    Returns a list of (P_b,3) tensors (and, with return_bins, lists of row and bin indices)."""
    lib = _capi.load_library()
    x, dt, u8, B, A, R = _scan_args(raw_scans)
    if not u8 and x.dtype != torch.float32:
        x = x.float()
    wdt = azimuth_angles.dtype if azimuth_angles.dtype in _DT else torch.float32
    p = _capi.KStrongParams()
    p.B, p.A, p.R, p.dtype, p.scan_u8, p.k = B, A, R, _DT[wdt], int(u8), int(kstrong)
    p.res, p.minr, p.maxr = float(res), float(minr), float(maxr)
    p.threshold2, p.threshold3, p.range_offset = float(threshold2), float(threshold3), float(range_offset)
    dev = x.device
    az = azimuth_angles.detach().to(device=dev, dtype=wdt).contiguous()
    with torch.cuda.device(dev):
        ws_bytes = lib.mmb_kstrong_workspace_bytes(p)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        totals = torch.empty(B, dtype=torch.int32, device=dev)
        _capi.check(lib.mmb_kstrong_select(p, _capi.ptr(x), _capi.ptr(ws), ws_bytes, _capi.ptr(totals),
                                           _capi.stream_ptr()), "mmb_kstrong_select")
        cap = max(A * int(kstrong), 1)                    # the exact upper bound: no read of the totals before the write
        pts = torch.empty((B, cap, 3), dtype=wdt, device=dev)
        rows = torch.empty((B, cap), dtype=torch.int32, device=dev) if return_bins else None
        bins = torch.empty((B, cap), dtype=torch.int32, device=dev) if return_bins else None
        _capi.check(lib.mmb_kstrong_write(p, _capi.ptr(az), _capi.ptr(ws), ws_bytes, cap, _capi.ptr(pts),
                                          _capi.ptr(rows), _capi.ptr(bins), _capi.stream_ptr()), "mmb_kstrong_write")
        tot = totals.cpu().tolist()                        # the one host sync, after both launches
    back = (lambda t: t) if raw_scans.is_cuda else (lambda t: t.to(raw_scans.device))
    out = [back(pts[b, :tot[b]]) for b in range(B)]
    if return_bins:
        return out, [back(rows[b, :tot[b]]) for b in range(B)], [back(bins[b, :tot[b]]) for b in range(B)]
    return out


# ---------------------------------------------------------------------------------------------
# K6: per-point weight sampling
# ---------------------------------------------------------------------------------------------
def extract_weights(mask, scan_pc):
    """Per-point bilinear sample of the weight mask, same contract as `radar_utils.py:108-140`.

    mask (N,640,640), scan_pc (N,n,3) -> (weights (N,n), diff_mean_num_non0, mean_num_non0, mean_w,
    max_w, min_w).  Points with x == 0 and y == 0 are batch padding and get weight 0.  The five
    statistics are small reductions over (N,n) and stay torch ops (the differentiable one keeps
    its autograd graph)."""
    _check_dtype(mask, "mask")
    dev = _device_of(mask)
    m = mask.to(dev)
    pts = scan_pc.detach().to(device=dev, dtype=m.dtype).contiguous()
    assert m.ndim == 3 and pts.ndim == 3 and m.shape[0] == pts.shape[0]
    weights = torch.ops.mmb.sample_weights(m, pts, 0.2384)
    fake = (pts[:, :, 0] == 0.0) & (pts[:, :, 1] == 0.0)
    real = weights[~fake]
    N = weights.shape[0]
    mean_num_non0 = (torch.sum(real > 0.05) / N).detach()
    mean_w, max_w, min_w = torch.mean(real).detach(), torch.max(real).detach(), torch.min(real).detach()
    diff_mean_num_non0 = torch.sum(0.5 * torch.tanh(5 * real) + 0.5) / N
    if not mask.is_cuda:
        weights = weights.to(mask.device)
    return weights, diff_mean_num_non0, mean_num_non0, mean_w, max_w, min_w


# ---------------------------------------------------------------------------------------------
# K1: polar -> Cartesian
# ---------------------------------------------------------------------------------------------
def radar_polar_to_cartesian_diff(fft_data, azimuths, radar_resolution, cart_resolution=0.2384, cart_pixel_width=640,
                                  interpolate_crossover=True, fix_wobble=True):
    """Polar scan batch (B,A,R) -> Cartesian (B,width,width), same contract as `radar_utils.py:258-336`
    (including its use of the DEFAULT Cartesian resolution for the pixel grid whatever
    `cart_resolution` is passed, `:276`).  fft_data may also be (B,A,R+11) uint8 raw Navtech rows.
    Differentiable in a float fft_data, like the reference (autograd through its `grid_sample`)."""
    assert fft_data.ndim == 3, "raw_scans must be 3D"
    dev = _device_of(fft_data)
    x = fft_data.to(dev).contiguous()
    if x.dtype == torch.uint8:
        dt, u8, B, A, R = _capi.F32, True, x.shape[0], x.shape[1], x.shape[2] - 11
    else:
        _check_dtype(x, "fft_data")
        dt, u8, B, A, R = _DT[x.dtype], False, x.shape[0], x.shape[1], x.shape[2]
    wdt = torch.float32 if dt == _capi.F32 else torch.float64
    az = azimuths.detach().to(device=dev, dtype=wdt).contiguous()
    rng, ang = form_cart_range_angle_grid(cart_pixel_width=cart_pixel_width, dtype=wdt, device=dev)
    out = torch.ops.mmb.polar_to_cart(x, az, rng.contiguous(), ang.contiguous(), int(cart_pixel_width),
                                      float(radar_resolution), bool(interpolate_crossover), bool(fix_wobble))
    return out if fft_data.is_cuda else out.to(fft_data.device)


# ---------------------------------------------------------------------------------------------
# N3: Cartesian -> polar
# ---------------------------------------------------------------------------------------------
def form_polar_range_grid(polar_resolution=0.2384, polar_pixel_shape=(400, 3360), dtype=None, device='cpu'):
    """`radar_utils.py:421-438`: range (m) of every polar pixel, one identical row per azimuth."""
    polar_range = (polar_pixel_shape[1] - 1) * polar_resolution
    kw = {} if dtype is None else {"dtype": dtype}
    range_coords = torch.linspace(0.0, polar_range, polar_pixel_shape[1], **kw).to(device)
    return range_coords.unsqueeze(0).expand(polar_pixel_shape[0], -1)


def radar_cartesian_to_polar(cart, azimuths, radar_resolution, cart_resolution=0.2384, polar_pixel_shape=(400, 3360)):
    """Cartesian image batch (B,H,W) -> polar (B,A,R), same contract as `radar_utils.py:338-372`,
    differentiable in `cart`.  The reference casts its sampling grid to float64 (`:370`), so it only
    runs for float64 images; float32 images are accepted here as well (all-float32 arithmetic)."""
    _check_dtype(cart, "cart")
    dev = _device_of(cart)
    c = cart.to(dev)
    az = azimuths.detach().to(device=dev, dtype=c.dtype).contiguous()
    assert c.ndim == 3 and az.ndim == 2 and az.shape[0] == c.shape[0]
    assert az.shape[1] == polar_pixel_shape[0]
    rng = form_polar_range_grid(radar_resolution, polar_pixel_shape, dtype=c.dtype, device=dev)[0].contiguous()
    out = torch.ops.mmb.cart_to_polar(c, az, rng, float(cart_resolution))
    return out if cart.is_cuda else out.to(cart.device)


# ---------------------------------------------------------------------------------------------
# N2: occupancy BEV of a point cloud
# ---------------------------------------------------------------------------------------------
def extract_bev_from_pts(pc, cart_pixel_width=640):
    """Point clouds (N,n,2|3) -> occupancy images (N,width,width), same contract as
    `radar_utils.py:142-165` (the BCE mask-loss target of `train_icp_weights.py:225`): the four
    floor/ceil pixels around every point are 1; out-of-range coordinates and padding points fall on
    the centre pixel, which is zeroed.  Not differentiable (index assignment in the reference)."""
    _check_dtype(pc, "pc")
    lib = _capi.load_library()
    dev = _device_of(pc)
    pts = pc.detach().to(dev).contiguous()
    N, n, cols = pts.shape
    with torch.cuda.device(dev):
        bev = torch.empty((N, cart_pixel_width, cart_pixel_width), dtype=pts.dtype, device=dev)
        _capi.check(lib.mmb_bev_from_points(N, n, cols, int(cart_pixel_width), _DT[pts.dtype], 0.2384, 320.0, _capi.ptr(pts),
                                            _capi.ptr(bev), _capi.stream_ptr()), "mmb_bev_from_points")
    return bev if pc.is_cuda else bev.to(pc.device)
