"""ctypes binding of the C ABI in include/mm_masking_b200.h.

The shared library is built in-tree by `__graft_entry__.build()` (plain nvcc, sm_100a) and
holds every CUDA kernel of this package.  There is no fallback: if the library is missing or
CUDA is unavailable the ops raise.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MMB_LIBRARY") or os.path.join(_HERE, "libmm_masking_b200.so")   # override: A/B builds

F32, F64 = 0, 1
PT2PL, PT2PT = 0, 1
LOSS_NONE, LOSS_HUBER, LOSS_CAUCHY = 0, 1, 2
NN_AUTO, NN_BRUTE, NN_GRID, NN_ROWS = 0, 1, 2, 3
PROFILE_CLASSES = 8


class IcpParams(C.Structure):
    """mirror of `mmb_icp_params`"""
    _fields_ = [
        ("N", C.c_int32), ("n", C.c_int32), ("m", C.c_int32), ("target_cols", C.c_int32),
        ("dtype", C.c_int32), ("icp_type", C.c_int32), ("dim", C.c_int32), ("max_iterations", C.c_int32),
        ("differentiable", C.c_int32), ("const_iter", C.c_int32), ("use_trim", C.c_int32),
        ("loss_kind", C.c_int32), ("has_weight", C.c_int32), ("save_for_backward", C.c_int32),
        ("nn_slices", C.c_int32), ("nn_backend", C.c_int32),
        ("tolerance", C.c_double), ("trim_dist", C.c_double), ("loss_metric", C.c_double),
        ("tanh_steepness", C.c_double), ("match_ratio_thresh", C.c_double),
    ]


class CfarParams(C.Structure):
    """mirror of `mmb_cfar_params`"""
    _fields_ = [("B", C.c_int32), ("A", C.c_int32), ("R", C.c_int32), ("dtype", C.c_int32), ("scan_u8", C.c_int32),
                ("width", C.c_int32), ("guard", C.c_int32), ("diff", C.c_int32),
                ("res", C.c_double), ("minr", C.c_double), ("maxr", C.c_double), ("a_thresh", C.c_double),
                ("b_thresh", C.c_double), ("steep_fact", C.c_double)]


class ExtractParams(C.Structure):
    """mirror of `mmb_extract_params`"""
    _fields_ = [("B", C.c_int32), ("A", C.c_int32), ("R", C.c_int32), ("dtype", C.c_int32), ("diff", C.c_int32),
                ("fused", C.c_int32), ("res", C.c_double), ("steep_fact", C.c_double), ("cfar", CfarParams)]


class KStrongParams(C.Structure):
    """mirror of `mmb_kstrong_params`"""
    _fields_ = [("B", C.c_int32), ("A", C.c_int32), ("R", C.c_int32), ("dtype", C.c_int32), ("scan_u8", C.c_int32),
                ("k", C.c_int32), ("res", C.c_double), ("minr", C.c_double), ("maxr", C.c_double),
                ("threshold2", C.c_double), ("threshold3", C.c_double), ("range_offset", C.c_double)]


# name -> (restype, argtypes); tests check that every one of these is exported.
_P = C.c_void_p
SIGNATURES = {
    "mmb_abi_version": (C.c_int, []),
    "mmb_status_string": (C.c_char_p, [C.c_int]),
    "mmb_last_cuda_error": (C.c_char_p, []),
    "mmb_launch_count": (C.c_longlong, []),
    "mmb_icp_workspace_bytes": (C.c_size_t, [C.POINTER(IcpParams)]),
    "mmb_icp_saved_bytes": (C.c_size_t, [C.POINTER(IcpParams)]),
    "mmb_icp_forward": (C.c_int, [C.POINTER(IcpParams)] + [_P] * 15 + [C.c_size_t, _P]),
    "mmb_icp_forward_profiled": (C.c_int, [C.POINTER(IcpParams)] + [_P] * 15 + [C.c_size_t, _P,
                                                                                 C.POINTER(C.c_float), C.POINTER(C.c_int32)]),
    "mmb_icp_backward": (C.c_int, [C.POINTER(IcpParams)] + [_P] * 11 + [C.c_size_t, _P]),
    "mmb_nn_workspace_bytes": (C.c_size_t, [C.c_int32] * 6),
    "mmb_nn_search": (C.c_int, [C.c_int32] * 8 + [_P] * 4 + [C.c_size_t, _P]),
    "mmb_nn_search_static": (C.c_int, [C.c_int32] * 8 + [_P] * 4),
    "mmb_sample_weights_fwd": (C.c_int, [C.c_int32] * 6 + [C.c_double, _P, _P, _P, _P]),
    "mmb_sample_weights_bwd": (C.c_int, [C.c_int32] * 6 + [C.c_double, _P, _P, _P, _P]),
    "mmb_sample_weights_bwd_det_workspace_bytes": (C.c_size_t, [C.c_int32] * 5),
    "mmb_sample_weights_bwd_det": (C.c_int, [C.c_int32] * 6 + [C.c_double, _P, _P, _P, _P, C.c_size_t, _P]),
    "mmb_polar_to_cart": (C.c_int, [C.c_int32] * 6 + [_P] * 4 + [C.c_double, C.c_int32, C.c_int32, _P, _P]),
    "mmb_polar_to_cart_bwd": (C.c_int, [C.c_int32] * 5 + [_P] * 4 + [C.c_double, C.c_int32, C.c_int32, _P, _P]),
    "mmb_cart_to_polar": (C.c_int, [C.c_int32] * 6 + [C.c_double] + [_P] * 5),
    "mmb_cart_to_polar_bwd": (C.c_int, [C.c_int32] * 6 + [C.c_double] + [_P] * 5),
    "mmb_bev_from_points": (C.c_int, [C.c_int32] * 5 + [C.c_double, C.c_double, _P, _P, _P]),
    "mmb_cfar_columns": (C.c_int, [C.POINTER(CfarParams)] + [C.POINTER(C.c_int32)] * 3),
    "mmb_cfar_mask": (C.c_int, [C.POINTER(CfarParams), _P, _P, _P]),
    "mmb_extract_workspace_bytes": (C.c_size_t, [C.POINTER(ExtractParams)]),
    "mmb_extract_count": (C.c_int, [C.POINTER(ExtractParams), _P, _P, C.c_size_t, _P, _P]),
    "mmb_extract_write": (C.c_int, [C.POINTER(ExtractParams), _P, _P, _P, C.c_size_t, C.c_int32, _P, _P,
                                    C.c_int32, _P, _P]),
    "mmb_kstrong_workspace_bytes": (C.c_size_t, [C.POINTER(KStrongParams)]),
    "mmb_kstrong_select": (C.c_int, [C.POINTER(KStrongParams), _P, _P, C.c_size_t, _P, _P]),
    "mmb_kstrong_write": (C.c_int, [C.POINTER(KStrongParams), _P, _P, C.c_size_t, C.c_int32, _P, _P, _P, _P]),
    "mmb_fp32_probe": (C.c_int, [C.c_int32] * 3 + [_P, _P]),
}

_lib = None


class MmbError(RuntimeError):
    pass


def load_library(path: str | None = None):
    """dlopen the kernel library and bind every symbol of the header."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or LIB_PATH
    if not os.path.exists(path):
        raise MmbError(
            f"{path} not found: build the CUDA library first (python -c 'import __graft_entry__ as g; g.build()'). "
            "mm_masking_b200 has no CPU fallback.")
    lib = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    if lib.mmb_abi_version() != 5:
        raise MmbError("libmm_masking_b200.so ABI version mismatch")
    _lib = lib
    return lib


def check(status: int, what: str):
    if status == 0:
        return
    lib = load_library()
    msg = lib.mmb_status_string(status).decode()
    if status == 6:
        msg += ": " + lib.mmb_last_cuda_error().decode()
    raise MmbError(f"{what} failed: {msg}")


_cuda_ok = False


def require_cuda():
    global _cuda_ok
    if _cuda_ok:
        return
    import torch
    if not torch.cuda.is_available():
        raise MmbError("mm_masking_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    _cuda_ok = True


def ptr(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def stream_ptr():
    """the current CUDA stream of the current device as a raw `cudaStream_t` (no Stream object is built)"""
    import torch
    return C.c_void_p(torch._C._cuda_getCurrentRawStream(torch.cuda.current_device()))


class _Here:
    """no-op context: the tensor's device is already the current one"""
    def __enter__(self):
        return self
    def __exit__(self, *exc):
        return False


_HERE = _Here()


def on_device(dev):
    """`torch.cuda.device(dev)` only when `dev` is not the current device (the context manager costs ~10 us a call)"""
    import torch
    return _HERE if dev.index is None or dev.index == torch.cuda.current_device() else torch.cuda.device(dev)
