"""CPU-side checks of the C-ABI boundary: the in-tree library loads, exports every symbol that
include/*.h declares, the ctypes signatures in `_capi.py` have the header's arity, and the
argument-validation paths (which run before any CUDA call) return the documented status codes.
No kernel is launched here."""
import ctypes as C
import glob
import os
import re

import pytest

from conftest import ROOT
from mm_masking_b200 import _capi


def _header_functions():
    """name -> number of parameters, parsed from the prototypes in include/*.h"""
    out = {}
    for path in glob.glob(os.path.join(ROOT, "include", "*.h")):
        text = open(path).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        text = re.sub(r"//[^\n]*", "", text)
        for m in re.finditer(r"\b(mmb_\w+)\s*\(([^;{}]*?)\)\s*;", text, flags=re.S):
            name, args = m.group(1), m.group(2).strip()
            out[name] = 0 if args in ("", "void") else args.count(",") + 1
    return out


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(_capi.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return _capi.load_library()


def test_every_declared_symbol_is_exported_and_bound(lib):
    decl = _header_functions()
    assert len(decl) >= 10
    raw = C.CDLL(_capi.LIB_PATH)
    for name, nargs in decl.items():
        assert hasattr(raw, name), f"{name} declared in include/ but not exported"
        assert name in _capi.SIGNATURES, f"{name} has no ctypes signature in _capi.py"
        assert len(_capi.SIGNATURES[name][1]) == nargs, f"{name}: ctypes arity differs from the header"
    for name in _capi.SIGNATURES:
        assert name in decl, f"{name} bound in _capi.py but not declared in include/"


def test_status_strings(lib):
    assert lib.mmb_abi_version() == 5
    assert lib.mmb_status_string(0) == b"ok"
    assert lib.mmb_status_string(4) == b"workspace too small"
    assert lib.mmb_status_string(99) == b"unknown status"


def _params(**kw):
    p = _capi.IcpParams()
    p.N, p.n, p.m, p.target_cols = 2, 100, 1000, 6
    p.dtype, p.icp_type, p.dim, p.max_iterations = _capi.F32, _capi.PT2PL, 2, 5
    p.differentiable, p.loss_kind, p.save_for_backward = 1, _capi.LOSS_CAUCHY, 1
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def test_icp_argument_validation_runs_before_cuda(lib):
    p = _params()
    assert lib.mmb_icp_workspace_bytes(p) > 0
    assert lib.mmb_icp_saved_bytes(p) > 0
    assert lib.mmb_icp_saved_bytes(_params(save_for_backward=0)) == 0
    assert lib.mmb_icp_workspace_bytes(_params(target_cols=5)) == 0
    null = [None] * 15
    assert lib.mmb_icp_forward(p, *null, 0, None) == 1                       # null pointer
    assert lib.mmb_icp_forward(_params(n=0), *null, 0, None) == 2            # invalid shape
    assert lib.mmb_icp_forward(_params(icp_type=_capi.PT2PL, target_cols=3), *null, 0, None) == 2
    assert lib.mmb_icp_forward(_params(dtype=7), *null, 0, None) == 3        # unsupported
    assert lib.mmb_icp_backward(p, *([None] * 11), 0, None) == 1
    # workspace too small / misaligned are reported before anything is enqueued
    fake = [C.c_void_p(256)] * 15
    assert lib.mmb_icp_forward(p, *fake, 16, None) == 4
    fake[14] = C.c_void_p(264)
    assert lib.mmb_icp_forward(p, *fake, 1 << 40, None) == 5


def test_nn_argument_validation(lib):
    assert lib.mmb_nn_workspace_bytes(1, 10, 10, 0, 0, 0) > 0
    assert lib.mmb_nn_workspace_bytes(0, 10, 10, 0, 0, 0) == 0
    assert lib.mmb_nn_search(1, 10, 10, 4, 0, 3, 0, 0, None, None, None, None, 0, None) == 2
    assert lib.mmb_nn_search(1, 10, 10, 3, 0, 3, 0, 0, None, None, None, None, 0, None) == 1
    assert lib.mmb_nn_search(1, 10, 10, 3, 5, 3, 0, 0, None, None, None, None, 0, None) == 3


def test_product_refuses_to_run_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from mm_masking_b200.dICP import ICP
    alg = ICP(icp_type="pt2pt", max_iterations=2)
    with pytest.raises(_capi.MmbError):
        alg.icp(torch.zeros(4, 3), torch.zeros(5, 3), torch.eye(4))
