"""GPU: `mmb_nn_search_static` -- the nearest-neighbour search with its scratch memory in a file-scope device array
of the library (no allocation inside the call; include/mm_masking_b200.h).  No counterpart in the reference,
so the check is against the existing entry point
`mmb_nn_search` (caller's workspace) on the same inputs, through the C ABI, on calls in a row that share the array.
Indices are integer work: the tolerance is bit-exact, as in tests/test_nn_gpu.py.

The file name makes pytest collect it last."""
import ctypes as C

import pytest
import torch

from mm_masking_b200 import _capi

pytestmark = pytest.mark.gpu

STATIC_BYTES = 16 << 20          # MMB_NN_STATIC_SCRATCH_BYTES


def _clouds(seed, N, n, m, dtype, dim):
    g = torch.Generator().manual_seed(seed)
    tgt = (torch.rand((N, m, 6), generator=g, dtype=torch.float64) * 150 - 75).to(dtype)
    qry = (torch.rand((N, n, 3), generator=g, dtype=torch.float64) * 170 - 85).to(dtype)     # some outside the map
    if dim == 2:
        tgt[:, :, 2] = 0
        qry[:, :, 2] = 0
    tgt[:, m // 2:m // 2 + 50] = tgt[:, :50]                  # exact ties: the lowest index wins in both
    return qry.cuda().contiguous(), tgt.cuda().contiguous()


def _dt(dtype):
    return _capi.F32 if dtype == torch.float32 else _capi.F64


def _with_workspace(lib, qry, tgt, dim, backend):
    N, n, m = qry.shape[0], qry.shape[1], tgt.shape[1]
    nbytes = lib.mmb_nn_workspace_bytes(N, n, m, _dt(qry.dtype), 0, backend)
    assert 0 < nbytes <= STATIC_BYTES
    ws = torch.empty(nbytes, dtype=torch.uint8, device=qry.device)
    idx = torch.full((N, n), -7, dtype=torch.int32, device=qry.device)
    rc = lib.mmb_nn_search(N, n, m, tgt.shape[2], _dt(qry.dtype), dim, 0, backend, _capi.ptr(qry), _capi.ptr(tgt),
                           _capi.ptr(idx), _capi.ptr(ws), nbytes, _capi.stream_ptr())
    _capi.check(rc, "mmb_nn_search")
    torch.cuda.synchronize()
    return idx


def _static(lib, qry, tgt, dim, backend):
    N, n, m = qry.shape[0], qry.shape[1], tgt.shape[1]
    idx = torch.full((N, n), -7, dtype=torch.int32, device=qry.device)
    rc = lib.mmb_nn_search_static(N, n, m, tgt.shape[2], _dt(qry.dtype), dim, 0, backend, _capi.ptr(qry),
                                  _capi.ptr(tgt), _capi.ptr(idx), _capi.stream_ptr())
    _capi.check(rc, "mmb_nn_search_static")
    torch.cuda.synchronize()
    return idx


# (seed, N, n, m, dtype, dim, backend): consecutive calls differ in shape, dtype and backend, so each one finds the
# array laid out by another problem (cell tables, sorted copies and partial minima at other offsets)
CALLS = [
    (1, 3, 1500, 30000, torch.float32, 2, _capi.NN_ROWS),
    (2, 2, 900, 12000, torch.float32, 3, _capi.NN_ROWS),
    (3, 2, 700, 9000, torch.float64, 2, _capi.NN_GRID),
    (4, 2, 300, 2500, torch.float32, 2, _capi.NN_BRUTE),
    (5, 1, 1200, 40000, torch.float64, 3, _capi.NN_AUTO),
    (1, 3, 1500, 30000, torch.float32, 2, _capi.NN_ROWS),     # the first problem again, over the others' leftovers
]


def test_static_scratch_matches_the_workspace_entry_point_on_calls_in_a_row():
    lib = _capi.load_library()
    for seed, N, n, m, dtype, dim, backend in CALLS:
        qry, tgt = _clouds(seed, N, n, m, dtype, dim)
        want = _with_workspace(lib, qry, tgt, dim, backend)
        got = _static(lib, qry, tgt, dim, backend)
        again = _static(lib, qry, tgt, dim, backend)           # straight after, over its own leftovers
        assert int(want.min()) >= 0 and int(want.max()) < m
        assert torch.equal(got, want), (seed, N, n, m, dtype, dim, backend)
        assert torch.equal(again, want), (seed, N, n, m, dtype, dim, backend)


def test_static_scratch_allocates_nothing_and_rejects_what_does_not_fit():
    lib = _capi.load_library()
    qry, tgt = _clouds(9, 2, 800, 20000, torch.float32, 2)
    _static(lib, qry, tgt, 2, _capi.NN_ROWS)                   # set-up call (symbol address looked up once per device)
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        _static(lib, qry, tgt, 2, _capi.NN_ROWS)
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0                                      # the driver handed out no memory during the calls

    # a problem whose workspace exceeds the array is refused before any launch: idx untouched, no CUDA error
    N, n, m = 8, 2000, 200000
    assert lib.mmb_nn_workspace_bytes(N, n, m, _capi.F32, 0, _capi.NN_ROWS) > STATIC_BYTES
    idx = torch.full((4,), -7, dtype=torch.int32, device="cuda")
    launches = lib.mmb_launch_count()
    rc = lib.mmb_nn_search_static(N, n, m, 6, _capi.F32, 2, 0, _capi.NN_ROWS, _capi.ptr(qry), _capi.ptr(tgt),
                                  _capi.ptr(idx), _capi.stream_ptr())
    assert rc == 7                                             # MMB_ERR_SCRATCH_LIMIT
    assert b"MMB_NN_STATIC_SCRATCH_BYTES" in lib.mmb_status_string(rc)
    assert lib.mmb_launch_count() == launches
    torch.cuda.synchronize()
    assert torch.all(idx == -7)
    with pytest.raises(_capi.MmbError):
        _capi.check(rc, "mmb_nn_search_static")

    # argument checks are those of mmb_nn_search
    assert lib.mmb_nn_search_static(0, 10, 10, 6, _capi.F32, 2, 0, 0, _capi.ptr(qry), _capi.ptr(tgt), _capi.ptr(idx),
                                    _capi.stream_ptr()) == 2
    assert lib.mmb_nn_search_static(1, 10, 10, 6, _capi.F32, 2, 0, 0, None, _capi.ptr(tgt), _capi.ptr(idx),
                                    _capi.stream_ptr()) == 1
