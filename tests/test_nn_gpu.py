"""GPU: nearest-neighbour search (K3) -- the reference's toy test, tie-breaking, and the oracle
on random clouds, through `mmb_nn_search`."""
import numpy as np
import pytest
import torch

from oracle import dicp_oracle as O

pytestmark = pytest.mark.gpu


def test_toy_points():
    """`external/dICP/tests/test_nn.py:9-41`"""
    from mm_masking_b200.dICP import nn
    pts = torch.tensor([(5.0, 4.0, 0.0), (2.0, 6.0, 0.0), (13.0, 3.0, 0.0), (8.0, 7.0, 0.0), (3.0, 1.0, 0.0)],
                       requires_grad=True)
    q = torch.tensor([[9.0, 4.0, 0.0]], requires_grad=True)
    finder = nn(differentiable=True, use_gumbel=False)
    near = finder.find_nn(q, pts)
    assert torch.all(near == torch.tensor((8.0, 7.0, 0.0)))
    near.sum().backward()
    assert pts.grad is not None and not torch.isnan(pts.grad).any()
    pts2 = torch.cat((pts.detach(), torch.tensor([[10.0, 2.0, 0.0]])))
    assert torch.all(finder.find_nn(q, pts2) == torch.tensor((10.0, 2.0, 0.0)))
    nondiff = nn(differentiable=False)
    assert torch.all(nondiff.find_nn(q, pts2) == torch.tensor((10.0, 2.0, 0.0)))


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("dim", [2, 3])
def test_exact_ties_take_the_lowest_index(dtype, dim):
    from mm_masking_b200.dICP.nn import nearest_index
    g = torch.Generator().manual_seed(5)
    N, n, m = 2, 700, 5000
    tgt = (torch.rand((N, m, 3), generator=g, dtype=torch.float64) * 160 - 80).to(dtype)
    if dim == 2:
        tgt[:, :, 2] = 0
    # every target row exists 4 times, at scattered positions
    tgt[:, 1250:2500] = tgt[:, :1250]
    tgt[:, 2500:3750] = tgt[:, :1250].flip(1)
    tgt[:, 3750:] = tgt[:, :1250]
    q = tgt[:, torch.randperm(m, generator=g)[:n]].clone()           # queries coincide with targets: distance 0
    q[:, ::2] += 0.01
    if dim == 2:
        q[:, :, 2] = 0
    for slices, backend in ((0, 1), (2, 1), (5, 1), (0, 2), (0, 3)):
        idx = nearest_index(q.cuda(), tgt.cuda(), dim=dim, nn_slices=slices, nn_backend=backend).cpu()
        ref = O.nearest_index(q.double(), tgt.double(), "exact")
        assert torch.equal(idx, ref)
        assert int(idx.max()) < 1250                                   # always the first copy


@pytest.mark.parametrize("shape", [(1, 1, 1), (1, 65, 65), (3, 513, 1025), (2, 2000, 7000), (1, 100, 50000)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_random_clouds_match_oracle(shape, dtype):
    from mm_masking_b200.dICP.nn import nearest_index
    N, n, m = shape
    g = torch.Generator().manual_seed(N * 1000 + n + m)
    tgt = (torch.rand((N, m, 6), generator=g, dtype=torch.float64) * 160 - 80).to(dtype)
    q = (torch.rand((N, n, 3), generator=g, dtype=torch.float64) * 160 - 80).to(dtype)
    if m > 10:
        tgt[:, -m // 10:, :] = 80000.0                                  # padding rows (ICP.py:446)
    for dim in (2, 3):
        qq, tt = q.double().clone(), tgt[:, :, :3].double().clone()
        if dim == 2:
            qq[:, :, 2] = 0; tt[:, :, 2] = 0
        ref = O.nearest_index(qq, tt, "exact")
        for backend in (0, 1, 2, 3):                                    # auto, brute force, Morton grid, row-major grid
            idx = nearest_index(q.cuda(), tgt.cuda(), dim=dim, nn_backend=backend).cpu()
            assert torch.equal(idx, ref), f"backend {backend}"


def test_clustered_near_ties_fp64_queries():
    """Queries that are not fp32-representable, next to tight clusters of targets: the fp32 filter
    must hand every candidate within its error band to the exact comparison."""
    from mm_masking_b200.dICP.nn import nearest_index
    g = torch.Generator().manual_seed(9)
    N, n, m = 1, 512, 4096
    centers = torch.rand((N, m // 8, 3), generator=g, dtype=torch.float64) * 150 - 75
    tgt = centers.repeat_interleave(8, dim=1) + 1e-6 * torch.randn((N, m, 3), generator=g, dtype=torch.float64)
    q = centers[:, :n] + 1e-3 * torch.randn((N, n, 3), generator=g, dtype=torch.float64)
    for backend in (1, 2, 3):
        idx = nearest_index(q.cuda(), tgt.cuda(), dim=3, nn_backend=backend).cpu()
        assert torch.equal(idx, O.nearest_index(q, tgt, "exact"))


def test_cold_start_bound_pass_keeps_every_tie():
    """m >= 8192 switches on the cold-start bound pass (strided subsample, fp32 only): its threshold
    must still let every exact tie and near-tie of the full set through, lowest index first."""
    from mm_masking_b200.dICP.nn import nearest_index
    g = torch.Generator().manual_seed(21)
    N, n, m = 2, 600, 20000
    centers = torch.rand((N, m // 8, 3), generator=g, dtype=torch.float64) * 150 - 75
    tgt = centers.repeat_interleave(8, dim=1) + 1e-6 * torch.randn((N, m, 3), generator=g, dtype=torch.float64)
    tgt[:, 8::16] = tgt[:, 0::16]                                        # exact duplicates at a higher index
    perm = torch.randperm(m, generator=g)
    tgt = tgt[:, perm]
    q = centers[:, :n] + 1e-3 * torch.randn((N, n, 3), generator=g, dtype=torch.float64)
    for dtype in (torch.float64, torch.float32):
        qq, tt = q.to(dtype), tgt.to(dtype)
        for slices, backend in ((0, 1), (3, 1), (0, 2), (0, 3)):
            idx = nearest_index(qq.cuda(), tt.cuda(), dim=3, nn_slices=slices, nn_backend=backend).cpu()
            assert torch.equal(idx, O.nearest_index(qq.double(), tt.double(), "exact"))


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_cell_grid_clamped_cells_and_far_queries(dtype):
    """Cell-grid backend corner cases: the grid covers the QUERY footprint, so map points far outside
    it (pad rows, a distant wall) are clamped into border cells and queries far from every target
    must climb the levels; the answer must still be the brute-force one."""
    from mm_masking_b200.dICP.nn import nearest_index
    g = torch.Generator().manual_seed(33)
    N, n, m = 3, 300, 9000
    q = (torch.rand((N, n, 3), generator=g, dtype=torch.float64) * 20 - 10)          # footprint 20 m
    tgt = torch.empty((N, m, 3), dtype=torch.float64)
    tgt[0] = torch.rand((m, 3), generator=g, dtype=torch.float64) * 400 - 200          # mostly outside the grid
    tgt[1] = torch.rand((m, 3), generator=g, dtype=torch.float64) * 2 + 300            # everything in one corner
    tgt[2] = torch.rand((m, 3), generator=g, dtype=torch.float64) * 30 - 15
    tgt[2, ::3] = 1.0e5                                                               # pad rows
    q[2, :10] = 5000.0                                                                # queries off the grid
    q[2, 10:20] = 0.0                                                                 # zero-padded source rows
    q, tgt = q.to(dtype), tgt.to(dtype)
    for dim in (2, 3):
        qq, tt = q.double().clone(), tgt.double().clone()
        if dim == 2:
            qq[:, :, 2] = 0; tt[:, :, 2] = 0
        ref = O.nearest_index(qq, tt, "exact")
        for backend in (2, 3):
            idx = nearest_index(q.cuda(), tgt.cuda(), dim=dim, nn_backend=backend).cpu()
            assert torch.equal(idx, ref), backend
