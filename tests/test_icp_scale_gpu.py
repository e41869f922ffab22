"""GPU parity at the sizes bench.py times (BASELINE.json configs[2] and configs[4]): the whole ICP call --
correspondence indices of EVERY iteration, pose, weight gradient -- for both nearest-neighbour backends
against the CPU oracle (`oracle/dicp_oracle.py`, fp64, `nn_mode="kdtree"`: the exhaustive search of
`external/dICP/dICP/nn.py:32-38` evaluated on k-d tree proposals; pinned to the exhaustive mode in
`tests/test_oracle_properties_cpu.py`).  Tolerances are BASELINE.json's: indices bit-exact, 1e-5 m / 1e-6 rad,
gradients 1e-4 relative (fp32 I/O).  Shapes follow `external/dICP/tests/test_ICP_inputs.py:105-110`
(batched call, per-pair weights)."""
import numpy as np
import pytest
import torch

from oracle import dicp_oracle as O

pytestmark = pytest.mark.gpu

CALL = dict(trim_dist=5.0, loss_fn={"name": "cauchy", "metric": 1.0})


def _pose_err(T, T_ref):
    T = np.asarray(T, dtype=np.float64); T_ref = np.asarray(T_ref, dtype=np.float64)
    dt = np.abs(T[:, :3, 3] - T_ref[:, :3, 3]).max()
    dR = np.einsum("bij,bkj->bik", T[:, :3, :3], T_ref[:, :3, :3])
    sk = np.stack([dR[:, 2, 1] - dR[:, 1, 2], dR[:, 0, 2] - dR[:, 2, 0], dR[:, 1, 0] - dR[:, 0, 1]], 1) / 2
    return dt, np.linalg.norm(sk, axis=1).max()


def _oracle(batch, dim, iters, const_iter, tol):
    t64 = lambda a: torch.tensor(np.asarray(a, dtype=np.float64))
    w = t64(batch["weight"]).requires_grad_(True)
    ref = O.dicp(t64(batch["source"]), t64(batch["target"]), t64(batch["T_init"]), w, icp_type="pt2pl",
                 max_iterations=iters, tolerance=tol, differentiable=True, dim=dim, const_iter=const_iter,
                 nn_mode="kdtree", return_indices=True, **CALL)
    ref["T"].sum().backward()
    return ref, w.grad


def _ours(batch, dim, iters, const_iter, tol, backend):
    from mm_masking_b200.dICP import ICP
    tt = lambda a: torch.tensor(a, dtype=torch.float32, device="cuda")
    w = tt(batch["weight"]).requires_grad_(True)
    alg = ICP(icp_type="pt2pl", differentiable=True, max_iterations=iters, tolerance=tol)
    alg.const_iter = const_iter
    alg.nn_backend = backend
    alg.record_indices = True
    out = alg.icp(tt(batch["source"]), tt(batch["target"]), tt(batch["T_init"]), weight=w, dim=dim, **CALL)
    out["T"].sum().backward()
    return out, w.grad, alg.last_indices


def _check(batch, dim, iters, const_iter=True, tol=1e-5, backends=None):
    from mm_masking_b200 import _capi
    ref, g_ref = _oracle(batch, dim, iters, const_iter, tol)
    done = ref["indices"].shape[0]
    for backend in backends or (_capi.NN_ROWS, _capi.NN_GRID, _capi.NN_BRUTE):
        out, g, idx = _ours(batch, dim, iters, const_iter, tol, backend)
        idx = idx[:done].cpu().numpy().astype(np.int64)
        ref_idx = ref["indices"].numpy()
        if not const_iter:
            # a pair that froze keeps no new correspondences (ICP.py:246-247 zeroes its weights): compare
            # only the iterations in which the pair was still being updated
            its = ref["stats"]["iterations"].numpy().astype(int)
            for b, kb in enumerate(its):
                idx[kb:, b] = ref_idx[kb:, b]
            np.testing.assert_array_equal(out["stats"]["iterations"].cpu().numpy(), ref["stats"]["iterations"].numpy())
        for k in range(done):
            mism = int((idx[k] != ref_idx[k]).sum())
            assert mism == 0, f"backend {backend}, iteration {k}: {mism} of {idx[k].size} indices differ"
        dt, dr = _pose_err(out["T"].detach().cpu().numpy(), ref["T"].detach().numpy())
        assert dt < 1e-5 and dr < 1e-6, (backend, dt, dr)
        scale = float(g_ref.abs().max())
        dg = float((g.double().cpu() - g_ref).abs().max())
        assert dg <= 1e-4 * scale, (backend, dg, scale)
        np.testing.assert_allclose(out["weights"].squeeze(-1).double().cpu().numpy(),
                                   ref["weights"].squeeze(-1).numpy(), atol=2e-6)


def test_c3_shape_both_backends():
    """BASELINE.json configs[2]: 2 000 x 50 000, dim 2, 10 iterations, train-style perturbation (2 m / 0.6 rad)."""
    from mm_masking_b200 import synth
    _check(synth.make_icp_batch(N=4, n=2000, m=50000, seed=0), dim=2, iters=10)


def test_c5_shape_both_backends():
    """BASELINE.json configs[4]: 200 000-point submaps, SE(3), 20 iterations (the 2-D cell grid searched with
    a 3-D radius)."""
    from mm_masking_b200 import synth
    _check(synth.make_icp_batch(N=2, n=2000, m=200000, seed=1, three_d=True), dim=3, iters=20)


def test_dense_map_forces_the_sub_cell_split():
    """200 000 targets in 2-D: level >= 2 cells hold far more than GRID_SPLIT targets, so the far (clutter)
    queries go through the 4 x 4 sub-cell culling."""
    from mm_masking_b200 import synth, _capi
    _check(synth.make_icp_batch(N=2, n=2000, m=200000, seed=2), dim=2, iters=6, backends=(_capi.NN_ROWS, _capi.NN_GRID))


def test_half_map_forces_the_outside_bucket():
    """The map covers only x < 10 m while the scan covers +-76 m: queries on the empty side have bounds of
    tens of metres, their search square leaves the grid and the outside bucket (incl. pad rows) is scanned."""
    from mm_masking_b200 import synth, _capi
    batch = synth.make_icp_batch(N=2, n=2000, m=50000, seed=3)
    tgt = batch["target"]
    gone = tgt[:, :, 0] > 10.0
    tgt[gone] = synth.MAP_PAD_VALUE
    _check(batch, dim=2, iters=6, backends=(_capi.NN_ROWS, _capi.NN_GRID))


def test_early_exit_and_freezing_at_scale():
    """`const_iter: False` (the reference default): pairs converge at different iterations, freeze, and the
    loop stops early; iterations, indices while active, pose and gradient against the oracle."""
    from mm_masking_b200 import synth
    _check(synth.make_icp_batch(N=3, n=2000, m=50000, seed=4, pos=0.5, rot=0.1), dim=2, iters=40,
           const_iter=False, tol=2e-3)


def test_map_much_larger_than_the_scan_footprint():
    """the cell grid is fitted to the scan footprint; with a map 2-3x as large most targets sit in the
    outside bucket and the queries near the border (or far from everything) scan it.  Exactness must not depend on
    where a target was binned; `const_iter: False`, so pairs also freeze at different iterations."""
    from mm_masking_b200 import synth, _capi
    batch = synth.make_icp_batch(N=3, n=2000, m=50000, seed=6, pos=1.0, rot=0.2, scan_range=30.0)
    _check(batch, dim=2, iters=25, const_iter=False, tol=2e-3, backends=(_capi.NN_ROWS, _capi.NN_GRID))
