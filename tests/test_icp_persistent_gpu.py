"""GPU: the persistent launch of the fused row-major backend in the corners its work queue has --
fewer work items than resident blocks (blocks of later iterations spin on the pair's `ready` flag from the start),
one query group per pair (an iteration is one item: every item waits for the one before), pairs that converge at
different iterations (their items are skipped and the finish kernel repeats their histories, `ICP.py:215-224,
246-247`), and the early stop once every pair has converged.  The reference for each case is the launch-per-iteration
path of the exhaustive backend on the same inputs: identical correspondences, so every output agrees to rounding."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(batch, backend, iters, tol, const_iter, dim, dtype=torch.float32):
    from mm_masking_b200.dICP import ICP
    tt = lambda a: torch.tensor(a, dtype=dtype, device="cuda")
    w = tt(batch["weight"]).requires_grad_(True)
    alg = ICP(icp_type="pt2pl", differentiable=True, max_iterations=iters, tolerance=tol)
    alg.const_iter = const_iter
    alg.nn_backend = backend
    alg.record_indices = True
    out = alg.icp(tt(batch["source"]), tt(batch["target"]), tt(batch["T_init"]), weight=w, trim_dist=5.0,
                  loss_fn={"name": "cauchy", "metric": 1.0}, dim=dim)
    out["T"].sum().backward()
    torch.cuda.synchronize()
    return out, w.grad, alg.last_indices


def _same(a, b, tol, what):
    assert a.shape == b.shape, (what, a.shape, b.shape)
    a, b = a.detach(), b.detach()
    scale = max(float(b.abs().max()), 1.0)
    assert float((a - b).abs().max()) <= tol * scale, (what, float((a - b).abs().max()), scale)


CASES = [
    # N, n, m, iterations, tolerance, const_iter, three_d
    (1, 40, 9000, 40, 1e-6, True, False),      # 2 query groups: 80 items for 740 resident blocks, all iterations run
    (1, 30, 9000, 60, 2e-3, False, False),     # 1 query group per iteration, converges on the way, early stop
    (5, 300, 12000, 30, 2e-3, False, False),   # pairs converge at different iterations
    (4, 200, 20000, 25, 2e-3, False, True),    # the same in SE(3)
    (37, 64, 8192, 12, 1e-6, True, False),     # more pairs than a 32-lane window, two groups each
]


@pytest.mark.parametrize("N,n,m,iters,tol,const_iter,three_d", CASES)
def test_persistent_launch_matches_launch_per_iteration(N, n, m, iters, tol, const_iter, three_d):
    from mm_masking_b200 import synth, _capi
    batch = synth.make_icp_batch(N=N, n=n, m=m, seed=100 + N + n, three_d=three_d, pos=1.0, rot=0.3)
    dim = 3 if three_d else 2
    ref, g_ref, idx_ref = _run(batch, _capi.NN_BRUTE, iters, tol, const_iter, dim)
    out, g, idx = _run(batch, _capi.NN_ROWS, iters, tol, const_iter, dim)
    again, g2, idx2 = _run(batch, _capi.NN_ROWS, iters, tol, const_iter, dim)

    it = ref["deltas"].shape[1]
    assert out["deltas"].shape[1] == it, "a different number of iterations was executed"
    if not const_iter and N > 1:
        conv_at = ref["stats"]["iterations"].cpu().numpy()
        assert len(np.unique(conv_at)) > 1 or it < iters, "the case is meant to have pairs converge at different iterations"
    # correspondences of the executed iterations of pairs that were still active are identical
    active = torch.arange(it, device="cuda")[:, None] < ref["stats"]["iterations"].to(torch.int64)[None, :]
    if const_iter:
        active[:] = True
    a, b = idx[:it][active], idx_ref[:it][active]
    assert torch.equal(a, b), int((a != b).sum())
    for key in ("T", "pc", "deltas", "weights", "costs"):
        _same(out[key], ref[key], 2e-5, key)
    for key in ("converged", "iterations"):
        assert torch.equal(out["stats"][key], ref["stats"][key]), key
    _same(out["stats"]["matched_ratio"], ref["stats"]["matched_ratio"], 2e-3, "matched_ratio")   # counts of w > 0.1
    _same(g, g_ref, 1e-4, "weight gradient")
    # run-to-run: bit for bit, whatever order the blocks took the items in
    for key in ("T", "pc", "deltas", "weights", "costs"):
        assert torch.equal(out[key], again[key]), key
    assert torch.equal(g, g2)
    assert torch.equal(idx[:it][active], idx2[:it][active])
