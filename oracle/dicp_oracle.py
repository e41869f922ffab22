"""CPU oracle for the weighted differentiable ICP path.  TEST INFRASTRUCTURE ONLY.

This file restates, in plain CPU torch (so that autograd yields the reference's
*unrolled* gradient), the algorithm of the reference's `external/dICP/dICP/ICP.py:48-293`
(`ICP.dICP`), `nn.py:23-40` (brute-force 1-NN) and `loss.py:21-56` (trim / Huber / Cauchy
weights).  Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline legs may
import it; the product (`mm_masking_b200/`) never does and fails loudly without its CUDA
library.

Pinning: `tests/test_oracle_golden.py` checks this restatement against vectors produced by
the unmodified reference (`oracle/make_golden.py`, run in the build container where
`/root/reference` is mounted): the reference's own 65-point fixture poses
(`external/dICP/tests/test_ICP.py:35-148`), seeded synthetic pairs in fp64 and fp32, per-iteration
correspondence indices and autograd gradients.

Conventions restated here (all file:line into `/root/reference/external/dICP/dICP/`):
  * dense inputs only: source (N,n,3), target (N,m,3|6), T_init (N,4,4), weight (N,n);
    ragged-list handling (`ICP.py:295-497`) is host logic tested separately.
  * dim == 2 zeroes source z, target z and normal z without renormalising (`ICP.py:105-114`).
  * pt2pt applies trim / loss weights per residual component (`ICP.py:147,155`).
  * w_sqrt = sqrt(w + 1e-10) - 1e-5, A = J^T W J + 1e-12 I, delta = -A^-1 J^T W e (`ICP.py:184-191`).
  * C <- exp(dphi^)^T C, r <- r - dr (`ICP.py:200-207`).
  * converged samples are frozen by zeroing their prior weights; the loop stops when all
    samples satisfy |delta| < tol, unless const_iter (`ICP.py:227-250`).
"""
from __future__ import annotations

import torch


def nearest_index(query: torch.Tensor, target_xyz: torch.Tensor, mode: str = "exact",
                  chunk: int = 256) -> torch.Tensor:
    """1-NN index of every query row in its pair's target, lowest index on ties.

    `nn.py:32-35` computes `torch.cdist(x, y[:, :, :3])` and `argmin(dim=2)`.  mode="cdist"
    calls exactly that (default compute mode: a matmul formulation above 25 rows, whose
    rounding flips some near-ties); mode="exact" evaluates squared differences
    directly, which is the index contract of this repo.
    """
    if mode == "cdist":
        return torch.argmin(torch.cdist(query, target_xyz, p=2), dim=2)
    if mode == "kdtree":
        return nearest_index_kdtree(query, target_xyz)
    N, n, _ = query.shape
    out = torch.empty((N, n), dtype=torch.long)
    for b in range(N):
        t = target_xyz[b]
        for s in range(0, n, chunk):
            q = query[b, s:s + chunk]
            d2 = ((q[:, None, :] - t[None, :, :]) ** 2).sum(dim=2)
            out[b, s:s + chunk] = torch.argmin(d2, dim=1)
    return out


_TREES: dict = {}


def nearest_index_kdtree(query: torch.Tensor, target_xyz: torch.Tensor, k: int = 8) -> torch.Tensor:
    """The same answer as mode="exact" at map sizes where the (n, m) distance table is too slow
    (c3: m = 50 000, c5: m = 200 000): a k-d tree (scipy) proposes the k nearest targets of every
    query, the squared distances of those k are then evaluated with the expression of mode="exact"
    and the winner is the lexicographic minimum of (d^2, index).  The proposal is exact unless more
    than k - 1 targets lie within rounding of the nearest distance (the synthetic maps duplicate
    1 % of their points pairwise, so k = 8 leaves ample room); when the k-th proposal still ties
    with the best, the query falls back to the full scan.  `tests/test_oracle_properties_cpu.py`
    checks it against mode="exact"."""
    import numpy as np
    from scipy.spatial import cKDTree
    N, n, _ = query.shape
    out = torch.empty((N, n), dtype=torch.long)
    for b in range(N):
        t = target_xyz[b].double()
        q = query[b].double()
        kk = min(k, t.shape[0])
        # the map does not move between iterations: keep the tree of the latest few maps
        key = (target_xyz.data_ptr(), tuple(target_xyz.shape), b, float(t.sum()), float(t[-1].sum()))
        tree = _TREES.get(key)
        if tree is None:
            if len(_TREES) >= 16:
                _TREES.clear()
            tree = _TREES[key] = cKDTree(t.numpy())
        _, cand = tree.query(q.numpy(), k=kk)
        cand = torch.from_numpy(np.asarray(cand).reshape(n, kk).astype(np.int64))
        d2 = ((q[:, None, :] - t[cand]) ** 2).sum(dim=2)                     # (n, kk), the "exact" expression
        best = d2.min(dim=1, keepdim=True).values
        tie = d2 == best
        idx = torch.where(tie, cand, torch.full_like(cand, t.shape[0])).min(dim=1).values
        unsure = tie.all(dim=1) & (kk < t.shape[0])                            # every proposal ties: scan everything
        for i in torch.nonzero(unsure).flatten().tolist():
            idx[i] = torch.argmin(((q[i][None, :] - t) ** 2).sum(dim=1))
        out[b] = idx
    return out


def trim_weight(rho: torch.Tensor, trim_dist: float, differentiable: bool, steep: float):
    """`loss.py:40-56`: soft (tanh) or hard gate on the residual norm."""
    if differentiable:
        return 0.5 * torch.tanh(steep * (trim_dist - rho) - 3.0) + 0.5
    return torch.where(rho > trim_dist, torch.zeros_like(rho), torch.ones_like(rho))


def loss_weight(err: torch.Tensor, name: str, metric: float, differentiable: bool):
    """`loss.py:21-38`: pseudo-Huber / hard Huber / Cauchy weights."""
    if name == "huber":
        if differentiable:
            return metric ** 2 / (metric ** 2 + err ** 2)
        return torch.where(err.abs() > metric, metric / err.abs(), torch.ones_like(err))
    if name == "cauchy":
        return 1.0 / (1.0 + (err / metric) ** 2)
    raise ValueError("Invalid loss name: {}".format(name))


def so3_exp(phi: torch.Tensor) -> torch.Tensor:
    """exp(phi^) for (N,3) rotation vectors; the reference calls `torch.matrix_exp` (`ICP.py:200`)."""
    zero = torch.zeros_like(phi[:, 0])
    K = torch.stack([zero, -phi[:, 2], phi[:, 1],
                     phi[:, 2], zero, -phi[:, 0],
                     -phi[:, 1], phi[:, 0], zero], dim=1).reshape(-1, 3, 3)
    return torch.matrix_exp(K)


def dicp(source, target, T_init, weight=None, *, icp_type="pt2pl", max_iterations=100,
         tolerance=1e-12, differentiable=True, trim_dist=None, loss_fn=None, dim=3,
         const_iter=False, tanh_steepness=5.0, match_ratio_thresh=0.1, nn_mode="exact",
         return_indices=False):
    """Batched Gauss-Newton ICP; returns the reference's result dict (`ICP.py:284-291`)."""
    N, n, _ = source.shape
    dtype = source.dtype
    src = source[:, :, :3]
    tgt = target if icp_type == "pt2pl" else target[:, :, :3]
    if icp_type == "pt2pl":
        assert target.shape[2] == 6
    w_init = torch.ones((N, n), dtype=dtype) if weight is None else weight
    if dim == 2:
        keep_s = torch.tensor([1.0, 1.0, 0.0], dtype=dtype)
        src = src * keep_s
        keep_t = torch.tensor([1.0, 1.0, 0.0, 1.0, 1.0, 0.0], dtype=dtype)[: tgt.shape[2]]
        tgt = tgt * keep_t
    C = T_init[:, :3, :3]
    r = T_init[:, :3, 3]
    cols = [2, 3, 4] if dim == 2 else [0, 1, 2, 3, 4, 5]
    d = len(cols)
    eye_d = torch.eye(d, dtype=dtype)

    deltas, weights, costs, indices = [], [], [], []
    converged = torch.zeros(N, dtype=torch.bool)
    num_iters = torch.zeros(N, dtype=dtype)
    match_ratio = torch.zeros(N, dtype=dtype)
    it = 0
    for it in range(max_iterations):
        a = src @ C.transpose(1, 2)                       # rotated source, (N,n,3)
        y = a + r[:, None, :]                             # in the target frame
        idx = nearest_index(y.detach(), tgt[:, :, :3].detach(), nn_mode)
        indices.append(idx)
        nbr = torch.gather(tgt, 1, idx[:, :, None].expand(-1, -1, tgt.shape[2]))
        diff = y - nbr[:, :, :3]                          # (N,n,3)

        if icp_type == "pt2pl":
            nrm = nbr[:, :, 3:6]
            err = (diff * nrm).sum(dim=2)                 # (N,n)
            rho = torch.linalg.norm(diff, dim=2)
            J = torch.cat([torch.cross(nrm, a, dim=2), -nrm], dim=2)     # (N,n,6)
            w0 = w_init
        else:
            err = diff.reshape(N, 3 * n)                  # rows (point, component)
            rho = err.abs()
            ax, ay, az = a[:, :, 0], a[:, :, 1], a[:, :, 2]
            o, one = torch.zeros_like(ax), torch.ones_like(ax)
            J = torch.stack([o, -az, ay, -one, o, o,
                             az, o, -ax, o, -one, o,
                             -ay, ax, o, o, o, -one], dim=2).reshape(N, 3 * n, 6)
            w0 = w_init.repeat_interleave(3, dim=1)
        w = w0
        if trim_dist is not None and trim_dist >= 0.0:
            w = w * trim_weight(rho, trim_dist, differentiable, tanh_steepness)
        if loss_fn is not None:
            w = w * loss_weight(err, loss_fn["name"], loss_fn["metric"], differentiable)

        J = J[:, :, cols]
        ws = torch.sqrt(w + 1.0e-10) - 1.0e-5
        e_w = (ws * err)[:, :, None]
        J_w = ws[:, :, None] * J
        A = J_w.transpose(1, 2) @ J_w + 1e-12 * eye_d
        step = -torch.linalg.inv(A) @ J_w.transpose(1, 2) @ e_w          # (N,d,1)
        step6 = torch.zeros((N, 6, 1), dtype=dtype)
        step6[:, cols] = step

        C = so3_exp(step6[:, 0:3, 0]).transpose(1, 2) @ C
        r = r - step6[:, 3:6, 0]

        deltas.append(step6.detach())
        w_save = w.detach()
        if weights:
            dead = (w_save.sum(dim=1) == 0.0)[:, None]
            w_save = torch.where(dead, weights[-1][:, :, 0], w_save)
        weights.append(w_save[:, :, None])
        cost = (e_w.transpose(1, 2) @ e_w).detach()
        if costs:
            cost = torch.where(cost == 0.0, costs[-1], cost)
        costs.append(cost)

        step_norm = torch.linalg.norm(step6.detach(), dim=1)[:, 0]
        small = step_norm < tolerance
        converged = converged | small
        if bool(small.any()) and not const_iter:
            num_iters = torch.where(small, num_iters + (it + 1) * (num_iters == 0), num_iters)
            matched = (w > match_ratio_thresh).sum(dim=1)
            at_start = (w0 > match_ratio_thresh).sum(dim=1)
            at_start[at_start == 0] = 1
            match_ratio = torch.where(small, match_ratio + matched / at_start * (match_ratio == 0),
                                      match_ratio)
            w_init = w_init * (~small).to(dtype)[:, None]
            w0 = w_init if icp_type == "pt2pl" else w_init.repeat_interleave(3, dim=1)
            if bool(small.all()):
                break

    num_iters = torch.where(num_iters == 0, torch.full_like(num_iters, it + 1), num_iters)
    matched = (w > match_ratio_thresh).sum(dim=1)
    at_start = (w0 > match_ratio_thresh).sum(dim=1)
    at_start[at_start == 0] = 1
    match_ratio = torch.where(match_ratio == 0, (matched / at_start).to(dtype), match_ratio)

    pc = src @ C.transpose(1, 2) + r[:, None, :]
    T = torch.diag_embed(torch.ones((N, 4), dtype=dtype))
    T[:, :3, :3] = C
    T[:, :3, 3] = r
    out = {
        "pc": pc, "T": T,
        "costs": torch.stack(costs, dim=1).squeeze(-1),
        "deltas": torch.stack(deltas, dim=1),
        "weights": torch.stack(weights, dim=1),
        "stats": {"converged": converged, "iterations": num_iters, "matched_ratio": match_ratio},
    }
    if return_indices:
        out["indices"] = torch.stack(indices, dim=0)      # (iters, N, n)
    return out


# --- SE(3) helpers standing in for pylgmath in the reference tests -----------------------
# (`external/dICP/tests/test_ICP.py:11-13`; lgmath convention: xi = [rho; phi], translation first)

def se3_exp(xi) -> torch.Tensor:
    xi = torch.as_tensor(xi, dtype=torch.float64).reshape(6)
    rho, phi = xi[:3], xi[3:]
    th = torch.linalg.norm(phi)
    K = torch.tensor([[0, -phi[2], phi[1]], [phi[2], 0, -phi[0]], [-phi[1], phi[0], 0]],
                     dtype=torch.float64)
    I = torch.eye(3, dtype=torch.float64)
    if th < 1e-12:
        R, V = I + K, I + 0.5 * K
    else:
        R = I + torch.sin(th) / th * K + (1 - torch.cos(th)) / th ** 2 * K @ K
        V = I + (1 - torch.cos(th)) / th ** 2 * K + (th - torch.sin(th)) / th ** 3 * K @ K
    T = torch.eye(4, dtype=torch.float64)
    T[:3, :3] = R
    T[:3, 3] = V @ rho
    return T


def se3_log(T) -> torch.Tensor:
    T = torch.as_tensor(T, dtype=torch.float64)
    R, t = T[:3, :3], T[:3, 3]
    cos = torch.clamp((torch.trace(R) - 1) / 2, -1.0, 1.0)
    th = torch.acos(cos)
    I = torch.eye(3, dtype=torch.float64)
    if th < 1e-10:
        K = 0.5 * (R - R.T)
        Vinv = I - 0.5 * K
    else:
        K = th / (2 * torch.sin(th)) * (R - R.T)
        half = th / 2
        Vinv = I - 0.5 * K + (1 - half / torch.tan(half)) / th ** 2 * K @ K
    phi = torch.stack([K[2, 1], K[0, 2], K[1, 0]])
    return torch.cat([Vinv @ t, phi])
