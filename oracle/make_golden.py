"""Generate tests/golden/*.npz by running the UNMODIFIED reference.  TEST INFRASTRUCTURE ONLY.

Run in the build container (where /root/reference is mounted):

    python -m oracle.make_golden            # writes tests/golden/

Each .npz holds the inputs and the reference's outputs for one case, so the GPU box (which
has no /root/reference) can check both the oracle restatement and the CUDA path against the
reference itself.  The 65-point clouds come from the reference's own test fixtures
(`external/dICP/tests/data/points_{scan,map}.npy`, used by `tests/test_ICP.py:21-33`).
"""
from __future__ import annotations

import os
import sys
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle.ref_loader import REFERENCE_ROOT, exact_cdist, load_reference  # noqa: E402
from oracle.dicp_oracle import se3_exp  # noqa: E402
from mm_masking_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def _config_path(const_iter: bool) -> str:
    """A dICP config identical to the reference's except for const_iter."""
    import yaml
    with open(os.path.join(REFERENCE_ROOT, "external", "dICP", "config", "dICP_config.yaml")) as f:
        cfg = yaml.safe_load(f)
    cfg["dICP"]["parameters"]["const_iter"] = bool(const_iter)
    fd, path = tempfile.mkstemp(suffix=".yaml")
    with os.fdopen(fd, "w") as f:
        yaml.safe_dump(cfg, f)
    return path


def run_reference(ICP, arrays, *, icp_type, differentiable, max_iterations, tolerance, trim_dist,
                  loss_fn, dim, const_iter, dtype, exact=True, grads=True, seed=0):
    """One reference call on dense inputs; returns a flat dict of numpy arrays."""
    tt = lambda a: torch.tensor(np.asarray(a), dtype=dtype)
    source = tt(arrays["source"]).requires_grad_(grads)
    target = tt(arrays["target"] if icp_type == "pt2pl" else arrays["target"][..., :3]).requires_grad_(grads)
    T_init = tt(arrays["T_init"]).requires_grad_(grads)
    weight = None
    if arrays.get("weight") is not None:
        weight = tt(arrays["weight"]).requires_grad_(grads)
    alg = ICP(config_path=_config_path(const_iter), icp_type=icp_type, differentiable=differentiable,
              max_iterations=max_iterations, tolerance=tolerance)
    with exact_cdist(exact) as rec:
        res = alg.icp(source, target, T_init, weight=weight, trim_dist=trim_dist, loss_fn=loss_fn, dim=dim)
    out = {
        "T": res["T"].detach().numpy(), "pc": res["pc"].detach().numpy(),
        "costs": res["costs"].detach().numpy(), "deltas": res["deltas"].detach().numpy(),
        "weights": res["weights"].detach().numpy(),
        "converged": res["stats"]["converged"].numpy(),
        "iterations": res["stats"]["iterations"].numpy(),
        "matched_ratio": res["stats"]["matched_ratio"].numpy(),
        "indices": torch.stack(rec.indices, 0).numpy().astype(np.int32),
    }
    if grads:
        g = torch.Generator().manual_seed(1234 + seed)
        G_T = torch.randn(res["T"].shape, generator=g, dtype=torch.float64).to(dtype)
        G_pc = 0.01 * torch.randn(res["pc"].shape, generator=g, dtype=torch.float64).to(dtype)
        G_T[:, 3, :] = 0
        loss = (res["T"] * G_T).sum() + (res["pc"] * G_pc).sum()
        loss.backward()
        out["G_T"], out["G_pc"] = G_T.numpy(), G_pc.numpy()
        out["grad_source"] = source.grad.numpy()
        out["grad_target"] = target.grad.numpy()
        out["grad_T_init"] = T_init.grad.numpy()
        if weight is not None:
            out["grad_weight"] = weight.grad.numpy()
    return out


def save(name, inputs, outputs, meta):
    blob = {}
    for k, v in inputs.items():
        if v is not None:
            blob["in_" + k] = v
    for k, v in outputs.items():
        blob["ref_" + k] = v
    blob["meta"] = np.array(repr(meta))
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **blob)
    print("wrote", name, {k: getattr(v, "shape", None) for k, v in outputs.items() if k in ("T", "indices")})


def fixture_clouds():
    base = os.path.join(REFERENCE_ROOT, "external", "dICP", "tests", "data")
    return np.load(os.path.join(base, "points_scan.npy")), np.load(os.path.join(base, "points_map.npy"))


def gen_fixture_cases(ICP):
    scan, mp = fixture_clouds()
    T_true = torch.linalg.inv(se3_exp([1.0, 1.0, 0, 0, 0, 0.1])).numpy()
    arrays = dict(source=scan[None, :, :3], target=mp[None], T_init=np.eye(4)[None], weight=None)
    cases = [
        ("fixture_pt2pt_diff", dict(icp_type="pt2pt", differentiable=True, loss_fn={"name": "huber", "metric": 1.0})),
        ("fixture_pt2pl_diff", dict(icp_type="pt2pl", differentiable=True, loss_fn={"name": "huber", "metric": 10.0})),
        ("fixture_pt2pt_nondiff", dict(icp_type="pt2pt", differentiable=False, loss_fn={"name": "huber", "metric": 10.0})),
        ("fixture_pt2pl_cauchy_nondiff", dict(icp_type="pt2pl", differentiable=False, loss_fn={"name": "cauchy", "metric": 0.5})),
    ]
    for name, kw in cases:
        meta = dict(max_iterations=100, tolerance=1e-10, trim_dist=5.0, dim=2, const_iter=False, **kw)
        out = run_reference(ICP, arrays, dtype=torch.float64, **meta)
        out["T_true"] = T_true
        save(name, arrays, out, meta)


def gen_ragged_cases(ICP):
    """The list-input cases of `external/dICP/tests/test_ICP_inputs.py:36-211`, with the random
    rows drawn from a fixed seed so they can be stored."""
    scan, mp = fixture_clouds()
    rng = np.random.default_rng(7)
    rand_pt = rng.random((1, 3)) * 1000
    src = [np.vstack([scan[:50, :3], rand_pt]), scan[:, :3], scan[:55, :3]]
    tgt = [mp[:55], mp, mp[:60]]
    alg = ICP(icp_type="pt2pl", differentiable=True, max_iterations=25, tolerance=1e-8)
    lf = {"name": "huber", "metric": 1.0}
    tt = lambda a: torch.tensor(a, dtype=torch.float64)
    res = alg.icp([tt(s) for s in src], [tt(t) for t in tgt], torch.eye(4, dtype=torch.float64).repeat(3, 1, 1),
                  trim_dist=5.0, loss_fn=lf, dim=2)
    loop_T, loop_mr, loop_it = [], [], []
    for s, t in zip(src, tgt):
        r1 = alg.icp(tt(s), tt(t), torch.eye(4, dtype=torch.float64), trim_dist=5.0, loss_fn=lf, dim=2)
        loop_T.append(r1["T"][0].numpy()); loop_mr.append(r1["stats"]["matched_ratio"][0].item())
        loop_it.append(r1["stats"]["iterations"][0].item())
    blob = {f"in_source_{i}": s for i, s in enumerate(src)}
    blob.update({f"in_target_{i}": t for i, t in enumerate(tgt)})
    blob.update(ref_T=res["T"].detach().numpy(), ref_matched_ratio=res["stats"]["matched_ratio"].numpy(),
                ref_iterations=res["stats"]["iterations"].numpy(), ref_converged=res["stats"]["converged"].numpy(),
                ref_deltas=res["deltas"].numpy(), ref_weights=res["weights"].numpy(), ref_costs=res["costs"].numpy(),
                ref_pc=res["pc"].detach().numpy(),
                ref_loop_T=np.stack(loop_T), ref_loop_matched_ratio=np.array(loop_mr), ref_loop_iterations=np.array(loop_it))
    np.savez_compressed(os.path.join(OUT, "ragged_input_types.npz"), **blob)
    print("wrote ragged_input_types")

    # weights given as None / ones / ones+zeros for 10 extra rows (`test_ICP_inputs.py:157-211`)
    extra = rng.random((10, 3))
    src = [scan[:, :3], scan[:, :3], np.vstack([scan[:, :3], extra])]
    wts = [None, np.ones(len(scan)), np.hstack([np.ones(len(scan)), np.zeros(10)])]
    res = alg.icp([tt(s) for s in src], [tt(mp)] * 3, torch.eye(4, dtype=torch.float64).repeat(3, 1, 1),
                  weight=[None if w is None else tt(w) for w in wts], trim_dist=5.0, loss_fn=lf, dim=2)
    blob = {f"in_source_{i}": s for i, s in enumerate(src)}
    blob.update(in_target=mp, in_weight_1=wts[1], in_weight_2=wts[2], ref_T=res["T"].detach().numpy(),
                ref_iterations=res["stats"]["iterations"].numpy(), ref_matched_ratio=res["stats"]["matched_ratio"].numpy())
    np.savez_compressed(os.path.join(OUT, "ragged_weight_inputs.npz"), **blob)
    print("wrote ragged_weight_inputs")


def gen_synthetic_cases(ICP):
    """Seeded synthetic pairs: reference run in fp64 on the fp32-valued inputs (the truth for
    poses and gradients), plus the reference's own fp32 pose for comparison."""
    variants = [
        # name, synth kwargs, icp kwargs
        ("synth_pt2pl_dim2_cauchy", dict(N=3, n=200, m=3000, seed=11, pos=0.5, rot=0.1),
         dict(icp_type="pt2pl", differentiable=True, max_iterations=10, tolerance=1e-5, trim_dist=5.0,
              loss_fn={"name": "cauchy", "metric": 1.0}, dim=2, const_iter=False)),
        ("synth_pt2pl_dim2_cauchy_hard", dict(N=4, n=256, m=4000, seed=12, pos=2.0, rot=0.6),
         dict(icp_type="pt2pl", differentiable=True, max_iterations=10, tolerance=1e-5, trim_dist=5.0,
              loss_fn={"name": "cauchy", "metric": 1.0}, dim=2, const_iter=True)),
        ("synth_pt2pl_dim2_nondiff", dict(N=3, n=200, m=3000, seed=13, pos=0.5, rot=0.1),
         dict(icp_type="pt2pl", differentiable=False, max_iterations=50, tolerance=1e-5, trim_dist=5.0,
              loss_fn={"name": "cauchy", "metric": 1.0}, dim=2, const_iter=False)),
        ("synth_pt2pt_dim2_cauchy", dict(N=2, n=150, m=2000, seed=14, pos=0.5, rot=0.1),
         dict(icp_type="pt2pt", differentiable=True, max_iterations=8, tolerance=1e-5, trim_dist=5.0,
              loss_fn={"name": "cauchy", "metric": 1.0}, dim=2, const_iter=True)),
        ("synth_pt2pl_dim3_huber", dict(N=2, n=200, m=3000, seed=15, pos=0.5, rot=0.1, three_d=True),
         dict(icp_type="pt2pl", differentiable=True, max_iterations=8, tolerance=1e-6, trim_dist=3.0,
              loss_fn={"name": "huber", "metric": 0.5}, dim=3, const_iter=True)),
        ("synth_pt2pt_dim3_plain", dict(N=2, n=150, m=2000, seed=16, pos=0.3, rot=0.05, three_d=True),
         dict(icp_type="pt2pt", differentiable=True, max_iterations=6, tolerance=1e-6, trim_dist=None,
              loss_fn=None, dim=3, const_iter=True)),
        ("synth_pt2pl_dim3_hardhuber", dict(N=2, n=150, m=2000, seed=17, pos=0.3, rot=0.05, three_d=True),
         dict(icp_type="pt2pl", differentiable=False, max_iterations=6, tolerance=1e-6, trim_dist=2.0,
              loss_fn={"name": "huber", "metric": 0.1}, dim=3, const_iter=True)),
    ]
    for name, skw, ikw in variants:
        batch = synth.make_icp_batch(**skw)
        arrays = dict(source=batch["source"], target=batch["target"], T_init=batch["T_init"], weight=batch["weight"])
        out = run_reference(ICP, arrays, dtype=torch.float64, exact=True, grads=True, **ikw)
        ref32 = run_reference(ICP, arrays, dtype=torch.float32, exact=False, grads=False, **ikw)
        out["T_fp32_default_cdist"] = ref32["T"]
        out["indices_fp32_default_cdist"] = ref32["indices"]
        save(name, arrays, out, dict(synth=skw, **ikw))


def main():
    os.makedirs(OUT, exist_ok=True)
    ICP, _nn, _loss, _ru = load_reference()
    torch.manual_seed(0)
    gen_fixture_cases(ICP)
    gen_ragged_cases(ICP)
    gen_synthetic_cases(ICP)


if __name__ == "__main__":
    main()
