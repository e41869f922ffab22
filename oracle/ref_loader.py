"""Import the UNMODIFIED reference in the build container.  TEST INFRASTRUCTURE ONLY.

`/root/reference` exists only where the goldens are generated (not on the GPU box), so
nothing that runs under `-m gpu`, `smoke()` or `bench.py` may call this.  The only thing
missing for the reference's hot path to import here is `matplotlib`, which
`external/dICP/dICP/ICP.py:8` and `mm_masking/radar_utils.py:3-5` import but never use on the
path; a two-module stub stands in for it.  No reference file is modified or copied.
"""
from __future__ import annotations

import os
import sys
import types

REFERENCE_ROOT = os.environ.get("MMB_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "external", "dICP", "dICP"))


def _stub_matplotlib():
    if "matplotlib" in sys.modules:
        return
    try:
        import matplotlib  # noqa: F401
        return
    except ImportError:
        pass
    mpl = types.ModuleType("matplotlib")
    mpl.use = lambda *a, **k: None
    plt = types.ModuleType("matplotlib.pyplot")
    mpl.pyplot = plt
    sys.modules["matplotlib"] = mpl
    sys.modules["matplotlib.pyplot"] = plt


def load_reference():
    """Returns (ICP class, nn class, loss class, radar_utils module) of the reference."""
    if not reference_available():
        raise RuntimeError("reference tree not found at " + REFERENCE_ROOT)
    _stub_matplotlib()
    for p in (os.path.join(REFERENCE_ROOT, "external", "dICP"),
              os.path.join(REFERENCE_ROOT, "mm_masking")):
        if p not in sys.path:
            sys.path.insert(0, p)
    from dICP.ICP import ICP
    from dICP.nn import nn
    from dICP.loss import loss
    import radar_utils
    return ICP, nn, loss, radar_utils


class exact_cdist:
    """Context manager: force `torch.cdist` to evaluate exact differences and
    record every `torch.argmin` result, without touching reference files."""

    def __init__(self, force_exact: bool = True):
        self.force_exact = force_exact
        self.indices = []

    def __enter__(self):
        import torch
        self._torch = torch
        self._cdist, self._argmin = torch.cdist, torch.argmin
        outer = self

        def cdist(x1, x2, p=2.0, compute_mode="use_mm_for_euclid_dist_if_necessary"):
            if outer.force_exact:
                compute_mode = "donot_use_mm_for_euclid_dist"
            return outer._cdist(x1, x2, p=p, compute_mode=compute_mode)

        def argmin(*a, **k):
            out = outer._argmin(*a, **k)
            outer.indices.append(out.detach().clone())
            return out

        torch.cdist, torch.argmin = cdist, argmin
        return self

    def __exit__(self, *exc):
        self._torch.cdist, self._torch.argmin = self._cdist, self._argmin
        return False


def load_reference_training():
    """(LearnICPWeightPolicy class, eval_training_loss, eval_validation_loss) of the reference.

    `mm_masking/icp_weight_policy.py:8` imports `neptune.types.File` (logging only, absent here): a
    stub module stands in.  `train_icp_weights.py` pulls in the whole dataset / Neptune stack at
    import time, so its two loss FUNCTIONS are compiled on their own from the unmodified source
    file (AST extraction, nothing copied into the repo) with the reference's `radar_utils` names in
    scope."""
    import ast
    _, _, _, ru = load_reference()
    if "neptune" not in sys.modules:
        nep = types.ModuleType("neptune"); nt = types.ModuleType("neptune.types")
        nt.File = object; nep.types = nt
        sys.modules["neptune"] = nep; sys.modules["neptune.types"] = nt
    plt = sys.modules["matplotlib.pyplot"]
    if not hasattr(plt, "figure"):
        plt.figure = lambda *a, **k: None
    import icp_weight_policy as ref_policy
    import torch
    path = os.path.join(REFERENCE_ROOT, "mm_masking", "train_icp_weights.py")
    tree = ast.parse(open(path).read(), filename=path)
    wanted = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in ("eval_training_loss", "eval_validation_loss")]
    ns = {"torch": torch, "extract_bev_from_pts": ru.extract_bev_from_pts,
          "radar_polar_to_cartesian_diff": ru.radar_polar_to_cartesian_diff}
    exec(compile(ast.Module(body=wanted, type_ignores=[]), path, "exec"), ns)
    return ref_policy.LearnICPWeightPolicy, ns["eval_training_loss"], ns["eval_validation_loss"]
