"""SURVEY.md 8f row N4 (the reachable part): `examples/localization_icp.cpp`, a ROS-free C++ caller of
include/mm_masking_b200.h that runs the localisation-style ICP of
`vtr_radar_lidar/src/modules/localization/localization_icp_module.cpp:276-420` (1-NN, distance gate, Cauchy
point-to-point Gauss-Newton, per-point weights) through the C ABI.  CPU: it is built and links the in-tree
library.  GPU: its pose equals the Python path's on the same pair."""
import os
import subprocess

import numpy as np
import pytest
import torch

from conftest import ROOT

BIN = os.path.join(ROOT, "examples", "localization_icp")


def _build():
    import __graft_entry__ as g
    g.build()


def test_example_is_built_and_links_the_library():
    _build()
    assert os.access(BIN, os.X_OK)
    out = subprocess.run(["ldd", BIN], capture_output=True, text=True).stdout
    assert "libmm_masking_b200.so" in out and "not found" not in out.split("libmm_masking_b200.so")[1].split("\n")[0]


@pytest.mark.gpu
def test_cpp_caller_matches_the_python_path(tmp_path):
    from mm_masking_b200 import synth
    from mm_masking_b200.dICP import ICP
    _build()
    b = synth.make_icp_batch(N=1, n=1500, m=20000, seed=12, pos=0.8, rot=0.15, three_d=True)
    src, tgt, w, T0 = b["source"][0], b["target"][0, :, :3].copy(), b["weight"][0], b["T_init"][0]
    path = tmp_path / "pair.bin"
    with open(path, "wb") as f:
        f.write(np.asarray([src.shape[0], tgt.shape[0]], np.int32).tobytes())
        for a in (src, tgt, w, T0):
            f.write(np.ascontiguousarray(a, np.float32).tobytes())
    out = subprocess.run([BIN, str(path), "30", "5.0", "1.0"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    lines = out.stdout.strip().split("\n")
    T_cpp = np.array([[float(x) for x in ln.split()] for ln in lines[1:5]])
    its = int(lines[5].split()[1])
    alg = ICP(icp_type="pt2pt", differentiable=False, max_iterations=30, tolerance=1e-6)
    tt = lambda a: torch.tensor(a, device="cuda")
    ref = alg.icp(tt(src)[None], tt(tgt)[None], tt(T0)[None], weight=tt(w)[None], trim_dist=5.0,
                  loss_fn={"name": "cauchy", "metric": 1.0}, dim=3)
    np.testing.assert_allclose(T_cpp, ref["T"][0].cpu().numpy(), atol=1e-6)
    assert its == int(ref["stats"]["iterations"][0])
    assert np.abs(T_cpp - T0).max() > 1e-3            # it moved
