"""oracle/frontend_oracle.py (numpy restatement of DAS3R's render() front-end) against the golden vectors produced by
the reference's own functions (tests/golden/make_frontend_golden.py): forward within fp32 rounding, the analytic
backward against the reference's autograd gradients."""
import glob
import os

import numpy as np
import pytest

import _harness as Hn
from oracle import frontend_oracle as fo

GOLDEN = sorted(glob.glob(os.path.join(Hn.ROOT, "tests", "golden", "frontend_*.npz")))


def load(path):
    z = np.load(path)
    m = {k[3:]: z[k] for k in z.files if k.startswith("in_") and k not in ("in_poses", "in_mask")}
    m["conf_index"] = np.nonzero(z["in_mask"])[0]
    return z, m, z["in_poses"]


def test_golden_files_exist():
    assert len(GOLDEN) >= 3


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_forward_matches_the_reference(path):
    z, m, poses = load(path)
    out = fo.forward(m, poses)
    for v in range(len(poses)):
        np.testing.assert_allclose(out["means3D"][v], z["out_means3D"][v], rtol=2e-6, atol=2e-6)
        np.testing.assert_allclose(out["rotations"][v], z["out_rotations"][v], rtol=2e-6, atol=2e-6)
    for k in ("opacity", "scales", "shs"):
        np.testing.assert_allclose(out[k], z[f"out_{k}"], rtol=2e-6, atol=1e-7)
    assert np.array_equal(out["shs"], z["out_shs"])  # a pure copy


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_backward_matches_the_reference_autograd(path):
    z, m, poses = load(path)
    g = dict(dL_dmeans3D=list(z["gin_dL_dmeans3D"]), dL_drotations=list(z["gin_dL_drotations"]),
             dL_dopacity=z["gin_dL_dopacity"], dL_dscales=z["gin_dL_dscales"], dL_dshs=z["gin_dL_dshs"])
    got = fo.backward(m, poses, g)
    pairs = [("dL_dxyz", "grad_xyz"), ("dL_drotation", "grad_rotation"), ("dL_dopacity", "grad_opacity"),
             ("dL_dscaling", "grad_scaling"), ("dL_dfeatures_dc", "grad_features_dc"),
             ("dL_dfeatures_rest", "grad_features_rest"), ("dL_dposes", "grad_poses")]
    for ours, ref in pairs:
        assert Hn.rel_err(got[ours], z[ref]) <= 2e-5, ours  # the reference side is fp32 autograd
    conf_grad = np.zeros_like(z["grad_conf_static"])
    conf_grad[m["conf_index"]] = got["dL_dconf"]
    assert Hn.rel_err(conf_grad, z["grad_conf_static"]) <= 2e-5
