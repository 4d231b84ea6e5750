"""The CUDA path against the committed golden vectors (outputs of the reference CUDA rasterizer, see
tests/golden/make_golden.py). Unlike test_gpu_parity.py this needs neither oracle/_ref nor the oracle at run time."""
import glob
import os

import numpy as np
import pytest
import torch

import _harness as Hn

pytestmark = pytest.mark.gpu

GOLDEN_FILES = sorted(p for p in glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "*.npz"))
                      if not os.path.basename(p).startswith(("frontend_", "loss_", "init_")))  # those have their own test files


@pytest.mark.skipif(not GOLDEN_FILES, reason="golden vectors not generated yet")
@pytest.mark.parametrize("path", GOLDEN_FILES, ids=[os.path.basename(p)[:-4] for p in GOLDEN_FILES])
def test_cuda_path_against_golden(path):
    dev = torch.device("cuda:0")
    z = np.load(path)
    vi = {}
    for k in z.files:
        if k.startswith("in_") and k != "in_dL":
            v = z[k]
            vi[k[3:]] = torch.from_numpy(v).to(dev) if v.ndim > 0 else v.item()
    for k in ("shs", "colors_precomp", "scales", "rotations", "cov3D_precomp"):
        vi.setdefault(k, None)
    dL = torch.from_numpy(z["in_dL"]).to(dev)
    ours = Hn.run_ours(vi, dL, debug=True, extras=False)
    assert ours["num_rendered"] == int(z["out_num_rendered"])
    assert np.array_equal(Hn.np_(ours["radii"]), z["out_radii"])
    assert np.array_equal(Hn.np_(ours["tiles_touched"]), z["out_tiles_touched"])
    assert np.array_equal(Hn.np_(ours["point_list"]), z["out_point_list"])
    assert np.array_equal(Hn.np_(ours["keys_sorted"]), z["out_keys_sorted"])
    assert np.array_equal(Hn.np_(ours["ranges"]), z["out_ranges"])
    assert np.array_equal(Hn.np_(ours["n_contrib"]), z["out_n_contrib"])
    assert float(np.abs(Hn.np_(ours["color"]) - z["out_color"]).max()) <= Hn.FWD_TOL
    assert float(np.abs(Hn.np_(ours["final_T"]) - z["out_final_T"]).max()) <= Hn.FWD_TOL
    gmap = {"dL_dmeans3D": "dL_dmeans3D", "dL_dmeans2D": "dL_dmeans2D", "dL_dopacity": "dL_dopacity", "dL_dsh": "dL_dsh",
            "dL_dcolors": "dL_dcolors", "dL_dscales": "dL_dscales", "dL_drotations": "dL_drotations",
            "dL_dcov3D": "dL_dcov3D"}
    checked = 0
    for k, gk in gmap.items():
        g = ours["grads"].get(k)
        ref = z["grad_" + gk]
        if g is None or ref.size == 0:
            continue
        assert Hn.rel_err(Hn.np_(g).reshape(ref.shape), ref) <= Hn.GRAD_RTOL, k
        checked += 1
    assert checked >= 3
    from simple_knn._C import distCUDA2

    np.testing.assert_allclose(distCUDA2(vi["means3D"].contiguous()).cpu().numpy(), z["out_knn"], rtol=1e-6)
