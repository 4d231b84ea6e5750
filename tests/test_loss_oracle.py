"""oracle/loss_oracle.py (numpy restatement of DAS3R's photometric loss) against golden vectors produced by the
reference's own utils/loss_utils.py / utils/image_utils.py (tests/golden/make_loss_golden.py)."""
import glob
import os

import numpy as np
import pytest

import _harness as Hn
from oracle import loss_oracle as lo

GOLDEN = sorted(glob.glob(os.path.join(Hn.ROOT, "tests", "golden", "loss_*.npz")))


def test_golden_files_exist():
    assert len(GOLDEN) >= 4


def test_window_is_the_reference_window():
    g = lo.gaussian_window(np.float32)
    assert g.shape == (11,) and abs(float(g.sum()) - 1.0) < 1e-6 and g[5] == g.max() and np.allclose(g, g[::-1])


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_forward_and_backward_match_the_reference(path):
    z = np.load(path)
    static = z["static"] if "static" in z.files else None
    lam, gs = float(z["lambda_dssim"]), float(z["grad_scale"])
    fw = lo.forward(z["image"], z["gt"], static, lam)
    assert abs(fw["loss"] - float(z["loss"])) <= 2e-6 * max(1.0, abs(float(z["loss"])))
    assert abs(fw["psnr"] - float(z["psnr"])) <= 1e-4
    np.testing.assert_allclose(fw["ssim_map"], z["ssim_map"], rtol=0, atol=5e-4)  # the reference forms E[x^2] - mu^2 in fp32
    bw = lo.backward(z["image"], z["gt"], static, lam, grad_loss=gs)
    assert Hn.rel_err(bw["dL_dimage"], z["grad_image"]) <= 2e-4
    assert Hn.rel_err(bw["dL_dgt"], z["grad_gt"]) <= 2e-4
    if static is not None:
        assert Hn.rel_err(bw["dL_dstatic"], z["grad_static"]) <= 2e-4
