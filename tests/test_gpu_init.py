"""The pointmap initialisation on the GPU (include/das3r_init.h, das3r_b200.init.create_from_pointmaps) against the
golden vectors of the reference's own functions, against the numpy oracle on a larger seeded case, and - at C4's 6 M
points - against torch's formulation of the same lines on the GPU with the reference's distCUDA2 (oracle/_ref)."""
import glob
import os

import numpy as np
import pytest
import torch

import _harness as Hn
import oracle
from oracle import init_oracle as io
from test_init_oracle import check_against_golden

pytestmark = pytest.mark.gpu
GOLDEN = sorted(glob.glob(os.path.join(Hn.ROOT, "tests", "golden", "init_*.npz")))
DEV = "cuda:0"


def _run(m, conf_thre):
    from das3r_b200.init import create_from_pointmaps

    t = lambda a: torch.as_tensor(np.ascontiguousarray(a), device=DEV)
    r = create_from_pointmaps(t(m["K"]), t(m["poses"]), t(m["depth"]), t(m["confidence"]), t(m["dyna_avg"]), t(m["rgb"]),
                              conf_thre)
    torch.cuda.synchronize()
    return r


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_against_reference_golden(path):
    z = np.load(path)
    m = {k[3:]: z[k] for k in z.files if k.startswith("in_") and k != "in_conf_thre"}
    r = _run(m, float(z["in_conf_thre"]))
    got = {k: Hn.np_(v) for k, v in r._asdict().items()}
    check_against_golden(got, z)


def test_against_the_oracle_on_a_larger_case():
    m = io.synth_pointmaps(V=6, H=96, W=160, seed=9)
    r = _run(m, 1.0)
    want = io.create_from_pointmaps(**m, conf_thre=1.0)
    got = {k: Hn.np_(v) for k, v in r._asdict().items()}
    assert np.array_equal(got["aggregated_mask"], want["aggregated_mask"])
    assert np.array_equal(got["conf_index"], want["conf_index"])
    np.testing.assert_allclose(got["xyz"], want["xyz"], rtol=2e-6, atol=2e-6)
    np.testing.assert_allclose(got["features_dc"], want["features_dc"], rtol=2e-6, atol=1e-6)
    np.testing.assert_allclose(got["dist2"], want["dist2"], rtol=1e-4, atol=1e-12)
    np.testing.assert_allclose(got["scaling"], want["scaling"], rtol=1e-5, atol=1e-5)
    assert np.array_equal(got["conf_static"], want["conf_static"])
    assert float(np.abs(got["features_rest"]).max()) == 0.0 and np.array_equal(got["rotation"], want["rotation"])
    np.testing.assert_allclose(got["opacity"], want["opacity"], rtol=1e-6)


def test_empty_selection_and_no_cpu_path():
    from das3r_b200.init import create_from_pointmaps

    m = io.synth_pointmaps(V=2, H=16, W=16, seed=1)
    r = _run(m, 1e6)  # nothing is that confident
    assert r.xyz.shape == (0, 3) and r.conf_index.shape == (0,) and not bool(r.aggregated_mask.any())
    assert r.conf_static.shape == (2, 16, 16)
    with pytest.raises(RuntimeError):
        t = lambda a: torch.as_tensor(a)
        create_from_pointmaps(t(m["K"]), t(m["poses"]), t(m["depth"]), t(m["confidence"]), t(m["dyna_avg"]), t(m["rgb"]))


def test_c4_sized_initialisation_against_torch_and_reference_knn():
    """50 views of 512x288 with ~81 % confident pixels = 6 M Gaussians (config C4's count): every output against the
    reference's formulation in torch on the GPU (same lines as create_from_cameras) with the reference's distCUDA2."""
    V, H, W = 50, 288, 512
    m = io.synth_pointmaps(V=V, H=H, W=W, seed=4)
    r = _run(m, 1.0)
    t = {k: torch.as_tensor(v, device=DEV) for k, v in m.items()}
    P = r.xyz.shape[0]
    assert 5_500_000 < P < 6_500_000
    jj, ii = torch.meshgrid(torch.arange(H, device=DEV), torch.arange(W, device=DEV), indexing="ij")
    grid = torch.stack([ii, jj], dim=-1)[None].expand(V, H, W, 2)                    # xy_grid
    pp = t["K"][:, :2, 2].unsqueeze(1)
    focal = t["K"][:, 0, 0][:, None, None]
    d = t["depth"].reshape(V, -1, 1)
    rel = torch.cat((d * (grid.reshape(V, -1, 2) - pp) / focal, d), dim=-1).reshape(V, H, W, 3)   # pose_utils.py:683
    p3d = torch.einsum("bij, bhwj -> bhwi", t["poses"][:, :3, :3], rel) + t["poses"][:, None, None, :3, 3]
    mask = t["confidence"].reshape(-1) > torch.tensor(1.0).log()
    assert torch.equal(r.aggregated_mask, mask)
    xyz = p3d.reshape(-1, 3)[mask]
    assert float((r.xyz - xyz).abs().max()) <= 4e-6
    col = t["rgb"].permute(0, 2, 3, 1).reshape(-1, 3)[mask]
    assert float((r.features_dc[:, 0] - (col - 0.5) / 0.28209479177387814).abs().max()) <= 2e-6
    assert torch.equal(r.conf_index, mask.nonzero().flatten())
    assert torch.equal(r.conf_static, 1 - t["dyna_avg"])
    if oracle.reference_available():
        _, K = oracle.load_reference()
        ref_d2 = K.distCUDA2(r.xyz.contiguous())
        assert torch.allclose(r.dist2, ref_d2, rtol=1e-6, atol=0)
        want = torch.log(torch.sqrt(torch.clamp_min(ref_d2, 0.0000001)))[..., None].repeat(1, 3)
        assert float((r.scaling - want).abs().max()) <= 2e-6
