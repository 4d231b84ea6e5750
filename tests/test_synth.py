import numpy as np
import torch

from das3r_b200.synth import knn_dist2_cpu, synth_scene, view_inputs


def test_scene_is_deterministic_and_shaped():
    a = synth_scene(P=500, V=4, W=64, H=48, seed=3)
    b = synth_scene(P=500, V=4, W=64, H=48, seed=3)
    c = synth_scene(P=500, V=4, W=64, H=48, seed=4)
    assert torch.equal(a.xyz, b.xyz) and torch.equal(a.shs, b.shs) and torch.equal(a.scales, b.scales)
    assert not torch.equal(a.xyz, c.xyz)
    assert a.xyz.shape == (500, 3) and a.shs.shape == (500, 16, 3) and a.rot.shape == (500, 4)
    vi = view_inputs(a, 1)
    assert vi["means3D"].shape == (500, 3) and not vi["means3D"].is_contiguous()  # transposed-slice view like DAS3R
    assert torch.equal(vi["viewmatrix"], torch.eye(4)) and float(vi["campos"].abs().sum()) == 0.0
    assert float(vi["means3D"][:, 2].min()) > 0.5  # in front of the camera
    assert abs(float(vi["projmatrix"][2, 3]) - 1.0) < 1e-6  # transposed projection: w = z


def test_knn_cpu_matches_bruteforce():
    g = torch.Generator().manual_seed(0)
    pts = torch.rand(300, 3, generator=g)
    pts[10] = pts[11]  # exact duplicate: distance 0 counts (simple_knn.cu:159,:178 skip only the index itself)
    d = knn_dist2_cpu(pts).numpy()
    p = pts.numpy().astype(np.float32)
    for i in (0, 10, 11, 150, 299):
        diff = p - p[i]
        d2 = (diff[:, 0] * diff[:, 0] + diff[:, 1] * diff[:, 1] + diff[:, 2] * diff[:, 2]).astype(np.float32)
        d2[i] = np.inf
        best = np.sort(d2)[:3]
        assert np.isclose(d[i], (best[0] + best[1] + best[2]) / np.float32(3), rtol=1e-6)
