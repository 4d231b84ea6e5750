"""Seeded small scenes shared by the CPU and GPU parity tests and by tests/golden/make_golden.py."""
from __future__ import annotations

import torch

from das3r_b200.synth import synth_scene, view_inputs, grad_seed


def _cov3d_from(scales, rotations, mod=1.0):
    """cov3D_precomp as scene/gaussian_model.py:build_covariance_from_scaling_rotation would hand it over (6 upper
    entries of R S S^T R^T), from normalised quaternions."""
    q = rotations / rotations.norm(dim=1, keepdim=True)
    r, x, y, z = q.unbind(1)
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                     2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                     2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=1).view(-1, 3, 3)
    L = R * (mod * scales)[:, None, :]
    S = L @ L.transpose(1, 2)
    return torch.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], dim=1).contiguous()


def make_case(name: str):
    """-> (view_inputs dict on CPU, dL_dout_color on CPU). Every case is deterministic in `name`."""
    if name == "sh3_small":            # SH degree 3, scales+rotations: DAS3R's default path
        sc = synth_scene(P=800, V=4, W=64, H=48, seed=11, sh_degree=3)
        vi = view_inputs(sc, 1)
    elif name == "sh1_bg_odd":          # degree 1, non-zero background, image not a multiple of 16, scale modifier
        sc = synth_scene(P=700, V=3, W=70, H=50, seed=12, sh_degree=1)
        vi = view_inputs(sc, 2)
        vi["bg"] = torch.tensor([0.2, 0.5, 0.9])
        vi["scale_modifier"] = 1.3
    elif name == "precomp":             # colors_precomp + cov3D_precomp (render_confidence / compute_cov3D_python)
        sc = synth_scene(P=600, V=4, W=64, H=64, seed=13, sh_degree=0)
        vi = view_inputs(sc, 0)
        g = torch.Generator().manual_seed(99)
        vi["colors_precomp"] = torch.rand(600, 3, generator=g)
        vi["cov3D_precomp"] = _cov3d_from(vi["scales"], vi["rotations"])
        vi["shs"] = None
        vi["scales"] = None
        vi["rotations"] = None
    elif name == "sh0_opaque":          # degree 0, high opacity: exercises the T < 1e-4 early termination
        sc = synth_scene(P=900, V=4, W=48, H=48, seed=14, sh_degree=0, opacity_views=1)
        vi = view_inputs(sc, 3)
        vi["scale_modifier"] = 2.0
    elif name == "behind_and_big":      # a third of the points behind the near plane, a few screen-filling Gaussians
        sc = synth_scene(P=500, V=2, W=80, H=48, seed=15, sh_degree=2)
        vi = view_inputs(sc, 0)
        m = vi["means3D"].clone()
        m[::3, 2] = -m[::3, 2]
        m[1, 2] = 0.001       # exactly on the patched near plane: culled (<=)
        m[4, 2] = 0.0011      # just in front of it
        vi["means3D"] = m
        s = vi["scales"].clone()
        s[5::50] = 0.4
        vi["scales"] = s
    else:
        raise KeyError(name)
    vi["means3D"] = vi["means3D"].contiguous()
    W, H = vi["image_width"], vi["image_height"]
    g = torch.Generator().manual_seed(1234)
    dL = torch.randn(3, H, W, generator=g) / (W * H)
    return vi, dL


CASE_NAMES = ["sh3_small", "sh1_bg_odd", "precomp", "sh0_opaque", "behind_and_big"]
