"""Generates tests/golden/frontend_*.npz: outputs and autograd gradients of the REFERENCE's own front-end code, run
where /root/reference exists (CPU, fp32). utils/pose_utils.py is imported from the reference as is
(get_camera_from_tensor, quadmultiply); the five lines of render() that glue them together
(gaussian_renderer/__init__.py:83-97,:106,:126 - they cannot be imported without the CUDA rasterizer and the scene
classes) are restated below with the same torch calls. Run: python tests/golden/make_frontend_golden.py"""
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
sys.path.insert(0, REF)
from utils.pose_utils import get_camera_from_tensor, quadmultiply  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def make(name, P, V, seed, n_pix=None):
    g = torch.Generator().manual_seed(seed)
    rn = lambda *s: torch.randn(*s, generator=g)
    n_pix = n_pix or 2 * P
    mask = torch.zeros(n_pix, dtype=torch.bool)
    mask[torch.randperm(n_pix, generator=g)[:P]] = True          # aggregated_mask (scene/gaussian_model.py:626)
    m = dict(xyz=rn(P, 3), rotation=torch.tensor([1.0, 0, 0, 0]) + 0.1 * rn(P, 4), opacity=rn(P, 1),
             scaling=-3 + 0.5 * rn(P, 3), features_dc=rn(P, 1, 3), features_rest=0.1 * rn(P, 15, 3),
             conf_static=torch.rand(n_pix, generator=g))
    poses = torch.cat([torch.tensor([1.0, 0, 0, 0]) + 0.2 * rn(V, 4), 0.3 * rn(V, 3)], dim=1)   # un-normalised quaternions
    leaves = {k: v.clone().requires_grad_(True) for k, v in m.items()}
    poses_l = poses.clone().requires_grad_(True)
    outs = dict(means3D=[], rotations=[])
    for v in range(V):
        camera_pose = poses_l[v]
        rel_w2c = get_camera_from_tensor(camera_pose)                                   # render():83
        gaussians_xyz = leaves["xyz"].clone()
        gaussians_rot = leaves["rotation"].clone()
        xyz_ones = torch.ones(gaussians_xyz.shape[0], 1).float()
        xyz_homo = torch.cat((gaussians_xyz, xyz_ones), dim=1)
        outs["means3D"].append((rel_w2c @ xyz_homo.T).T[:, :3])                         # :90
        outs["rotations"].append(quadmultiply(camera_pose[:4], gaussians_rot))          # :91
    opacity = torch.sigmoid(leaves["opacity"])                                          # get_opacity
    outs["opacity"] = opacity * leaves["conf_static"].reshape(-1, 1)[mask]              # :97
    outs["scales"] = torch.exp(leaves["scaling"])                                       # get_scaling
    outs["shs"] = torch.cat((leaves["features_dc"], leaves["features_rest"]), dim=1)    # get_features
    gin = dict(dL_dmeans3D=[rn(P, 3) for _ in range(V)], dL_drotations=[rn(P, 4) for _ in range(V)],
               dL_dopacity=rn(P, 1), dL_dscales=rn(P, 3), dL_dshs=rn(P, 16, 3))
    loss = sum((o * w).sum() for o, w in zip(outs["means3D"], gin["dL_dmeans3D"]))
    loss = loss + sum((o * w).sum() for o, w in zip(outs["rotations"], gin["dL_drotations"]))
    loss = loss + (outs["opacity"] * gin["dL_dopacity"]).sum() + (outs["scales"] * gin["dL_dscales"]).sum()
    loss = loss + (outs["shs"] * gin["dL_dshs"]).sum()
    loss.backward()
    save = {f"in_{k}": v.numpy() for k, v in m.items()}
    save["in_poses"] = poses.numpy()
    save["in_mask"] = mask.numpy()
    for k in ("opacity", "scales", "shs"):
        save[f"out_{k}"] = outs[k].detach().numpy()
    save["out_means3D"] = np.stack([o.detach().numpy() for o in outs["means3D"]])
    save["out_rotations"] = np.stack([o.detach().numpy() for o in outs["rotations"]])
    save["gin_dL_dmeans3D"] = np.stack([t.numpy() for t in gin["dL_dmeans3D"]])
    save["gin_dL_drotations"] = np.stack([t.numpy() for t in gin["dL_drotations"]])
    for k in ("dL_dopacity", "dL_dscales", "dL_dshs"):
        save[f"gin_{k}"] = gin[k].numpy()
    for k, v in leaves.items():
        save[f"grad_{k}"] = v.grad.numpy()
    save["grad_poses"] = poses_l.grad.numpy()
    np.savez_compressed(os.path.join(HERE, f"frontend_{name}.npz"), **save)
    print(name, "P", P, "V", V, "saved")


if __name__ == "__main__":
    make("small", P=700, V=3, seed=0)
    make("one_view", P=257, V=1, seed=1)
    make("many_views", P=300, V=16, seed=2)
