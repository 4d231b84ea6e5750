"""Synthetic DAS3R-shaped scenes: no datasets or checkpoints exist on the box, so tests and bench use
`synth_scene(P, V, W, H, seed)`, which mimics GaussianModel.create_from_cameras (reference scene/gaussian_model.py:
573-659): one small isotropic Gaussian per sampled pointmap pixel of each of V views, scale = 3-NN distance, opacity
= 1/V times a static-confidence factor, and `view_inputs(scene, v)`, which builds the rasterizer inputs exactly like
the render glue does (reference gaussian_renderer/__init__.py:57-108: identity view matrix, Gaussians moved into the
camera frame, quaternion product left un-normalised).

Everything is generated on the CPU with a seeded torch.Generator so CPU oracle, reference and CUDA path see
bit-identical inputs; nothing here touches the native extension.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np
import torch

SH_C0 = 0.28209479177387814


@dataclass
class Scene:
    P: int
    V: int
    W: int
    H: int
    sh_degree: int
    focal: float
    xyz: torch.Tensor        # [P,3] world
    rot: torch.Tensor        # [P,4] (r,x,y,z), near identity, NOT normalised
    scales: torch.Tensor     # [P,3] activated
    opacity: torch.Tensor    # [P,1] activated, already multiplied by static confidence
    shs: torch.Tensor        # [P,16,3]
    poses: torch.Tensor      # [V,7] w2c as quaternion (r,x,y,z) + translation
    seed: int
    extras: dict = field(default_factory=dict)

    @property
    def tanfovx(self) -> float:
        return self.W / (2.0 * self.focal)

    @property
    def tanfovy(self) -> float:
        return self.H / (2.0 * self.focal)


def projection_matrix(znear: float, zfar: float, fovx: float, fovy: float) -> torch.Tensor:
    """reference utils/graphics_utils.py:80-100 (getProjectionMatrix), returned un-transposed."""
    tan_y = math.tan(fovy / 2)
    tan_x = math.tan(fovx / 2)
    top = tan_y * znear
    bottom = -top
    right = tan_x * znear
    left = -right
    Pm = torch.zeros(4, 4)
    Pm[0, 0] = 2.0 * znear / (right - left)
    Pm[1, 1] = 2.0 * znear / (top - bottom)
    Pm[0, 2] = (right + left) / (right - left)
    Pm[1, 2] = (top + bottom) / (top - bottom)
    Pm[3, 2] = 1.0
    Pm[2, 2] = zfar / (zfar - znear)
    Pm[2, 3] = -(zfar * znear) / (zfar - znear)
    return Pm


def quat_to_rot(q: torch.Tensor) -> torch.Tensor:
    """reference utils/pose_utils.py:34-55 (quad2rotation, normalises), single quaternion -> [3,3]."""
    q = q / q.norm()
    r, x, y, z = [float(v) for v in q]
    return torch.tensor([
        [1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y)],
        [2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x)],
        [2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)],
    ], dtype=torch.float32)


def quat_mul(q1: torch.Tensor, q2: torch.Tensor) -> torch.Tensor:
    """reference utils/pose_utils.py:86-104 (quadmultiply)."""
    w1, x1, y1, z1 = q1.unbind(dim=-1)
    w2, x2, y2, z2 = q2.unbind(dim=-1)
    return torch.stack([
        w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2,
        w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
        w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
        w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2,
    ], dim=-1)


def _camera_arc(V: int, n_poses: int | None = None) -> torch.Tensor:
    """Cameras on a smooth arc: yaw within +-15 degrees, centre offset <= 0.3 scene units. Returns w2c [n,7]."""
    n = n_poses or V
    out = torch.zeros(n, 7)
    for i in range(n):
        s = 0.0 if n == 1 else (2.0 * i / (n - 1) - 1.0)
        yaw = math.radians(-15.0 * s)
        c = torch.tensor([0.3 * s, 0.05 * math.sin(math.pi * s), 0.02 * math.cos(math.pi * s) - 0.02])
        # c2w rotation = R_y(yaw); w2c quaternion = conjugate
        half = -yaw / 2.0
        q = torch.tensor([math.cos(half), 0.0, math.sin(half), 0.0])
        R = quat_to_rot(q)
        t = -(R @ c)
        out[i, :4] = q
        out[i, 4:] = t
    return out


def knn_dist2_cpu(xyz: torch.Tensor) -> torch.Tensor:
    """Exact mean squared distance to the 3 nearest other points (semantics of simple_knn.cu:148-184), on the CPU
    with a k-d tree; squared distances and the mean are formed in float32 with the reference's expression order."""
    from scipy.spatial import cKDTree

    pts = xyz.detach().cpu().numpy().astype(np.float32)
    n = pts.shape[0]
    k = min(4, n)
    tree = cKDTree(pts.astype(np.float64))
    _, idx = tree.query(pts.astype(np.float64), k=k, workers=-1)
    idx = idx.reshape(n, k)
    # drop self (first column is normally self; with exact duplicates any of the zero-distance hits may be self)
    rows = np.arange(n)
    d2 = np.full((n, 3), np.finfo(np.float32).max, dtype=np.float32)
    if k == 4:
        self_pos = np.argmax(idx == rows[:, None], axis=1)
        has_self = (idx == rows[:, None]).any(axis=1)
        self_pos = np.where(has_self, self_pos, 3)  # duplicates pushed self out of the top 4: drop the farthest
        keep = np.ones((n, 4), dtype=bool)
        keep[rows, self_pos] = False
        nb = idx[keep].reshape(n, 3)
        d = pts[nb] - pts[:, None, :]
        dx, dy, dz = d[..., 0], d[..., 1], d[..., 2]
        d2 = (dx * dx + dy * dy + dz * dz).astype(np.float32)
        d2.sort(axis=1)
    else:
        for r in range(n):
            others = [j for j in range(n) if j != r]
            vals = sorted(float(np.float32(((pts[j] - pts[r]) ** 2).sum())) for j in others)
            for j, v in enumerate(vals[:3]):
                d2[r, j] = v
    mean = ((d2[:, 0] + d2[:, 1]) + d2[:, 2]) / np.float32(3.0)
    return torch.from_numpy(mean.astype(np.float32))


def synth_scene(P: int, V: int, W: int, H: int, seed: int, sh_degree: int = 3, n_poses: int | None = None,
                focal_mul: float = 1.1, opacity_views: int | None = None, depth_noise: float = 7e-4,
                dist2_fn=None) -> Scene:
    """dist2_fn: optional replacement for the CPU k-d tree that gives the 3-NN mean squared distances (e.g. distCUDA2
    on the GPU for the multi-million-point configs, where the tree takes minutes); xyz [P,3] float32 CPU -> [P] CPU."""
    g = torch.Generator().manual_seed(seed)
    focal = focal_mul * W
    poses = _camera_arc(V)
    per = [P // V + (1 if v < P % V else 0) for v in range(V)]
    assert max(per) <= W * H, "more Gaussians per view than pixels"
    ph = torch.rand(3, generator=g) * 2 * math.pi

    def zfield(u, w_):
        # depth of the ONE shared scene surface as seen from the middle ("reference") camera, values in [1.0, 2.0]
        return 1.5 + 0.5 * (0.5 * torch.sin(2 * math.pi * (1.3 * u) + ph[0]) +
                            0.3 * torch.sin(2 * math.pi * (2.1 * w_) + ph[1]) +
                            0.2 * torch.sin(2 * math.pi * (3.0 * u + 2.0 * w_) + ph[2]))

    ref = V // 2
    R_ref = quat_to_rot(poses[ref, :4]).double()
    t_ref = poses[ref, 4:].double()
    chunks = []
    for v in range(V):
        n = per[v]
        pix = torch.randperm(W * H, generator=g)[:n].sort().values
        py = (pix // W).double()
        px = (pix % W).double()
        # pixel-aligned rays of view v (pixel centres in the rasterizer's ndc2Pix convention, auxiliary.h:41-44)
        dirs = torch.stack([(px + 0.5 - W / 2.0) / focal, (py + 0.5 - H / 2.0) / focal, torch.ones_like(px)], dim=1)
        R = quat_to_rot(poses[v, :4]).double()
        t = poses[v, 4:].double()
        # intersect each ray with the shared surface by fixed-point iteration on the depth along the ray, so that all
        # V pointmaps lie on the same surface (as aligned DUSt3R/MonST3R pointmaps do) instead of V unrelated ones
        zc = torch.full((n,), 1.5, dtype=torch.float64)
        for _ in range(8):
            world = (dirs * zc[:, None] - t) @ R            # R^T (cam - t), row-vector form
            pr = world @ R_ref.T + t_ref                      # into the reference camera
            u = (pr[:, 0] / pr[:, 2] * focal + W / 2.0) / W
            w_ = (pr[:, 1] / pr[:, 2] * focal + H / 2.0) / H
            zc = zc * (zfield(u, w_) / pr[:, 2])
        zc = zc * (1.0 + depth_noise * torch.randn(n, generator=g).double())
        world = (dirs * zc[:, None] - t) @ R
        chunks.append(world.float())
    xyz = torch.cat(chunks, dim=0).contiguous().float()
    dist2 = (dist2_fn(xyz) if dist2_fn is not None else knn_dist2_cpu(xyz)).clamp_min(1e-7)
    scales = torch.sqrt(dist2)[:, None].repeat(1, 3).contiguous()
    rot = torch.zeros(P, 4)
    rot[:, 0] = 1.0
    rot = rot + 0.01 * torch.randn(P, 4, generator=g)
    ov = opacity_views if opacity_views is not None else max(V, 4)
    conf = 0.5 + 0.5 * torch.rand(P, 1, generator=g)
    opacity = (1.0 / ov) * conf
    shs = torch.zeros(P, 16, 3)
    shs[:, 0, :] = (torch.rand(P, 3, generator=g) - 0.5) / SH_C0
    shs[:, 1:, :] = 0.05 * torch.randn(P, 15, 3, generator=g)
    if n_poses is not None and n_poses != V:
        poses = _camera_arc(V, n_poses)
    return Scene(P=P, V=poses.shape[0], W=W, H=H, sh_degree=sh_degree, focal=focal, xyz=xyz, rot=rot.contiguous(),
                 scales=scales, opacity=opacity.contiguous(), shs=shs.contiguous(), poses=poses, seed=seed)


def view_inputs(scene: Scene, v: int, device: str | torch.device = "cpu", sh_degree: int | None = None) -> dict:
    """Rasterizer inputs for view v, built like reference gaussian_renderer/__init__.py:57-108."""
    dev = torch.device(device)
    pose = scene.poses[v].to(dev)
    xyz = scene.xyz.to(dev)
    rel = torch.eye(4, device=dev)
    rel[:3, :3] = quat_to_rot(scene.poses[v, :4]).to(dev)
    rel[:3, 3] = pose[4:]
    ones = torch.ones(xyz.shape[0], 1, device=dev)
    means3D = (rel @ torch.cat((xyz, ones), dim=1).T).T[:, :3]  # a transposed-slice view, like the reference
    rotations = quat_mul(pose[:4], scene.rot.to(dev))
    fovx = 2 * math.atan(scene.tanfovx)
    fovy = 2 * math.atan(scene.tanfovy)
    proj = projection_matrix(0.01, 100.0, fovx, fovy).transpose(0, 1).to(dev)
    w2c = torch.eye(4, device=dev)
    projmatrix = (w2c.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0)
    campos = w2c.inverse()[3, :3]
    return dict(
        means3D=means3D, rotations=rotations, scales=scene.scales.to(dev), opacities=scene.opacity.to(dev),
        shs=scene.shs.to(dev), viewmatrix=w2c, projmatrix=projmatrix, campos=campos,
        bg=torch.zeros(3, device=dev), tanfovx=scene.tanfovx, tanfovy=scene.tanfovy,
        image_height=scene.H, image_width=scene.W,
        sh_degree=scene.sh_degree if sh_degree is None else sh_degree,
    )


def grad_seed(scene: Scene, seed_offset: int = 1) -> torch.Tensor:
    """dL/d(out_color): N(0,1)/Npix."""
    g = torch.Generator().manual_seed(scene.seed + seed_offset)
    return torch.randn(3, scene.H, scene.W, generator=g) / (scene.W * scene.H)


# BASELINE.json configs: name -> kwargs of synth_scene
CONFIGS = {
    "C1": dict(P=100_000, V=4, W=512, H=288, seed=0, sh_degree=3),
    "C2": dict(P=1_000_000, V=50, W=512, H=288, seed=1, sh_degree=3),
    "C3": dict(P=2_000_000, V=80, W=512, H=288, seed=2, sh_degree=3),
    "C4": dict(P=6_000_000, V=50, W=1024, H=436, seed=3, sh_degree=3),
    "C5": dict(P=2_000_000, V=50, W=854, H=480, seed=4, sh_degree=3, n_poses=500),
}
