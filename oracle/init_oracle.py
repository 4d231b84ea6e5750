"""TEST INFRASTRUCTURE - NOT PRODUCT CODE.

numpy (float32) restatement of the pointmap initialisation, reference scene/gaussian_model.py:573-659 with
utils/pose_utils.py:572-583 (depth_to_pts3d), :672-683 (_fast_depthmap_to_pts3d), :609-659 (geotrf),
utils/sh_utils.py:114-115 (RGB2SH) and utils/general_utils.py:18-19 (inverse_sigmoid). The 3-NN distances
(simple_knn.cu:148-184 semantics: exact mean of the 3 smallest squared distances to other indices) come from a k-d tree.

Parity pin: tests/golden/init_*.npz hold the outputs of the reference's own depth_to_pts3d / RGB2SH / inverse_sigmoid
(imported from /root/reference by tests/golden/make_init_golden.py) glued as in create_from_cameras;
tests/test_init_oracle.py checks this file against them. Only tests/ and __graft_entry__.smoke() may import it.
"""
from __future__ import annotations

import numpy as np

C0 = 0.28209479177387814


def knn_mean_dist2(pts: np.ndarray) -> np.ndarray:
    from scipy.spatial import cKDTree

    n = pts.shape[0]
    if n < 4:
        raise ValueError("the oracle needs at least 4 points")
    tree = cKDTree(pts.astype(np.float64))
    _, idx = tree.query(pts.astype(np.float64), k=4, workers=-1)
    rows = np.arange(n)
    is_self = idx == rows[:, None]
    self_pos = np.where(is_self.any(axis=1), np.argmax(is_self, axis=1), 3)
    keep = np.ones((n, 4), dtype=bool)
    keep[rows, self_pos] = False
    nb = idx[keep].reshape(n, 3)
    d = pts[nb] - pts[:, None, :]
    d2 = (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1] + d[..., 2] * d[..., 2]).astype(np.float32)
    d2.sort(axis=1)
    return (((d2[:, 0] + d2[:, 1]) + d2[:, 2]) / np.float32(3.0)).astype(np.float32)


def create_from_pointmaps(K, poses, depth, confidence, dyna_avg, rgb, conf_thre=1.0, dist2=None) -> dict:
    K, poses, depth = (np.asarray(a, np.float32) for a in (K, poses, depth))
    confidence, dyna_avg, rgb = (np.asarray(a, np.float32) for a in (confidence, dyna_avg, rgb))
    V, H, W = depth.shape
    focal = K[:, 0, 0][:, None, None]                       # pose_utils.py:575
    pp = K[:, :2, 2]
    jj, ii = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")   # xy_grid: [...,0] = column i, [...,1] = row j
    gx = ii[None].astype(np.float32) - pp[:, 0][:, None, None]
    gy = jj[None].astype(np.float32) - pp[:, 1][:, None, None]
    cx = (depth * gx / focal).astype(np.float32)            # :683
    cy = (depth * gy / focal).astype(np.float32)
    cam = np.stack([cx, cy, depth], axis=-1)                # [V,H,W,3]
    R, t = poses[:, :3, :3], poses[:, :3, 3]
    world = (np.einsum("bij,bhwj->bhwi", R, cam) + t[:, None, None, :]).astype(np.float32)   # geotrf :636
    mask = confidence.reshape(-1) > np.float32(np.log(np.float32(conf_thre)))                # gaussian_model.py:624
    xyz = world.reshape(-1, 3)[mask]
    colour = rgb.transpose(0, 2, 3, 1).reshape(-1, 3)[mask]                                 # :615, :630
    f_dc = ((colour - np.float32(0.5)) / np.float32(C0)).astype(np.float32)                 # RGB2SH
    P = xyz.shape[0]
    if dist2 is None:
        dist2 = knn_mean_dist2(xyz)
    s = np.log(np.sqrt(np.maximum(dist2.astype(np.float32), np.float32(1e-7)))).astype(np.float32)   # :641-642
    x = np.float32(1.0 / V)
    with np.errstate(divide="ignore"):
        opa = np.log(x / (np.float32(1.0) - x)).astype(np.float32)                              # :646-647
    rot = np.zeros((P, 4), np.float32)
    rot[:, 0] = 1
    return dict(xyz=xyz, features_dc=f_dc[:, None, :], features_rest=np.zeros((P, 15, 3), np.float32),
                scaling=np.repeat(s[:, None], 3, axis=1), rotation=rot, opacity=np.full((P, 1), opa, np.float32),
                conf_static=(np.float32(1.0) - dyna_avg).astype(np.float32), aggregated_mask=mask,
                conf_index=np.nonzero(mask)[0].astype(np.int64), dist2=dist2.astype(np.float32))


def synth_pointmaps(V, H, W, seed, conf_ratio=0.8):
    """Seeded DAS3R-shaped inputs: smooth depth, DUSt3R-like log-confidence around 0, camera arc."""
    r = np.random.default_rng(seed)
    f = np.float32(1.1 * W)
    K = np.tile(np.array([[f, 0, W / 2], [0, f, H / 2], [0, 0, 1]], np.float32), (V, 1, 1))
    poses = np.tile(np.eye(4, dtype=np.float32), (V, 1, 1))
    for v in range(V):
        a = np.float32(0.3 * (v / max(V - 1, 1) - 0.5))
        poses[v, :3, :3] = np.array([[np.cos(a), 0, np.sin(a)], [0, 1, 0], [-np.sin(a), 0, np.cos(a)]], np.float32)
        poses[v, :3, 3] = np.array([0.3 * a, 0.02 * v, 0.0], np.float32)
    yy, xx = np.meshgrid(np.linspace(0, 1, H, dtype=np.float32), np.linspace(0, 1, W, dtype=np.float32), indexing="ij")
    depth = np.stack([1.5 + 0.4 * np.sin(6 * xx + 0.3 * v) * np.cos(5 * yy) + 0.01 * r.standard_normal((H, W)).astype(np.float32)
                      for v in range(V)]).astype(np.float32)
    conf = (r.standard_normal((V, H, W)).astype(np.float32) + np.float32(0.8416 if conf_ratio == 0.8 else 0.0))
    dyna = r.random((V, H, W)).astype(np.float32) ** 3
    rgb = r.random((V, 3, H, W)).astype(np.float32)
    return dict(K=K, poses=poses, depth=depth, confidence=conf.astype(np.float32), dyna_avg=dyna.astype(np.float32), rgb=rgb)
