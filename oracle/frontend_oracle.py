"""CPU restatement (numpy) of DAS3R's render() front-end and its backward - TEST INFRASTRUCTURE, like everything under
oracle/: only tests/, __graft_entry__.smoke() and bench.py's baseline legs may import it; the product path
(das3r_b200/) never does.

What it restates, with the reference lines each step follows (paths relative to /root/reference):
  means3D   = (get_camera_from_tensor(pose) @ [xyz | 1]^T)^T[:, :3]   gaussian_renderer/__init__.py:83-90,
                                                                      utils/pose_utils.py:57-85 (pose -> 4x4),
                                                                      utils/pose_utils.py:36-55 (quad2rotation: the
                                                                      quaternion is NORMALISED for the matrix)
  rotations = quadmultiply(pose[:4], _rotation)                       gaussian_renderer/__init__.py:91,
                                                                      utils/pose_utils.py:86-104 (RAW pose quaternion)
  opacity   = sigmoid(_opacity) * _conf_static.reshape(-1,1)[mask]    gaussian_renderer/__init__.py:95-97,
                                                                      scene/gaussian_model.py:193-194
  scales    = exp(_scaling)                                           gaussian_renderer/__init__.py:106,
                                                                      scene/gaussian_model.py:103-105
  shs       = cat(_features_dc, _features_rest, dim=1)                gaussian_renderer/__init__.py:126,
                                                                      scene/gaussian_model.py:186-190
The backward is the analytic adjoint of those five lines (the reference gets it from autograd).

Parity pinned: tests/golden/frontend_*.npz hold outputs and autograd gradients of the reference's own functions
(utils/pose_utils.py imported from /root/reference by tests/golden/make_frontend_golden.py, CPU fp32);
tests/test_frontend_oracle.py checks this file against them.
"""
from __future__ import annotations

import numpy as np


def quad2rotation(q: np.ndarray) -> np.ndarray:
    """utils/pose_utils.py:36-55, one quaternion (r, x, y, z) -> 3x3."""
    q = q / np.sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3])
    r, x, y, z = q
    one, two = q.dtype.type(1), q.dtype.type(2)
    return np.array([[one - two * (y * y + z * z), two * (x * y - r * z), two * (x * z + r * y)],
                     [two * (x * y + r * z), one - two * (x * x + z * z), two * (y * z - r * x)],
                     [two * (x * z - r * y), two * (y * z + r * x), one - two * (x * x + y * y)]], dtype=q.dtype)


def quadmultiply(q1: np.ndarray, q2: np.ndarray) -> np.ndarray:
    """utils/pose_utils.py:86-104; q1 [4], q2 [P,4]."""
    w1, x1, y1, z1 = q1
    w2, x2, y2, z2 = q2[:, 0], q2[:, 1], q2[:, 2], q2[:, 3]
    return np.stack([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2,
                     w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                     w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
                     w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2], axis=-1)


def _conf(m):
    c = m["conf_static"].reshape(-1)
    return c[m["conf_index"]] if m.get("conf_index") is not None else c[:m["xyz"].shape[0]]


def forward(m: dict, poses: np.ndarray, dtype=np.float32) -> dict:
    """m: xyz [P,3], rotation [P,4], opacity [P,1], scaling [P,3], features_dc [P,1,3], features_rest [P,15,3],
    conf_static [...], conf_index [P] or None. poses [V,7]."""
    f = lambda a: np.asarray(a, dtype=dtype)
    xyz, rot = f(m["xyz"]), f(m["rotation"])
    poses = f(poses)
    means, rots = [], []
    for p in poses:
        R, t = quad2rotation(p[:4]), p[4:]
        means.append(xyz @ R.T + t)                     # gaussian_renderer/__init__.py:87-89
        rots.append(quadmultiply(p[:4], rot))           # :91
    sig = dtype(1) / (dtype(1) + np.exp(-f(m["opacity"])))
    return dict(means3D=means, rotations=rots, opacity=sig * f(_conf(m)).reshape(-1, 1),
                scales=np.exp(f(m["scaling"])), shs=np.concatenate([f(m["features_dc"]), f(m["features_rest"])], axis=1))


def backward(m: dict, poses: np.ndarray, g: dict, dtype=np.float64) -> dict:
    """g: dL_dmeans3D / dL_drotations (lists, one per view), dL_dopacity [P,1], dL_dscales [P,3], dL_dshs [P,16,3]."""
    f = lambda a: np.asarray(a, dtype=dtype)
    xyz, rot, poses = f(m["xyz"]), f(m["rotation"]), f(poses)
    d_xyz, d_rot = np.zeros_like(xyz), np.zeros_like(rot)
    d_poses = np.zeros((len(poses), 7), dtype)
    for v, p in enumerate(poses):
        gm, d = f(g["dL_dmeans3D"][v]), f(g["dL_drotations"][v])
        n = np.sqrt((p[:4] ** 2).sum())
        r, x, y, z = p[:4] / n
        R = quad2rotation(p[:4])
        d_xyz += gm @ R                                   # (R^T g)^T
        M = gm.T @ xyz                                    # dL/dR[a][b] = sum_i g_a X_b
        dR = [np.array([[0, -2 * z, 2 * y], [2 * z, 0, -2 * x], [-2 * y, 2 * x, 0]]),
              np.array([[0, 2 * y, 2 * z], [2 * y, -4 * x, -2 * r], [2 * z, 2 * r, -4 * x]]),
              np.array([[-4 * y, 2 * x, 2 * r], [2 * x, 0, 2 * z], [-2 * r, 2 * z, -4 * y]]),
              np.array([[-4 * z, -2 * r, 2 * x], [2 * r, -4 * z, 2 * y], [2 * x, 2 * y, 0]])]
        dq_hat = np.array([(M * D).sum() for D in dR])
        qh = np.array([r, x, y, z])
        d_poses[v, :4] = (dq_hat - qh * (qh @ dq_hat)) / n      # through q / |q|
        d_poses[v, 4:] = gm.sum(0)
        w1, x1, y1, z1 = p[:4]
        d0, d1, d2, d3 = d[:, 0], d[:, 1], d[:, 2], d[:, 3]
        w2, x2, y2, z2 = rot[:, 0], rot[:, 1], rot[:, 2], rot[:, 3]
        d_rot += np.stack([d0 * w1 + d1 * x1 + d2 * y1 + d3 * z1, -d0 * x1 + d1 * w1 + d2 * z1 - d3 * y1,
                           -d0 * y1 - d1 * z1 + d2 * w1 + d3 * x1, -d0 * z1 + d1 * y1 - d2 * x1 + d3 * w1], axis=-1)
        d_poses[v, 0] += (d0 * w2 + d1 * x2 + d2 * y2 + d3 * z2).sum()
        d_poses[v, 1] += (-d0 * x2 + d1 * w2 - d2 * z2 + d3 * y2).sum()
        d_poses[v, 2] += (-d0 * y2 + d1 * z2 + d2 * w2 - d3 * x2).sum()
        d_poses[v, 3] += (-d0 * z2 - d1 * y2 + d2 * x2 + d3 * w2).sum()
    sig = 1 / (1 + np.exp(-f(m["opacity"])))
    conf = f(_conf(m)).reshape(-1, 1)
    go = f(g["dL_dopacity"])
    dsh = f(g["dL_dshs"])
    return dict(dL_dxyz=d_xyz, dL_drotation=d_rot, dL_dopacity=go * conf * sig * (1 - sig), dL_dconf=(go * sig).reshape(-1),
                dL_dscaling=f(g["dL_dscales"]) * np.exp(f(m["scaling"])), dL_dfeatures_dc=dsh[:, :1].copy(),
                dL_dfeatures_rest=dsh[:, 1:].copy(), dL_dposes=d_poses)
