"""The training and rendering drivers (das3r_b200/train.py, das3r_b200/render_all.py; reference train_gui.py:499-641,
render.py:74-124) on a small pointmap scene built by the pointmap initialisation:

  * the SAME loop - random view stack, SH degree schedule, learning-rate schedules, PSNR-gated pose step - run in two
    arms from the same parameters: this repository's fused components, and the reference's formulation (torch glue of
    gaussian_renderer/__init__.py:83-126 + the reference CUDA rasterizer from oracle/_ref + utils/loss_utils.py in torch +
    torch.optim.Adam), for 260 iterations across three SH-degree switches; loss curves, gate decisions and final
    parameters are compared;
  * the all-frames render against the per-view drop-in rasterizer, and its frame sharding.
"""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import _harness as Hn
import oracle
from oracle import init_oracle as io

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
E = torch.Tensor([])


def _rot_to_quat(R):
    """(r,x,y,z) of a rotation matrix (numpy, trace > -1 here)."""
    t = np.trace(R)
    r = math.sqrt(max(1.0 + t, 1e-12)) / 2
    return np.array([r, (R[2, 1] - R[1, 2]) / (4 * r), (R[0, 2] - R[2, 0]) / (4 * r), (R[1, 0] - R[0, 1]) / (4 * r)], np.float32)


def _scene(V=6, H=96, W=160, seed=3):
    from das3r_b200.init import create_from_pointmaps
    from das3r_b200.synth import projection_matrix
    from das3r_b200.train import Camera, ViewGeometry, state_from_init

    m = io.synth_pointmaps(V=V, H=H, W=W, seed=seed)
    yy, xx = np.meshgrid(np.linspace(0, 1, H, dtype=np.float32), np.linspace(0, 1, W, dtype=np.float32), indexing="ij")
    r = np.random.default_rng(seed)
    rgb = np.stack([np.stack([0.5 + 0.4 * np.sin(5 * xx + c + 0.2 * v) * np.cos(4 * yy - c) for c in range(3)])
                    for v in range(V)]).astype(np.float32)
    m["rgb"] = rgb
    t = lambda a: torch.as_tensor(np.ascontiguousarray(a), device=DEV)
    init = create_from_pointmaps(t(m["K"]), t(m["poses"]), t(m["depth"]), t(m["confidence"]), t(m["dyna_avg"]), t(m["rgb"]))
    w2c = []
    for v in range(V):
        c2w = m["poses"][v].astype(np.float64)
        Rw = c2w[:3, :3].T
        tw = -Rw @ c2w[:3, 3]
        w2c.append(np.concatenate([_rot_to_quat(Rw), tw.astype(np.float32)]))
    poses_w2c = t(np.stack(w2c).astype(np.float32))
    f = float(m["K"][0, 0, 0])
    tanx, tany = W / (2 * f), H / (2 * f)
    proj = projection_matrix(0.01, 100.0, 2 * math.atan(tanx), 2 * math.atan(tany)).transpose(0, 1).to(DEV).contiguous()
    geom = ViewGeometry(image_height=H, image_width=W, tanfovx=tanx, tanfovy=tany, projmatrix=proj, bg=torch.zeros(3, device=DEV))
    # ground truth: the views' own colour maps; the odd views get noise on top (their PSNR stays under the gate)
    gts = []
    for v in range(V):
        g = rgb[v].copy()
        if v % 2 == 1:
            g = np.clip(g + 0.25 * r.standard_normal(g.shape).astype(np.float32), 0, 1)
        gts.append(t(g))
    cams = [Camera(uid=v, original_image=gts[v]) for v in range(V)]
    return init, poses_w2c, geom, cams, (V, H, W)


def _reference_hooks():
    """render / loss / optimizer of the reference's formulation."""
    C, _ = oracle.load_reference()

    class Raster(torch.autograd.Function):
        @staticmethod
        def forward(ctx, means3D, shs, opacities, scales, rotations, geom, D):
            out = C.rasterize_gaussians(geom.bg, means3D, E, opacities, scales, rotations, 1.0, E, geom.viewmatrix,
                                        geom.projmatrix, geom.tanfovx, geom.tanfovy, geom.image_height, geom.image_width, shs,
                                        D, geom.campos, False, False)
            R, color, radii, gbuf, bbuf, ibuf = out[:6]
            ctx.geom, ctx.R, ctx.D = geom, R, D
            ctx.save_for_backward(means3D, shs, scales, rotations, radii, gbuf, bbuf, ibuf)
            return color

        @staticmethod
        def backward(ctx, g):
            means3D, shs, scales, rotations, radii, gbuf, bbuf, ibuf = ctx.saved_tensors
            geom = ctx.geom
            (_m2, _dc, d_op, d_m3, _cov, d_sh, d_sc, d_rot) = C.rasterize_gaussians_backward(
                geom.bg, means3D, radii, E, scales, rotations, 1.0, E, geom.viewmatrix, geom.projmatrix, geom.tanfovx,
                geom.tanfovy, g.contiguous(), shs, ctx.D, geom.campos, gbuf, ctx.R, bbuf, ibuf, False)
            return d_m3, d_sh, d_op, d_sc, d_rot, None, None

    def render_fn(st, geom, pose, filtering=None):
        # gaussian_renderer/__init__.py:83-97,:106,:126 with utils/pose_utils.py:36-104
        q = pose[:4]
        r, x, y, z = q / torch.sqrt((q * q).sum())
        Rm = torch.stack([torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y)]),
                          torch.stack([2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x)]),
                          torch.stack([2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)])])
        w2c = torch.cat([torch.cat([Rm, pose[4:].reshape(3, 1)], dim=1), torch.eye(4, device=pose.device)[3:]], dim=0)
        xyz, rot = st.xyz.clone(), st.rotation.clone()
        homo = torch.cat((xyz, torch.ones(xyz.shape[0], 1, device=xyz.device)), dim=1)
        means = (w2c @ homo.T).T[:, :3].contiguous()
        w1, x1, y1, z1 = q.unbind(-1)
        w2, x2, y2, z2 = rot.unbind(-1)
        rots = torch.stack([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                            w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2], dim=-1)
        opacity = torch.sigmoid(st.opacity) * st.conf_static.reshape(-1, 1)[st.conf_index]
        shs = torch.cat((st.features_dc, st.features_rest), dim=1)
        return Raster.apply(means, shs, opacity, torch.exp(st.scaling), rots, geom, st.active_sh_degree)

    x = np.arange(11)
    w1 = np.exp(-(x - 5) ** 2 / (2 * 1.5 ** 2)).astype(np.float32)
    w1 = torch.tensor(w1 / w1.sum(), device=DEV)
    window = (w1[:, None] @ w1[None, :])[None, None].expand(3, 1, 11, 11).contiguous()

    def loss_fn(image, gt, static, lam):
        # train_gui.py:567-576 with utils/loss_utils.py:36-67 and utils/image_utils.py:17-19
        im, g = image * static, gt * static
        conv = lambda t: F.conv2d(t, window, padding=5, groups=3)
        mu1, mu2 = conv(im), conv(g)
        s1, s2, s12 = conv(im * im) - mu1.pow(2), conv(g * g) - mu2.pow(2), conv(im * g) - mu1 * mu2
        C1, C2 = 0.01 ** 2, 0.03 ** 2
        ssim_map = ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1.pow(2) + mu2.pow(2) + C1) * (s1 + s2 + C2))
        mse = ((im - g) ** 2).view(3, -1).mean(1, keepdim=True)
        psnr = (20 * torch.log10(1.0 / torch.sqrt(mse))).mean()
        return ((1.0 - lam) * torch.abs(im - g) + lam * (1.0 - ssim_map)).mean(), psnr.detach()

    return render_fn, loss_fn, torch.optim.Adam


def test_training_loop_against_the_reference_formulation():
    if not oracle.reference_available():
        pytest.skip("oracle/_ref not built")
    from das3r_b200.train import OptimizationParams, Trainer, state_from_init

    init, poses_w2c, geom, cams, (V, H, W) = _scene()
    opt = OptimizationParams(iterations=260, sh_increase_interval=80, psnr_threshold=20.0)
    ours = Trainer(state_from_init(init, V, H, W, poses_w2c), cams, geom, opt, seed=5)
    rf, lf, adam = _reference_hooks()
    ref = Trainer(state_from_init(init, V, H, W, poses_w2c), cams, geom, opt, seed=5, render_fn=rf, loss_fn=lf, adam_cls=adam)
    ha, hb = ours.train(260), ref.train(260)
    torch.cuda.synchronize()
    assert ours.state.active_sh_degree == 3 and ref.state.active_sh_degree == 3  # switched at 80, 160, 240
    la, lb = np.array([h[2] for h in ha]), np.array([h[2] for h in hb])
    pa, pb = np.array([h[3] for h in ha]), np.array([h[3] for h in hb])
    assert [h[1] for h in ha] == [h[1] for h in hb]  # same views in the same order
    rel = np.abs(la - lb) / lb
    print("loss rel diff: first 20 max", rel[:20].max(), "all max", rel.max(), "final losses", la[-1], lb[-1],
          "psnr range", pa.min(), pa.max(), "pose steps", sum(h[4] for h in ha), sum(h[4] for h in hb))
    assert la[-20:].mean() < 0.8 * la[:20].mean(), "the loop does not train"
    assert rel[:20].max() <= 1e-4, rel[:20].max()
    assert rel.max() <= 2e-2, rel.max()
    gates_a, gates_b = [h[4] for h in ha], [h[4] for h in hb]
    assert 0 < sum(gates_a) < len(gates_a), "the PSNR gate never switched in this test"
    assert sum(a != b for a, b in zip(gates_a, gates_b)) <= 3
    # parameters after 260 optimizer steps: the arms moved the same way
    for name in ("xyz", "features_dc", "opacity", "scaling", "rotation", "conf_static", "Q", "T"):
        a, b = getattr(ours.state, name).detach(), getattr(ref.state, name).detach()
        start = getattr(state_from_init(init, V, H, W, poses_w2c), name)
        moved = float((b - start).abs().mean())
        diff = float((a - b).abs().mean())
        print(name, "mean |update|", moved, "mean |ours - ref|", diff)
        assert diff <= 0.05 * moved + 1e-7, (name, diff, moved)


def test_render_all_frames_matches_the_per_view_rasterizer_and_shards():
    from das3r_b200.rasterizer import GaussianRasterizer
    from das3r_b200.render_all import render_frames, render_my_frames
    from das3r_b200.train import render_view, state_from_init

    init, poses_w2c, geom, cams, (V, H, W) = _scene(V=4, H=64, W=96, seed=8)
    st = state_from_init(init, V, H, W, poses_w2c)
    st.active_sh_degree = 3
    with torch.no_grad():
        st.features_rest.copy_(0.05 * torch.randn(st.features_rest.shape, generator=torch.Generator().manual_seed(1)).to(DEV))
    # a trajectory of 11 interpolated poses (render.py renders the interpolated path)
    F_ = 11
    a = torch.linspace(0, 1, F_, device=DEV)[:, None]
    traj = (poses_w2c[0][None] * (1 - a) + poses_w2c[-1][None] * a).contiguous()
    frames = render_frames(st, geom, traj, frames_per_batch=4)
    assert len(frames) == F_
    for f in (0, 5, 10):
        with torch.no_grad():
            want = render_view(st, geom, traj[f])
        assert torch.equal(frames[f], want)
    got = {}
    for rank in range(3):
        mine, imgs = render_my_frames(st, geom, traj, rank, 3, frames_per_batch=2)
        assert all(i % 3 == rank for i in mine)
        got.update(dict(zip(mine, imgs)))
    assert sorted(got) == list(range(F_))
    for f in range(F_):
        assert torch.equal(got[f], frames[f])
    # per-Gaussian static confidence, as after load_ply (render_test, gaussian_renderer/__init__.py:226)
    st2 = state_from_init(init, V, H, W, poses_w2c)
    st2.active_sh_degree = 3
    st2.conf_static = st.conf_static.reshape(-1)[st.conf_index].clone()
    st2.conf_index = None
    with torch.no_grad():
        st2.features_rest.copy_(st.features_rest)
    again = render_frames(st2, geom, traj, frames_per_batch=8, frames=[0, 10])
    assert torch.equal(again[0], frames[0]) and torch.equal(again[1], frames[10])
