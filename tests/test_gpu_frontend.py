"""The fused DAS3R render front-end (include/das3r_frontend.h) on the GPU against
  * the golden vectors made with the reference's own functions (tests/golden/frontend_*.npz),
  * oracle/frontend_oracle.py on a larger seeded case,
  * the reference's torch formulation run on the GPU (same five lines as gaussian_renderer/__init__.py:83-126), through
    autograd, including the pose gradients,
and chained with MultiViewRasterizer. fp32 tolerances: forward 2e-6 relative (a pure copy for the SH rows), gradients
1e-4 relative to the tensor's largest entry (sums over up to 10^5 Gaussians in a different order)."""
import glob
import os

import numpy as np
import pytest
import torch

import _harness as Hn
from das3r_b200 import load_native
from oracle import frontend_oracle as fo

pytestmark = pytest.mark.gpu
GOLDEN = sorted(glob.glob(os.path.join(Hn.ROOT, "tests", "golden", "frontend_*.npz")))
DEV = "cuda:0"


def _t(a, dtype=torch.float32):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype, device=DEV)


def _run(m, poses, gin=None):
    C = load_native()
    ci = _t(m["conf_index"], torch.int64) if m.get("conf_index") is not None else torch.empty(0, dtype=torch.int64)
    args = [_t(m[k]) for k in ("xyz", "rotation", "opacity", "scaling", "features_dc", "features_rest")]
    conf = _t(m["conf_static"]).reshape(-1)
    opac, scales, shs, means, rots = C.frontend_forward(*args, conf, ci, _t(poses))
    out = dict(opacity=opac, scales=scales, shs=shs, means3D=means, rotations=rots)
    if gin is None:
        return out, None
    g = C.frontend_backward(args[0], args[1], args[2], args[3], conf, ci, _t(poses), _t(gin["dL_dopacity"]),
                            _t(gin["dL_dscales"]), _t(gin["dL_dshs"]), [_t(x) for x in gin["dL_dmeans3D"]],
                            [_t(x) for x in gin["dL_drotations"]])
    names = ("dL_dxyz", "dL_drotation", "dL_dopacity", "dL_dscaling", "dL_dfeatures_dc", "dL_dfeatures_rest", "dL_dconf",
             "dL_dposes")
    torch.cuda.synchronize()
    return out, dict(zip(names, g))


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_against_reference_golden(path):
    z = np.load(path)
    m = {k[3:]: z[k] for k in z.files if k.startswith("in_") and k not in ("in_poses", "in_mask")}
    m["conf_index"] = np.nonzero(z["in_mask"])[0]
    gin = dict(dL_dmeans3D=list(z["gin_dL_dmeans3D"]), dL_drotations=list(z["gin_dL_drotations"]),
               dL_dopacity=z["gin_dL_dopacity"], dL_dscales=z["gin_dL_dscales"], dL_dshs=z["gin_dL_dshs"])
    out, g = _run(m, z["in_poses"], gin)
    for v in range(len(z["in_poses"])):
        np.testing.assert_allclose(Hn.np_(out["means3D"][v]), z["out_means3D"][v], rtol=2e-6, atol=2e-6)
        np.testing.assert_allclose(Hn.np_(out["rotations"][v]), z["out_rotations"][v], rtol=2e-6, atol=2e-6)
    np.testing.assert_allclose(Hn.np_(out["opacity"]), z["out_opacity"], rtol=2e-6, atol=1e-7)
    np.testing.assert_allclose(Hn.np_(out["scales"]), z["out_scales"], rtol=2e-6, atol=1e-7)
    assert np.array_equal(Hn.np_(out["shs"]), z["out_shs"])
    for ours, ref in (("dL_dxyz", "grad_xyz"), ("dL_drotation", "grad_rotation"), ("dL_dopacity", "grad_opacity"),
                      ("dL_dscaling", "grad_scaling"), ("dL_dfeatures_dc", "grad_features_dc"),
                      ("dL_dfeatures_rest", "grad_features_rest"), ("dL_dposes", "grad_poses")):
        assert Hn.rel_err(g[ours], z[ref]) <= 1e-4, ours
    conf_grad = np.zeros_like(z["grad_conf_static"])
    conf_grad[m["conf_index"]] = Hn.np_(g["dL_dconf"])
    assert Hn.rel_err(conf_grad, z["grad_conf_static"]) <= 1e-4


def _case(P, V, seed, with_index=True):
    r = np.random.default_rng(seed)
    f = lambda *s: r.standard_normal(s).astype(np.float32)
    n_pix = 2 * P
    m = dict(xyz=f(P, 3), rotation=(np.array([1, 0, 0, 0], np.float32) + 0.1 * f(P, 4)), opacity=f(P, 1),
             scaling=-3 + 0.5 * f(P, 3), features_dc=f(P, 1, 3), features_rest=0.1 * f(P, 15, 3),
             conf_static=r.random(n_pix if with_index else P).astype(np.float32),
             conf_index=np.sort(r.choice(n_pix, P, replace=False)) if with_index else None)
    poses = np.concatenate([np.array([1, 0, 0, 0], np.float32) + 0.2 * f(V, 4), 0.3 * f(V, 3)], axis=1)
    gin = dict(dL_dmeans3D=[f(P, 3) for _ in range(V)], dL_drotations=[f(P, 4) for _ in range(V)], dL_dopacity=f(P, 1),
               dL_dscales=f(P, 3), dL_dshs=f(P, 16, 3))
    return m, poses, gin


@pytest.mark.parametrize("P,V,with_index", [(100_003, 8, True), (4_097, 16, False), (1, 1, True)])
def test_against_the_oracle(P, V, with_index):
    m, poses, gin = _case(P, V, seed=P + V, with_index=with_index)
    out, g = _run(m, poses, gin)
    want = fo.forward(m, poses)
    for v in range(V):
        np.testing.assert_allclose(Hn.np_(out["means3D"][v]), want["means3D"][v], rtol=3e-6, atol=3e-6)
        np.testing.assert_allclose(Hn.np_(out["rotations"][v]), want["rotations"][v], rtol=3e-6, atol=3e-6)
    np.testing.assert_allclose(Hn.np_(out["opacity"]), want["opacity"], rtol=3e-6, atol=1e-7)
    np.testing.assert_allclose(Hn.np_(out["scales"]), want["scales"], rtol=3e-6, atol=1e-7)
    assert np.array_equal(Hn.np_(out["shs"]), want["shs"])
    wg = fo.backward(m, poses, gin)  # float64
    for k in wg:
        assert Hn.rel_err(g[k], wg[k]) <= 1e-4, k


def test_autograd_matches_the_reference_formulation_on_the_gpu():
    """Same five lines as render() with torch ops, on the GPU, against the fused op - through autograd, with the
    rasterizer-side gradients replaced by random weights."""
    from das3r_b200.frontend import render_inputs

    P, V = 30_000, 5
    m, poses, gin = _case(P, V, seed=3)
    names = ("xyz", "rotation", "opacity", "scaling", "features_dc", "features_rest", "conf_static")

    def leaves():
        return {k: _t(m[k]).requires_grad_(True) for k in names}, _t(poses).requires_grad_(True)

    ci = _t(m["conf_index"], torch.int64)
    w = {k: ([_t(x) for x in v] if isinstance(v, list) else _t(v)) for k, v in gin.items()}

    def loss_of(means, rots, opac, scales, shs):
        l = sum((a * b).sum() for a, b in zip(means, w["dL_dmeans3D"])) + sum(
            (a * b).sum() for a, b in zip(rots, w["dL_drotations"]))
        return l + (opac * w["dL_dopacity"]).sum() + (scales * w["dL_dscales"]).sum() + (shs * w["dL_dshs"]).sum()

    # reference formulation (gaussian_renderer/__init__.py:83-97,:106,:126; utils/pose_utils.py:36-104)
    A, pa = leaves()
    means, rots = [], []
    for v in range(V):
        q = pa[v, :4]
        qn = q / torch.sqrt((q * q).sum())
        r, x, y, z = qn
        R = torch.stack([torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y)]),
                         torch.stack([2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x)]),
                         torch.stack([2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)])])
        means.append(A["xyz"] @ R.T + pa[v, 4:])
        w1, x1, y1, z1 = q
        w2, x2, y2, z2 = A["rotation"].unbind(-1)
        rots.append(torch.stack([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                                 w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2], dim=-1))
    opac = torch.sigmoid(A["opacity"]) * A["conf_static"].reshape(-1, 1)[ci]
    loss_of(means, rots, opac, torch.exp(A["scaling"]), torch.cat((A["features_dc"], A["features_rest"]), dim=1)).backward()

    B, pb = leaves()
    fe = render_inputs(B["xyz"], B["rotation"], B["opacity"], B["scaling"], B["features_dc"], B["features_rest"],
                       B["conf_static"], ci, pb)
    loss_of(fe.means3D, fe.rotations, fe.opacity, fe.scales, fe.shs).backward()
    torch.cuda.synchronize()
    for k in names:
        assert Hn.rel_err(B[k].grad, A[k].grad) <= 1e-4, k
    assert Hn.rel_err(pb.grad, pa.grad) <= 1e-4, "poses"


def test_chained_with_the_multi_view_rasterizer():
    """parameters + poses -> fused front-end -> MultiViewRasterizer -> loss; gradients reach the raw parameters and
    the poses and agree with the unfused torch front-end feeding the per-view rasterizer."""
    from das3r_b200.frontend import render_inputs
    from das3r_b200.multiview import MultiViewRasterizer
    from das3r_b200.rasterizer import GaussianRasterizationSettings, GaussianRasterizer
    from das3r_b200.synth import synth_scene, view_inputs

    P, V, W, H = 8_000, 3, 128, 80
    sc = synth_scene(P=P, V=V, W=W, H=H, seed=9, sh_degree=3)
    vi = view_inputs(sc, 0)
    dev = torch.device(DEV)
    g = torch.Generator().manual_seed(0)
    base = dict(xyz=vi["means3D"].clone(), rotation=vi["rotations"].clone(),
                opacity=torch.logit(vi["opacities"].clamp(1e-4, 1 - 1e-4)), scaling=torch.log(vi["scales"]),
                features_dc=vi["shs"][:, :1].clone(), features_rest=vi["shs"][:, 1:].clone(),
                conf_static=0.5 + 0.5 * torch.rand(P, generator=g))
    poses = torch.cat([torch.tensor([1.0, 0, 0, 0]) + 0.01 * torch.randn(V, 4, generator=g),
                       0.01 * torch.randn(V, 3, generator=g)], dim=1)
    dLs = [(torch.randn(3, H, W, generator=g) / (W * H)).to(dev) for _ in range(V)]
    settings = [GaussianRasterizationSettings(
        image_height=H, image_width=W, tanfovx=vi["tanfovx"], tanfovy=vi["tanfovy"], bg=vi["bg"].to(dev),
        scale_modifier=1.0, viewmatrix=vi["viewmatrix"].to(dev), projmatrix=vi["projmatrix"].to(dev), sh_degree=3,
        campos=vi["campos"].to(dev), prefiltered=False, debug=False) for _ in range(V)]

    def leaves():
        return {k: v.to(dev).contiguous().requires_grad_(True) for k, v in base.items()}, poses.to(dev).requires_grad_(True)

    A, pa = leaves()
    fe = render_inputs(A["xyz"], A["rotation"], A["opacity"], A["scaling"], A["features_dc"], A["features_rest"],
                       A["conf_static"], None, pa)
    colors, _ = MultiViewRasterizer(streams=2)(settings, fe.means3D, fe.rotations, fe.opacity, fe.shs, fe.scales)
    sum((c * d).sum() for c, d in zip(colors, dLs)).backward()

    B, pb = leaves()
    loss = 0
    for v in range(V):
        q = pb[v, :4]
        r, x, y, z = q / torch.sqrt((q * q).sum())
        R = torch.stack([torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y)]),
                         torch.stack([2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x)]),
                         torch.stack([2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)])])
        means = B["xyz"] @ R.T + pb[v, 4:]
        w1, x1, y1, z1 = q
        w2, x2, y2, z2 = B["rotation"].unbind(-1)
        rots = torch.stack([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                            w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2], dim=-1)
        opac = torch.sigmoid(B["opacity"]) * B["conf_static"].reshape(-1, 1)
        color, _ = GaussianRasterizer(settings[v])(
            means3D=means, means2D=torch.zeros(P, 3, device=dev, requires_grad=True), opacities=opac,
            shs=torch.cat((B["features_dc"], B["features_rest"]), dim=1), scales=torch.exp(B["scaling"]), rotations=rots)
        loss = loss + (color * dLs[v]).sum()
    loss.backward()
    torch.cuda.synchronize()
    for k in base:
        assert Hn.rel_err(A[k].grad, B[k].grad) <= Hn.GRAD_RTOL, k
    assert Hn.rel_err(pa.grad, pb.grad) <= Hn.GRAD_RTOL, "poses"
