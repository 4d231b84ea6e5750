"""Fused DAS3R render front-end: what gaussian_renderer/__init__.py:83-108,:126 does to the model parameters between
`GaussianModel` and the rasterizer call - pose transform of means and rotations, opacity and scale activations, SH
concatenation - for all views of a step in one kernel, and the backward (parameter AND pose gradients) in one more
(include/das3r_frontend.h, das3r_b200/csrc/frontend.cu).

    fe = render_inputs(xyz, rotation, opacity, scaling, features_dc, features_rest, conf_static, conf_index, poses)
    colors, radii = MultiViewRasterizer()(settings, fe.means3D, fe.rotations, fe.opacity, fe.shs, fe.scales)

`poses` is [V,7] (quaternion r,x,y,z + translation per view = `camera_pose` of render()); `conf_index` is
`aggregated_mask.reshape(-1).nonzero()` (int64, [P]) or None when conf_static already has one entry per Gaussian. Like
the reference, the matrix that moves the means uses the normalised pose quaternion and the quaternion product the raw
one. features_rest is always [P,15,3] (max_sh_degree 3); the active degree selects what is copied. At most 16 views per call.
"""
from __future__ import annotations

from typing import NamedTuple, Sequence

import torch

from . import load_native

_C = load_native()
_EMPTY_I64 = torch.empty(0, dtype=torch.int64)


class RenderInputs(NamedTuple):
    means3D: Sequence[torch.Tensor]    # per view [P,3]
    rotations: Sequence[torch.Tensor]  # per view [P,4]
    opacity: torch.Tensor              # [P,1]
    scales: torch.Tensor               # [P,3]
    shs: torch.Tensor                  # [P,16,3]


class _Frontend(torch.autograd.Function):
    @staticmethod
    def forward(ctx, xyz, rotation, opacity, scaling, features_dc, features_rest, conf_static, poses, conf_index,
                sh_degree, filter_index):
        ci = conf_index if conf_index is not None else _EMPTY_I64
        fi = filter_index if filter_index is not None else _EMPTY_I64
        args = [t.contiguous() for t in (xyz, rotation, opacity, scaling, features_dc, features_rest, conf_static)]
        poses_c = poses.contiguous()
        opac, scales, shs, means, rots = _C.frontend_forward_ex(*args, ci, poses_c, int(sh_degree), fi)
        ctx.save_for_backward(args[0], args[1], args[2], args[3], args[6], poses_c, ci, fi)
        ctx.n_views = len(means)
        ctx.conf_shape = conf_static.shape
        ctx.sh_degree = int(sh_degree)
        return (opac, scales, shs, *means, *rots)

    @staticmethod
    def backward(ctx, d_opac, d_scales, d_shs, *rest):
        xyz, rotation, opacity, scaling, conf_static, poses, ci, fi = ctx.saved_tensors
        V = ctx.n_views
        P = fi.numel() if fi.numel() else xyz.shape[0]  # rows the forward produced
        z = lambda t, *shape: t if t is not None else torch.zeros(*shape, device=xyz.device)
        d_means = [z(rest[v], P, 3) for v in range(V)]
        d_rots = [z(rest[V + v], P, 4) for v in range(V)]
        dxyz, drot, dop, dsc, ddc, drest, dconf, dposes = _C.frontend_backward_ex(
            xyz, rotation, opacity, scaling, conf_static, ci, poses, z(d_opac, P, 1), z(d_scales, P, 3),
            z(d_shs, P, 16, 3), d_means, d_rots, ctx.sh_degree, fi, [], torch.empty(0, dtype=torch.uint8), 0, -1, True)
        d_conf_static = torch.zeros(conf_static.numel(), device=xyz.device)
        if ci.numel():
            d_conf_static[ci] = dconf  # the indices are unique (positions of a boolean mask's true entries)
        else:
            d_conf_static[:P] = dconf
        return dxyz, drot, dop, dsc, ddc, drest, d_conf_static.view(ctx.conf_shape), dposes, None, None, None


def render_inputs(xyz, rotation, opacity, scaling, features_dc, features_rest, conf_static, conf_index, poses,
                  sh_degree: int = 3, filtering=None) -> RenderInputs:
    """sh_degree: the model's active SH degree (only its (D+1)^2 coefficients are copied into the joint rows - the
    rasterizer reads nothing else). filtering: render()'s optional boolean mask over the Gaussians
    (gaussian_renderer/__init__.py:38-46) or an int64 index tensor; the outputs then have one row per selected Gaussian,
    and `conf_static` / `conf_index` are indexed by selected row, like the reference's `opacity * _conf_static[...]`."""
    if poses.dim() != 2 or poses.shape[1] != 7:
        raise ValueError("poses must be [V,7]: quaternion (r,x,y,z) and translation per view")
    if not xyz.is_cuda:
        raise RuntimeError("xyz must be a CUDA tensor: the front-end has no CPU path")
    fi = None
    if filtering is not None:
        fi = filtering.nonzero().flatten() if filtering.dtype == torch.bool else filtering.to(torch.int64)
        fi = fi.contiguous()
        if fi.numel() == xyz.shape[0] and filtering.dtype == torch.bool:
            fi = None  # all-true mask (render()'s default): identity
    out = _Frontend.apply(xyz, rotation, opacity, scaling, features_dc, features_rest, conf_static, poses, conf_index,
                          sh_degree, fi)
    V = poses.shape[0]
    return RenderInputs(means3D=list(out[3:3 + V]), rotations=list(out[3 + V:3 + 2 * V]), opacity=out[0], scales=out[1],
                        shs=out[2])
