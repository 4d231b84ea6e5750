"""Several views of one set of Gaussians in one autograd node - the shape of a data-parallel DAS3R step.

DAS3R renders one view per iteration (train_gui.py:555-559) and poses the Gaussians for that view itself
(gaussian_renderer/__init__.py:83-92): every view gets its own `means3D` and `rotations`, while opacities, SH
coefficients and scales are shared. A step that renders n views (this repository's multi-GPU plan: each rank takes n
views of the step) therefore runs every stage ONCE for all views on the caller's stream:

  * `das3r_raster_forward_preprocess`: one kernel projects all views (the shared SH rows are read once),
  * `das3r_raster_forward_render_batch`: one host read-back for all views, then binning, sort, rect lists and the
    blend as one launch per stage with grid.y = view - multi-wave grids that fill the 148 SMs, instead of n small ones,
  * `das3r_raster_backward_blend_batch`: likewise for the backward blend,
  * `das3r_raster_backward_finish`: turns the n screen-space gradient buffers into Gaussian gradients in ONE pass that
    reads the shared SH rows once and writes dL_dsh / dL_dscales / dL_dopacity once, summed over the views, plus every
    view's own dL_dmeans3D / dL_drotations.

No host threads and no side streams are involved. This has no counterpart in the reference package; `GaussianRasterizer`
(rasterizer.py) remains the drop-in per-view surface and gives the same numbers (tests/test_gpu_multiview.py). Only the
input variant DAS3R trains with is supported here: SH colours + scale/rotation; use the per-view surface for anything
else.
"""
from __future__ import annotations

from typing import Sequence

import torch

from . import load_native
from .rasterizer import GaussianRasterizationSettings

_C = load_native()
_EMPTY = torch.Tensor([])


class _RasterizeViews(torch.autograd.Function):
    @staticmethod
    def forward(ctx, settings, opacities, shs, scales, *mr):
        n = len(settings)
        means = mr[0::2]
        rots = mr[1::2]
        rs0 = settings[0]
        pre = _C.rasterize_gaussians_forward_preprocess(
            [(rs.bg, means[i], rots[i], rs.viewmatrix, rs.projmatrix, float(rs.tanfovx), float(rs.tanfovy),
              int(rs.image_height), int(rs.image_width), rs.campos) for i, rs in enumerate(settings)],
            opacities, scales, shs, rs0.sh_degree, float(rs0.scale_modifier), bool(rs0.prefiltered), bool(rs0.debug))
        res = _C.rasterize_gaussians_forward_render_batch(
            [(settings[i].bg, pre[i][0], pre[i][1], pre[i][2], pre[i][3]) for i in range(n)], bool(rs0.debug))
        ctx.settings, ctx.n = settings, n
        ctx.num_rendered = [r[0] for r in res]
        colors = [p[0] for p in pre]
        radii = [p[1] for p in pre]
        saved = [opacities, shs, scales, *mr]
        for p, r in zip(pre, res):
            saved += [p[1], p[2], r[1], p[3]]  # radii, geom, binning, img
        ctx.save_for_backward(*saved)
        ctx.mark_non_differentiable(*radii)
        return (*colors, *radii)

    @staticmethod
    def backward(ctx, *grads):
        n, settings = ctx.n, ctx.settings
        sv = ctx.saved_tensors
        opacities, shs, scales = sv[0], sv[1], sv[2]
        mr = sv[3:3 + 2 * n]
        st = sv[3 + 2 * n:]
        dev = opacities.device
        P = opacities.shape[0]
        M = shs.shape[1]
        bviews = []
        for i, rs in enumerate(settings):
            radii, geom, binning, img = st[4 * i:4 * i + 4]
            g = grads[i]
            if g is None:
                g = torch.zeros(3, rs.image_height, rs.image_width, device=dev)
            bviews.append((rs.bg, g.contiguous(), geom, ctx.num_rendered[i], binning, img))
        accs = _C.rasterize_gaussians_backward_blend_batch(bviews, P, bool(settings[0].debug))
        o = dict(device=dev, dtype=torch.float32)
        # every element is written by the finish half (fused pass or per-view fallback): no zero fills
        d_opac, d_shs, d_scales = torch.empty(P, 1, **o), torch.empty(P, M, 3, **o), torch.empty(P, 3, **o)
        d_means_sum, d_rots_sum = torch.empty(P, 3, **o), torch.empty(P, 4, **o)  # unused: every view has its own
        d_mr = []
        views = []
        for i, rs in enumerate(settings):
            radii, geom, binning, img = st[4 * i:4 * i + 4]
            dm, dr = torch.empty(P, 3, **o), torch.empty(P, 4, **o)
            d_mr += [dm, dr]
            views.append((mr[2 * i], radii, scales, mr[2 * i + 1], rs.viewmatrix, rs.projmatrix, float(rs.tanfovx),
                          float(rs.tanfovy), rs.campos, geom, accs[i], int(rs.image_height), int(rs.image_width), dm,
                          dr))
        _C.rasterize_gaussians_backward_finish(views, shs, settings[0].sh_degree, float(settings[0].scale_modifier),
                                               [d_opac, d_means_sum, d_shs, d_scales, d_rots_sum], False,
                                               bool(settings[0].debug))
        return (None, d_opac, d_shs, d_scales, *d_mr)


class MultiViewRasterizer:
    """`MultiViewRasterizer()(settings, means3D, rotations, opacities, shs, scales)` -> (colors, radii), two lists with
    one entry per view; differentiable w.r.t. opacities, shs, scales and every view's means3D / rotations.

    settings   sequence of GaussianRasterizationSettings, one per view (same sh_degree and scale_modifier)
    means3D    sequence of [P,3] tensors, the Gaussians as posed for each view
    rotations  sequence of [P,4] tensors, likewise
    opacities [P,1], shs [P,M,3], scales [P,3]   shared by all views
    """

    def __init__(self, streams: int = 1):
        # `streams` is accepted for compatibility with round-1 callers; every stage is now one batched launch on the
        # caller's stream
        self.n_streams = 1

    def __call__(self, settings: Sequence[GaussianRasterizationSettings], means3D: Sequence[torch.Tensor],
                 rotations: Sequence[torch.Tensor], opacities: torch.Tensor, shs: torch.Tensor, scales: torch.Tensor):
        n = len(settings)
        if n == 0 or len(means3D) != n or len(rotations) != n:
            raise ValueError("settings, means3D and rotations need one entry per view")
        if not opacities.is_cuda:
            raise RuntimeError("opacities must be a CUDA tensor: this rasterizer has no CPU path")
        if shs is None or shs.dim() != 3 or scales is None:
            raise Exception("MultiViewRasterizer takes SH colours and scale/rotation inputs; use GaussianRasterizer "
                            "for precomputed colours or covariances")
        if any(rs.sh_degree != settings[0].sh_degree or rs.scale_modifier != settings[0].scale_modifier or
               rs.prefiltered != settings[0].prefiltered or rs.debug != settings[0].debug for rs in settings):
            raise ValueError("all views must use the same sh_degree, scale_modifier, prefiltered and debug flags")
        mr = []
        for m, r in zip(means3D, rotations):
            mr += [m, r]
        out = _RasterizeViews.apply(tuple(settings), opacities, shs, scales, *mr)
        return list(out[:n]), list(out[n:])
