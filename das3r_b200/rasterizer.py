"""The Python surface of the rasterizer, call-compatible with the reference package
submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py:

  GaussianRasterizationSettings   (:157-169)  per-call camera/config bundle (NamedTuple, same field order)
  GaussianRasterizer              (:171-220)  nn.Module with .forward(...) -> (color, radii) and .markVisible
  _RasterizeGaussians             (:44-155)   autograd.Function; backward returns the 9 gradients in input order
  rasterize_gaussians             (:21-42)    functional entry

Additive surface (BASELINE.json north_star "color/radii/depth/alpha"): `GaussianRasterizer.forward(...,
return_depth_alpha=True)` returns `(color, radii, depth, alpha)`; the default stays the reference's 2-tuple so DAS3R's
five call sites (gaussian_renderer/__init__.py:131,:259,:390,:499) keep working unchanged.
"""
from __future__ import annotations

from typing import NamedTuple

import torch
import torch.nn as nn

from . import load_native

_C = load_native()


def cpu_deep_copy_tuple(input_tuple):
    """Debug helper (reference :17-19): host copies of every tensor argument, taken before the call can corrupt them."""
    return tuple(item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple)


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def _forward_args(raster_settings, means3D, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, sh):
    # argument order of _C.rasterize_gaussians (reference :60-80, rasterize_points.h:18-38)
    return (
        raster_settings.bg,
        means3D,
        colors_precomp,
        opacities,
        scales,
        rotations,
        raster_settings.scale_modifier,
        cov3Ds_precomp,
        raster_settings.viewmatrix,
        raster_settings.projmatrix,
        raster_settings.tanfovx,
        raster_settings.tanfovy,
        raster_settings.image_height,
        raster_settings.image_width,
        sh,
        raster_settings.sh_degree,
        raster_settings.campos,
        raster_settings.prefiltered,
        raster_settings.debug,
    )


def _call_with_snapshot(fn, args, debug, dump_name, what):
    """Reference debug behaviour (:83-92, :132-141): on failure in debug mode dump the CPU copies of the arguments."""
    if not debug:
        return fn(*args)
    cpu_args = cpu_deep_copy_tuple(args)
    try:
        return fn(*args)
    except Exception as ex:
        torch.save(cpu_args, dump_name)
        print(f"\nAn error occured in {what}. Please forward {dump_name} for debugging.")
        raise ex


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings, *rest):
        # `rest` is () for the reference's 9-argument call and (extras,) for the additive depth/alpha surface
        extras = bool(rest[0]) if rest else False
        ctx.n_rest = len(rest)
        args = _forward_args(raster_settings, means3D, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, sh)
        if extras:
            num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer, depth, alpha = _call_with_snapshot(
                _C.rasterize_gaussians_ex, args, raster_settings.debug, "snapshot_fw.dump", "forward")
        else:
            num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer = _call_with_snapshot(
                _C.rasterize_gaussians, args, raster_settings.debug, "snapshot_fw.dump", "forward")
        ctx.raster_settings = raster_settings
        ctx.num_rendered = num_rendered
        ctx.extras = extras
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer,
                              binningBuffer, imgBuffer)
        if extras:
            ctx.mark_non_differentiable(radii, depth, alpha)
            return color, radii, depth, alpha
        return color, radii

    @staticmethod
    def backward(ctx, grad_out_color, *_unused):
        num_rendered = ctx.num_rendered
        raster_settings = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer, binningBuffer,
         imgBuffer) = ctx.saved_tensors
        # argument order of _C.rasterize_gaussians_backward (reference :109-129, rasterize_points.h:40-62)
        args = (
            raster_settings.bg,
            means3D,
            radii,
            colors_precomp,
            scales,
            rotations,
            raster_settings.scale_modifier,
            cov3Ds_precomp,
            raster_settings.viewmatrix,
            raster_settings.projmatrix,
            raster_settings.tanfovx,
            raster_settings.tanfovy,
            grad_out_color,
            sh,
            raster_settings.sh_degree,
            raster_settings.campos,
            geomBuffer,
            num_rendered,
            binningBuffer,
            imgBuffer,
            raster_settings.debug,
        )
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh, grad_scales,
         grad_rotations) = _call_with_snapshot(_C.rasterize_gaussians_backward, args, raster_settings.debug,
                                               "snapshot_bw.dump", "backward")
        # one gradient per forward input, in input order (reference :143-155); settings (and extras) get None
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities, grad_scales, grad_rotations,
                grad_cov3Ds_precomp, None) + (None,) * ctx.n_rest


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings, extras=False):
    if extras:
        return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                         cov3Ds_precomp, raster_settings, True)
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """Boolean mask of points in front of the near plane (reference :176-185; DAS3R near plane 0.001)."""
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None, return_depth_alpha=False):
        rs = self.raster_settings
        # same validation and messages as the reference (:191-195)
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (
                (scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        # absent inputs travel as empty CPU tensors -> NULL pointers below the boundary (:197-207)
        if shs is None:
            shs = torch.Tensor([])
        if colors_precomp is None:
            colors_precomp = torch.Tensor([])
        if scales is None:
            scales = torch.Tensor([])
        if rotations is None:
            rotations = torch.Tensor([])
        if cov3D_precomp is None:
            cov3D_precomp = torch.Tensor([])
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                                   rs, extras=bool(return_depth_alpha))
