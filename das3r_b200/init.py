"""Pointmap initialisation: what `GaussianModel.create_from_cameras` (reference scene/gaussian_model.py:573-659) computes
from the stacked per-view maps - unprojection of every pixel (`depth_to_pts3d`, utils/pose_utils.py:572-583), the
confidence mask, the compaction, RGB2SH colours, `distCUDA2` scales, identity rotations, `inverse_sigmoid(1/V)` opacities
and `_conf_static = 1 - dyna_avg` - as two passes over the pixels plus the 3-NN search (include/das3r_init.h,
das3r_b200/csrc/init.cu).

    init = create_from_pointmaps(K, poses, depth, confidence, dyna_avg, rgb, conf_thre=1.0)
    init.xyz, init.features_dc, init.features_rest, init.scaling, init.rotation, init.opacity   # the nn.Parameters
    init.conf_static [V,H,W], init.aggregated_mask [V*H*W] bool, init.conf_index [P] int64

`conf_index` (= aggregated_mask.nonzero()) is what `das3r_b200.frontend.render_inputs` takes to look up a Gaussian's
static confidence. No CPU path.
"""
from __future__ import annotations

import math
from typing import NamedTuple

import torch

from . import load_native

_C = load_native()


class PointmapInit(NamedTuple):
    xyz: torch.Tensor              # [P,3]
    features_dc: torch.Tensor      # [P,1,3]
    features_rest: torch.Tensor    # [P,15,3]
    scaling: torch.Tensor          # [P,3]  log scales
    rotation: torch.Tensor         # [P,4]
    opacity: torch.Tensor          # [P,1]  pre-sigmoid
    conf_static: torch.Tensor      # [V,H,W]
    aggregated_mask: torch.Tensor  # [V*H*W] bool
    conf_index: torch.Tensor       # [P] int64
    dist2: torch.Tensor            # [P] mean squared distance to the 3 nearest neighbours (before the 1e-7 clamp)


def create_from_pointmaps(K: torch.Tensor, poses: torch.Tensor, depth: torch.Tensor, confidence: torch.Tensor,
                          dyna_avg: torch.Tensor, rgb: torch.Tensor, conf_thre: float = 1.0) -> PointmapInit:
    """K [V,3,3] (fx == fy as the reference asserts), poses [V,4,4] camera-to-world, depth / confidence / dyna_avg
    [V,H,W], rgb [V,3,H,W]; all float32 CUDA tensors."""
    if not depth.is_cuda:
        raise RuntimeError("depth must be a CUDA tensor: the initialisation has no CPU path")
    if not bool((K[:, 0, 0] == K[:, 1, 1]).all()):
        raise AssertionError("fx != fy")  # utils/pose_utils.py:574
    log_thre = float(torch.tensor(conf_thre).log())  # as the reference forms it (gaussian_model.py:624), in fp32
    if math.isnan(log_thre):
        raise ValueError("conf_thre must be positive")
    return PointmapInit(*_C.pointmap_init(K, poses, depth, confidence, dyna_avg, rgb, log_thre))
