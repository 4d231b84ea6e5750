"""DAS3R's training loop on tensors (reference train_gui.py:499-641 without the GUI, the dataset readers and the ply
writer): the per-iteration semantics are the reference's -

  * `update_learning_rate(iteration)`: log-linear decay of the xyz, pose and static-confidence learning rates
    (scene/gaussian_model.py:301-324, utils/general_utils.py:29-62),
  * `oneupSHdegree()` every 3000 iterations (train_gui.py:541-543, gaussian_model.py:199-201): degree 0 for iterations
    1-2999, 1 afterwards in the 4000-iteration runs of scripts/testing_psnr_sintel.sh:46,
  * one random view per iteration from a stack that is refilled when empty (`viewpoint_stack.pop(randint(...))`,
    :545-555),
  * render with the view's optimisable pose (:556-559), masked L1 + SSIM loss and the frame's PSNR (:565-576),
  * `optimizer.step()` for the seven Gaussian groups, and `optimizer_cam.step()` ONLY when the frame's PSNR exceeds
    `psnr_threshold` = 26 (:580-586); both are zeroed every iteration,

while every heavy step is one of this repository's fused components: `render_inputs` (front-end), the rasterizer,
`photometric_loss`, `FusedAdam`. The three hooks (`render_fn`, `loss_fn`, `adam_cls`) exist so that the tests can run the
SAME loop with the reference's formulation (torch glue + the reference CUDA rasterizer + torch loss + torch.optim.Adam)
and compare loss curves and parameters. No CPU path.
"""
from __future__ import annotations

import math
import random
from dataclasses import dataclass, field
from typing import Callable, Sequence

import numpy as np
import torch

from .frontend import render_inputs
from .loss import photometric_loss
from .optim import FusedAdam
from .rasterizer import GaussianRasterizationSettings, GaussianRasterizer


@dataclass
class OptimizationParams:
    """arguments/__init__.py:71-90 (defaults) with the values DAS3R's scripts override (--iter 4000)."""
    iterations: int = 4000
    position_lr_init: float = 0.00016
    position_lr_final: float = 0.0000016
    position_lr_delay_mult: float = 0.01
    position_lr_max_steps: int = 30_000
    feature_lr: float = 0.0025
    opacity_lr: float = 0.05
    scaling_lr: float = 0.005
    rotation_lr: float = 0.001
    lambda_dssim: float = 0.2
    conf_lr_init: float = 3e-3       # gaussian_model.py:233-234
    conf_lr_final: float = 3e-4
    cam_lr_init: float = 0.00003     # :246-249
    cam_lr_final: float = 0.000003
    psnr_threshold: float = 26.0     # train_gui.py:734
    sh_increase_interval: int = 3000  # train_gui.py:542
    spatial_lr_scale: float = 1.0
    optim_pose: bool = True


def get_expon_lr_func(lr_init, lr_final, lr_delay_steps=0, lr_delay_mult=1.0, max_steps=1000000) -> Callable[[int], float]:
    """utils/general_utils.py:29-62."""

    def helper(step):
        if step < 0 or (lr_init == 0.0 and lr_final == 0.0):
            return 0.0
        if lr_delay_steps > 0:
            delay_rate = lr_delay_mult + (1 - lr_delay_mult) * np.sin(0.5 * np.pi * np.clip(step / lr_delay_steps, 0, 1))
        else:
            delay_rate = 1.0
        t = np.clip(step / max_steps, 0, 1)
        log_lerp = np.exp(np.log(lr_init) * (1 - t) + np.log(lr_final) * t)
        return delay_rate * log_lerp

    return helper


@dataclass
class Camera:
    """What the loop needs of scene/cameras.py:Camera: the ground-truth frame and the view's index into the per-view
    parameters (Q, T, _conf_static)."""
    uid: int
    original_image: torch.Tensor  # [3,H,W]


@dataclass
class GaussianState:
    """The optimisable tensors of scene/gaussian_model.py:GaussianModel (leaf tensors with requires_grad)."""
    xyz: torch.Tensor            # [P,3]
    features_dc: torch.Tensor    # [P,1,3]
    features_rest: torch.Tensor  # [P,15,3]
    opacity: torch.Tensor        # [P,1] pre-sigmoid
    scaling: torch.Tensor        # [P,3] log
    rotation: torch.Tensor       # [P,4]
    conf_static: torch.Tensor    # [V,H,W] (or [P] after load_ply)
    Q: torch.Tensor              # [V,4] pose quaternions (r,x,y,z)
    T: torch.Tensor              # [V,3] pose translations
    conf_index: torch.Tensor | None = None  # [P] int64 = aggregated_mask.nonzero(); None: conf_static is per Gaussian
    active_sh_degree: int = 0
    max_sh_degree: int = 3

    def parameters(self):
        return dict(xyz=self.xyz, f_dc=self.features_dc, f_rest=self.features_rest, opacity=self.opacity,
                    scaling=self.scaling, rotation=self.rotation, conf_static=self.conf_static)

    def oneupSHdegree(self):  # gaussian_model.py:199-201
        if self.active_sh_degree < self.max_sh_degree:
            self.active_sh_degree += 1

    def get_RT(self, idx):  # :170-174
        return torch.cat((self.Q[idx], self.T[idx]), dim=0)


@dataclass
class ViewGeometry:
    """Per-run camera constants of render() (gaussian_renderer/__init__.py:52-77): identity view matrix, the camera's
    projection matrix, camera position at the origin."""
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    projmatrix: torch.Tensor  # [4,4] as the rasterizer takes it (transposed, row-vector convention)
    bg: torch.Tensor          # [3]
    viewmatrix: torch.Tensor = field(default=None)
    campos: torch.Tensor = field(default=None)

    def __post_init__(self):
        dev = self.projmatrix.device
        if self.viewmatrix is None:
            self.viewmatrix = torch.eye(4, device=dev)
        if self.campos is None:
            self.campos = torch.zeros(3, device=dev)

    def settings(self, sh_degree: int, debug: bool = False) -> GaussianRasterizationSettings:
        return GaussianRasterizationSettings(
            image_height=int(self.image_height), image_width=int(self.image_width), tanfovx=self.tanfovx,
            tanfovy=self.tanfovy, bg=self.bg, scale_modifier=1.0, viewmatrix=self.viewmatrix, projmatrix=self.projmatrix,
            sh_degree=sh_degree, campos=self.campos, prefiltered=False, debug=debug)


def render_view(state: GaussianState, geom: ViewGeometry, pose: torch.Tensor, filtering=None) -> torch.Tensor:
    """gaussian_renderer/__init__.py:23-149 (`render`) for one view: fused front-end + rasterizer. -> image [3,H,W]."""
    fe = render_inputs(state.xyz, state.rotation, state.opacity, state.scaling, state.features_dc, state.features_rest,
                       state.conf_static.reshape(-1), state.conf_index, pose.reshape(1, 7),
                       sh_degree=state.active_sh_degree, filtering=filtering)
    P = fe.opacity.shape[0]
    means2D = torch.zeros(P, 3, device=fe.opacity.device, requires_grad=fe.opacity.requires_grad)
    color, _radii = GaussianRasterizer(geom.settings(state.active_sh_degree))(
        means3D=fe.means3D[0], means2D=means2D, opacities=fe.opacity, shs=fe.shs, scales=fe.scales,
        rotations=fe.rotations[0])
    return color


def masked_loss(image, gt, static, lambda_dssim):
    """train_gui.py:565-576: -> (loss, psnr_frame)."""
    return photometric_loss(image, gt, static, lambda_dssim)


class Trainer:
    """trainer = Trainer(state, cameras, geom, opt, seed); trainer.train(n) or trainer.step() -> loss (float).

    `history` collects (iteration, view uid, loss, psnr, pose step taken) per iteration."""

    def __init__(self, state: GaussianState, cameras: Sequence[Camera], geom: ViewGeometry,
                 opt: OptimizationParams | None = None, seed: int = 0, render_fn=render_view, loss_fn=masked_loss,
                 adam_cls=FusedAdam):
        self.state, self.cameras, self.geom = state, list(cameras), geom
        self.opt = opt or OptimizationParams()
        self.render_fn, self.loss_fn = render_fn, loss_fn
        self.rng = random.Random(seed)
        self.iteration = 0
        self.viewpoint_stack: list[Camera] = []
        self.history: list[tuple] = []
        o = self.opt
        p = state.parameters()
        for t in list(p.values()) + [state.Q, state.T]:
            if not t.is_cuda:
                raise RuntimeError("the trainer needs CUDA tensors: there is no CPU path")
            t.requires_grad_(True)
        # scene/gaussian_model.py:236-244
        groups = [
            {"params": [p["xyz"]], "lr": o.position_lr_init * o.spatial_lr_scale, "name": "xyz"},
            {"params": [p["f_dc"]], "lr": o.feature_lr, "name": "f_dc"},
            {"params": [p["f_rest"]], "lr": o.feature_lr / 20.0, "name": "f_rest"},
            {"params": [p["opacity"]], "lr": o.opacity_lr, "name": "opacity"},
            {"params": [p["scaling"]], "lr": o.scaling_lr, "name": "scaling"},
            {"params": [p["rotation"]], "lr": o.rotation_lr, "name": "rotation"},
            {"params": [p["conf_static"]], "lr": o.conf_lr_init, "name": "conf_static"},
        ]
        cam_groups = [{"params": [state.Q], "lr": o.cam_lr_init, "name": "pose_Q"},
                      {"params": [state.T], "lr": o.cam_lr_init, "name": "pose_T"}]
        self.optimizer = adam_cls(groups, lr=0.0, eps=1e-15)          # :258
        self.optimizer_cam = adam_cls(cam_groups, lr=0.0, eps=1e-15)  # :260
        self.xyz_sched = get_expon_lr_func(o.position_lr_init * o.spatial_lr_scale, o.position_lr_final * o.spatial_lr_scale,
                                           lr_delay_mult=o.position_lr_delay_mult, max_steps=o.position_lr_max_steps)
        self.cam_sched = get_expon_lr_func(o.cam_lr_init, o.cam_lr_final, lr_delay_mult=o.position_lr_delay_mult,
                                           max_steps=1000)
        self.conf_sched = get_expon_lr_func(o.conf_lr_init, o.conf_lr_final, lr_delay_mult=o.position_lr_delay_mult,
                                            max_steps=o.iterations)

    def update_learning_rate(self, iteration: int) -> None:  # gaussian_model.py:301-324
        for g in self.optimizer_cam.param_groups:
            if g["name"] in ("pose_Q", "pose_T"):
                g["lr"] = self.cam_sched(iteration)
        for g in self.optimizer.param_groups:
            if g["name"] == "xyz":
                g["lr"] = self.xyz_sched(iteration)
            if g["name"] == "conf_static":
                g["lr"] = self.conf_sched(iteration)

    def step(self) -> float:
        self.iteration += 1
        it, o, st = self.iteration, self.opt, self.state
        self.update_learning_rate(it)
        if not o.optim_pose:
            st.Q.requires_grad_(False)
            st.T.requires_grad_(False)
        if it % o.sh_increase_interval == 0:
            st.oneupSHdegree()
        if not self.viewpoint_stack:
            self.viewpoint_stack = self.cameras.copy()
        cam = self.viewpoint_stack.pop(self.rng.randint(0, len(self.viewpoint_stack) - 1))
        pose = st.get_RT(cam.uid)
        image = self.render_fn(st, self.geom, pose)
        static = st.conf_static[cam.uid] if st.conf_index is not None else None
        loss, psnr_frame = self.loss_fn(image, cam.original_image, static, o.lambda_dssim)
        loss.backward()
        with torch.no_grad():
            self.optimizer.step()
            self.optimizer.zero_grad(set_to_none=True)
            took_pose_step = bool(psnr_frame > o.psnr_threshold)  # train_gui.py:584 (a host read, as `loss.item()` is)
            if took_pose_step and o.optim_pose:
                self.optimizer_cam.step()
            self.optimizer_cam.zero_grad(set_to_none=True)
        lv = float(loss.item())
        self.history.append((it, cam.uid, lv, float(psnr_frame), took_pose_step))
        return lv

    def train(self, iterations: int | None = None) -> list[tuple]:
        n = self.opt.iterations if iterations is None else iterations
        for _ in range(n):
            self.step()
        return self.history


def state_from_init(init, V: int, H: int, W: int, poses_w2c: torch.Tensor, max_sh_degree: int = 3) -> GaussianState:
    """GaussianState from `das3r_b200.init.create_from_pointmaps` (what Scene.__init__ -> create_from_cameras +
    init_RT_seq do): poses_w2c [V,7] = (quaternion, translation) of every view's world-to-camera transform."""
    return GaussianState(xyz=init.xyz.clone(), features_dc=init.features_dc.clone(), features_rest=init.features_rest.clone(),
                         opacity=init.opacity.clone(), scaling=init.scaling.clone(), rotation=init.rotation.clone(),
                         conf_static=init.conf_static.clone().view(V, H, W), Q=poses_w2c[:, :4].clone().contiguous(),
                         T=poses_w2c[:, 4:].clone().contiguous(), conf_index=init.conf_index, active_sh_degree=0,
                         max_sh_degree=max_sh_degree)
