"""Generates tests/golden/init_*.npz: outputs of the REFERENCE's own pointmap-initialisation code, run where
/root/reference exists (CPU, fp32). utils/pose_utils.py (depth_to_pts3d), utils/sh_utils.py (RGB2SH) and
utils/general_utils.py (inverse_sigmoid) are imported from the reference as they are; the lines of
GaussianModel.create_from_cameras that glue them together (scene/gaussian_model.py:614-659 - the class cannot be
imported here: it needs plyfile and the CUDA simple_knn) are restated below with the same torch calls. distCUDA2 is
CUDA-only: the exact 3-NN mean squared distance comes from a k-d tree (das3r_b200.synth.knn_dist2_cpu, whose semantics
the GPU tests pin to the reference's simple_knn). Run: python tests/golden/make_init_golden.py"""
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
sys.path.insert(0, REF)
from utils.general_utils import inverse_sigmoid  # noqa: E402
from utils.pose_utils import depth_to_pts3d  # noqa: E402
from utils.sh_utils import RGB2SH  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from das3r_b200.synth import knn_dist2_cpu  # noqa: E402
from oracle.init_oracle import synth_pointmaps  # noqa: E402


def make(name, V, H, W, seed, conf_thre):
    m = {k: torch.from_numpy(v) for k, v in synth_pointmaps(V, H, W, seed).items()}
    K, poses, depth_maps, confidences, dyna_avg, rgbs = m["K"], m["poses"], m["depth"], m["confidence"], m["dyna_avg"], m["rgb"]
    p3d = depth_to_pts3d(K, poses, W, H, depth_maps).float()                       # :614
    p3d_color = rgbs.permute(0, 2, 3, 1).reshape(-1, 3)                            # :615
    pts_4_3dgs = p3d.reshape(-1, 3)
    conf_static = 1 - dyna_avg                                                     # :620
    confidence = confidences.reshape(-1)
    confidence_masks = confidence > torch.tensor(conf_thre).log()                  # :624
    pts_4_3dgs = pts_4_3dgs[confidence_masks]                                      # :629
    color_4_3dgs = p3d_color.reshape(-1, 3)[confidence_masks]
    fused_color = RGB2SH(color_4_3dgs)                                             # :634
    features = torch.zeros((fused_color.shape[0], 3, 16)).float()
    features[:, :3, 0] = fused_color
    dist2_raw = knn_dist2_cpu(pts_4_3dgs)
    dist2 = torch.clamp_min(dist2_raw, 0.0000001)                                  # :641
    scales = torch.log(torch.sqrt(dist2))[..., None].repeat(1, 3)
    rots = torch.zeros((pts_4_3dgs.shape[0], 4))
    rots[:, 0] = 1
    opa = 1 / V
    opacities = inverse_sigmoid(opa * torch.ones((pts_4_3dgs.shape[0], 1), dtype=torch.float))
    save = {f"in_{k}": v.numpy() for k, v in m.items()}
    save["in_conf_thre"] = np.float32(conf_thre)
    save.update(out_xyz=pts_4_3dgs.numpy(), out_features_dc=features[:, :, 0:1].transpose(1, 2).contiguous().numpy(),
                out_features_rest=features[:, :, 1:].transpose(1, 2).contiguous().numpy(), out_scaling=scales.numpy(),
                out_rotation=rots.numpy(), out_opacity=opacities.numpy(), out_conf_static=conf_static.numpy(),
                out_aggregated_mask=confidence_masks.numpy(), out_dist2=dist2_raw.numpy())
    np.savez_compressed(os.path.join(HERE, f"init_{name}.npz"), **save)
    print(name, "P =", pts_4_3dgs.shape[0], "of", confidence.numel())


if __name__ == "__main__":
    make("small", V=3, H=24, W=40, seed=1, conf_thre=1.0)
    make("thre3", V=5, H=18, W=32, seed=2, conf_thre=3.0)
    make("one_view", V=1, H=30, W=30, seed=3, conf_thre=1.0)
