"""`diff_gaussian_rasterization._C` — same three functions as the reference pybind module (ext.cpp:15-18)."""
from das3r_b200 import load_native as _load

_n = _load()
rasterize_gaussians = _n.rasterize_gaussians
rasterize_gaussians_backward = _n.rasterize_gaussians_backward
mark_visible = _n.mark_visible
rasterize_gaussians_ex = _n.rasterize_gaussians_ex  # additive: also returns depth and alpha
