"""Drop-in package name: `from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer`
(DAS3R gaussian_renderer/__init__.py:14-17) resolves to the B200-native implementation in das3r_b200."""
from das3r_b200.rasterizer import (  # noqa: F401
    GaussianRasterizationSettings,
    GaussianRasterizer,
    _RasterizeGaussians,
    cpu_deep_copy_tuple,
    rasterize_gaussians,
)
from . import _C  # noqa: F401
