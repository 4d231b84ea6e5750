"""Host logic of the drop-in Python surface (no GPU): same names, fields, validation and messages as the reference
package (diff_gaussian_rasterization/__init__.py:157-220), and loud failure instead of a CPU fallback."""
import pytest
import torch


def test_dropin_module_names():
    import diff_gaussian_rasterization as dgr
    from simple_knn._C import distCUDA2  # noqa: F401

    for n in ("GaussianRasterizationSettings", "GaussianRasterizer", "_RasterizeGaussians", "rasterize_gaussians",
              "cpu_deep_copy_tuple"):
        assert hasattr(dgr, n)
    for n in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible"):
        assert hasattr(dgr._C, n)


def test_settings_fields_match_the_reference_order():
    from diff_gaussian_rasterization import GaussianRasterizationSettings as S

    assert S._fields == ("image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix",
                         "projmatrix", "sh_degree", "campos", "prefiltered", "debug")


def _settings():
    from diff_gaussian_rasterization import GaussianRasterizationSettings as S

    return S(image_height=32, image_width=32, tanfovx=0.5, tanfovy=0.5, bg=torch.zeros(3), scale_modifier=1.0,
             viewmatrix=torch.eye(4), projmatrix=torch.eye(4), sh_degree=0, campos=torch.zeros(3), prefiltered=False,
             debug=False)


def test_input_validation_messages():
    from diff_gaussian_rasterization import GaussianRasterizer

    r = GaussianRasterizer(_settings())
    m = torch.zeros(4, 3)
    with pytest.raises(Exception, match="Please provide excatly one of either SHs or precomputed colors!"):
        r(means3D=m, means2D=m, opacities=torch.ones(4, 1), scales=torch.ones(4, 3), rotations=torch.ones(4, 4))
    with pytest.raises(Exception, match="Please provide excatly one of either SHs or precomputed colors!"):
        r(means3D=m, means2D=m, opacities=torch.ones(4, 1), shs=torch.zeros(4, 16, 3), colors_precomp=torch.zeros(4, 3),
          scales=torch.ones(4, 3), rotations=torch.ones(4, 4))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair or precomputed 3D covariance"):
        r(means3D=m, means2D=m, opacities=torch.ones(4, 1), shs=torch.zeros(4, 16, 3))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair or precomputed 3D covariance"):
        r(means3D=m, means2D=m, opacities=torch.ones(4, 1), shs=torch.zeros(4, 16, 3), scales=torch.ones(4, 3),
          rotations=torch.ones(4, 4), cov3D_precomp=torch.zeros(4, 6))


def test_no_cpu_fallback():
    """CPU tensors must raise: the product path never falls back to the oracle or to PyTorch."""
    from diff_gaussian_rasterization import GaussianRasterizer
    from simple_knn._C import distCUDA2

    r = GaussianRasterizer(_settings())
    m = torch.zeros(4, 3)
    with pytest.raises(RuntimeError, match="no CPU path"):
        r(means3D=m, means2D=m, opacities=torch.ones(4, 1), shs=torch.zeros(4, 16, 3), scales=torch.ones(4, 3),
          rotations=torch.ones(4, 4))
    with pytest.raises(RuntimeError, match="no CPU path"):
        distCUDA2(torch.zeros(8, 3))
    with pytest.raises(RuntimeError, match="no CPU path"):
        r.markVisible(m)
    with pytest.raises(Exception, match="means3D must have dimensions"):
        r(means3D=torch.zeros(4, 2), means2D=m, opacities=torch.ones(4, 1), shs=torch.zeros(4, 16, 3),
          scales=torch.ones(4, 3), rotations=torch.ones(4, 4))


def test_product_never_imports_the_oracle():
    import os
    import re

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for d in ("das3r_b200", "diff_gaussian_rasterization", "simple_knn"):
        for dirpath, _, files in os.walk(os.path.join(root, d)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                    text = open(os.path.join(dirpath, f)).read()
                    assert not re.search(r"^\s*(import|from)\s+oracle\b", text, flags=re.M), os.path.join(dirpath, f)
                    assert "liboracle" not in text and "raster_oracle" not in text, os.path.join(dirpath, f)
