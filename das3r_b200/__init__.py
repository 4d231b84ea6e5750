"""das3r_b200 — B200-native (sm_100a) drop-in for DAS3R's differentiable Gaussian-splat rasterizer
(submodules/diff-gaussian-rasterization) and 3-NN initialiser (submodules/simple-knn).

Layout:
  csrc/            hand-written CUDA kernels + the C ABI (include/das3r_raster.h, include/das3r_frontend.h, include/das3r_loss.h, include/das3r_adam.h) + the thin
                   torch binding
  rasterizer.py    the Python surface DAS3R imports (GaussianRasterizationSettings, GaussianRasterizer, ...)
  multiview.py     MultiViewRasterizer: all views of a step in one autograd node (streams + fused gradient pass)
  frontend.py      render_inputs: DAS3R's render() front-end (pose transform, activations, SH concat) as one fused op
  loss.py          photometric_loss: the masked L1 + SSIM training loss (and PSNR) as one fused op
  optim.py         FusedAdam: the Adam step of all parameter groups in one launch
  knn.py           distCUDA2
  synth.py         synthetic DAS3R-shaped scenes for tests and bench (no datasets on the box)
  sharding.py      view/frame sharding + the one gradient all-reduce of the multi-GPU plan

There is no CPU path and no fallback: importing the native extension fails loudly if it has not been built
(`python -m das3r_b200.build`), and every op raises on non-CUDA tensors.
"""
from __future__ import annotations

__version__ = "0.1.0"


def load_native():
    """Import and return the torch extension `_C`; raise a clear error when it is missing."""
    try:
        from . import _C  # noqa: F401
    except ImportError as ex:  # pragma: no cover - exercised only on a broken build
        raise ImportError(
            "das3r_b200._C (the sm_100a CUDA extension) is not built or failed to load: "
            f"{ex}. Build it in-tree with `python -m das3r_b200.build`; there is no CPU or PyTorch fallback."
        ) from ex
    return _C
