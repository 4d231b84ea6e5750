"""distCUDA2 — mean squared distance to the 3 nearest other points (reference simple-knn/spatial.cu:15-25)."""
from __future__ import annotations

from . import load_native

_C = load_native()


def distCUDA2(points):
    return _C.distCUDA2(points)
