"""b200splat — B200-native (sm_100a) differentiable Gaussian-splatting rasterizer.

Drop-in behind the reference's Python rasterizer API:
  neuralgaussiansplatting_b200.diff_gaussian_rasterization  (GaussianRasterizationSettings, GaussianRasterizer)
  neuralgaussiansplatting_b200.rasterizer2                  (GaussianRasterizationSettings2, GaussianRasterizer2)
  neuralgaussiansplatting_b200.renderer                     (render, render_idxmap, render_batch)
All compute is hand-written CUDA in libb200splat.so behind the C ABI of include/b200splat.h.
"""
from . import build as build_tools  # noqa: F401

__version__ = "0.1.0"


def library_path():
    from . import _lib
    return _lib.lib_path()
