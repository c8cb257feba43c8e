"""Drop-in replacement for the reference package `rasterizer2`
(submodules/rasterizer2/rasterizer2/__init__.py): the "idxmap" extra channels.

  GaussianRasterizationSettings2   same 12 fields as the reference NamedTuple      (:140-152)
  GaussianRasterizer2(raster_settings)
      .forward(means3D, normal, means2D, feature_vector, shs=None, scales=None, rotations=None)
          -> (idxmap[H,W] i32 (-1 = no point), colmap[H,W,3], depthmap[H,W], featuremap[H,W,64])  (:155-179)
  rasterize_gaussians2(...)                                                         (:7-28)

Gradient contract copied from the reference (:74-138): only `feature_vector` receives a gradient
(channels 25..63 scatter-added from grad_featuremap); every other input gets zeros.
The reference's scatter is a racy `+=` (rasterizer_impl.cu:844-846); ours accumulates atomically.

rasterizer2 also exports the stripped alpha-compositing pair at C level (ext.cpp:16-17, no Python
wrapper): `_C.rasterize_gaussians` / `_C.rasterize_gaussians_backward` with the precomp arguments
removed. `alpha_forward` / `alpha_backward` below expose exactly those 17-/19-argument calls.
"""
from __future__ import annotations

from typing import NamedTuple

import torch
import torch.nn as nn

from .. import _C

__all__ = ["GaussianRasterizationSettings2", "GaussianRasterizer2", "rasterize_gaussians2",
           "alpha_forward", "alpha_backward"]


class GaussianRasterizationSettings2(NamedTuple):
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


class _GetIdxMap2(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, normal, means2D, feature_vector, sh, scales, rotations, raster_settings):
        s = raster_settings
        num_rendered, idxmap, colmap, depthmap, featuremap = _C.get_idxmap2(
            s.bg, means3D, normal, scales, rotations, s.viewmatrix, s.projmatrix, s.tanfovx, s.tanfovy,
            s.image_height, s.image_width, feature_vector, sh, s.campos, s.debug)
        ctx.raster_settings = s
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(means3D, normal, scales, rotations, feature_vector, sh, idxmap)
        ctx.mark_non_differentiable(idxmap)
        return idxmap, colmap, depthmap, featuremap

    @staticmethod
    def backward(ctx, _grad_idx, _grad_colmap, _grad_depth, grad_featuremap):
        s = ctx.raster_settings
        means3D, normal, scales, rotations, feature_vector, sh, idxmap = ctx.saved_tensors
        (grad_feature_vector,) = _C.get_idxmap_backward2(means3D, s.image_height, s.image_width, s.campos, idxmap,
                                                         grad_featuremap, s.debug)
        z = torch.zeros_like
        return (z(means3D), z(normal), z(means3D), grad_feature_vector,
                z(sh) if sh is not None else None, z(scales) if scales is not None else None,
                z(rotations) if rotations is not None else None, None)


def rasterize_gaussians2(means3D, normal, means2D, feature_vector, sh, scales, rotations, raster_settings):
    return _GetIdxMap2.apply(means3D, normal, means2D, feature_vector, sh, scales, rotations, raster_settings)


class GaussianRasterizer2(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def forward(self, means3D, normal, means2D, feature_vector, shs=None, scales=None, rotations=None):
        return rasterize_gaussians2(means3D, normal, means2D, feature_vector, shs, scales, rotations,
                                    self.raster_settings)


_EMPTY = torch.Tensor([])


def alpha_forward(background, means3D, opacity, scales, rotations, scale_modifier, viewmatrix, projmatrix, tan_fovx,
                  tan_fovy, image_height, image_width, sh, degree, campos, prefiltered, debug):
    """rasterizer2's `_C.rasterize_gaussians` (rasterize_points.h:8-27): 17 arguments, no precomp inputs."""
    return _C.rasterize_gaussians(background, means3D, _EMPTY, opacity, scales, rotations, scale_modifier, _EMPTY,
                                  viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree,
                                  campos, prefiltered, debug)


def alpha_backward(background, means3D, radii, scales, rotations, scale_modifier, viewmatrix, projmatrix, tan_fovx,
                   tan_fovy, dL_dout_color, sh, degree, campos, geomBuffer, R, binningBuffer, imageBuffer, debug):
    """rasterizer2's `_C.rasterize_gaussians_backward` (rasterize_points.h:29-49): returns the reference's 6-tuple
    (dL_dmeans2D, dL_dopacity, dL_dmeans3D, dL_dsh, dL_dscales, dL_drotations) (rasterize_points.cu:177)."""
    g2d, _gc, gop, g3d, _gcov, gsh, gs, gr = _C.rasterize_gaussians_backward(
        background, means3D, radii, _EMPTY, scales, rotations, scale_modifier, _EMPTY, viewmatrix, projmatrix,
        tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos, geomBuffer, R, binningBuffer, imageBuffer, debug)
    return g2d, gop, g3d, gsh, gs, gr
