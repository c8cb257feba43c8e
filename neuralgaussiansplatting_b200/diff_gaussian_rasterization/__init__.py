"""Drop-in replacement for the reference package `diff_gaussian_rasterization`
(submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py).

Public surface kept identical (names, argument order, keyword names, return shapes/dtypes):
  GaussianRasterizationSettings  NamedTuple, same 12 fields in the same order   (:157-169)
  GaussianRasterizer(raster_settings)                                           (:171-220)
      .forward(means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
               rotations=None, cov3D_precomp=None) -> (color[3,H,W] f32, radii[P] i32)
      .markVisible(positions) -> bool[P]
  rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                      cov3Ds_precomp, raster_settings)                          (:21-42)
The kernels behind `_C` are the B200-native ones (libb200splat.so); there is no CPU path.
"""
from __future__ import annotations

from typing import NamedTuple

import torch
import torch.nn as nn

from .. import _C

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians"]


class GaussianRasterizationSettings(NamedTuple):
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


def _snapshot(args):
    """CPU deep copy of a call's arguments, taken before the call can corrupt them (:17-19)."""
    return tuple(a.detach().cpu().clone() if isinstance(a, torch.Tensor) else a for a in args)


def _call_with_dump(fn, args, debug: bool, dump_path: str, what: str):
    """debug=True: snapshot the arguments and, if the native call raises, write them to
    `dump_path` with torch.save before re-raising (reference :83-90 and :132-139)."""
    if not debug:
        return fn(*args)
    saved = _snapshot(args)
    try:
        return fn(*args)
    except Exception:
        torch.save(saved, dump_path)
        print(f"\nAn error occured in {what}. Writing {dump_path} for debugging.\n")
        raise


class _RasterizeGaussians(torch.autograd.Function):
    """Autograd binding; argument and gradient order follow the reference (:44-155)."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        s = raster_settings
        native_args = (s.bg, means3D, colors_precomp, opacities, scales, rotations, s.scale_modifier,
                       cov3Ds_precomp, s.viewmatrix, s.projmatrix, s.tanfovx, s.tanfovy, s.image_height,
                       s.image_width, sh, s.sh_degree, s.campos, s.prefiltered, s.debug)
        num_rendered, color, radii, geom, binning, img = _call_with_dump(
            _C.rasterize_gaussians, native_args, s.debug, "snapshot_fw.dump", "forward")
        ctx.raster_settings = s
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning, img)
        ctx.mark_non_differentiable(radii)
        return color, radii

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii):
        s = ctx.raster_settings
        colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning, img = ctx.saved_tensors
        native_args = (s.bg, means3D, radii, colors_precomp, scales, rotations, s.scale_modifier, cov3Ds_precomp,
                       s.viewmatrix, s.projmatrix, s.tanfovx, s.tanfovy, grad_out_color, sh, s.sh_degree, s.campos,
                       geom, ctx.num_rendered, binning, img, s.debug)
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh, grad_scales,
         grad_rotations) = _call_with_dump(_C.rasterize_gaussians_backward, native_args, s.debug,
                                           "snapshot_bw.dump", "backward")
        # same positional order as the forward inputs; the settings tuple gets no gradient
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities, grad_scales,
                grad_rotations, grad_cov3Ds_precomp, None)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """Boolean near-plane frustum mask per point (reference :176-185)."""
        with torch.no_grad():
            s = self.raster_settings
            return _C.mark_visible(positions, s.viewmatrix, s.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        # same validation and messages as the reference (:191-195)
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        # absent inputs travel as empty CPU tensors, which the native side maps to NULL (:197-207)
        empty = torch.Tensor([])
        shs = empty if shs is None else shs
        colors_precomp = empty if colors_precomp is None else colors_precomp
        scales = empty if scales is None else scales
        rotations = empty if rotations is None else rotations
        cov3D_precomp = empty if cov3D_precomp is None else cov3D_precomp
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                   cov3D_precomp, self.raster_settings)
