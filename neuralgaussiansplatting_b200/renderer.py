"""The caller either side of the rasterizer: `render()` with the reference's signature and return
dict (gaussian_renderer/__init__.py:19-101), and the view-batched variant the north star names.

`render_batch` launches B views back to back on the current stream. With the library's
device-side num_rendered there is no host round trip inside a view (the reference blocks on a
4-byte D2H per view, rasterizer_impl.cu:281); the only synchronisation is the one header read
per view that reports num_rendered, issued after the whole view is enqueued.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence

import torch

from .diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
from .rasterizer2 import GaussianRasterizationSettings2, GaussianRasterizer2
from .sh_utils import eval_sh


class PipelineParams:
    """arguments/__init__.py:64-69"""

    def __init__(self, convert_SHs_python: bool = False, compute_cov3D_python: bool = False, debug: bool = False):
        self.convert_SHs_python = convert_SHs_python
        self.compute_cov3D_python = compute_cov3D_python
        self.debug = debug


def _settings(cls, cam, pc, pipe, bg_color, scaling_modifier):
    return cls(
        image_height=int(cam.image_height), image_width=int(cam.image_width),
        tanfovx=math.tan(cam.FoVx * 0.5), tanfovy=math.tan(cam.FoVy * 0.5),
        bg=bg_color, scale_modifier=scaling_modifier,
        viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform,
        sh_degree=pc.active_sh_degree, campos=cam.camera_center,
        prefiltered=False, debug=pipe.debug)


def render(viewpoint_camera, pc, pipe, bg_color: torch.Tensor, scaling_modifier: float = 1.0,
           override_color: Optional[torch.Tensor] = None) -> Dict[str, torch.Tensor]:
    """Same contract as gaussian_renderer.render(): returns
    {"render": [3,H,W], "viewspace_points": [P,3] (grad sink), "visibility_filter": bool[P], "radii": int32[P]}."""
    xyz = pc.get_xyz
    screenspace_points = torch.zeros_like(xyz, dtype=xyz.dtype, requires_grad=True, device=xyz.device) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    rasterizer = GaussianRasterizer(raster_settings=_settings(GaussianRasterizationSettings, viewpoint_camera, pc,
                                                              pipe, bg_color, scaling_modifier))
    scales = rotations = cov3D_precomp = None
    if pipe.compute_cov3D_python:
        cov3D_precomp = pc.get_covariance(scaling_modifier)
    else:
        scales, rotations = pc.get_scaling, pc.get_rotation
    shs = colors_precomp = None
    if override_color is None:
        if pipe.convert_SHs_python:
            feats = pc.get_features
            shs_view = feats.transpose(1, 2).view(-1, 3, (pc.max_sh_degree + 1) ** 2)
            dir_pp = xyz - viewpoint_camera.camera_center.repeat(feats.shape[0], 1)
            dir_pp = dir_pp / dir_pp.norm(dim=1, keepdim=True)
            colors_precomp = torch.clamp_min(eval_sh(pc.active_sh_degree, shs_view, dir_pp) + 0.5, 0.0)
        else:
            shs = pc.get_features
    else:
        colors_precomp = override_color
    rendered_image, radii = rasterizer(means3D=xyz, means2D=screenspace_points, shs=shs, colors_precomp=colors_precomp,
                                       opacities=pc.get_opacity, scales=scales, rotations=rotations,
                                       cov3D_precomp=cov3D_precomp)
    return {"render": rendered_image, "viewspace_points": screenspace_points, "visibility_filter": radii > 0,
            "radii": radii}


def render_idxmap(viewpoint_camera, pc, pipe, bg_color: torch.Tensor, scaling_modifier: float = 1.0):
    """The rasterizer2 call render1/2/3 make (gaussian_renderer/__init__.py:175,285,397), without the
    neural heads: returns idxmap / colmap / depthmap / featuremap and the reference's bookkeeping keys
    (`visibility_filter = idxmap > 0`, `radii = ones[P]`, :299-302)."""
    xyz = pc.get_xyz
    screenspace_points = torch.zeros_like(xyz, dtype=xyz.dtype, requires_grad=True, device=xyz.device) + 0
    rasterizer = GaussianRasterizer2(raster_settings=_settings(GaussianRasterizationSettings2, viewpoint_camera, pc,
                                                               pipe, bg_color, scaling_modifier))
    idxmap, colmap, depthmap, featuremap = rasterizer(
        means3D=xyz, normal=pc.get_normal, means2D=screenspace_points, feature_vector=pc.get_feature_vector,
        shs=pc.get_features, scales=pc.get_scaling, rotations=pc.get_rotation)
    return {"idxmap": idxmap, "colmap": colmap, "depthmap": depthmap, "featuremap": featuremap,
            "viewspace_points": screenspace_points, "visibility_filter": idxmap > 0,
            "radii": torch.ones(xyz.shape[0], device=xyz.device)}


def render_batch(cameras: Sequence, pc, pipe, bg_color: torch.Tensor, scaling_modifier: float = 1.0,
                 out: Optional[torch.Tensor] = None) -> Dict[str, object]:
    """Render a list of views back to back (the loop of render.py:24-35, gaussian_renderer.render() per view).
    Returns {"render": [B,3,H,W], "radii": [B,P], "num_rendered": [B ints], "views": [dict, ...]}.

    Without gradients (torch.no_grad(), or no input requires grad) and with equal image sizes this is ONE batched
    call: `_C.rasterize_gaussians_batch` shares one geometry / image / binning workspace between the views, writes
    every image straight into `out` / the returned [B,3,H,W] tensor and synchronises the host ONCE per batch.
    When gradients are needed every view keeps its own state buffers for its backward, so the views go through
    `render()` one by one (their forward still never blocks mid-pipeline)."""
    cameras = list(cameras)
    sizes = {(int(c.image_height), int(c.image_width)) for c in cameras}
    needs_grad = torch.is_grad_enabled() and any(
        t is not None and t.requires_grad for t in (pc.get_xyz, pc.get_features, pc.get_opacity, pc.get_scaling,
                                                    pc.get_rotation))
    if cameras and len(sizes) == 1 and not needs_grad and not pipe.debug and not pipe.convert_SHs_python:
        from . import _C
        H, W = next(iter(sizes))
        E = torch.Tensor([])
        xyz = pc.get_xyz.detach()
        colors, shs, cov = E, E, E
        scales, rots = E, E
        if pipe.compute_cov3D_python:
            cov = pc.get_covariance(scaling_modifier).detach()
        else:
            scales, rots = pc.get_scaling.detach(), pc.get_rotation.detach()
        shs = pc.get_features.detach()
        nums, img, radii = _C.rasterize_gaussians_batch(
            bg_color, xyz, colors, pc.get_opacity.detach(), scales, rots, scaling_modifier, cov,
            [(c.world_view_transform, c.full_proj_transform, c.camera_center) for c in cameras],
            [math.tan(c.FoVx * 0.5) for c in cameras], [math.tan(c.FoVy * 0.5) for c in cameras], H, W, shs,
            pc.active_sh_degree, False, out_color=out)
        views = [{"render": img[b], "viewspace_points": None, "visibility_filter": radii[b] > 0, "radii": radii[b]}
                 for b in range(len(cameras))]
        return {"render": img, "radii": radii, "num_rendered": nums, "views": views}
    views: List[Dict[str, torch.Tensor]] = [render(c, pc, pipe, bg_color, scaling_modifier) for c in cameras]
    res: Dict[str, object] = {"views": views}
    if len(sizes) == 1 and views:
        if out is not None:
            torch.stack([v["render"] for v in views], out=out)
            res["render"] = out
        else:
            res["render"] = torch.stack([v["render"] for v in views])
        res["radii"] = torch.stack([v["radii"] for v in views])
    return res
