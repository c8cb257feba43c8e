"""Shared helpers for the parity tests: run the product (through the reference-shaped `_C` surface over the
C ABI) and the reference-compiled oracle (tests/refbind.py) on the same seeded device inputs."""
from __future__ import annotations

import math
from typing import Dict, Optional

import numpy as np
import torch

from neuralgaussiansplatting_b200 import _C
from neuralgaussiansplatting_b200.synthetic import CONFIGS, SceneConfig, make_camera, make_grad_image, make_scene

EMPTY = torch.Tensor([])


def scene_on(cfg_name: str, device, seed: int = 0, P: Optional[int] = None, view: int = 0):
    cfg = CONFIGS[cfg_name] if isinstance(cfg_name, str) else cfg_name
    sc = make_scene(cfg, seed=seed, device=device, P=P)
    cam = make_camera(cfg, view, device=device)
    return cfg, sc, cam


def fwd_args(cfg: SceneConfig, sc: Dict[str, torch.Tensor], cam, bg: torch.Tensor, degree: int = 3,
             colors_precomp=None, cov3D_precomp=None, scale_modifier: float = 1.0, debug: bool = False,
             prefiltered: bool = False):
    """Positional argument list of _C.rasterize_gaussians (reference rasterize_points.h:18-38)."""
    use_sh = colors_precomp is None
    use_sr = cov3D_precomp is None
    return (bg, sc["means3D"], EMPTY if use_sh else colors_precomp, sc["opacities"],
            sc["scales"] if use_sr else EMPTY, sc["rotations"] if use_sr else EMPTY, scale_modifier,
            EMPTY if use_sr else cov3D_precomp, cam.world_view_transform, cam.full_proj_transform,
            math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5), cam.image_height, cam.image_width,
            sc["shs"] if use_sh else EMPTY, degree, cam.camera_center, prefiltered, debug)


def ours_forward(cfg, sc, cam, bg, names=None, **kw):
    args = fwd_args(cfg, sc, cam, bg, **kw)
    out, state = _C.forward_with_state(*args, names=names)
    return args, out, state


def ours_backward(args, out, dL_dpix):
    (bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix,
     tan_fovx, tan_fovy, H, W, sh, degree, campos, prefiltered, debug) = args
    num_rendered, color, radii, geom, binning, img = out
    g = _C.rasterize_gaussians_backward(bg, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp,
                                        viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dpix, sh, degree, campos, geom,
                                        num_rendered, binning, img, debug)
    names = ("means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations")
    return dict(zip(names, g))


def ref_forward(ref, cfg, sc, cam, bg, degree: int = 3, colors_precomp=None, cov3D_precomp=None,
                scale_modifier: float = 1.0):
    use_sh = colors_precomp is None
    use_sr = cov3D_precomp is None
    kw = dict(bg=bg, means3D=sc["means3D"], opacities=sc["opacities"], viewmatrix=cam.world_view_transform,
              projmatrix=cam.full_proj_transform, campos=cam.camera_center, tan_fovx=math.tan(cam.FoVx * 0.5),
              tan_fovy=math.tan(cam.FoVy * 0.5), W=cam.image_width, H=cam.image_height, degree=degree,
              scale_modifier=scale_modifier,
              shs=sc["shs"] if use_sh else None, colors_precomp=None if use_sh else colors_precomp,
              scales=sc["scales"] if use_sr else None, rotations=sc["rotations"] if use_sr else None,
              cov3D_precomp=None if use_sr else cov3D_precomp)
    return kw, ref.forward(**kw)


def ref_backward(ref, kw, radii, dL_dpix):
    keys = ("bg", "means3D", "viewmatrix", "projmatrix", "campos", "tan_fovx", "tan_fovy", "shs", "colors_precomp",
            "scales", "rotations", "cov3D_precomp", "scale_modifier")
    return ref.backward(radii=radii, dL_dpix=dL_dpix, **{k: kw[k] for k in keys})


def rel_l2(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.double().flatten(), b.double().flatten()
    d = (a - b).norm().item()
    n = b.norm().item()
    return d / n if n > 0 else d


def np_of(t: torch.Tensor) -> np.ndarray:
    return t.detach().cpu().numpy()


def compare_forward_state(state: Dict[str, torch.Tensor], ref, radii_ours, radii_ref) -> Dict[str, object]:
    """Mismatch counts for the bit-exact set (north_star): radii, tiles_touched, sorted keys (+values),
    per-tile ranges, n_contrib; plus max-abs of the float state for visible Gaussians."""
    res: Dict[str, object] = {}
    r_o, r_r = np_of(radii_ours), np_of(radii_ref)
    res["radii"] = int((r_o != r_r).sum())
    vis = r_r > 0
    tt_r = ref.fetch("tiles_touched")
    res["tiles_touched"] = int((np_of(state["tiles_touched"]).view(np.uint32) != tt_r).sum())
    res["point_offsets"] = int((np_of(state["point_offsets"]).view(np.uint32) != ref.fetch("point_offsets")).sum())
    k_r = ref.fetch("point_list_keys")
    k_o = np_of(state["point_list_keys"]).view(np.uint64)
    res["num_rendered"] = (int(k_o.shape[0]), int(k_r.shape[0]))
    if k_o.shape == k_r.shape:
        res["point_list_keys"] = int((k_o != k_r).sum())
        res["point_list"] = int((np_of(state["point_list"]).view(np.uint32) != ref.fetch("point_list")).sum())
    else:
        res["point_list_keys"] = res["point_list"] = -1
    res["ranges"] = int((np_of(state["ranges"]).view(np.uint32).reshape(-1) != ref.fetch("ranges")).sum())
    res["n_contrib"] = int((np_of(state["n_contrib"]).view(np.uint32) != ref.fetch("n_contrib")).sum())
    # float state (visible Gaussians only: the reference leaves culled rows uninitialised)
    d_r = ref.fetch("depths")
    res["depth_bits"] = int((np_of(state["depths"]).view(np.uint32)[vis] != d_r.view(np.uint32)[vis]).sum())
    m_r = ref.fetch("means2D").reshape(-1, 2)
    res["means2D_bits"] = int((np_of(state["means2D"]).view(np.uint32)[vis] != m_r.view(np.uint32)[vis]).sum())
    c_r = ref.fetch("conic_opacity").reshape(-1, 4)
    c_o = np_of(state["conic_opacity"])
    res["conic_bits"] = int((c_o.view(np.uint32)[vis] != c_r.view(np.uint32)[vis]).sum())
    res["conic_maxrel"] = float(np.max(np.abs(c_o[vis] - c_r[vis]) / (np.abs(c_r[vis]) + 1e-12))) if vis.any() else 0.0
    g_r = ref.fetch("rgb").reshape(-1, 3)
    g_o = np_of(state["rgb"])
    res["rgb_maxabs"] = float(np.max(np.abs(g_o[vis] - g_r[vis]))) if vis.any() else 0.0
    cl_r = ref.fetch("clamped").reshape(-1, 3)
    res["clamped"] = int((np_of(state["clamped"])[vis] != cl_r[vis]).sum())
    res["final_T_maxabs"] = float(np.max(np.abs(np_of(state["final_T"]) - ref.fetch("accum_alpha"))))
    return res


def cpu_covariance_and_colors(sc, cam, degree: int = 3):
    """The reference's Python-side toggles (gaussian_renderer/__init__.py:63-81) on whatever device sc lives."""
    from neuralgaussiansplatting_b200.sh_utils import build_covariance_from_scaling_rotation, eval_sh
    cov = build_covariance_from_scaling_rotation(sc["scales"], 1.0, sc["rotations"])
    shs_view = sc["shs"].transpose(1, 2).reshape(-1, 3, 16)
    d = sc["means3D"] - cam.camera_center[None, :]
    d = d / d.norm(dim=1, keepdim=True)
    col = torch.clamp_min(eval_sh(degree, shs_view, d) + 0.5, 0.0)
    return cov.contiguous(), col.contiguous()
