"""Seeded synthetic Gaussians and cameras of the shapes BASELINE.json names (no datasets offline).

Distributions follow SURVEY.md §8(d). Camera matrices follow the reference's conventions exactly:
  world_view_transform = getWorld2View2(R, T).T                 (utils/graphics_utils.py:38-49, scene/cameras.py:54)
  projection_matrix    = getProjectionMatrix(znear, zfar, fovX, fovY).T   (graphics_utils.py:51-71, cameras.py:55)
  full_proj_transform  = world_view_transform @ projection_matrix          (cameras.py:56)
  camera_center        = world_view_transform.inverse()[3, :3]            (cameras.py:57)
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, Optional

import numpy as np
import torch


@dataclass(frozen=True)
class SceneConfig:
    name: str
    P: int
    width: int
    height: int
    fovx: float
    shape: str          # "nerf" (uniform cube) or "360" (gaussian core + far shell)
    cam_radius: float
    cam_height: float
    log_scale_mu: float  # calibrated so that tiles / visible Gaussian ~= 6 at camera 0
    views: int = 1

    @property
    def fovy(self) -> float:
        if self.width == self.height:
            return self.fovx
        return 2.0 * math.atan(math.tan(self.fovx * 0.5) * self.height / self.width)


# log_scale_mu values come from tools/calibrate_scales.py (bisection on the CPU oracle's tiles_touched at
# camera 0 so that num_rendered / visible Gaussians = 6.0); achieved values are in DESIGN.md.
CONFIGS: Dict[str, SceneConfig] = {
    "C1": SceneConfig("C1", 50_000, 800, 800, 0.6911112, "nerf", 4.031, 0.0, -4.86),
    "C2": SceneConfig("C2", 300_000, 800, 800, 0.6911112, "nerf", 4.031, 0.0, -4.86),
    "C3": SceneConfig("C3", 3_000_000, 1344, 896, 1.0, "360", 4.0, 1.0, -4.91),
    "C4": SceneConfig("C4", 6_000_000, 1920, 1080, 1.2, "360", 4.0, 1.0, -5.07, views=64),
    "C5": SceneConfig("C5", 3_000_000, 1344, 896, 1.0, "360", 4.0, 1.0, -4.91, views=8),
    # tiny shapes for tests
    "T0": SceneConfig("T0", 2_000, 160, 96, 0.9, "nerf", 4.031, 0.5, -2.6),
    "T1": SceneConfig("T1", 20_000, 400, 304, 0.8, "360", 4.0, 1.0, -2.9),
}


def make_scene(cfg: SceneConfig, seed: int = 0, device="cpu", P: Optional[int] = None) -> Dict[str, torch.Tensor]:
    """Gaussians in the layouts the rasterizer takes: means3D[P,3], scales[P,3] (activated),
    rotations[P,4] (unit), opacities[P,1] (activated), shs[P,16,3], plus the idxmap inputs
    feature_vector[P,64] and normal[P,3]."""
    g = torch.Generator().manual_seed(seed)
    P = int(P if P is not None else cfg.P)
    if cfg.shape == "nerf":
        means = (torch.rand(P, 3, generator=g) * 2.0 - 1.0) * 1.3
    else:
        core = (torch.randn(P, 3, generator=g) * 1.5).clamp_(-6.0, 6.0)
        d = torch.randn(P, 3, generator=g)
        d = d / d.norm(dim=1, keepdim=True).clamp_min(1e-8)
        r = 6.0 + 24.0 * torch.rand(P, 1, generator=g)
        shell = d * r
        pick = torch.rand(P, 1, generator=g) < 0.3
        means = torch.where(pick, shell, core)
        # far-shell Gaussians are proportionally larger so their footprint stays a few pixels
    scales = torch.exp(torch.randn(P, 3, generator=g) * 0.6 + cfg.log_scale_mu)
    if cfg.shape == "360":
        dist = means.norm(dim=1, keepdim=True).clamp_min(1.0)
        scales = scales * torch.where(dist > 6.0, dist / 4.0, torch.ones_like(dist))
    rot = torch.randn(P, 4, generator=g)
    rot = rot / rot.norm(dim=1, keepdim=True).clamp_min(1e-8)
    opac = torch.sigmoid(torch.randn(P, 1, generator=g) * 2.0)
    shs = torch.randn(P, 16, 3, generator=g)
    shs[:, 1:, :] *= 0.15
    feat = torch.randn(P, 64, generator=g)
    nrm = torch.randn(P, 3, generator=g)
    nrm = nrm / nrm.norm(dim=1, keepdim=True).clamp_min(1e-8)
    out = dict(means3D=means, scales=scales, rotations=rot, opacities=opac, shs=shs, feature_vector=feat, normal=nrm)
    return {k: v.float().contiguous().to(device) for k, v in out.items()}


def _projection_matrix(znear: float, zfar: float, fovx: float, fovy: float) -> np.ndarray:
    # getProjectionMatrix (graphics_utils.py:51-71)
    tan_y, tan_x = math.tan(fovy / 2), math.tan(fovx / 2)
    top, right = tan_y * znear, tan_x * znear
    bottom, left = -top, -right
    Pm = np.zeros((4, 4), dtype=np.float64)
    Pm[0, 0] = 2.0 * znear / (right - left)
    Pm[1, 1] = 2.0 * znear / (top - bottom)
    Pm[0, 2] = (right + left) / (right - left)
    Pm[1, 2] = (top + bottom) / (top - bottom)
    Pm[3, 2] = 1.0
    Pm[2, 2] = zfar / (zfar - znear)
    Pm[2, 3] = -(zfar * znear) / (zfar - znear)
    return Pm


class SyntheticCamera:
    """The attributes gaussian_renderer.render() reads from scene.cameras.Camera (cameras.py:17-57)."""

    def __init__(self, width: int, height: int, fovx: float, fovy: float, position, target=(0.0, 0.0, 0.0),
                 up=(0.0, 0.0, 1.0), znear: float = 0.01, zfar: float = 100.0, device="cpu", uid: int = 0):
        self.uid = uid
        self.image_width, self.image_height = int(width), int(height)
        self.FoVx, self.FoVy = float(fovx), float(fovy)
        self.znear, self.zfar = znear, zfar
        c = np.asarray(position, dtype=np.float64)
        fwd = np.asarray(target, dtype=np.float64) - c
        fwd /= np.linalg.norm(fwd)
        upv = np.asarray(up, dtype=np.float64)
        right = np.cross(fwd, upv)
        if np.linalg.norm(right) < 1e-8:
            right = np.cross(fwd, np.array([0.0, 1.0, 0.0]))
        right /= np.linalg.norm(right)
        down = np.cross(fwd, right)           # camera y points down, z forward (COLMAP convention)
        Rw2c = np.stack([right, down, fwd], axis=0)
        W2V = np.eye(4)
        W2V[:3, :3] = Rw2c
        W2V[:3, 3] = -Rw2c @ c
        wvt = torch.tensor(np.float32(W2V)).transpose(0, 1).contiguous()
        proj = torch.tensor(np.float32(_projection_matrix(znear, zfar, fovx, fovy))).transpose(0, 1).contiguous()
        self.world_view_transform = wvt.to(device)
        self.projection_matrix = proj.to(device)
        self.full_proj_transform = (wvt.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0).contiguous().to(device)
        self.camera_center = wvt.inverse()[3, :3].contiguous().to(device)

    def to(self, device):
        for k in ("world_view_transform", "projection_matrix", "full_proj_transform", "camera_center"):
            setattr(self, k, getattr(self, k).to(device))
        return self


def make_camera(cfg: SceneConfig, view_id: int = 0, device="cpu") -> SyntheticCamera:
    """View `view_id` of the config's orbit (seed 1000 + view_id fixes the small jitter)."""
    n = max(cfg.views, 8)
    rng = np.random.default_rng(1000 + view_id)
    ang = 2.0 * math.pi * (view_id % n) / n + 0.05 * rng.standard_normal()
    if cfg.shape == "nerf":
        elev = 0.35 + 0.1 * rng.standard_normal()
        pos = cfg.cam_radius * np.array([math.cos(ang) * math.cos(elev), math.sin(ang) * math.cos(elev), math.sin(elev)])
    else:
        pos = np.array([cfg.cam_radius * math.cos(ang), cfg.cam_radius * math.sin(ang), cfg.cam_height])
    return SyntheticCamera(cfg.width, cfg.height, cfg.fovx, cfg.fovy, pos, device=device, uid=view_id)


def make_grad_image(cfg: SceneConfig, view_id: int = 0, device="cpu") -> torch.Tensor:
    """dL/dout_color ~ N(0,1)*1e-3, seed 2000 + view_id."""
    g = torch.Generator().manual_seed(2000 + view_id)
    return (torch.randn(3, cfg.height, cfg.width, generator=g) * 1e-3).to(device)


def make_grad_featuremap(cfg: SceneConfig, view_id: int = 0, device="cpu") -> torch.Tensor:
    g = torch.Generator().manual_seed(3000 + view_id)
    return (torch.randn(cfg.height, cfg.width, 64, generator=g) * 1e-3).to(device)


class SyntheticGaussians:
    """The GaussianModel getters gaussian_renderer.render() uses (scene/gaussian_model.py:113-144)."""

    def __init__(self, scene: Dict[str, torch.Tensor], sh_degree: int = 3, requires_grad: bool = False):
        self.max_sh_degree = 3
        self.active_sh_degree = sh_degree
        self._t = {k: v.clone().requires_grad_(requires_grad and v.is_floating_point()) for k, v in scene.items()}

    get_xyz = property(lambda self: self._t["means3D"])
    get_scaling = property(lambda self: self._t["scales"])
    get_rotation = property(lambda self: self._t["rotations"])
    get_opacity = property(lambda self: self._t["opacities"])
    get_features = property(lambda self: self._t["shs"])
    get_feature_vector = property(lambda self: self._t["feature_vector"])
    get_normal = property(lambda self: self._t["normal"])

    def get_covariance(self, scaling_modifier: float = 1.0):
        from .sh_utils import build_covariance_from_scaling_rotation
        return build_covariance_from_scaling_rotation(self.get_scaling, scaling_modifier, self.get_rotation)

    def parameters(self):
        return [v for v in self._t.values() if v.requires_grad]
