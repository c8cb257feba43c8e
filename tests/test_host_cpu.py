"""Host-side logic that needs no GPU: the drop-in Python surface (names, argument validation, error behaviour),
the synthetic generators, the torch helpers behind the reference's pipeline toggles, and that the product path
fails loudly (no CPU fallback) when handed CPU tensors."""
import math

import pytest
import torch

import helpers as Hh
from neuralgaussiansplatting_b200.diff_gaussian_rasterization import (GaussianRasterizationSettings,
                                                                      GaussianRasterizer)
from neuralgaussiansplatting_b200.rasterizer2 import GaussianRasterizationSettings2, GaussianRasterizer2
from neuralgaussiansplatting_b200.sh_utils import build_covariance_from_scaling_rotation, build_rotation, eval_sh
from neuralgaussiansplatting_b200.synthetic import CONFIGS, make_camera, make_scene


def _settings(cls, cfg, cam):
    return cls(cfg.height, cfg.width, 0.5, 0.5, torch.zeros(3), 1.0, cam.world_view_transform,
               cam.full_proj_transform, 3, cam.camera_center, False, False)


def test_settings_are_the_reference_namedtuples():
    fields = ("image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix",
              "projmatrix", "sh_degree", "campos", "prefiltered", "debug")
    assert GaussianRasterizationSettings._fields == fields
    assert GaussianRasterizationSettings2._fields == fields


def test_forward_signature_matches_reference():
    import inspect
    sig = inspect.signature(GaussianRasterizer.forward)
    assert list(sig.parameters) == ["self", "means3D", "means2D", "opacities", "shs", "colors_precomp", "scales",
                                    "rotations", "cov3D_precomp"]
    sig2 = inspect.signature(GaussianRasterizer2.forward)
    assert list(sig2.parameters) == ["self", "means3D", "normal", "means2D", "feature_vector", "shs", "scales",
                                     "rotations"]


def test_validation_and_no_cpu_fallback():
    cfg, sc, cam = Hh.scene_on("T0", "cpu", P=16)
    r = GaussianRasterizer(_settings(GaussianRasterizationSettings, cfg, cam))
    m2d = torch.zeros_like(sc["means3D"])
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(sc["means3D"], m2d, sc["opacities"], shs=sc["shs"], colors_precomp=torch.zeros(16, 3),
          scales=sc["scales"], rotations=sc["rotations"])
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):
        r(sc["means3D"], m2d, sc["opacities"], shs=sc["shs"], scales=sc["scales"], rotations=sc["rotations"],
          cov3D_precomp=torch.zeros(16, 6))
    with pytest.raises(RuntimeError, match="CUDA-only"):
        r(sc["means3D"], m2d, sc["opacities"], shs=sc["shs"], scales=sc["scales"], rotations=sc["rotations"])
    r2 = GaussianRasterizer2(_settings(GaussianRasterizationSettings2, cfg, cam))
    with pytest.raises(RuntimeError, match="CUDA-only"):
        r2(sc["means3D"], sc["normal"], m2d, sc["feature_vector"], sc["shs"], sc["scales"], sc["rotations"])
    with pytest.raises(RuntimeError, match="CUDA-only"):
        r.markVisible(sc["means3D"])


def test_synthetic_scene_is_seeded_and_shaped():
    cfg = CONFIGS["T1"]
    a, b = make_scene(cfg, seed=0), make_scene(cfg, seed=0)
    c = make_scene(cfg, seed=1)
    for k in a:
        assert torch.equal(a[k], b[k])
    assert not torch.equal(a["means3D"], c["means3D"])
    assert a["means3D"].shape == (cfg.P, 3) and a["shs"].shape == (cfg.P, 16, 3)
    assert a["opacities"].shape == (cfg.P, 1) and a["feature_vector"].shape == (cfg.P, 64)
    assert torch.allclose(a["rotations"].norm(dim=1), torch.ones(cfg.P), atol=1e-5)
    assert float(a["opacities"].min()) > 0 and float(a["opacities"].max()) < 1


def test_camera_follows_reference_conventions():
    cfg = CONFIGS["C1"]
    cam = make_camera(cfg, 0)
    # full_proj = world_view @ projection (scene/cameras.py:56); camera_center = inverse(world_view)[3,:3] (:57)
    assert torch.allclose(cam.full_proj_transform, cam.world_view_transform @ cam.projection_matrix, atol=1e-6)
    assert torch.allclose(cam.camera_center, cam.world_view_transform.inverse()[3, :3], atol=1e-5)
    assert abs(float(cam.camera_center.norm()) - cfg.cam_radius) < 1e-3
    # the origin projects to the image centre with positive depth
    p = torch.tensor([0.0, 0.0, 0.0, 1.0]) @ cam.full_proj_transform
    assert abs(float(p[0] / p[3])) < 1e-5 and abs(float(p[1] / p[3])) < 1e-5
    z = (torch.tensor([0.0, 0.0, 0.0, 1.0]) @ cam.world_view_transform)[2]
    assert float(z) > 0.2
    assert abs(CONFIGS["C3"].fovy - 2 * math.atan(math.tan(0.5) * 896 / 1344)) < 1e-9


def test_eval_sh_degree_gating_and_covariance():
    torch.manual_seed(0)
    sh = torch.randn(10, 3, 16)
    d = torch.nn.functional.normalize(torch.randn(10, 3), dim=1)
    full = eval_sh(3, sh, d)
    sh0 = sh.clone(); sh0[..., 1:] = 0
    assert torch.allclose(eval_sh(0, sh, d), eval_sh(3, sh0, d), atol=1e-6)
    assert not torch.allclose(full, eval_sh(2, sh, d))
    q = torch.randn(10, 4); s = torch.rand(10, 3) + 0.1
    R = build_rotation(q)
    assert torch.allclose(R @ R.transpose(1, 2), torch.eye(3).expand(10, 3, 3), atol=1e-5)
    cov = build_covariance_from_scaling_rotation(s, 2.0, q)
    L = R * (2.0 * s)[:, None, :]
    full_cov = L @ L.transpose(1, 2)
    assert torch.allclose(cov[:, 0], full_cov[:, 0, 0], atol=1e-6) and torch.allclose(cov[:, 4], full_cov[:, 1, 2], atol=1e-6)
