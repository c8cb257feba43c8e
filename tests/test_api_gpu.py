"""The drop-in Python surface on the GPU: GaussianRasterizer / GaussianRasterizer2 modules with autograd, the
render() caller, the equivalence relations the reference's own pipeline toggles define
(gaussian_renderer/__init__.py:63-81), the idxmap extra channels against the reference rasterizer2, and the
committed golden vectors."""
import json
import math
from pathlib import Path

import numpy as np
import pytest
import torch

import helpers as Hh
import oracle_cpu
import refbind
from neuralgaussiansplatting_b200 import _C
from neuralgaussiansplatting_b200.renderer import PipelineParams, render, render_batch, render_idxmap
from neuralgaussiansplatting_b200.synthetic import (CONFIGS, SyntheticGaussians, make_camera, make_grad_featuremap,
                                                    make_grad_image, make_scene)

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).resolve().parent / "golden"


def test_top_level_import_names_resolve_to_this_package():
    import diff_gaussian_rasterization as dgr
    import rasterizer2 as r2
    assert dgr.GaussianRasterizer.__module__.startswith("neuralgaussiansplatting_b200")
    assert r2.GaussianRasterizer2.__module__.startswith("neuralgaussiansplatting_b200")
    assert dgr.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug")


def test_render_contract_and_autograd(dev):
    cfg, sc, cam = Hh.scene_on("T1", dev)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    pc = SyntheticGaussians(sc, requires_grad=True)
    pkg = render(cam, pc, PipelineParams(), bg)
    assert set(pkg) == {"render", "viewspace_points", "visibility_filter", "radii"}
    assert pkg["render"].shape == (3, cfg.height, cfg.width) and pkg["render"].dtype == torch.float32
    assert pkg["radii"].shape == (cfg.P,) and pkg["radii"].dtype == torch.int32
    assert pkg["visibility_filter"].dtype == torch.bool
    dL = make_grad_image(cfg, 0, device=dev)
    (pkg["render"] * dL).sum().backward()
    # same numbers as the raw _C backward
    args, out, _ = Hh.ours_forward(cfg, sc, cam, bg, names=[])
    g = Hh.ours_backward(args, out, dL)
    for k, name in (("means3D", "means3D"), ("opacities", "opacity"), ("shs", "sh"), ("scales", "scales"),
                    ("rotations", "rotations")):
        assert Hh.rel_l2(pc._t[k].grad.reshape(g[name].shape), g[name]) < 1e-5, k
    vs = pkg["viewspace_points"].grad
    assert vs is not None and vs.shape == (cfg.P, 3) and float(vs[:, 2].abs().max()) == 0.0
    assert Hh.rel_l2(vs, g["means2D"]) < 1e-5
    assert pc._t["feature_vector"].grad is None


def test_python_toggles_are_equivalent(dev):
    """convert_SHs_python / compute_cov3D_python must render what the kernels render."""
    cfg, sc, cam = Hh.scene_on("T1", dev)
    bg = torch.tensor([0.0, 0.0, 0.0], device=dev)
    pc = SyntheticGaussians(sc)
    a = render(cam, pc, PipelineParams(), bg)
    b = render(cam, pc, PipelineParams(convert_SHs_python=True), bg)
    c = render(cam, pc, PipelineParams(compute_cov3D_python=True), bg)
    assert torch.equal(a["radii"], b["radii"])
    assert float((a["render"] - b["render"]).abs().max()) < 1e-4
    assert float((a["radii"] != c["radii"]).float().mean()) < 2e-3    # torch vs kernel cov3D differ by ulps
    assert float((a["render"] - c["render"]).abs().max()) < 5e-3


def test_argument_validation_matches_reference_messages(dev):
    from neuralgaussiansplatting_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    cfg, sc, cam = Hh.scene_on("T0", dev)
    s = GaussianRasterizationSettings(cfg.height, cfg.width, 0.5, 0.5, torch.zeros(3, device=dev), 1.0,
                                      cam.world_view_transform, cam.full_proj_transform, 3, cam.camera_center,
                                      False, False)
    r = GaussianRasterizer(s)
    m2d = torch.zeros_like(sc["means3D"])
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(sc["means3D"], m2d, sc["opacities"], scales=sc["scales"], rotations=sc["rotations"])
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair or precomputed 3D covariance"):
        r(sc["means3D"], m2d, sc["opacities"], shs=sc["shs"])
    with pytest.raises(RuntimeError, match=r"means3D must have dimensions \(num_points, 3\)"):
        r(sc["means3D"][:, :2], m2d, sc["opacities"], shs=sc["shs"], scales=sc["scales"], rotations=sc["rotations"])
    with pytest.raises(RuntimeError, match="no CPU path"):
        r(sc["means3D"].cpu(), m2d.cpu(), sc["opacities"].cpu(), shs=sc["shs"].cpu(), scales=sc["scales"].cpu(),
          rotations=sc["rotations"].cpu())
    vis = r.markVisible(sc["means3D"])
    assert vis.dtype == torch.bool and vis.shape == (cfg.P,)


def test_non_default_stream_and_noncontiguous_inputs(dev):
    cfg, sc, cam = Hh.scene_on("T0", dev)
    bg = torch.zeros(3, device=dev)
    base = _C.rasterize_gaussians(*Hh.fwd_args(cfg, sc, cam, bg))
    sc2 = dict(sc)
    sc2["means3D"] = torch.cat([sc["means3D"], sc["means3D"]], dim=1)[:, :3]     # strided view
    sc2["shs"] = sc["shs"].transpose(1, 2).contiguous().transpose(1, 2)           # get_features-style view
    st = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(st):
        out = _C.rasterize_gaussians(*Hh.fwd_args(cfg, sc2, cam, bg))
    st.synchronize()
    assert out[0] == base[0] and torch.equal(out[1], base[1]) and torch.equal(out[2], base[2])


def test_render_batch_matches_single_views(dev):
    cfg, sc, _ = Hh.scene_on("T1", dev)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    pc = SyntheticGaussians(sc)
    cams = [make_camera(cfg, v, device=dev) for v in range(4)]
    res = render_batch(cams, pc, PipelineParams(), bg)
    assert res["render"].shape == (4, 3, cfg.height, cfg.width) and res["radii"].shape == (4, cfg.P)
    for v, cam in enumerate(cams):
        single = render(cam, pc, PipelineParams(), bg)
        assert torch.equal(res["render"][v], single["render"]) and torch.equal(res["radii"][v], single["radii"])


def test_render_batch_shares_one_workspace_and_synchronises_once(dev):
    """B views back to back, one workspace, images written into the caller's [B,3,H,W] tensor,
    exactly ONE host synchronisation per batch once the capacity estimate for the shape exists."""
    cfg, sc, _ = Hh.scene_on("T1", dev)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    pc = SyntheticGaussians(sc)
    cams = [make_camera(cfg, v, device=dev) for v in range(6)]
    with torch.no_grad():
        render_batch(cams[:2], pc, PipelineParams(), bg)                  # learns the capacity estimate
        out = torch.full((6, 3, cfg.height, cfg.width), float("nan"), device=dev)
        s0, o0 = _C.sync_count, _C.overflow_count
        res = render_batch(cams, pc, PipelineParams(), bg, out=out)
        assert _C.sync_count - s0 == 1 and _C.overflow_count == o0
    assert res["render"].data_ptr() == out.data_ptr() and torch.isfinite(out).all()
    assert len(res["num_rendered"]) == 6 and all(n > 0 for n in res["num_rendered"])
    for v, cam in enumerate(cams):
        single = render(cam, pc, PipelineParams(), bg)
        assert torch.equal(out[v], single["render"]) and torch.equal(res["radii"][v], single["radii"])
        assert torch.equal(res["views"][v]["visibility_filter"], single["visibility_filter"])


def test_render_batch_recovers_from_a_too_small_capacity_estimate(dev):
    cfg, sc, _ = Hh.scene_on("T1", dev)
    bg = torch.zeros(3, device=dev)
    pc = SyntheticGaussians(sc)
    cams = [make_camera(cfg, v, device=dev) for v in range(3)]
    key = (dev.index, cfg.P, cfg.width, cfg.height)
    with torch.no_grad():
        render_batch(cams, pc, PipelineParams(), bg)
        _C._capacity_hint[key] = 64                                        # far too small: every view overflows
        o0 = _C.overflow_count
        res = render_batch(cams, pc, PipelineParams(), bg)
        assert _C.overflow_count - o0 == 3
    for v, cam in enumerate(cams):
        single = render(cam, pc, PipelineParams(), bg)
        assert torch.equal(res["render"][v], single["render"])
    assert _C._capacity_hint[key] >= max(res["num_rendered"])


def test_capacity_hint_is_a_running_maximum(dev):
    """A loop that alternates small and large views must not overflow on every large one."""
    cfg, sc, _ = Hh.scene_on("T1", dev)
    bg = torch.zeros(3, device=dev)
    small = {k: (v[:2000].contiguous() if v.shape[0] == cfg.P else v) for k, v in sc.items()}
    cam = make_camera(cfg, 0, device=dev)
    far = make_camera(cfg, 0, device=dev)
    _C._capacity_hint.clear()
    a_big = Hh.fwd_args(cfg, sc, cam, bg)
    big = _C.rasterize_gaussians(*a_big)[0]
    o0 = _C.overflow_count
    # same shape key (P, W, H), much smaller view: shrink the scales so few tiles are touched
    sc_small = dict(sc); sc_small["scales"] = sc["scales"] * 0.05
    for _ in range(3):
        n_small = _C.rasterize_gaussians(*Hh.fwd_args(cfg, sc_small, cam, bg))[0]
        n_big = _C.rasterize_gaussians(*a_big)[0]
    assert n_small < 0.7 * big and n_big == big
    assert _C.overflow_count == o0
    del small, far


# ------------------------------------------------------------------ rasterizer2 extra channels
def _idxmap_ours(cfg, sc, cam, dev):
    return _C.get_idxmap2(torch.zeros(3, device=dev), sc["means3D"], sc["normal"], sc["scales"], sc["rotations"],
                          cam.world_view_transform, cam.full_proj_transform, math.tan(cam.FoVx * 0.5),
                          math.tan(cam.FoVy * 0.5), cfg.height, cfg.width, sc["feature_vector"], sc["shs"],
                          cam.camera_center, False)


@pytest.mark.skipif(not refbind.available("ref_r2"), reason="oracle/_ref not built")
@pytest.mark.parametrize("cfg_name", ["T0", "T1", "C1"])
def test_idxmap_matches_reference_rasterizer2(dev, cfg_name):
    cfg, sc, cam = Hh.scene_on(cfg_name, dev)
    n, idx, col, dep, feat = _idxmap_ours(cfg, sc, cam, dev)
    r2 = refbind.RefR2()
    R, ridx, rcol, rdep, rfeat, rrad = r2.getidxmap(
        means3D=sc["means3D"], normal=sc["normal"], scales=sc["scales"], rotations=sc["rotations"],
        viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform, tan_fovx=math.tan(cam.FoVx * 0.5),
        tan_fovy=math.tan(cam.FoVy * 0.5), feature_vector=sc["feature_vector"], shs=sc["shs"],
        campos=cam.camera_center, W=cfg.width, H=cfg.height)
    assert n == R
    assert torch.equal(idx, ridx)                                   # bit-exact nearest-point index map
    assert torch.equal(dep.view(torch.int32), rdep.view(torch.int32))
    assert float((col - rcol).abs().max()) < 1e-6
    assert float((feat[..., :25] - rfeat[..., :25]).abs().max()) < 1e-5   # depth + positional encoding
    assert torch.equal(feat[..., 25:], rfeat[..., 25:])              # gathered learned channels
    # backward: the reference's scatter is a racy += (rasterizer_impl.cu:844-846); compare where each point is
    # hit by exactly one pixel, and against the race-free CPU oracle everywhere
    gf = make_grad_featuremap(cfg, 0, device=dev)
    (g,) = _C.get_idxmap_backward2(sc["means3D"], cfg.height, cfg.width, cam.camera_center, idx, gf, False)
    gr = r2.getidxmap_backward(means3D=sc["means3D"], campos=cam.camera_center, out_idx=ridx, grad_feature=gf,
                               W=cfg.width, H=cfg.height)
    hits = torch.bincount(idx[idx >= 0].long().flatten(), minlength=cfg.P)
    once = hits == 1
    assert torch.equal(g[once], gr[once])
    assert float(g[hits == 0].abs().max()) == 0.0 and float(g[:, :25].abs().max()) == 0.0
    o = oracle_cpu.idxmap(means3D=sc["means3D"], viewmatrix=cam.world_view_transform,
                          projmatrix=cam.full_proj_transform, campos=cam.camera_center,
                          feature_vector=sc["feature_vector"], W=cfg.width, H=cfg.height, grad_feature=gf)
    assert np.array_equal(o["idx"], idx.cpu().numpy())
    assert Hh.rel_l2(g.cpu(), torch.from_numpy(o["dL_dfeature_vector"])) < 1e-6
    r2.close()


@pytest.mark.skipif(not refbind.available("ref_r2"), reason="oracle/_ref not built")
@pytest.mark.parametrize("cfg_name", ["T1", "C1"])
def test_rasterizer2_alpha_exports_match_reference(dev, cfg_name):
    """rasterizer2's C-level alpha path — the 17-argument `rasterize_gaussians` and the
    6-tuple `rasterize_gaussians_backward` (rasterizer2/ext.cpp:16-17, rasterize_points.cu:27,105) — against the
    reference rasterizer2 compiled for sm_100 (oracle/ref_shim_r2.cu: ref_r2_forward / ref_r2_backward)."""
    from neuralgaussiansplatting_b200 import rasterizer2 as r2mod
    from neuralgaussiansplatting_b200.synthetic import make_grad_image
    cfg, sc, cam = Hh.scene_on(cfg_name, dev)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    tx, ty = math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5)
    out = r2mod.alpha_forward(bg, sc["means3D"], sc["opacities"], sc["scales"], sc["rotations"], 1.0,
                              cam.world_view_transform, cam.full_proj_transform, tx, ty, cfg.height, cfg.width,
                              sc["shs"], 3, cam.camera_center, False, False)
    assert len(out) == 6                                              # (R, color, radii, geom, binning, img)
    R, color, radii, geom, binning, img = out
    r2 = refbind.RefR2()
    rR, rcolor, rradii = r2.forward(bg=bg, means3D=sc["means3D"], opacities=sc["opacities"], scales=sc["scales"],
                                    rotations=sc["rotations"], viewmatrix=cam.world_view_transform,
                                    projmatrix=cam.full_proj_transform, campos=cam.camera_center, tan_fovx=tx,
                                    tan_fovy=ty, W=cfg.width, H=cfg.height, shs=sc["shs"])
    assert R == rR and torch.equal(radii, rradii)
    assert float((color - rcolor).abs().max()) < 1e-5
    dL = make_grad_image(cfg, 0, device=dev)
    g = r2mod.alpha_backward(bg, sc["means3D"], radii, sc["scales"], sc["rotations"], 1.0, cam.world_view_transform,
                             cam.full_proj_transform, tx, ty, dL, sc["shs"], 3, cam.camera_center, geom, R, binning,
                             img, False)
    gr = r2.backward(bg=bg, means3D=sc["means3D"], radii=rradii, shs=sc["shs"], scales=sc["scales"],
                     rotations=sc["rotations"], viewmatrix=cam.world_view_transform,
                     projmatrix=cam.full_proj_transform, campos=cam.camera_center, tan_fovx=tx, tan_fovy=ty,
                     dL_dpix=dL)
    assert len(g) == 6 and [tuple(t.shape) for t in g] == [tuple(t.shape) for t in gr]
    for name, a, b in zip(("means2D", "opacity", "means3D", "sh", "scales", "rotations"), g, gr):
        assert Hh.rel_l2(a, b) < 1e-4, name
    r2.close()


def test_idxmap_module_autograd_contract(dev):
    cfg, sc, cam = Hh.scene_on("T0", dev)
    pc = SyntheticGaussians(sc, requires_grad=True)
    out = render_idxmap(cam, pc, PipelineParams(), torch.zeros(3, device=dev))
    assert out["idxmap"].shape == (cfg.height, cfg.width) and out["idxmap"].dtype == torch.int32
    assert out["colmap"].shape == (cfg.height, cfg.width, 3)
    assert out["depthmap"].shape == (cfg.height, cfg.width)
    assert out["featuremap"].shape == (cfg.height, cfg.width, 64)
    assert out["radii"].shape == (cfg.P,) and float(out["radii"].min()) == 1.0
    (out["featuremap"] * make_grad_featuremap(cfg, 0, device=dev)).sum().backward()
    fv = pc._t["feature_vector"].grad
    assert fv is not None and float(fv[:, :25].abs().max()) == 0.0 and float(fv.abs().max()) > 0.0
    for k in ("means3D", "scales", "rotations", "shs", "normal"):     # zero gradients, as in the reference
        g = pc._t[k].grad
        assert g is None or float(g.abs().max()) == 0.0


# ------------------------------------------------------------------ committed golden vectors
@pytest.mark.parametrize("variant", ["sh_sr", "precomp"])
def test_cuda_path_matches_committed_reference_golden(dev, variant):
    g = np.load(GOLD / f"ref_T0_{variant}.npz")
    t = lambda k: torch.from_numpy(g[k]).to(dev)  # noqa: E731
    tan = g["in_tanfov"]
    E = Hh.EMPTY
    pre = variant == "precomp"
    args = (t("bg"), t("in_means3D"), t("in_colors_precomp") if pre else E, t("in_opacities"),
            E if pre else t("in_scales"), E if pre else t("in_rotations"), 1.0, t("in_cov3D_precomp") if pre else E,
            t("in_viewmatrix"), t("in_projmatrix"), float(tan[0]), float(tan[1]), 96, 160, E if pre else t("in_shs"),
            3, t("in_campos"), False, False)
    out, state = _C.forward_with_state(*args)
    assert out[0] == int(g["num_rendered"])
    assert np.array_equal(out[2].cpu().numpy(), g["radii"])
    assert np.array_equal(state["tiles_touched"].cpu().numpy().view(np.uint32), g["tiles_touched"])
    assert np.array_equal(state["point_list_keys"].cpu().numpy().view(np.uint64), g["point_list_keys"])
    assert np.array_equal(state["point_list"].cpu().numpy().view(np.uint32), g["point_list"])
    assert np.array_equal(state["ranges"].cpu().numpy().view(np.uint32).reshape(-1), g["ranges"])
    assert np.array_equal(state["n_contrib"].cpu().numpy().view(np.uint32), g["n_contrib"])
    assert np.abs(out[1].cpu().numpy() - g["color"]).max() < 1e-5
    grads = Hh.ours_backward(args, out, t("in_dL_dpix"))
    names = ["means2D", "opacity", "means3D", "colors", "cov3D"] + ([] if pre else ["sh", "scales", "rotations"])
    for k in names:
        assert Hh.rel_l2(grads[k].cpu().reshape(g["g_" + k].shape), torch.from_numpy(g["g_" + k])) < 1e-4, k


@pytest.mark.gpu
def test_peer_allreduce_two_ranks_exact():
    """b200s_peer_allreduce_sum_f32 over NVLink peer memory, 2 ranks: exact fixed-order sum, identical replicas,
    copy path for non-aliasing gradients (tools/peer_check.py --quick). Needs two GPUs on the box."""
    import subprocess
    import sys
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = Path(__file__).resolve().parent.parent
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533", str(root / "tools" / "peer_check.py"),
                        "--quick"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    rep = json.loads(r.stdout.strip().splitlines()[-1])
    assert rep["stray_copy_path"] and all(rep[f"small_trial{i}"]["exact_fixed_order_sum"] for i in range(3))
    assert all(rep[f"factored_trial{i}"]["exact_fixed_order_sum"] and rep[f"factored_trial{i}"]["replicas_identical"]
               for i in range(3))
