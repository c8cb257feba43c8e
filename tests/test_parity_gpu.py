"""Parity of the CUDA path (through the reference-shaped `_C` surface over the C ABI) against the reference
rasterizer compiled for sm_100 and run on the same inputs in the same process (north_star):
  bit-exact : radii, tiles_touched, sorted keys (+values), per-tile ranges, n_contrib
  forward   : image within 1e-5 max-abs (fp32)
  gradients : within 1e-4 relative L2; bit-reproducible in deterministic mode
and against the CPU oracle where the reference library is not present."""
import math

import numpy as np
import pytest
import torch

import helpers as Hh
import oracle_cpu
import refbind
from neuralgaussiansplatting_b200 import _C
from neuralgaussiansplatting_b200.synthetic import make_grad_image

pytestmark = pytest.mark.gpu

IMG_TOL = 1e-5        # north_star: forward image max-abs, fp32
GRAD_TOL = 1e-4       # north_star: gradients, relative L2
MAX_REF_SPREAD = 5e-5 # largest run-to-run rel-L2 of the reference's own atomics seen (4e-5, profiles/r02j_grad_noise.jsonl)
BITEXACT = ("radii", "tiles_touched", "point_offsets", "point_list_keys", "point_list", "ranges", "n_contrib",
            "depth_bits", "means2D_bits", "conic_bits")

needs_ref = pytest.mark.skipif(not refbind.available("ref_dgr"), reason="oracle/_ref not built")


@pytest.fixture(scope="module")
def ref():
    r = refbind.RefDGR()
    yield r
    r.close()


def check_against_reference(ref, cfg_name, dev, P=None, view=0, degree=3, precomp=False, scale_modifier=1.0,
                            bg=(0.1, 0.2, 0.3), grads=True, sh_coeffs=None):
    cfg, sc, cam = Hh.scene_on(cfg_name, dev, P=P, view=view)
    if sh_coeffs is not None:
        sc = dict(sc, shs=sc["shs"][:, :sh_coeffs].contiguous())
    bgt = torch.tensor(bg, device=dev, dtype=torch.float32)
    kwv = {}
    if precomp:
        cov, col = Hh.cpu_covariance_and_colors(sc, cam, degree)
        kwv = dict(colors_precomp=col, cov3D_precomp=cov)
    args, out, state = Hh.ours_forward(cfg, sc, cam, bgt, degree=degree, scale_modifier=scale_modifier, **kwv)
    kw, (R, rcolor, rradii) = Hh.ref_forward(ref, cfg, sc, cam, bgt, degree=degree, scale_modifier=scale_modifier,
                                             **kwv)
    assert out[0] == R
    res = Hh.compare_forward_state(state, ref, out[2], rradii)
    for k in BITEXACT:
        assert res[k] == 0, (k, res)
    if not precomp:
        assert res["clamped"] == 0 and res["rgb_maxabs"] == 0.0, res
    assert res["final_T_maxabs"] < IMG_TOL
    assert float((out[1] - rcolor).abs().max()) < IMG_TOL
    if grads:
        dL = make_grad_image(cfg, view, device=dev)
        g_o = Hh.ours_backward(args, out, dL)
        g_r = Hh.ref_backward(ref, kw, rradii, dL)
        g_r2 = Hh.ref_backward(ref, kw, rradii, dL)
        # the gradients the autograd binding hands back to the caller for this input variant
        # (dL/dcov3D and dL/dcolors are internal intermediates when scales/rotations/SHs are given)
        names = ["means2D", "opacity", "means3D"]
        names += ["colors", "cov3D"] if precomp else ["sh", "scales", "rotations"]
        for k in names:
            # The reference sums its per-pixel contributions with float atomics, so it only defines its gradients
            # up to its own run-to-run spread (up to 4e-5 on dL/dscales / dL/drotations of the big-Gaussian T1
            # views, profiles/r02j_grad_noise.jsonl; this repo's own spread is 3x smaller). That spread, measured
            # here from a second reference run, is added to the 1e-4 of the north star: without it the comparison
            # fails about one run in ten on T1 view 3 (seen at 1.01e-4).
            spread = Hh.rel_l2(g_r2[k], g_r[k])
            assert spread <= MAX_REF_SPREAD, (k, spread)     # the allowance cannot widen silently
            assert Hh.rel_l2(g_o[k].reshape(g_r[k].shape), g_r[k]) < GRAD_TOL + spread, (k, spread)
    return res


@needs_ref
@pytest.mark.parametrize("cfg_name", ["T0", "T1", "C1"])
def test_forward_backward_parity(ref, dev, cfg_name):
    check_against_reference(ref, cfg_name, dev)


@needs_ref
def test_parity_C2_training_shape(ref, dev):
    check_against_reference(ref, "C2", dev)


@needs_ref
@pytest.mark.parametrize("degree", [0, 1, 2])
def test_lower_sh_degrees(ref, dev, degree):
    check_against_reference(ref, "T1", dev, degree=degree)


@needs_ref
@pytest.mark.parametrize("P", [257, None])
@pytest.mark.parametrize("degree", [0, 1, 2])
def test_sh_tensors_narrower_than_16_coefficients(ref, dev, degree, P):
    """Models trained at --sh_degree 0 / 1 / 2 carry [P, 1 / 4 / 9, 3] SH tensors (reference
    scene/gaussian_model.py:128-131, rasterize_points.cu:68-72 takes M from sh.size(1)). The backward must write
    exactly M rows per Gaussian: anything more lands on the next Gaussian's dL/dsh (and past the end of the
    tensor for the last one). P = 257 leaves a partial block."""
    check_against_reference(ref, "T1", dev, degree=degree, P=P, sh_coeffs=(degree + 1) ** 2)


@needs_ref
def test_precomputed_colors_and_covariance(ref, dev):
    check_against_reference(ref, "T1", dev, precomp=True)


@needs_ref
@pytest.mark.parametrize("mod", [0.5, 2.0])
def test_scale_modifier(ref, dev, mod):
    check_against_reference(ref, "T0", dev, scale_modifier=mod)


@needs_ref
@pytest.mark.parametrize("view", [1, 3, 6])
def test_other_views(ref, dev, view):
    check_against_reference(ref, "T1", dev, view=view)


@needs_ref
@pytest.mark.parametrize("P", [1, 31, 257, 4097])
def test_ragged_point_counts(ref, dev, P):
    check_against_reference(ref, "T0", dev, P=P)


@needs_ref
def test_image_size_not_multiple_of_tile(ref, dev):
    import dataclasses
    from neuralgaussiansplatting_b200.synthetic import CONFIGS
    cfg = dataclasses.replace(CONFIGS["T0"], width=157, height=83)
    check_against_reference(ref, cfg, dev)


@needs_ref
@pytest.mark.parametrize("wh", [(173, 99), (401, 303), (16, 16), (17, 1)])
def test_ragged_image_sizes_forward_and_backward(ref, dev, wh):
    """Image sides that are not multiples of the 16-pixel tile: partial tiles in BOTH blend kernels (out-of-image
    lanes, `inside` masking, block-last over missing pixels), forward state bit-exact and gradients against the
    reference, in the atomic and in the deterministic backward."""
    import dataclasses
    from neuralgaussiansplatting_b200.synthetic import CONFIGS
    cfg = dataclasses.replace(CONFIGS["T1" if wh[0] > 200 else "T0"], width=wh[0], height=wh[1])
    check_against_reference(ref, cfg, dev)
    try:
        _C.config["deterministic"] = True
        check_against_reference(ref, cfg, dev)
    finally:
        _C.config["deterministic"] = False


@needs_ref
@pytest.mark.parametrize("wh", [(4112, 48), (48, 4112)])
def test_grids_wider_than_256_tiles_use_tile_id_keys(ref, dev, wh):
    """More than 256 tile columns (or rows): the (x, y) sort keys do not fit 8 + 8 bits, so the binning falls back to
    tile-id keys on the default path (csrc/binning.cu: launch_binning). Forward state bit-exact, gradients as usual."""
    import dataclasses
    from neuralgaussiansplatting_b200.synthetic import CONFIGS
    cfg = dataclasses.replace(CONFIGS["T0"], width=wh[0], height=wh[1])
    res = check_against_reference(ref, cfg, dev)
    assert res["num_rendered"][0] > 0


@needs_ref
def test_huge_gaussians_cover_every_tile(ref, dev):
    """A few Gaussians whose rectangle is the whole tile grid (the load-balanced expansion path)."""
    cfg, sc, cam = Hh.scene_on("T0", dev, P=64)
    sc["scales"] = sc["scales"] * 60.0
    bg = torch.zeros(3, device=dev)
    args, out, state = Hh.ours_forward(cfg, sc, cam, bg)
    kw, (R, rcolor, rradii) = Hh.ref_forward(ref, cfg, sc, cam, bg)
    res = Hh.compare_forward_state(state, ref, out[2], rradii)
    assert all(res[k] == 0 for k in BITEXACT), res
    assert float((out[1] - rcolor).abs().max()) < IMG_TOL
    assert int(state["tiles_touched"].max()) == 60   # 10 x 6 tiles


@needs_ref
def test_long_tile_runs(ref, dev):
    """60 tiles x ~5,000 entries each: tile lists many blend batches long, several onesweep tiles per digit."""
    res = check_against_reference(ref, "T0", dev, P=40_000, grads=False)
    assert res["num_rendered"][0] > 60 * 4096


def test_empty_input(dev):
    cfg, sc, cam = Hh.scene_on("T0", dev, P=0)
    bg = torch.tensor([0.5, 0.25, 0.125], device=dev)
    out = _C.rasterize_gaussians(*Hh.fwd_args(cfg, sc, cam, bg))
    assert out[0] == 0 and out[1].shape == (3, cfg.height, cfg.width) and out[2].shape == (0,)
    assert float(out[1].abs().max()) == 0.0     # RasterizeGaussiansCUDA returns the zero image when P == 0


def test_everything_culled_gives_background(dev):
    cfg, sc, cam = Hh.scene_on("T0", dev, P=500)
    sc["means3D"] = sc["means3D"] * 0.1 + cam.camera_center[None, :] * 3.0   # all behind the camera
    bg = torch.tensor([0.5, 0.25, 0.125], device=dev)
    args = Hh.fwd_args(cfg, sc, cam, bg)
    out = _C.rasterize_gaussians(*args)
    assert out[0] == 0 and int((out[2] != 0).sum()) == 0
    assert torch.equal(out[1], bg[:, None, None].expand_as(out[1]))
    g = Hh.ours_backward(args, out, make_grad_image(cfg, 0, device=dev))
    assert all(float(v.abs().max()) == 0.0 for v in g.values() if v is not None)


def test_against_cpu_oracle(dev):
    """Same gates against the C restatement (covers boxes where oracle/_ref was not built)."""
    cfg, sc, cam = Hh.scene_on("T1", dev)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    dL = make_grad_image(cfg, 0, device=dev)
    args, out, state = Hh.ours_forward(cfg, sc, cam, bg)
    g = Hh.ours_backward(args, out, dL)
    o = oracle_cpu.render_scene(cfg, {k: v.cpu() for k, v in sc.items()}, cam.to("cpu"), bg.cpu(), dL_dpix=dL.cpu())
    cam.to(dev)
    assert int((o["radii"] != out[2].cpu()).sum()) == 0
    assert int((o["tiles_touched"] != state["tiles_touched"].cpu()).sum()) == 0
    assert o["num_rendered"] == out[0]
    assert int((o["point_list_keys"] != state["point_list_keys"].cpu()).sum()) == 0
    assert int((o["point_list"] != state["point_list"].cpu()).sum()) == 0
    assert int((o["ranges"] != state["ranges"].cpu()).sum()) == 0
    # libm expf vs device expf: isolated threshold pixels may differ
    assert float((o["n_contrib"].reshape(-1) != state["n_contrib"].cpu()).float().mean()) <= 1e-4
    assert float((o["color"] - out[1].cpu()).abs().max()) < IMG_TOL
    for k in ("means2D", "opacity", "means3D", "sh", "scales", "rotations"):
        assert Hh.rel_l2(g[k].cpu().reshape(o["grads"][k].shape), o["grads"][k]) < GRAD_TOL, k


@pytest.mark.parametrize("cfg_name", ["T0", "T1", "C1", "C2"])
def test_deterministic_mode_against_the_order_free_oracle(dev, cfg_name):
    """SURVEY.md §8(c) asks for the deterministic mode against "our own deterministic oracle". A CPU mirror of the
    kernel's summation ORDER cannot be bit-exact: every addend contains the device's expf (range reduction + MUFU.EX2)
    and MUFU.RCP, whose low bits no host libm reproduces. The oracle therefore removes order from the question instead:
    oracle/oracle_cpu.c accumulates every per-Gaussian sum in float64 (error ~1e-16 of the sum, i.e. order-free at
    float32 resolution), and the fixed-order float32 result must sit within DET_TOL = 2e-5 relative L2 of it — 5x
    tighter than the tolerance of the atomic mode, whose order changes run to run."""
    DET_TOL = 2e-5
    cfg, sc, cam = Hh.scene_on(cfg_name, dev)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    dL = make_grad_image(cfg, 0, device=dev)
    args, out, _ = Hh.ours_forward(cfg, sc, cam, bg, names=[])
    try:
        _C.config["deterministic"] = True
        g = Hh.ours_backward(args, out, dL)
    finally:
        _C.config["deterministic"] = False
    o = oracle_cpu.render_scene(cfg, {k: v.cpu() for k, v in sc.items()}, cam.to("cpu"), bg.cpu(), dL_dpix=dL.cpu())
    cam.to(dev)
    for k in ("means2D", "opacity", "means3D", "sh", "scales", "rotations"):
        assert Hh.rel_l2(g[k].cpu().reshape(o["grads"][k].shape), o["grads"][k]) < DET_TOL, k


def test_deterministic_mode_is_bit_reproducible(dev):
    cfg, sc, cam = Hh.scene_on("T1", dev)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    dL = make_grad_image(cfg, 0, device=dev)
    args, out, _ = Hh.ours_forward(cfg, sc, cam, bg, names=[])
    try:
        _C.config["deterministic"] = False
        g_atomic = Hh.ours_backward(args, out, dL)
        _C.config["deterministic"] = True
        runs = [Hh.ours_backward(args, out, dL) for _ in range(3)]
    finally:
        _C.config["deterministic"] = False
    for k, v in runs[0].items():
        if v is None:
            continue
        for r in runs[1:]:
            assert torch.equal(v, r[k]), k
        assert Hh.rel_l2(v, g_atomic[k]) < GRAD_TOL, k


def test_forward_is_bit_reproducible(dev):
    cfg, sc, cam = Hh.scene_on("T1", dev)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    a = _C.rasterize_gaussians(*Hh.fwd_args(cfg, sc, cam, bg))
    b = _C.rasterize_gaussians(*Hh.fwd_args(cfg, sc, cam, bg))
    assert a[0] == b[0] and torch.equal(a[1], b[1]) and torch.equal(a[2], b[2])


def test_optimistic_capacity_overflow_falls_back(dev):
    """The no-host-sync forward sizes the binning workspace from the previous call; a much larger second
    view must still produce the exact result (overflow flag -> re-run with the exact size)."""
    cfg, sc, cam = Hh.scene_on("T0", dev)
    bg = torch.zeros(3, device=dev)
    small = dict(sc); small["scales"] = sc["scales"] * 0.05
    _C._capacity_hint.clear()
    r_small = _C.rasterize_gaussians(*Hh.fwd_args(cfg, small, cam, bg))
    r_big = _C.rasterize_gaussians(*Hh.fwd_args(cfg, sc, cam, bg))       # hint too small -> overflow path
    _C._capacity_hint.clear()
    r_exact = _C.rasterize_gaussians(*Hh.fwd_args(cfg, sc, cam, bg))     # two-stage exact path
    assert r_small[0] < r_big[0] == r_exact[0]
    assert torch.equal(r_big[1], r_exact[1]) and torch.equal(r_big[2], r_exact[2])


@needs_ref
def test_mark_visible(ref, dev):
    cfg, sc, cam = Hh.scene_on("T1", dev)
    ours = _C.mark_visible(sc["means3D"], cam.world_view_transform, cam.full_proj_transform)
    theirs = ref.mark_visible(sc["means3D"], cam.world_view_transform, cam.full_proj_transform)
    assert ours.dtype == torch.bool and torch.equal(ours, theirs)
    assert 0 < int(ours.sum()) < ours.numel()


def test_linearity_of_backward_in_dL(dev):
    """Size-independent property: the backward is linear in dL/dout_color."""
    cfg, sc, cam = Hh.scene_on("C1", dev)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    args, out, _ = Hh.ours_forward(cfg, sc, cam, bg, names=[])
    d1, d2 = make_grad_image(cfg, 0, device=dev), make_grad_image(cfg, 1, device=dev)
    _C.config["deterministic"] = True
    try:
        g1, g2, g12 = (Hh.ours_backward(args, out, d) for d in (d1, d2, d1 + 2.0 * d2))
    finally:
        _C.config["deterministic"] = False
    for k in ("means3D", "opacity", "sh", "scales", "rotations", "means2D"):
        assert Hh.rel_l2(g1[k] + 2.0 * g2[k], g12[k]) < 1e-5, k
