"""Factored SH gradient (row (e) of SURVEY.md §8): b200s_backward's 3-float factors + b200s_sh_grad_reconstruct must
reproduce, BIT FOR BIT, the fixed-order sum over views of the dL/dsh rows the dense backward writes (the quantity the
two-shot all-reduce of the materialised rows delivers). Runs on one GPU: the views of a multi-GPU step are emulated by
filling the gather slots locally, so the driver's 1-GPU test lease covers the arithmetic of the collective."""
import ctypes as C

import pytest
import torch

import helpers as Hh
from neuralgaussiansplatting_b200 import _C, _lib
from neuralgaussiansplatting_b200.synthetic import make_camera, make_grad_image

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda:0")


def _views_backward(cfg_name, dev, views, degree, factored_slots=None, slot_floats=0):
    cfg, sc, _ = Hh.scene_on(cfg_name, dev)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    outs = []
    for i, v in enumerate(views):
        cam = make_camera(cfg, v, device=dev)
        dL = make_grad_image(cfg, v, device=dev)
        args = Hh.fwd_args(cfg, sc, cam, bg, degree=degree)
        out = _C.rasterize_gaussians(*args)
        if factored_slots is not None:
            P = sc["means3D"].shape[0]
            slot = factored_slots[i * slot_floats:(i + 1) * slot_floats]
            _C.config["grad_out"] = lambda slot=slot: {"sh_factors": slot}
        try:
            g = Hh.ours_backward(args, out, dL)
        finally:
            _C.config["grad_out"] = None
        outs.append(g)
    return cfg, sc, outs


@pytest.mark.parametrize("cfg_name,degree,views", [("T1", 3, [0, 3, 5]), ("T0", 3, [0, 1]), ("T1", 1, [2, 6]),
                                                    ("C1", 3, [0, 1, 2, 3, 4, 5, 6, 7])])
def test_reconstruct_equals_fixed_order_sum_of_dense_rows(dev, cfg_name, degree, views):
    lib = _lib.load()
    _C.config["deterministic"] = True         # bit-reproducible local gradients: both backward runs see equal dL/dRGB
    try:
        cfg, sc, dense = _views_backward(cfg_name, dev, views, degree)
        P = sc["means3D"].shape[0]
        slot = -(-(3 * P + 4) // 64) * 64
        factors = torch.full((len(views) * slot,), float("nan"), device=dev)
        _, _, fact = _views_backward(cfg_name, dev, views, degree, factors, slot)
    finally:
        _C.config["deterministic"] = False
    # every other gradient is unaffected by the factored output
    for a, b in zip(dense, fact):
        for k in ("means2D", "opacity", "means3D", "scales", "rotations"):
            assert torch.equal(a[k], b[k]), k
    expect = dense[0]["sh"].clone()
    for g in dense[1:]:
        expect += g["sh"]
    got = torch.full_like(expect, float("nan"))
    a = _lib.ShGradArgs()
    a.P, a.D, a.M, a.views = P, degree, 16, len(views)
    a.means3D, a.factors, a.stride_floats, a.dL_dsh = sc["means3D"].data_ptr(), factors.data_ptr(), slot, got.data_ptr()
    _lib.check(lib.b200s_sh_grad_reconstruct(C.byref(a), torch.cuda.current_stream().cuda_stream), "reconstruct")
    torch.cuda.synchronize()
    assert torch.isfinite(got).all()
    assert float(expect.abs().max()) > 0
    assert torch.equal(got, expect), float((got - expect).abs().max())
    # fast form (fused multiply-adds, rsqrt direction; what the in-switch collective uses): same sum to rounding
    fastv = torch.full_like(expect, float("nan"))
    a.fast, a.dL_dsh = 1, fastv.data_ptr()
    _lib.check(lib.b200s_sh_grad_reconstruct(C.byref(a), torch.cuda.current_stream().cuda_stream), "reconstruct fast")
    torch.cuda.synchronize()
    assert torch.isfinite(fastv).all()
    assert float((fastv - expect).abs().max()) <= 2e-6 * float(expect.abs().max()), float((fastv - expect).abs().max())
    assert torch.equal(fastv == 0, expect == 0)          # culled rows stay exactly zero
    # the factor block carries the view's camera centre behind the 3P factors
    for i, v in enumerate(views):
        cam = make_camera(cfg, v, device=dev)
        assert torch.equal(factors[i * slot + 3 * P:i * slot + 3 * P + 3], cam.camera_center)


def test_reconstruct_argument_checks(dev):
    lib = _lib.load()
    a = _lib.ShGradArgs()
    a.P, a.D, a.M, a.views = 8, 3, 16, 0
    assert lib.b200s_sh_grad_reconstruct(C.byref(a), None) == -1
    a.views = 1
    assert lib.b200s_sh_grad_reconstruct(C.byref(a), None) == -1        # NULL pointers
    t = torch.zeros(64, device=dev)
    a.means3D, a.factors, a.dL_dsh, a.stride_floats = t.data_ptr(), t.data_ptr(), t.data_ptr(), 8
    assert lib.b200s_sh_grad_reconstruct(C.byref(a), None) == -1        # stride < 3P + 4
