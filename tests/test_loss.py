"""Fused L1 + SSIM loss (csrc/loss.cu) against the reference's loss restated with torch ops on the CPU
(oracle/loss_ref.py). Tolerances: |value| 2e-6 against the float64 oracle (the float32 oracle itself sits ~1e-6 from
it), gradient 1e-4 relative L2 against float64."""
import ctypes as C
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "oracle"))
import loss_ref  # noqa: E402
from neuralgaussiansplatting_b200 import _lib  # noqa: E402
from neuralgaussiansplatting_b200 import loss as L  # noqa: E402


def _pair(C_, H, W, seed, close=True):
    g = torch.Generator().manual_seed(seed)
    gt = torch.rand(C_, H, W, generator=g)
    if close:   # a render that resembles its target, as in training
        img = (gt + 0.1 * torch.randn(C_, H, W, generator=g)).clamp(0, 1)
        img[:, : H // 3] = gt[:, : H // 3].roll(1, -1)
    else:
        img = torch.rand(C_, H, W, generator=g)
    return img, gt


def test_oracle_identities():
    img, gt = _pair(3, 40, 52, 0)
    assert abs(float(loss_ref.ssim(gt[None], gt[None])) - 1.0) < 1e-6            # ssim(x, x) == 1
    assert float(loss_ref.l1(gt, gt)) == 0.0
    l32 = loss_ref.loss_and_grad(img, gt, 0.2, torch.float32)
    l64 = loss_ref.loss_and_grad(img, gt, 0.2, torch.float64)
    assert abs(float(l32[0]) - float(l64[0])) < 5e-6
    w = loss_ref.window(3)
    assert w.shape == (3, 1, 11, 11) and abs(float(w[0].sum()) - 1.0) < 1e-6 and torch.equal(w[0], w[2])


GOLDEN = ROOT / "tests" / "golden" / "ref_loss.npz"      # produced by the reference's own utils/loss_utils.py


def test_oracle_reproduces_the_reference_golden():
    import numpy as np
    z = np.load(GOLDEN)
    for name in ("a", "b", "c"):
        img, gt, lam = torch.tensor(z[f"{name}_img"]), torch.tensor(z[f"{name}_gt"]), float(z[f"{name}_lambda"])
        loss, l1, ssim, grad = loss_ref.loss_and_grad(img, gt, lam, torch.float32)
        assert abs(float(l1) - float(z[f"{name}_l1"])) < 1e-7 and abs(float(ssim) - float(z[f"{name}_ssim"])) < 1e-6
        assert abs(float(loss) - float(z[f"{name}_loss"])) < 1e-6
        ref_g = torch.tensor(z[f"{name}_grad"])
        assert float((grad - ref_g).norm() / ref_g.norm()) < 1e-5, name


@pytest.mark.gpu
def test_fused_loss_matches_the_reference_golden():
    import numpy as np
    z = np.load(GOLDEN)
    dev = torch.device("cuda:0")
    for name in ("a", "b", "c"):
        img, gt, lam = torch.tensor(z[f"{name}_img"]), torch.tensor(z[f"{name}_gt"]), float(z[f"{name}_lambda"])
        out, grad = L.fused_l1_ssim(img.to(dev), gt.to(dev), lam)
        assert abs(float(out[0]) - float(z[f"{name}_loss"])) < 5e-6 and abs(float(out[2]) - float(z[f"{name}_ssim"])) < 5e-6
        ref_g = torch.tensor(z[f"{name}_grad"])
        assert float((grad.cpu() - ref_g).norm() / ref_g.norm()) < 1e-4, name


def test_oracle_gradient_matches_finite_differences():
    img, gt = _pair(1, 12, 14, 1)
    _, _, _, g = loss_ref.loss_and_grad(img, gt, 0.2, torch.float64)
    x = img.double()
    for (c, i, j) in [(0, 0, 0), (0, 5, 7), (0, 11, 13)]:
        e = torch.zeros_like(x); e[c, i, j] = 1e-6
        f = lambda t: float((1 - 0.2) * loss_ref.l1(t, gt.double()) + 0.2 * (1 - loss_ref.ssim(t[None], gt.double()[None])))  # noqa: E731
        fd = (f(x + e) - f(x - e)) / 2e-6
        assert abs(fd - float(g[c, i, j])) < 1e-6 * max(1.0, abs(fd) * 1e3)


def test_loss_abi_argument_checks():
    lib = _lib.load()
    assert lib.b200s_loss_workspace_size(0, 4, 4) == 0
    n = lib.b200s_loss_workspace_size(3, 100, 200)
    assert n >= 3 * 3 * 100 * 200 * 4 and n % 256 == 0
    assert lib.b200s_l1_ssim_loss(None, None) == -1
    a = _lib.LossArgs()
    a.C, a.H, a.W = 3, 16, 16
    assert lib.b200s_l1_ssim_loss(C.byref(a), None) == -1                      # NULL images
    a.image = a.gt = a.loss_out = a.ws = 4096
    a.ws_bytes = 16
    assert lib.b200s_l1_ssim_loss(C.byref(a), None) == -2                      # workspace too small
    with pytest.raises(RuntimeError, match="CUDA"):
        L.fused_l1_ssim(torch.zeros(3, 8, 8), torch.zeros(3, 8, 8))


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(3, 16, 32), (3, 37, 53), (1, 5, 7), (3, 200, 333)])
@pytest.mark.parametrize("lam", [0.2, 0.0, 1.0])
def test_fused_loss_matches_reference(shape, lam):
    dev = torch.device("cuda:0")
    img, gt = _pair(*shape, seed=sum(shape))
    out, grad = L.fused_l1_ssim(img.to(dev), gt.to(dev), lam)
    torch.cuda.synchronize()
    loss64, l1_64, ssim64, g64 = loss_ref.loss_and_grad(img, gt, lam, torch.float64)
    assert abs(float(out[0]) - float(loss64)) < 2e-6
    assert abs(float(out[1]) - float(l1_64)) < 2e-6 and abs(float(out[2]) - float(ssim64)) < 2e-6
    rel = float((grad.cpu().double() - g64).norm() / g64.norm().clamp_min(1e-30))
    assert rel < 1e-4, rel
    _, _, _, g32 = loss_ref.loss_and_grad(img, gt, lam, torch.float32)           # no further from truth than torch fp32
    rel32 = float((g32.double() - g64).norm() / g64.norm().clamp_min(1e-30))
    assert rel < max(10 * rel32, 1e-5)


@pytest.mark.gpu
def test_fused_loss_autograd_and_reproducibility():
    dev = torch.device("cuda:0")
    img, gt = _pair(3, 64, 80, 5)
    x = img.to(dev).requires_grad_(True)
    loss = L.l1_ssim_loss(x, gt.to(dev), 0.2)
    (3.0 * loss).backward()
    _, _, _, g64 = loss_ref.loss_and_grad(img, gt, 0.2, torch.float64)
    assert float((x.grad.cpu().double() - 3.0 * g64).norm() / (3.0 * g64).norm()) < 1e-4
    a = L.fused_l1_ssim(x.detach(), gt.to(dev))[0]
    b = L.fused_l1_ssim(x.detach(), gt.to(dev))[0]
    assert torch.equal(a, b)                                                     # fixed-order reduction
    assert abs(float(L.ssim(gt.to(dev), gt.to(dev))) - 1.0) < 1e-6 and float(L.l1_loss(gt.to(dev), gt.to(dev))) == 0.0


def Hh_rel_l2(a, b):
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


@pytest.mark.gpu
def test_reference_style_composition_is_differentiable(dev):
    """train.py:91-92 writes the loss as (1 - l) * l1_loss(image, gt) + l * (1 - ssim(image, gt)) and calls
    .backward(): the drop-in l1_loss / ssim must carry gradients, and agree with the fused call."""
    g = torch.Generator().manual_seed(5)
    img = torch.rand(3, 57, 83, generator=g).to(dev).requires_grad_(True)
    gt = torch.rand(3, 57, 83, generator=g).to(dev)
    lam = 0.2
    loss = (1.0 - lam) * L.l1_loss(img, gt) + lam * (1.0 - L.ssim(img, gt))
    loss.backward()
    out, grad = L.fused_l1_ssim(img.detach(), gt, lam)
    assert abs(float(loss) - float(out[0])) < 2e-6
    assert img.grad is not None and Hh_rel_l2(img.grad, grad) < 1e-5
