"""ORACLE (test infrastructure only): the reference's photometric loss restated on the CPU with the same torch ops.

  l1            utils/loss_utils.py:17-18
  window        utils/loss_utils.py:23-31   (float32 1-D Gaussian, sigma 1.5, normalised; outer product; one copy per channel)
  ssim          utils/loss_utils.py:34-63   (depthwise conv2d, padding 5; C1 = 0.01^2, C2 = 0.03^2; mean of the map)
  combination   train.py:91-92
`dtype=torch.float64` gives the high-precision value the fp32 implementations are both compared with."""
from math import exp

import torch
import torch.nn.functional as F


def window(channels: int, size: int = 11, sigma: float = 1.5, dtype=torch.float32) -> torch.Tensor:
    g = torch.tensor([exp(-(i - size // 2) ** 2 / float(2 * sigma ** 2)) for i in range(size)], dtype=torch.float32)
    g = (g / g.sum()).unsqueeze(1)
    w2 = g.mm(g.t()).float()
    return w2.expand(channels, 1, size, size).contiguous().to(dtype)


def ssim(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    ch = a.size(-3)
    w = window(ch, dtype=a.dtype).to(a.device)
    conv = lambda t: F.conv2d(t, w, padding=5, groups=ch)  # noqa: E731
    mu_a, mu_b = conv(a), conv(b)
    var_a = conv(a * a) - mu_a.pow(2)
    var_b = conv(b * b) - mu_b.pow(2)
    cov = conv(a * b) - mu_a * mu_b
    c1, c2 = 0.01 ** 2, 0.03 ** 2
    m = ((2 * mu_a * mu_b + c1) * (2 * cov + c2)) / ((mu_a.pow(2) + mu_b.pow(2) + c1) * (var_a + var_b + c2))
    return m.mean()


def l1(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    return (a - b).abs().mean()


def loss_and_grad(image: torch.Tensor, gt: torch.Tensor, lambda_dssim: float = 0.2, dtype=torch.float32):
    """-> (loss, l1, ssim, dloss/dimage) on the CPU in `dtype`."""
    x = image.detach().cpu().to(dtype).requires_grad_(True)
    y = gt.detach().cpu().to(dtype)
    l, s = l1(x, y), ssim(x, y)
    loss = (1.0 - lambda_dssim) * l + lambda_dssim * (1.0 - s)
    loss.backward()
    return loss.detach(), l.detach(), s.detach(), x.grad.detach()
