"""Fused photometric loss of the training step, drop-in for the expression
`(1.0 - opt.lambda_dssim) * l1_loss(image, gt_image) + opt.lambda_dssim * (1.0 - ssim(image, gt_image))`
(train.py:91-92, utils/loss_utils.py:17-63). One call computes the value and dL/dimage with two kernels of
libb200splat.so (b200s_l1_ssim_loss, csrc/loss.cu) instead of five cuDNN depthwise convolutions, ~15 elementwise
kernels and their autograd mirror. CUDA float32 only; there is no CPU path."""
from __future__ import annotations

import ctypes as C
from typing import Tuple

import torch

from . import _lib


def fused_l1_ssim(image: torch.Tensor, gt: torch.Tensor, lambda_dssim: float = 0.2, want_grad: bool = True,
                  grad_scale: float = 1.0) -> Tuple[torch.Tensor, torch.Tensor]:
    """-> (float32[3] = {loss, l1, ssim}, dL/dimage [C,H,W] or None). No host synchronisation."""
    if not image.is_cuda or not gt.is_cuda:
        raise RuntimeError("b200splat loss kernels are CUDA-only (sm_100a); there is no CPU path")
    if image.dtype != torch.float32 or gt.dtype != torch.float32 or image.shape != gt.shape or image.dim() != 3:
        raise RuntimeError("image and gt must be float32 tensors of the same [C,H,W] shape")
    lib = _lib.load()
    dev = image.device
    Cn, H, W = image.shape
    with torch.cuda.device(dev):
        x, y = image.detach().contiguous(), gt.detach().contiguous()
        out = torch.empty(3, dtype=torch.float32, device=dev)
        grad = torch.empty_like(x) if want_grad else None
        nbytes = lib.b200s_loss_workspace_size(Cn, H, W)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        a = _lib.LossArgs()
        a.C, a.H, a.W = Cn, H, W
        a.lambda_dssim, a.grad_scale = float(lambda_dssim), float(grad_scale)
        a.image, a.gt, a.loss_out = x.data_ptr(), y.data_ptr(), out.data_ptr()
        a.dL_dimage = grad.data_ptr() if want_grad else None
        a.ws, a.ws_bytes = ws.data_ptr(), nbytes
        _lib.check(lib.b200s_l1_ssim_loss(C.byref(a), torch.cuda.current_stream(dev).cuda_stream),
                   "b200s_l1_ssim_loss")
    return out, grad


class _L1SSIM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, gt, lambda_dssim):
        out, grad = fused_l1_ssim(image, gt, lambda_dssim, want_grad=image.requires_grad)
        ctx.save_for_backward(grad)
        return out[0]

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return grad * g, None, None


def l1_ssim_loss(image: torch.Tensor, gt: torch.Tensor, lambda_dssim: float = 0.2) -> torch.Tensor:
    """Scalar loss tensor with autograd to `image` (gt is treated as a constant, as in train.py)."""
    return _L1SSIM.apply(image, gt, float(lambda_dssim))


class _Component(torch.autograd.Function):
    """One term of the photometric loss as a differentiable scalar: which = 1 -> L1, which = 2 -> SSIM. The fused
    kernels emit d((1-lambda) L1 + lambda (1-SSIM))/dimage; lambda = 0 gives dL1/dimage, lambda = 1 gives -dSSIM/dimage."""

    @staticmethod
    def forward(ctx, image, gt, which):
        out, grad = fused_l1_ssim(image, gt, 0.0 if which == 1 else 1.0, want_grad=image.requires_grad)
        ctx.save_for_backward(grad)
        ctx.sign = 1.0 if which == 1 else -1.0
        return out[which]

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return grad * (g * ctx.sign), None, None


def l1_loss(network_output: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    """utils/loss_utils.py:17-18, differentiable w.r.t. the first argument like the reference's (the second is a
    constant, as in train.py:91). For the training step prefer l1_ssim_loss: one fused call for both terms."""
    return _Component.apply(network_output, gt, 1)


def ssim(img1: torch.Tensor, img2: torch.Tensor) -> torch.Tensor:
    """utils/loss_utils.py:34-43 with the defaults window_size=11, size_average=True; differentiable w.r.t. img1."""
    return _Component.apply(img1, img2, 2)
