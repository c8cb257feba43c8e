"""Generates tests/golden/ref_loss.npz by IMPORTING the reference's own utils/loss_utils.py (pure torch, runs on the CPU
in the build container where /root/reference is mounted). The file pins oracle/loss_ref.py — and through it the CUDA
kernel — to outputs of the reference itself. Run: python tests/golden/make_loss_golden.py"""
import importlib.util
import sys
from pathlib import Path

import numpy as np
import torch

REF = Path("/root/reference/utils/loss_utils.py")
spec = importlib.util.spec_from_file_location("ref_loss_utils", REF)
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

out = {}
for name, (C, H, W, seed, lam) in {"a": (3, 24, 31, 11, 0.2), "b": (3, 40, 16, 12, 0.2), "c": (1, 9, 13, 13, 1.0)}.items():
    g = torch.Generator().manual_seed(seed)
    gt = torch.rand(C, H, W, generator=g)
    img = (gt + 0.15 * torch.randn(C, H, W, generator=g)).clamp(0, 1).requires_grad_(True)
    l1 = ref.l1_loss(img, gt)
    ss = ref.ssim(img, gt)
    loss = (1.0 - lam) * l1 + lam * (1.0 - ss)                 # train.py:92
    loss.backward()
    out.update({f"{name}_img": img.detach().numpy(), f"{name}_gt": gt.numpy(), f"{name}_lambda": np.float32(lam),
                f"{name}_l1": l1.detach().numpy(), f"{name}_ssim": ss.detach().numpy(), f"{name}_loss": loss.detach().numpy(),
                f"{name}_grad": img.grad.numpy()})
np.savez_compressed(Path(__file__).resolve().parent / "ref_loss.npz", **out)
print("wrote ref_loss.npz", {k: v.shape for k, v in out.items() if k.endswith("_grad")})
