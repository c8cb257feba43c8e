"""ORACLE (test infrastructure only — imported by tests/ and nothing else): the reference's optimizer path for the
Gaussian tensors, restated with the reference's own building blocks on the CPU.

It follows, line by line:
  * raw parameters and activations           scene/gaussian_model.py:35-43 (setup_functions), :113-141 (get_*)
  * torch.cat of f_dc / f_rest (get_features) scene/gaussian_model.py:130-133
  * param groups + Adam(lr=0.0, eps=1e-15)    scene/gaussian_model.py:257-279
  * loss.backward(); optimizer.step(); optimizer.zero_grad(set_to_none=True)   train.py:93,127-128
The "loss" is replaced by its vector-Jacobian product: the rasterizer's gradients w.r.t. the ACTIVATED tensors are
fed to torch.autograd.backward, which is exactly what loss.backward() propagates at that point of the graph.
Pinned: this module IS torch autograd + torch.optim.Adam, i.e. the reference's implementation, not a rewrite.
"""
from __future__ import annotations

from typing import Dict

import torch


class ReferenceGaussianOptimizer:
    def __init__(self, xyz, features, opacity, scaling, rotation, lrs: Dict[str, float], betas=(0.9, 0.999),
                 eps: float = 1e-15):
        P = torch.nn.Parameter
        self._xyz = P(xyz.detach().clone().cpu())
        f = features.detach().clone().cpu()
        self._features_dc = P(f[:, :1].contiguous())
        self._features_rest = P(f[:, 1:].contiguous())
        self._opacity = P(opacity.detach().clone().cpu())
        self._scaling = P(scaling.detach().clone().cpu())
        self._rotation = P(rotation.detach().clone().cpu())
        groups = [
            {"params": [self._xyz], "lr": lrs["xyz"], "name": "xyz"},
            {"params": [self._features_dc], "lr": lrs["f_dc"], "name": "f_dc"},
            {"params": [self._features_rest], "lr": lrs["f_rest"], "name": "f_rest"},
            {"params": [self._opacity], "lr": lrs["opacity"], "name": "opacity"},
            {"params": [self._scaling], "lr": lrs["scaling"], "name": "scaling"},
            {"params": [self._rotation], "lr": lrs["rotation"], "name": "rotation"},
        ]
        self.optimizer = torch.optim.Adam(groups, lr=0.0, eps=eps, betas=betas)

    # gaussian_model.py:113-141
    def activated(self) -> Dict[str, torch.Tensor]:
        return {
            "means3D": self._xyz,
            "shs": torch.cat((self._features_dc, self._features_rest), dim=1),
            "opacities": torch.sigmoid(self._opacity),
            "scales": torch.exp(self._scaling),
            "rotations": torch.nn.functional.normalize(self._rotation),
        }

    def set_lr(self, name: str, lr: float) -> None:           # gaussian_model.py:285-291
        for g in self.optimizer.param_groups:
            if g["name"] == name:
                g["lr"] = lr

    def step(self, grads: Dict[str, torch.Tensor]) -> None:
        act = self.activated()
        names = list(act)
        torch.autograd.backward([act[k] for k in names], [grads[k].detach().cpu().reshape(act[k].shape) for k in names])
        self.optimizer.step()
        self.optimizer.zero_grad(set_to_none=True)

    def raw(self) -> Dict[str, torch.Tensor]:
        return {"xyz": self._xyz.detach(), "features": torch.cat((self._features_dc, self._features_rest), 1).detach(),
                "opacity": self._opacity.detach(), "scaling": self._scaling.detach(),
                "rotation": self._rotation.detach()}

    def moments(self) -> Dict[str, Dict[str, torch.Tensor]]:
        st = self.optimizer.state
        cat = lambda key: torch.cat((st[self._features_dc][key], st[self._features_rest][key]), 1)  # noqa: E731
        out = {"features": {"exp_avg": cat("exp_avg"), "exp_avg_sq": cat("exp_avg_sq")}}
        for name, p in (("xyz", self._xyz), ("opacity", self._opacity), ("scaling", self._scaling),
                        ("rotation", self._rotation)):
            out[name] = {"exp_avg": st[p]["exp_avg"], "exp_avg_sq": st[p]["exp_avg_sq"]}
        return out
