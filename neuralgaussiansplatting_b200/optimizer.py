"""Fused optimizer step for the replicated Gaussian parameters.

Reference behaviour being replaced (for the Gaussian tensors only; the neural heads keep torch.optim):
  * raw parameters `_xyz, _features_dc, _features_rest, _opacity, _scaling, _rotation` and their activations
    (scene/gaussian_model.py:113-141: identity / cat / sigmoid / exp / F.normalize);
  * `torch.optim.Adam(l, lr=0.0, eps=1e-15)` with one param group per tensor (gaussian_model.py:257-279),
    stepped once per iteration (train.py:127-128), `update_learning_rate` rescheduling the xyz group
    (gaussian_model.py:285-291).

Here the five tensors live in the layout the rasterizer consumes (SH coefficients concatenated as [P,16,3], so
`get_features`' torch.cat disappears) and ONE kernel (b200s_adam_step, csrc/optimizer.cu) chains the rasterizer's
gradients through the activations, applies Adam and writes the activated tensors for the next forward.
There is no CPU path: tensors must be CUDA float32.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Dict, Optional

import torch

from . import _lib

# name of the activated tensor (rasterizer argument) -> (raw parameter name, activation mode)
TENSORS = {
    "means3D": ("xyz", "identity"),
    "shs": ("features", "sh"),
    "opacities": ("opacity", "sigmoid"),
    "scales": ("scaling", "exp"),
    "rotations": ("rotation", "normalize"),
}
# reference defaults (arguments/__init__.py OptimizationParams): position_lr_init * spatial_lr_scale is scheduled
DEFAULT_LRS = {"xyz": 0.00016, "f_dc": 0.0025, "f_rest": 0.0025 / 20.0, "opacity": 0.05, "scaling": 0.005,
               "rotation": 0.001}


def activate(name: str, raw: torch.Tensor) -> torch.Tensor:
    """The reference activations (gaussian_model.py:35-43), used once at construction."""
    mode = TENSORS[name][1]
    if mode == "exp":
        return torch.exp(raw)
    if mode == "sigmoid":
        return torch.sigmoid(raw)
    if mode == "normalize":
        return torch.nn.functional.normalize(raw)
    return raw


class FusedGaussianAdam:
    """raw: {"xyz"[P,3], "features"[P,16,3], "opacity"[P,1], "scaling"[P,3], "rotation"[P,4]} CUDA float32 tensors,
    updated in place. `activated` holds the tensors to hand to the rasterizer (means3D aliases xyz and shs
    aliases features: identity activations are not copied)."""

    def __init__(self, raw: Dict[str, torch.Tensor], lrs: Optional[Dict[str, float]] = None, betas=(0.9, 0.999),
                 eps: float = 1e-15):
        self._lib = _lib.load()
        self.raw = {}
        for act_name, (raw_name, _) in TENSORS.items():
            t = raw[raw_name]
            if not t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous():
                raise RuntimeError(f"{raw_name}: FusedGaussianAdam needs contiguous CUDA float32 tensors "
                                   "(b200splat has no CPU path)")
            self.raw[raw_name] = t
        P = self.raw["xyz"].shape[0]
        if tuple(self.raw["features"].shape[1:]) != (16, 3):
            raise RuntimeError("features must be [P,16,3] (f_dc and f_rest concatenated, SH degree 3)")
        for raw_name, width in (("opacity", 1), ("scaling", 3), ("rotation", 4)):
            if tuple(self.raw[raw_name].shape) != (P, width):
                raise RuntimeError(f"{raw_name} must be [{P},{width}]")
        self.lrs = dict(DEFAULT_LRS)
        if lrs:
            self.lrs.update(lrs)
        self.betas, self.eps = (float(betas[0]), float(betas[1])), float(eps)
        self.step_count = 0
        self.exp_avg = {k: torch.zeros_like(v) for k, v in self.raw.items()}
        self.exp_avg_sq = {k: torch.zeros_like(v) for k, v in self.raw.items()}
        self.activated = {}
        for act_name, (raw_name, mode) in TENSORS.items():
            self.activated[act_name] = (self.raw[raw_name] if mode in ("identity", "sh")
                                        else activate(act_name, self.raw[raw_name]).contiguous())

    def set_lr(self, group: str, lr: float) -> None:
        """update_learning_rate (gaussian_model.py:285-291) sets the xyz group; any group may be set."""
        if group not in self.lrs:
            raise KeyError(group)
        self.lrs[group] = float(lr)

    def step(self, grads: Dict[str, torch.Tensor]) -> None:
        """One optimizer.step() (train.py:127) from the rasterizer's gradients w.r.t. the ACTIVATED tensors
        (keys means3D, shs, opacities, scales, rotations — e.g. the views PeerGradientAllReduce.grad_out() handed
        to the backward, after wait())."""
        self.step_count += 1
        b1, b2 = self.betas
        bc1 = 1.0 - b1 ** self.step_count                      # python doubles, as torch.optim.adam does
        bc2_sqrt = math.sqrt(1.0 - b2 ** self.step_count)
        a = _lib.AdamArgs()
        a.count = len(TENSORS)
        a.beta1, a.beta2 = b1, b2
        a.one_minus_beta1, a.one_minus_beta2 = 1.0 - b1, 1.0 - b2
        a.eps, a.bias_correction2_sqrt = self.eps, bc2_sqrt
        dev = self.raw["xyz"].device
        keep = []
        for i, (act_name, (raw_name, mode)) in enumerate(TENSORS.items()):
            g = grads[act_name]
            p = self.raw[raw_name]
            if g.device != dev or g.dtype != torch.float32 or g.numel() != p.numel():
                raise RuntimeError(f"gradient {act_name}: expected float32 on {dev} with {p.numel()} elements")
            g = g.contiguous()
            if g.data_ptr() % 16:
                g = g.clone()
            keep.append(g)
            t = a.t[i]
            t.param, t.exp_avg, t.exp_avg_sq = p.data_ptr(), self.exp_avg[raw_name].data_ptr(), self.exp_avg_sq[raw_name].data_ptr()
            t.grad = g.data_ptr()
            act = self.activated[act_name]
            t.activated = None if act.data_ptr() == p.data_ptr() else act.data_ptr()
            t.n = p.numel()
            t.mode = _lib.ADAM_MODES[mode]
            if mode == "sh":
                t.neg_step, t.neg_step_rest = -(self.lrs["f_dc"] / bc1), -(self.lrs["f_rest"] / bc1)
            else:
                t.neg_step = t.neg_step_rest = -(self.lrs[raw_name] / bc1)
        with torch.cuda.device(dev):
            _lib.check(self._lib.b200s_adam_step(C.byref(a), torch.cuda.current_stream(dev).cuda_stream),
                       "b200s_adam_step")
        del keep

    # ---- adaptive density control (densify.py): hand the tensors out as a GaussianState, take a resized one back
    def densify_state(self, grad_accum: torch.Tensor, denom: torch.Tensor, max_radii2D: Optional[torch.Tensor] = None,
                      extras: Optional[Dict[str, torch.Tensor]] = None):
        from .densify import GaussianState
        P = self.raw["xyz"].shape[0]
        if max_radii2D is None:
            max_radii2D = torch.zeros(P, device=self.raw["xyz"].device)
        return GaussianState(dict(self.raw), dict(self.exp_avg), dict(self.exp_avg_sq), grad_accum, denom, max_radii2D,
                             dict(extras or {}))

    def load_densified(self, state) -> None:
        """Install the tensors of a GaussianState whose P may differ (clone / split / prune, reset_opacity): raw
        parameters and moments are replaced, the activated tensors recomputed. step_count is kept: the reference keeps
        each survivor's Adam `step` and new rows join with zero moments (cat_tensors_to_optimizer)."""
        for name in self.raw:
            self.raw[name] = state.params[name].contiguous()
            self.exp_avg[name] = state.exp_avg[name].contiguous()
            self.exp_avg_sq[name] = state.exp_avg_sq[name].contiguous()
        for act_name, (raw_name, mode) in TENSORS.items():
            self.activated[act_name] = (self.raw[raw_name] if mode in ("identity", "sh")
                                        else activate(act_name, self.raw[raw_name]).contiguous())

    # ---- interchange with the reference's layout (checkpoints: gaussian_model.py capture()/restore())
    def reference_tensors(self) -> Dict[str, torch.Tensor]:
        """{_xyz, _features_dc[P,1,3], _features_rest[P,15,3], _opacity, _scaling, _rotation} copies."""
        f = self.raw["features"]
        return {"_xyz": self.raw["xyz"].clone(), "_features_dc": f[:, :1].clone(), "_features_rest": f[:, 1:].clone(),
                "_opacity": self.raw["opacity"].clone(), "_scaling": self.raw["scaling"].clone(),
                "_rotation": self.raw["rotation"].clone()}

    def reference_state(self) -> Dict[str, Dict[str, torch.Tensor]]:
        """Adam state per reference param-group name (xyz, f_dc, f_rest, opacity, scaling, rotation)."""
        out = {}
        for raw_name in ("xyz", "opacity", "scaling", "rotation"):
            out[raw_name] = {"step": self.step_count, "exp_avg": self.exp_avg[raw_name].clone(),
                             "exp_avg_sq": self.exp_avg_sq[raw_name].clone()}
        m, v = self.exp_avg["features"], self.exp_avg_sq["features"]
        out["f_dc"] = {"step": self.step_count, "exp_avg": m[:, :1].clone(), "exp_avg_sq": v[:, :1].clone()}
        out["f_rest"] = {"step": self.step_count, "exp_avg": m[:, 1:].clone(), "exp_avg_sq": v[:, 1:].clone()}
        return out
