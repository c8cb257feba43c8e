"""Adaptive density control on this repo's parameter layout (the step that changes P between
iterations). Pure, device-agnostic torch functions over a `GaussianState`; FusedGaussianAdam / ViewParallelTrainer
install the result.

Reference (scene/gaussian_model.py): densify_and_prune :535-548, densify_and_clone :517-533, densify_and_split
:490-515, densification_postfix :462-488, cat_tensors_to_optimizer :441-460, prune_points / _prune_optimizer
:407-439, reset_opacity :331-334, build_rotation utils/general_utils.py:78-99. Followed to the letter, including:
  * clone runs first and its postfix ZEROES xyz_gradient_accum / denom / max_radii2D for every point, so the split
    that follows pads the OLD gradients with zeros for the clones, and the screen-size prune test (max_radii2D >
    max_screen_size) can never fire inside densify_and_prune;
  * new points get zero Adam moments, pruned points lose theirs, survivors keep theirs;
  * split children: mean + R(q) (std * N(0,1)), scale / (0.8 N), everything else copied; the parents are pruned.
Replicas of a view-parallel job must take identical decisions: pass the all-reduced statistics
(`parallel.DensificationStats.allreduce`) and the same `unit_normals` / identically seeded generator on every rank.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Optional

import torch

PARAMS = ("xyz", "features", "opacity", "scaling", "rotation")


@dataclass
class GaussianState:
    """Raw parameters + Adam moments + per-Gaussian statistics, all with the Gaussian index as dim 0.
    `extras`: per-Gaussian arrays that are not optimised here (normals, the fork's feature_vector): copied / pruned
    alongside."""
    params: Dict[str, torch.Tensor]
    exp_avg: Dict[str, torch.Tensor]
    exp_avg_sq: Dict[str, torch.Tensor]
    grad_accum: torch.Tensor            # [P,1] xyz_gradient_accum
    denom: torch.Tensor                 # [P,1]
    max_radii2D: torch.Tensor           # [P]
    extras: Dict[str, torch.Tensor] = field(default_factory=dict)

    @property
    def P(self) -> int:
        return self.params["xyz"].shape[0]


def rotation_matrices(q: torch.Tensor) -> torch.Tensor:
    """build_rotation (general_utils.py:78-99): normalises q = (r, x, y, z) and returns [K,3,3]."""
    n = torch.sqrt(q[:, 0] * q[:, 0] + q[:, 1] * q[:, 1] + q[:, 2] * q[:, 2] + q[:, 3] * q[:, 3])
    u = q / n[:, None]
    r, x, y, z = u[:, 0], u[:, 1], u[:, 2], u[:, 3]
    R = torch.zeros((q.size(0), 3, 3), dtype=q.dtype, device=q.device)
    R[:, 0, 0] = 1 - 2 * (y * y + z * z)
    R[:, 0, 1] = 2 * (x * y - r * z)
    R[:, 0, 2] = 2 * (x * z + r * y)
    R[:, 1, 0] = 2 * (x * y + r * z)
    R[:, 1, 1] = 1 - 2 * (x * x + z * z)
    R[:, 1, 2] = 2 * (y * z - r * x)
    R[:, 2, 0] = 2 * (x * z - r * y)
    R[:, 2, 1] = 2 * (y * z + r * x)
    R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


def _append(s: GaussianState, new: Dict[str, torch.Tensor], new_extras: Dict[str, torch.Tensor]) -> GaussianState:
    """densification_postfix + cat_tensors_to_optimizer: concatenate, zero moments for the new rows, reset the stats."""
    params = {k: torch.cat((s.params[k], new[k]), dim=0) for k in PARAMS}
    m = {k: torch.cat((s.exp_avg[k], torch.zeros_like(new[k])), dim=0) for k in PARAMS}
    v = {k: torch.cat((s.exp_avg_sq[k], torch.zeros_like(new[k])), dim=0) for k in PARAMS}
    extras = {k: torch.cat((t, new_extras[k]), dim=0) for k, t in s.extras.items()}
    P = params["xyz"].shape[0]
    z = lambda *shape: torch.zeros(shape, dtype=s.grad_accum.dtype, device=s.grad_accum.device)  # noqa: E731
    return GaussianState(params, m, v, z(P, 1), z(P, 1), z(P), extras)


def prune(s: GaussianState, mask: torch.Tensor) -> GaussianState:
    """prune_points(mask): drop the rows where mask is True (parameters, moments, statistics, extras)."""
    keep = ~mask
    return GaussianState({k: t[keep] for k, t in s.params.items()}, {k: t[keep] for k, t in s.exp_avg.items()},
                         {k: t[keep] for k, t in s.exp_avg_sq.items()}, s.grad_accum[keep], s.denom[keep],
                         s.max_radii2D[keep], {k: t[keep] for k, t in s.extras.items()})


def clone_small(s: GaussianState, grads: torch.Tensor, grad_threshold: float, scene_extent: float,
                percent_dense: float) -> GaussianState:
    """densify_and_clone: under-reconstructed (high view-space gradient) AND small -> duplicate in place."""
    sel = torch.where(torch.norm(grads, dim=-1) >= grad_threshold, True, False)
    sel = torch.logical_and(sel, torch.max(torch.exp(s.params["scaling"]), dim=1).values <= percent_dense * scene_extent)
    return _append(s, {k: s.params[k][sel] for k in PARAMS}, {k: t[sel] for k, t in s.extras.items()})


def split_large(s: GaussianState, grads: torch.Tensor, grad_threshold: float, scene_extent: float,
                percent_dense: float, N: int = 2, unit_normals: Optional[torch.Tensor] = None,
                generator: Optional[torch.Generator] = None) -> GaussianState:
    """densify_and_split: high gradient AND large -> N children sampled inside the parent, parent removed.
    `grads` may be shorter than the state (it predates the clone): the tail counts as zero gradient.
    unit_normals: [N*K,3] standard-normal draws to use instead of the generator (tests, replica consistency)."""
    n0 = s.P
    padded = torch.zeros((n0,), dtype=grads.dtype, device=grads.device)
    padded[:grads.shape[0]] = grads.squeeze()
    scaling = torch.exp(s.params["scaling"])
    sel = torch.where(padded >= grad_threshold, True, False)
    sel = torch.logical_and(sel, torch.max(scaling, dim=1).values > percent_dense * scene_extent)
    stds = scaling[sel].repeat(N, 1)
    if unit_normals is None:
        unit_normals = torch.randn(stds.shape, dtype=stds.dtype, device=stds.device, generator=generator)
    samples = unit_normals.to(stds.dtype) * stds                                  # torch.normal(mean=0, std=stds)
    rots = rotation_matrices(s.params["rotation"][sel]).repeat(N, 1, 1)
    new = {
        "xyz": torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1) + s.params["xyz"][sel].repeat(N, 1),
        "scaling": torch.log(scaling[sel].repeat(N, 1) / (0.8 * N)),
        "rotation": s.params["rotation"][sel].repeat(N, 1),
        "features": s.params["features"][sel].repeat(N, 1, 1),
        "opacity": s.params["opacity"][sel].repeat(N, 1),
    }
    extras = {k: t[sel].repeat(N, *([1] * (t.dim() - 1))) for k, t in s.extras.items()}
    out = _append(s, new, extras)
    parents = torch.cat((sel, torch.zeros(N * int(sel.sum()), dtype=torch.bool, device=sel.device)))
    return prune(out, parents)


def densify_and_prune(s: GaussianState, max_grad: float, min_opacity: float, extent: float,
                      max_screen_size: Optional[float], percent_dense: float = 0.01, N: int = 2,
                      unit_normals: Optional[torch.Tensor] = None,
                      generator: Optional[torch.Generator] = None) -> GaussianState:
    """densify_and_prune (gaussian_model.py:535-548; train.py:119-121 calls it with max_grad =
    densify_grad_threshold, min_opacity 0.005, extent = scene.cameras_extent)."""
    grads = s.grad_accum / s.denom
    grads[grads.isnan()] = 0.0
    s = clone_small(s, grads, max_grad, extent, percent_dense)
    s = split_large(s, grads, max_grad, extent, percent_dense, N, unit_normals, generator)
    mask = (torch.sigmoid(s.params["opacity"]) < min_opacity).squeeze(-1)
    if max_screen_size:
        big_vs = s.max_radii2D > max_screen_size           # all zeros here: the clone's postfix reset them (see module doc)
        big_ws = torch.exp(s.params["scaling"]).max(dim=1).values > 0.1 * extent
        mask = torch.logical_or(torch.logical_or(mask, big_vs), big_ws)
    return prune(s, mask)


def reset_opacity(s: GaussianState) -> GaussianState:
    """reset_opacity (gaussian_model.py:331-334) + replace_tensor_to_optimizer (:393-405): opacity capped at 0.01
    (stored as its logit), the opacity moments zeroed."""
    a = torch.sigmoid(s.params["opacity"])
    capped = torch.min(a, torch.ones_like(a) * 0.01)
    params = dict(s.params); m = dict(s.exp_avg); v = dict(s.exp_avg_sq)
    params["opacity"] = torch.log(capped / (1 - capped))                          # inverse_sigmoid, general_utils.py:18
    m["opacity"] = torch.zeros_like(params["opacity"])
    v["opacity"] = torch.zeros_like(params["opacity"])
    return GaussianState(params, m, v, s.grad_accum, s.denom, s.max_radii2D, s.extras)
