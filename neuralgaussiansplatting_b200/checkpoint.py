"""`chkpnt<iteration>.pth` in the reference's format (the second data format next to the path).

Reference: train.py:132 saves `(gaussians.capture(), iteration)`; GaussianModel.capture / restore
(scene/gaussian_model.py:75-111) define the 14-tuple
    (active_sh_degree, _xyz, _normals, _features_dc, _features_rest, _features, _scaling, _rotation, _opacity,
     max_radii2D, xyz_gradient_accum, denom, optimizer.state_dict(), spatial_lr_scale)
and the optimizer is torch.optim.Adam over the param groups of training_setup (:257-265) in the order
    xyz, f_dc, f_rest, features, opacity, scaling, rotation   (+ the neural heads, which this repo does not own).
`to_reference_tuple` / `from_reference_tuple` convert between that tuple and this repo's layout (SH coefficients
concatenated [P,16,3], moments per tensor), so a run can be handed over in either direction. Pure torch, any device.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

GROUP_ORDER = ("xyz", "f_dc", "f_rest", "features", "opacity", "scaling", "rotation")


def _adam_group_template(lr: float, name: str, index: int, betas, eps: float) -> Dict:
    """One param_groups entry exactly as the installed torch.optim.Adam serialises it."""
    probe = torch.optim.Adam([{"params": [torch.nn.Parameter(torch.zeros(1))], "lr": lr, "name": name}], lr=0.0,
                             eps=eps, betas=betas)
    g = dict(probe.state_dict()["param_groups"][0])
    g["params"] = [index]
    return g


def to_reference_tuple(raw: Dict[str, torch.Tensor], exp_avg: Dict[str, torch.Tensor], exp_avg_sq: Dict[str, torch.Tensor],
                       step: int, lrs: Dict[str, float], grad_accum: torch.Tensor, denom: torch.Tensor,
                       max_radii2D: Optional[torch.Tensor] = None, normals: Optional[torch.Tensor] = None,
                       feature_vector: Optional[torch.Tensor] = None, active_sh_degree: int = 3,
                       spatial_lr_scale: float = 1.0, betas=(0.9, 0.999), eps: float = 1e-15) -> Tuple:
    """capture() (gaussian_model.py:75-91) from this repo's tensors. `raw` / `exp_avg` / `exp_avg_sq`: dicts keyed
    xyz, features[P,16,3], opacity, scaling, rotation (FusedGaussianAdam.raw / .exp_avg / .exp_avg_sq).
    `lrs`: xyz, f_dc, f_rest, opacity, scaling, rotation (+ optional "features" for the 64-d descriptor group)."""
    P = raw["xyz"].shape[0]
    dev = raw["xyz"].device
    P_ = torch.nn.Parameter
    f = raw["features"]
    normals = torch.zeros(P, 3, device=dev) if normals is None else normals
    fv = torch.zeros(P, 64, device=dev) if feature_vector is None else feature_vector
    tensors = {"xyz": raw["xyz"], "f_dc": f[:, :1].contiguous(), "f_rest": f[:, 1:].contiguous(), "features": fv,
               "opacity": raw["opacity"], "scaling": raw["scaling"], "rotation": raw["rotation"]}
    def group_moments(d: Dict[str, torch.Tensor], group: str):
        if group == "f_dc":
            return d["features"][:, :1].contiguous()
        if group == "f_rest":
            return d["features"][:, 1:].contiguous()
        if group == "features":                              # the fork's 64-d descriptor: not optimised by this repo
            return None
        return d[group]

    state, groups = {}, []
    for i, name in enumerate(GROUP_ORDER):
        groups.append(_adam_group_template(float(lrs.get(name, lrs.get("f_dc", 0.0))), name, i, betas, eps))
        m, v = group_moments(exp_avg, name), group_moments(exp_avg_sq, name)
        if m is not None and step > 0:                      # torch keeps no state for a parameter that never stepped
            state[i] = {"step": torch.tensor(float(step)), "exp_avg": m.clone(), "exp_avg_sq": v.clone()}
    opt_dict = {"state": state, "param_groups": groups}
    max_radii2D = torch.zeros(P, device=dev) if max_radii2D is None else max_radii2D
    return (int(active_sh_degree), P_(tensors["xyz"].clone()), P_(normals.clone()), P_(tensors["f_dc"].clone()),
            P_(tensors["f_rest"].clone()), P_(fv.clone()), P_(tensors["scaling"].clone()), P_(tensors["rotation"].clone()),
            P_(tensors["opacity"].clone()), max_radii2D.clone(), grad_accum.clone(), denom.clone(), opt_dict,
            float(spatial_lr_scale))


def from_reference_tuple(model_args: Tuple) -> Dict:
    """restore() (gaussian_model.py:93-111) into this repo's layout:
    {"raw", "exp_avg", "exp_avg_sq" (dicts keyed xyz/features/opacity/scaling/rotation), "step", "lrs", "normals",
     "feature_vector", "max_radii2D", "grad_accum", "denom", "active_sh_degree", "spatial_lr_scale"}."""
    (sh_degree, xyz, normals, f_dc, f_rest, fv, scaling, rotation, opacity, max_radii2D, grad_accum, denom, opt_dict,
     spatial_lr_scale) = model_args
    d = lambda t: t.detach().clone()  # noqa: E731
    raw = {"xyz": d(xyz), "features": torch.cat((d(f_dc), d(f_rest)), dim=1).contiguous(), "opacity": d(opacity),
           "scaling": d(scaling), "rotation": d(rotation)}
    by_name = {g["name"]: g for g in opt_dict["param_groups"] if "name" in g}
    st = opt_dict["state"]

    def moments(name, like, key):
        g = by_name.get(name)
        s = st.get(g["params"][0]) if g is not None else None
        return d(s[key]) if s is not None else torch.zeros_like(like)

    exp_avg, exp_avg_sq, step = {}, {}, 0
    for name in ("xyz", "opacity", "scaling", "rotation"):
        exp_avg[name] = moments(name, raw[name], "exp_avg")
        exp_avg_sq[name] = moments(name, raw[name], "exp_avg_sq")
    for key, out in (("exp_avg", exp_avg), ("exp_avg_sq", exp_avg_sq)):
        out["features"] = torch.cat((moments("f_dc", f_dc, key), moments("f_rest", f_rest, key)), dim=1).contiguous()
    for name in ("xyz", "f_dc", "opacity"):
        g = by_name.get(name)
        if g is not None and g["params"][0] in st:
            step = max(step, int(float(st[g["params"][0]]["step"])))
    lrs = {name: float(g["lr"]) for name, g in by_name.items() if name in GROUP_ORDER}
    return {"raw": raw, "exp_avg": exp_avg, "exp_avg_sq": exp_avg_sq, "step": step, "lrs": lrs, "normals": d(normals),
            "feature_vector": d(fv), "max_radii2D": d(max_radii2D), "grad_accum": d(grad_accum), "denom": d(denom),
            "active_sh_degree": int(sh_degree), "spatial_lr_scale": float(spatial_lr_scale)}


def save_checkpoint(path: str, model_args: Tuple, iteration: int) -> None:
    """train.py:132."""
    torch.save((model_args, int(iteration)), path)


def load_checkpoint(path: str, map_location="cpu") -> Tuple[Tuple, int]:
    """train.py:38 (weights_only=False: the tuple holds nn.Parameters and an optimizer state dict)."""
    model_args, iteration = torch.load(path, map_location=map_location, weights_only=False)
    return model_args, int(iteration)
