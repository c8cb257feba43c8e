"""ORACLE (test infrastructure only): the reference's adaptive density control driven through a REAL torch.optim.Adam,
as scene/gaussian_model.py does it, on the CPU.

  model tensors / optimizer groups     gaussian_model.py:46-62, :257-279   (f_dc and f_rest are separate groups)
  optimizer-state surgery              :393-405 (replace), :407-425 (prune), :441-460 (cat)
  prune_points / postfix               :427-439, :462-488
  split / clone / densify_and_prune    :490-515, :517-533, :535-548
  reset_opacity                        :331-334
The only liberty: the split's Gaussian noise is injected (`unit_normals`, multiplied by the standard deviations — what
torch.normal(mean=0, std=stds) computes) so that the implementation under test can be given the same draws.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch
from torch import nn

GROUPS = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


def _rot(q):                                                # utils/general_utils.py:78-99
    n = torch.sqrt(q[:, 0] * q[:, 0] + q[:, 1] * q[:, 1] + q[:, 2] * q[:, 2] + q[:, 3] * q[:, 3])
    u = q / n[:, None]
    R = torch.zeros((q.size(0), 3, 3))
    r, x, y, z = u[:, 0], u[:, 1], u[:, 2], u[:, 3]
    R[:, 0, 0] = 1 - 2 * (y * y + z * z); R[:, 0, 1] = 2 * (x * y - r * z); R[:, 0, 2] = 2 * (x * z + r * y)
    R[:, 1, 0] = 2 * (x * y + r * z); R[:, 1, 1] = 1 - 2 * (x * x + z * z); R[:, 1, 2] = 2 * (y * z - r * x)
    R[:, 2, 0] = 2 * (x * z - r * y); R[:, 2, 1] = 2 * (y * z + r * x); R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


class ReferenceDensifier:
    def __init__(self, xyz, features, opacity, scaling, rotation, normals, percent_dense=0.01):
        f = features.detach().clone()
        self.t = {"xyz": nn.Parameter(xyz.detach().clone()), "f_dc": nn.Parameter(f[:, :1].contiguous()),
                  "f_rest": nn.Parameter(f[:, 1:].contiguous()), "opacity": nn.Parameter(opacity.detach().clone()),
                  "scaling": nn.Parameter(scaling.detach().clone()), "rotation": nn.Parameter(rotation.detach().clone())}
        self.normals = normals.detach().clone()
        self.percent_dense = percent_dense
        self.optimizer = torch.optim.Adam([{"params": [self.t[k]], "lr": 1e-3, "name": k} for k in GROUPS], lr=0.0,
                                          eps=1e-15)
        P = xyz.shape[0]
        self.accum, self.denom, self.max_radii2D = torch.zeros(P, 1), torch.zeros(P, 1), torch.zeros(P)

    # a few Adam steps so that the optimizer state exists and is non-trivial
    def warm(self, grads_by_group: Dict[str, torch.Tensor]):
        for k in GROUPS:
            self.t[k].grad = grads_by_group[k].clone()
        self.optimizer.step()
        self.optimizer.zero_grad(set_to_none=True)

    def _swap(self, group, new_tensor, state):
        old = group["params"][0]
        if old in self.optimizer.state:
            del self.optimizer.state[old]
        group["params"][0] = nn.Parameter(new_tensor.requires_grad_(True))
        if state is not None:
            self.optimizer.state[group["params"][0]] = state
        self.t[group["name"]] = group["params"][0]

    def _cat(self, ext: Dict[str, torch.Tensor]):            # :441-460
        for g in self.optimizer.param_groups:
            e = ext[g["name"]]
            st = self.optimizer.state.get(g["params"][0], None)
            if st is not None:
                st["exp_avg"] = torch.cat((st["exp_avg"], torch.zeros_like(e)), dim=0)
                st["exp_avg_sq"] = torch.cat((st["exp_avg_sq"], torch.zeros_like(e)), dim=0)
            self._swap(g, torch.cat((g["params"][0].detach(), e), dim=0), st)

    def _prune(self, mask):                                  # :407-439
        keep = ~mask
        for g in self.optimizer.param_groups:
            st = self.optimizer.state.get(g["params"][0], None)
            if st is not None:
                st["exp_avg"], st["exp_avg_sq"] = st["exp_avg"][keep], st["exp_avg_sq"][keep]
            self._swap(g, g["params"][0].detach()[keep], st)
        self.normals = self.normals[keep]
        self.accum, self.denom, self.max_radii2D = self.accum[keep], self.denom[keep], self.max_radii2D[keep]

    def _postfix(self, ext, new_normals):                    # :462-488
        self._cat(ext)
        self.normals = torch.cat((self.normals, new_normals), dim=0)
        P = self.t["xyz"].shape[0]
        self.accum, self.denom, self.max_radii2D = torch.zeros(P, 1), torch.zeros(P, 1), torch.zeros(P)

    def _scaling(self):
        return torch.exp(self.t["scaling"].detach())

    def clone(self, grads, thr, extent):                     # :517-533
        sel = torch.where(torch.norm(grads, dim=-1) >= thr, True, False)
        sel = torch.logical_and(sel, torch.max(self._scaling(), dim=1).values <= self.percent_dense * extent)
        self._postfix({k: self.t[k].detach()[sel] for k in GROUPS}, self.normals[sel])

    def split(self, grads, thr, extent, unit_normals, N=2):  # :490-515
        n0 = self.t["xyz"].shape[0]
        padded = torch.zeros(n0)
        padded[:grads.shape[0]] = grads.squeeze()
        sel = torch.where(padded >= thr, True, False)
        sel = torch.logical_and(sel, torch.max(self._scaling(), dim=1).values > self.percent_dense * extent)
        stds = self._scaling()[sel].repeat(N, 1)
        samples = unit_normals * stds
        rots = _rot(self.t["rotation"].detach()[sel]).repeat(N, 1, 1)
        ext = {"xyz": torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1) + self.t["xyz"].detach()[sel].repeat(N, 1),
               "scaling": torch.log(self._scaling()[sel].repeat(N, 1) / (0.8 * N)),
               "rotation": self.t["rotation"].detach()[sel].repeat(N, 1),
               "f_dc": self.t["f_dc"].detach()[sel].repeat(N, 1, 1),
               "f_rest": self.t["f_rest"].detach()[sel].repeat(N, 1, 1),
               "opacity": self.t["opacity"].detach()[sel].repeat(N, 1)}
        self._postfix(ext, self.normals[sel].repeat(N, 1))
        self._prune(torch.cat((sel, torch.zeros(N * int(sel.sum()), dtype=torch.bool))))

    def split_count(self, grads, thr, extent) -> int:
        """How many parents the split after a clone would select (to size `unit_normals`)."""
        n0 = self.t["xyz"].shape[0]
        padded = torch.zeros(n0); padded[:grads.shape[0]] = grads.squeeze()
        sel = torch.logical_and(padded >= thr, torch.max(self._scaling(), dim=1).values > self.percent_dense * extent)
        return int(sel.sum())

    def densify_and_prune(self, max_grad, min_opacity, extent, max_screen_size, unit_normals_fn):   # :535-548
        grads = self.accum / self.denom
        grads[grads.isnan()] = 0.0
        self.clone(grads, max_grad, extent)
        self.split(grads, max_grad, extent, unit_normals_fn(self.split_count(grads, max_grad, extent)))
        mask = (torch.sigmoid(self.t["opacity"].detach()) < min_opacity).squeeze()
        if max_screen_size:
            big_vs = self.max_radii2D > max_screen_size
            big_ws = self._scaling().max(dim=1).values > 0.1 * extent
            mask = torch.logical_or(torch.logical_or(mask, big_vs), big_ws)
        self._prune(mask)

    def reset_opacity(self):                                 # :331-334 + :393-405
        a = torch.sigmoid(self.t["opacity"].detach())
        new = torch.log(torch.min(a, torch.ones_like(a) * 0.01) / (1 - torch.min(a, torch.ones_like(a) * 0.01)))
        for g in self.optimizer.param_groups:
            if g["name"] == "opacity":
                st = self.optimizer.state.get(g["params"][0], None)
                st["exp_avg"], st["exp_avg_sq"] = torch.zeros_like(new), torch.zeros_like(new)
                self._swap(g, new, st)

    # ---- views in the layout of the implementation under test
    def params(self) -> Dict[str, torch.Tensor]:
        d = {k: self.t[k].detach() for k in ("xyz", "opacity", "scaling", "rotation")}
        d["features"] = torch.cat((self.t["f_dc"].detach(), self.t["f_rest"].detach()), dim=1)
        return d

    def moments(self, key: str) -> Dict[str, torch.Tensor]:
        st = self.optimizer.state
        d = {k: st[self.t[k]][key] for k in ("xyz", "opacity", "scaling", "rotation")}
        d["features"] = torch.cat((st[self.t["f_dc"]][key], st[self.t["f_rest"]][key]), dim=1)
        return d
