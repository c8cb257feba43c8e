"""View-parallel training iteration built from this repo's kernels only (the analogue of
train.py:75-128 for the Gaussian tensors, one process per GPU).

One step on every rank, for the view(s) that rank owns (`parallel.shard_views`):

    forward          _C.rasterize_gaussians on the activated tensors the optimizer maintains        (K1-K6)
    loss + dL/dimage loss.fused_l1_ssim                                   (train.py:91-92, two kernels)
    backward         _C.rasterize_gaussians_backward, gradients written into the peer-mapped buffer  (K7-K9)
    all-reduce       parallel.PeerGradientAllReduce (one kernel over NVLink; NCCL if unavailable)    (§8e)
    optimizer        optimizer.FusedGaussianAdam.step on the summed gradients                        (train.py:127)

Gradients of the ranks' views are SUMMED (the all-reduce the north star names); every rank then applies the identical
Adam update to its replica, so replicas stay bit-identical (the peer kernel delivers the same bits to every rank) —
`replicas_in_sync()` checks it. Densification statistics are accumulated per view (`parallel.DensificationStats`).
The neural heads, dataset readers and densification of the reference are not part of this path.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch
import torch.distributed as dist

from . import _C, parallel
from .loss import fused_l1_ssim
from .optimizer import FusedGaussianAdam

_GRAD_SLOTS = {"means3D": 3, "opacities": 2, "shs": 5, "scales": 6, "rotations": 7}   # slots of the backward's 8-tuple


class ViewParallelTrainer:
    def __init__(self, raw: Dict[str, torch.Tensor], lrs: Optional[Dict[str, float]] = None, lambda_dssim: float = 0.2,
                 background: Optional[torch.Tensor] = None, sh_degree: int = 3, group=None,
                 collective: str = "peer"):
        """raw: {"xyz","features"[P,16,3],"opacity","scaling","rotation"} CUDA float32 (updated in place)."""
        self.opt = FusedGaussianAdam(raw, lrs)
        self.dev = raw["xyz"].device
        self.P = raw["xyz"].shape[0]
        self.lambda_dssim, self.sh_degree = float(lambda_dssim), int(sh_degree)
        self.bg = background if background is not None else torch.zeros(3, device=self.dev)
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.reducer = None
        self.collective = "none"
        if self.world > 1 and collective != "none":        # "none": a private replica (e.g. rendering targets)
            if collective == "peer":
                try:
                    self.reducer = parallel.PeerGradientAllReduce(self.P, 16, self.dev, group=group)
                    self.collective = "peer"
                except Exception:  # noqa: BLE001 — no peer mappings on this box
                    self.reducer = None
                ok = torch.tensor([1 if self.reducer is not None else 0], device=self.dev)
                dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
                if int(ok.item()) == 0:
                    self.reducer = None
            if self.reducer is None:
                self.reducer = parallel.GradientAllReduce(group)
                self.collective = "nccl"
        self.stats = parallel.DensificationStats(self.P, self.dev)
        self.iteration = 0
        self._E = torch.Tensor([])
        # the rasterizer runs at high priority, the collective's side stream at normal priority (parallel.py)
        self.compute_stream = torch.cuda.Stream(device=self.dev, priority=-1) if self.world > 1 else None

    # ------------------------------------------------------------------ one view: forward, loss, backward
    def _view(self, camera, gt_image, tan_fovx, tan_fovy):
        a = self.opt.activated
        fa = (self.bg, a["means3D"], self._E, a["opacities"], a["scales"], a["rotations"], 1.0, self._E,
              camera.world_view_transform, camera.full_proj_transform, tan_fovx, tan_fovy, int(camera.image_height),
              int(camera.image_width), a["shs"], self.sh_degree, camera.camera_center, False, False)
        out = _C.rasterize_gaussians(*fa)
        loss3, dL = fused_l1_ssim(out[1], gt_image, self.lambda_dssim)
        g = _C.rasterize_gaussians_backward(fa[0], fa[1], out[2], fa[2], fa[4], fa[5], fa[6], fa[7], fa[8], fa[9],
                                            fa[10], fa[11], dL, fa[14], fa[15], fa[16], out[3], out[0], out[4], out[5],
                                            False)
        radii = out[2]
        self.stats.add_view(g[0], radii > 0)                    # add_densification_stats, gaussian_model.py:550-552
        return loss3, {k: g[i] for k, i in _GRAD_SLOTS.items()}, radii, out[1]

    def step(self, cameras, gt_images) -> torch.Tensor:
        """cameras / gt_images: the views THIS rank rasterizes this iteration (usually one). Returns the float32[3]
        {loss, l1, ssim} of the last view (device tensor, no synchronisation)."""
        import math
        if not isinstance(cameras, (list, tuple)):
            cameras, gt_images = [cameras], [gt_images]
        if self.compute_stream is not None and torch.cuda.current_stream(self.dev) != self.compute_stream:
            self.compute_stream.wait_stream(torch.cuda.current_stream(self.dev))
            with torch.cuda.stream(self.compute_stream):
                out = self.step(cameras, gt_images)
            torch.cuda.current_stream(self.dev).wait_stream(self.compute_stream)
            return out
        self.iteration += 1
        peer = isinstance(self.reducer, parallel.PeerGradientAllReduce)
        if peer:
            self._agree_on_view_count(len(cameras))
        total, loss3 = None, None
        for i, (cam, gt) in enumerate(zip(cameras, gt_images)):
            last = i == len(cameras) - 1
            # only the last view of the step writes straight into the peer buffer; earlier ones are accumulated into it
            _C.config["grad_out"] = self.reducer.grad_out if (peer and last and total is None) else None
            try:
                loss3, grads, radii, _ = self._view(cam, gt, math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5))
            finally:
                _C.config["grad_out"] = None
            if total is None:
                total = grads
            else:
                parallel.accumulate_(total, grads)
        if self.reducer is not None:
            if peer:
                # the collective's layout (factored SH gradient or dense rows) has to be the same on every rank: it is
                # derived from the agreed view count, never from what this rank's backward happened to do
                self.reducer.start(total, means3D=self.opt.activated["means3D"], sh_degree=self.sh_degree,
                                   factored=self.reducer.factored and len(cameras) == 1)
            else:
                self.reducer.start(total)
            self.reducer.wait()
        self.opt.step(total)
        return loss3

    def _agree_on_view_count(self, n: int) -> None:
        """Every rank of a peer-memory group must rasterize the same number of views per step (one view: factored SH
        gradient; several: dense rows). The first step checks it across ranks (one small all-reduce + host read);
        afterwards a rank that deviates from the agreed count fails here, before it can join the collective with a
        different buffer layout than its peers."""
        agreed = getattr(self, "_views_per_step", None)
        if agreed is None:
            t = torch.tensor([n, -n], device=self.dev, dtype=torch.int32)
            dist.all_reduce(t, op=dist.ReduceOp.MIN, group=self.group)
            lo, hi = int(t[0].item()), -int(t[1].item())
            if lo != hi:
                raise RuntimeError(f"view-parallel step: ranks rasterize between {lo} and {hi} views this step; "
                                   "every rank must pass the same number of cameras")
            self._views_per_step = n
        elif n != agreed:
            raise RuntimeError(f"view-parallel step: {n} views on rank {self.rank}, but the group agreed on {agreed} "
                               "per rank at its first step")

    # ------------------------------------------------------------------ adaptive density control
    def _make_reducer(self):
        if self.world > 1 and self.collective == "peer":
            self.reducer = None                                   # release the old peer buffers first
            torch.cuda.empty_cache()
            self.reducer = parallel.PeerGradientAllReduce(self.P, 16, self.dev, group=self.group)

    def densify_and_prune(self, max_grad: float = 0.0002, min_opacity: float = 0.005, extent: float = 1.0,
                          max_screen_size: Optional[float] = None, percent_dense: float = 0.01) -> int:
        """train.py:113-121 for a view-parallel job: the per-view statistics are summed over ranks first, the split's
        noise comes from a generator seeded with the iteration number, so every replica takes the same decisions and
        ends with bit-identical tensors. Returns the new number of Gaussians."""
        from .densify import densify_and_prune
        self.stats.allreduce(self.group)
        gen = torch.Generator(device=self.dev).manual_seed(1_000_003 * self.iteration + 17)
        st = self.opt.densify_state(self.stats.accum, self.stats.denom)
        st = densify_and_prune(st, max_grad, min_opacity, extent, max_screen_size, percent_dense, generator=gen)
        self.opt.load_densified(st)
        self.P = st.P
        self.stats = parallel.DensificationStats(self.P, self.dev)
        self._make_reducer()
        return self.P

    def reset_opacity(self) -> None:
        from .densify import reset_opacity
        st = reset_opacity(self.opt.densify_state(self.stats.accum, self.stats.denom))
        self.opt.load_densified(st)

    # ------------------------------------------------------------------ checks / export
    def replicas_in_sync(self) -> bool:
        return parallel.replicas_in_sync(self.opt.raw.values(), self.group)

    @torch.no_grad()
    def render(self, camera) -> torch.Tensor:
        import math
        a = self.opt.activated
        out = _C.rasterize_gaussians(self.bg, a["means3D"], self._E, a["opacities"], a["scales"], a["rotations"], 1.0,
                                     self._E, camera.world_view_transform, camera.full_proj_transform,
                                     math.tan(camera.FoVx * 0.5), math.tan(camera.FoVy * 0.5), int(camera.image_height),
                                     int(camera.image_width), a["shs"], self.sh_degree, camera.camera_center, False, False)
        return out[1]
