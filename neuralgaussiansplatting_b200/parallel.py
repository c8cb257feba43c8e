"""View-parallel multi-GPU plumbing (one process per GPU, torch.distributed: NCCL on the B200 box, gloo in the
CPU tests). The reference is single-GPU (utils/general_utils.py:133); the only axis along which the rasterizer
shards naturally is camera views:

  * Gaussians (59 trainable floats each) are replicated on every rank;
  * view v of a step goes to rank v mod world (render-only: static round-robin, no communication);
  * training: every rank rasterizes its views forward+backward, accumulates the per-Gaussian gradients locally and
    one SUM all-reduce per step makes them identical everywhere (236 B/Gaussian: xyz 3 + f_dc 3 + f_rest 45 +
    opacity 1 + scaling 3 + rotation 4 floats). Buckets are issued on a dedicated communication stream in
    parameter order, the big f_rest tensor split into chunks, so that the optimizer on early buckets (and, with
    more views than ranks, the next view's rasterization) overlaps the rest.
Densification statistics are a per-view norm of the 2-D mean gradient (scene/gaussian_model.py:550-552), so they
are accumulated per view locally and only summed across ranks when densification runs.
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional, Sequence

import torch
import torch.distributed as dist

GRAD_FLOATS_PER_GAUSSIAN = 59          # 3 + 48 + 1 + 3 + 4
PARAM_ORDER = ("means3D", "opacities", "scales", "rotations", "shs")


def shard_views(num_views: int, rank: int, world: int) -> List[int]:
    """Round-robin view ownership: view v -> rank v mod world."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    return list(range(rank, num_views, world))


def owner_of_view(view: int, world: int) -> int:
    return view % world


def allreduce_payload_bytes(P: int, sh_coeffs: int = 16) -> int:
    return 4 * P * (3 + 3 * sh_coeffs + 1 + 3 + 4)


class GradientAllReduce:
    """Bucketed SUM all-reduce of the per-Gaussian gradients on a side stream.

    start(grads) enqueues every bucket asynchronously (after the producing stream's work, via an event);
    wait() makes the current stream wait for all of them. With the gloo backend (CPU tests) the same calls run
    synchronously.

    Two ways to use it in a view loop:
      blocking    start(g_k); wait()                      — the reduce is exposed after every step;
      overlapped  wait(); start(g_k)  (+ a final wait())  — the reduce of view k runs on the NCCL stream while
                  view k+1 is rasterized (the gradients of a backward are fresh tensors, so view k+1 does not
                  touch the buffers in flight); the reduced gradients of view k are consumed one view late."""

    def __init__(self, group: Optional[dist.ProcessGroup] = None, max_bucket_bytes: int = 160 << 20):
        self.group = group
        self.max_bucket_bytes = int(max_bucket_bytes)
        self._works: list = []
        self._held: list = []          # buckets being reduced: kept alive until wait()
        self._stream: Optional[torch.cuda.Stream] = None
        self._done: Optional[torch.cuda.Event] = None

    def buckets(self, grads: Dict[str, torch.Tensor]) -> List[torch.Tensor]:
        """Flat views of the gradient tensors in PARAM_ORDER, large tensors split into row chunks."""
        out: List[torch.Tensor] = []
        for name in PARAM_ORDER:
            g = grads.get(name)
            if g is None:
                continue
            if not g.is_contiguous():
                raise ValueError(f"gradient {name} must be contiguous (it is reduced in place)")
            nbytes = g.numel() * g.element_size()
            if nbytes <= self.max_bucket_bytes or g.shape[0] <= 1:
                out.append(g.view(-1))
                continue
            nchunks = -(-nbytes // self.max_bucket_bytes)
            rows = -(-g.shape[0] // nchunks)
            for r0 in range(0, g.shape[0], rows):
                out.append(g[r0:r0 + rows].reshape(-1))
        return out

    def start(self, grads: Dict[str, torch.Tensor]) -> None:
        if not dist.is_initialized() or dist.get_world_size(self.group) == 1:
            return
        bks = self.buckets(grads)
        self._held = bks
        if bks and bks[0].is_cuda:
            dev = bks[0].device
            if self._stream is None:
                self._stream = torch.cuda.Stream(device=dev)
            ready = torch.cuda.Event()
            ready.record(torch.cuda.current_stream(dev))
            self._stream.wait_event(ready)
            with torch.cuda.stream(self._stream):
                for b in bks:
                    dist.all_reduce(b, op=dist.ReduceOp.SUM, group=self.group)
                self._done = torch.cuda.Event()
                self._done.record(self._stream)
        else:
            for b in bks:
                self._works.append(dist.all_reduce(b, op=dist.ReduceOp.SUM, group=self.group, async_op=True))

    def wait(self) -> None:
        for w in self._works:
            w.wait()
        self._works.clear()
        if self._done is not None:
            torch.cuda.current_stream().wait_event(self._done)
            self._done = None
        self._held = []

    def pending(self) -> bool:
        return self._done is not None or bool(self._works)


def accumulate_(total: Dict[str, torch.Tensor], part: Dict[str, torch.Tensor]) -> None:
    for k, v in part.items():
        if v is None:
            continue
        if k in total and total[k] is not None:
            total[k].add_(v.reshape(total[k].shape))
        else:
            total[k] = v.clone()


class DensificationStats:
    """xyz_gradient_accum / denom (gaussian_model.py:550-552): per-view norm of dL/dmean2D[:, :2] for visible
    Gaussians, accumulated locally; all-reduced only when densification needs them."""

    def __init__(self, P: int, device):
        self.accum = torch.zeros(P, 1, device=device)
        self.denom = torch.zeros(P, 1, device=device)

    def add_view(self, viewspace_grad: torch.Tensor, visibility: torch.Tensor) -> None:
        self.accum[visibility] += torch.norm(viewspace_grad[visibility, :2], dim=-1, keepdim=True)
        self.denom[visibility] += 1

    def allreduce(self, group=None) -> None:
        if dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(self.accum, group=group)
            dist.all_reduce(self.denom, group=group)


def replicas_in_sync(params: Iterable[torch.Tensor], group=None) -> bool:
    """True when every rank holds bit-identical parameters (checksum all-gather)."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return True
    sums = torch.stack([p.detach().double().sum() for p in params]).to(torch.float64)
    gathered = [torch.zeros_like(sums) for _ in range(dist.get_world_size(group))]
    dist.all_gather(gathered, sums, group=group)
    return all(torch.equal(gathered[0], g) for g in gathered[1:])
