"""Host-side view pipelining for the callers of the rasterizer (the train.py / render.py view
loops, reference train.py:75-93 and render.py:24-35).

The reference touches the host once per view: it picks a camera, uploads / reads the ground-truth image, renders,
computes the loss and reads it back (`loss.item()` for the progress bar, train.py:99) — each step waits for the
previous one. With a 2 ms rasterizer that serialisation IS the step time, so the drop-in ships the two pieces a
caller needs to keep the GPU busy:

  * ViewPrefetcher  double-buffered H2D staging of a view's host inputs (camera block, ground-truth image) on a
                    copy stream, so view k+1 uploads while view k rasterizes;
  * DeferredScalar  device->host read-back of a per-step scalar (the loss) that is consumed one step late, so the
                    host never blocks on the step it has just enqueued.

Both are plain CUDA stream/event plumbing through PyTorch; no computation happens here.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch


class ViewPrefetcher:
    """Ring of `depth` device slots; submit() copies a dict of pinned host tensors into the next slot on a copy
    stream, acquire() hands the oldest submitted slot to the current stream."""

    def __init__(self, device, spec: Dict[str, Tuple[Tuple[int, ...], torch.dtype]], depth: int = 2):
        if depth < 2:
            raise ValueError("ViewPrefetcher needs at least two slots")
        self.device = torch.device(device)
        self.depth = depth
        self.slots: List[Dict[str, torch.Tensor]] = [
            {k: torch.empty(shape, dtype=dt, device=self.device) for k, (shape, dt) in spec.items()}
            for _ in range(depth)]
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self._ready: List[Optional[torch.cuda.Event]] = [None] * depth      # copy finished
        self._free: List[Optional[torch.cuda.Event]] = [None] * depth       # consumer finished
        self._head = 0       # next slot to fill
        self._tail = 0       # next slot to hand out
        self._inflight = 0
        self._held: Optional[int] = None

    def submit(self, host: Dict[str, torch.Tensor]) -> None:
        if self._inflight >= self.depth - (1 if self._held is not None else 0):
            raise RuntimeError("ViewPrefetcher: all slots are in flight")
        i = self._head
        with torch.cuda.stream(self.copy_stream):
            if self._free[i] is not None:
                self.copy_stream.wait_event(self._free[i])      # the slot's last consumer is done
            for k, dst in self.slots[i].items():
                src = host[k]
                if not src.is_pinned():
                    raise ValueError(f"ViewPrefetcher: host tensor {k!r} must be pinned for an asynchronous copy")
                dst.copy_(src, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        self._ready[i] = ev
        self._head = (i + 1) % self.depth
        self._inflight += 1

    def acquire(self) -> Dict[str, torch.Tensor]:
        """Oldest submitted view; the current stream waits for its copy. Releases the previously acquired slot."""
        if self._inflight == 0:
            raise RuntimeError("ViewPrefetcher: nothing submitted")
        self.release()
        i = self._tail
        torch.cuda.current_stream(self.device).wait_event(self._ready[i])
        self._tail = (i + 1) % self.depth
        self._inflight -= 1
        self._held = i
        return self.slots[i]

    def release(self) -> None:
        """Marks the acquired slot reusable once the work enqueued so far on the current stream has finished."""
        if self._held is None:
            return
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self._free[self._held] = ev
        self._held = None


class DeferredScalar:
    """push(t) enqueues an asynchronous D2H copy of a 1-element tensor; pop() returns the value pushed `lag`
    calls earlier (blocking only on THAT copy). drain() returns everything still in flight."""

    def __init__(self, lag: int = 1, dtype=torch.float32):
        self.lag = lag
        self._bufs = [torch.zeros(1, dtype=dtype).pin_memory() for _ in range(lag + 1)]
        self._events: List[Optional[torch.cuda.Event]] = [None] * (lag + 1)
        self._n = 0

    def push(self, t: torch.Tensor) -> Optional[float]:
        i = self._n % (self.lag + 1)
        self._bufs[i].copy_(t.detach().reshape(1), non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(t.device))
        self._events[i] = ev
        self._n += 1
        if self._n > self.lag:
            j = (self._n - 1 - self.lag) % (self.lag + 1)
            self._events[j].synchronize()
            return float(self._bufs[j][0])
        return None

    def drain(self) -> List[float]:
        out = []
        first = max(self._n - self.lag, 0)
        for k in range(first, self._n):
            j = k % (self.lag + 1)
            self._events[j].synchronize()
            out.append(float(self._bufs[j][0]))
        self._n = 0
        return out
