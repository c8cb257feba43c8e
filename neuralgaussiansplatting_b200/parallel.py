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


class FlatGradientAllReduce:
    """ONE ncclAllReduce per step over a flat buffer that holds the five gradient tensors (gradient_layout) — the
    strongest plain-NCCL formulation (no per-tensor buckets, no staging copy: the backward writes into views of the
    buffer through `_C.config["grad_out"]`, or a reference binding zero-fills and accumulates into them). Buffers
    rotate so the reduce of view k can overlap view k+1. Same start()/wait() protocol as GradientAllReduce; used by
    both bench arms as the NCCL comparison / reference-arm collective."""

    def __init__(self, P: int, sh_coeffs: int = 16, device=None, group=None, n_buffers: int = 2):
        self.group = group
        self.layout, self.n_floats = gradient_layout(P, sh_coeffs)
        self.device = torch.device(device if device is not None else "cpu")
        self._flat = [torch.zeros(self.n_floats, dtype=torch.float32, device=self.device) for _ in range(n_buffers)]
        self.n_buffers = int(n_buffers)
        self._cur = -1
        self._stream: Optional[torch.cuda.Stream] = None
        self._done: Optional[torch.cuda.Event] = None
        self._work = None
        self._copy_back = None

    def _views(self, i: int) -> Dict[str, torch.Tensor]:
        out = {}
        for name, (off, shape) in self.layout.items():
            n = 1
            for d in shape:
                n *= d
            out[name] = self._flat[i][off:off + n].view(shape)
        return out

    def grad_out(self) -> Dict[str, torch.Tensor]:
        self._cur = (self._cur + 1) % self.n_buffers
        return self._views(self._cur)

    def start(self, grads: Optional[Dict[str, torch.Tensor]] = None, **_unused) -> None:
        if not dist.is_initialized() or dist.get_world_size(self.group) == 1:
            return
        if self._cur < 0:
            self._cur = 0
        views, flat = self._views(self._cur), self._flat[self._cur]
        self._copy_back = None
        if grads is not None:
            stray = {k: g for k, g in grads.items() if g is not None and k in views
                     and g.data_ptr() != views[k].data_ptr()}
            if stray:
                for k, g in stray.items():
                    views[k].copy_(g.reshape(views[k].shape))
                self._copy_back = (stray, views)
        if flat.is_cuda:
            if self._stream is None:
                self._stream = torch.cuda.Stream(device=flat.device)
            ready = torch.cuda.Event()
            ready.record(torch.cuda.current_stream(flat.device))
            self._stream.wait_event(ready)
            with torch.cuda.stream(self._stream):
                dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group)
                self._done = torch.cuda.Event()
                self._done.record(self._stream)
        else:
            self._work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)

    def wait(self) -> None:
        if self._work is not None:
            self._work.wait()
            self._work = None
        if self._done is not None:
            torch.cuda.current_stream().wait_event(self._done)
            self._done = None
        if self._copy_back is not None:
            stray, views = self._copy_back
            for k, g in stray.items():
                g.copy_(views[k].reshape(g.shape))
            self._copy_back = None

    def pending(self) -> bool:
        return self._done is not None or self._work is not None


def gradient_layout(P: int, sh_coeffs: int = 16, align_floats: int = 64):
    """Offsets (in floats) of the five gradient tensors inside one flat peer buffer, each sub-tensor starting on a
    256-byte boundary (preprocess_bwd stores dL/dsh rows as 16-byte vectors). Returns ({name: (offset, shape)},
    total_floats) with total_floats a multiple of `align_floats`."""
    shapes = {"means3D": (P, 3), "opacities": (P, 1), "scales": (P, 3), "rotations": (P, 4), "shs": (P, sh_coeffs, 3)}
    out, off = {}, 0
    for name in PARAM_ORDER:
        n = 1
        for d in shapes[name]:
            n *= d
        out[name] = (off, shapes[name])
        off = -(-(off + n) // align_floats) * align_floats
    return out, off


def peer_owner_slices(n_float4: int, world: int):
    """[(begin, end)] of the float4 elements each rank reduces in b200s_peer_allreduce_sum_f32 (csrc/peer_reduce.cu:
    per = ceil(n4 / world); rank r owns [min(n4, r*per), min(n4, r*per + per)))."""
    per = -(-n_float4 // world)
    return [(min(n_float4, r * per), min(n_float4, r * per + per)) for r in range(world)]


def peer_sparse_chunks(rows: int, world: int, ctas: int, chunk: int = 1024):
    """{(rank, cta): [(row_begin, row_end), ...]} — the 1024-row work units of the visibility-sparse tail: rank r owns
    rows [r*rows_per, (r+1)*rows_per) with rows_per = ceil(ceil(rows/world)/1024)*1024; its CTA b takes chunks
    b, b+ctas, ... and stops at the first chunk starting at or beyond `rows`."""
    rows_per = -(-(-(-rows // world)) // chunk) * chunk
    out = {}
    for r in range(world):
        for b in range(ctas):
            lst = []
            for c in range(b, rows_per // chunk, ctas):
                row0 = r * rows_per + c * chunk
                if row0 >= rows:
                    break
                lst.append((row0, min(rows, row0 + chunk)))
            out[(r, b)] = lst
    return out


class PeerMemory:
    """`nbytes` of device memory on every rank of the group, each rank's block mapped into every other rank's
    address space over NVLink (plumbing only: the kernels that use the mappings are in libb200splat.so).

    backend "symm": torch.distributed._symmetric_memory (cuMem + fabric/fd handle exchange; also yields the
    NVSwitch multicast mapping when the box has one). backend "ipc": legacy CUDA IPC handles of a plain torch
    allocation exchanged through the process group. "auto" tries symm first; all ranks agree on the outcome."""

    def __init__(self, nbytes: int, device, group=None, backend: str = "auto"):
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.nbytes = int(nbytes)
        self.device = torch.device(device)
        self.multicast_ptr = 0
        self.backend = None
        self._keep = []
        err = None
        if backend in ("auto", "symm"):
            try:
                self._init_symm()
            except Exception as e:  # noqa: BLE001 - any failure means "not available on this box"
                err = e
            ok = torch.tensor([1 if self.backend == "symm" else 0], device=self.device, dtype=torch.int32)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
            if int(ok.item()) == 0:
                self.backend = None
                self.multicast_ptr = 0
                if backend == "symm":
                    raise RuntimeError(f"symmetric memory is not available on every rank: {err}")
        if self.backend is None:
            self._init_ipc()

    def _init_symm(self):
        import torch.distributed._symmetric_memory as symm_mem
        grp = self.group if self.group is not None else dist.group.WORLD
        t = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=self.device)
        hdl = symm_mem.rendezvous(t, grp)
        self.local = t
        self.ptrs = [int(x) for x in hdl.buffer_ptrs]
        if getattr(hdl, "has_multicast_support", False):
            self.multicast_ptr = int(hdl.multicast_ptr or 0)
        self._keep = [hdl]
        self.backend = "symm"

    def _init_ipc(self):
        t = torch.empty(self.nbytes, dtype=torch.uint8, device=self.device)
        meta = t.untyped_storage()._share_cuda_()
        gathered = [None] * self.world
        dist.all_gather_object(gathered, (meta, t.storage_offset()), group=self.group)
        self.local = t
        self.ptrs = []
        for r, (m, off) in enumerate(gathered):
            if r == self.rank:
                self.ptrs.append(t.data_ptr())
                continue
            m = list(m)
            m[0] = self.device.index if self.device.index is not None else torch.cuda.current_device()
            with torch.cuda.device(self.device):
                st = torch.UntypedStorage._new_shared_cuda(*m)
            self._keep.append(st)
            self.ptrs.append(st.data_ptr() + off)
        self.backend = "ipc"
        dist.barrier(group=self.group)


class PeerGradientAllReduce:
    """The gradient all-reduce as ONE kernel of libb200splat.so over NVLink peer memory
    (b200s_peer_allreduce_sum_f32, csrc/peer_reduce.cu) instead of NCCL calls.

    The rasterizer's backward writes its five gradient tensors straight into a peer-mapped buffer
    (`grad_out()` hands `_C.config["grad_out"]` fresh views; buffers rotate so the reduce of view k can run while
    view k+1 is written). start(grads) launches the collective on a high-priority side stream after the producing
    stream's work; wait() makes the current stream wait for it. Sums are taken in rank order by a single owner
    per element, so every replica receives bit-identical values and results are reproducible run to run.
    Same start()/wait() protocol as GradientAllReduce. Gradients that do not alias the current buffer (autograd
    cloned them) are copied in and back out.

    Stream priorities: the collective's side stream has NORMAL priority (`priority=0`); run the rasterizer on a
    high-priority stream (`torch.cuda.Stream(priority=-1)`) so that the collective's CTAs only take the SM slots the
    rasterizer leaves free — measured on 2x B200: 1.93 ms/step vs 2.06 with the priorities the other way round
    (1.89 ms on one GPU)."""

    # CTAs of 512 threads that saturate NVLink per world size (measured on 8x B200, tools/peer_check.py): each
    # thread keeps 16 loads in flight of which (world-1)/world are remote; more CTAs only take SMs from the
    # rasterizer that runs concurrently.
    AUTO_CTAS = {2: 16, 3: 16, 4: 12}
    AUTO_CTAS_FACTORED = {2: 16, 3: 16, 4: 16}     # default 16: the payload is 2.5x smaller but latency-bound

    def __init__(self, P: int, sh_coeffs: int = 16, device=None, group=None, n_buffers: int = 2, ctas: int = 0,
                 multimem: Optional[bool] = None, backend: str = "auto", factored: bool = True,
                 gather_push: bool = True, priority: int = 0, fast_reconstruct: bool = True):
        from . import _lib
        self._libmod = _lib
        self.fast_reconstruct = bool(fast_reconstruct)
        self._lib = _lib.load()
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if self.world > _lib.MAX_PEERS:
            raise ValueError(f"at most {_lib.MAX_PEERS} peers")
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.layout, self.n_floats = gradient_layout(P, sh_coeffs)
        # Factored SH gradient (default): within one view dL/dsh is the outer product basis(dir) x dL/dRGB, so the
        # backward emits 3 floats per Gaussian (+ its camera centre) into this rank's slot of a gather region behind
        # the five tensors; the collective sums only the 11 dense floats, gathers the slots, and
        # b200s_sh_grad_reconstruct rebuilds the summed 48-float rows locally — bit-identical to reducing the rows,
        # with 2.5x fewer NVLink bytes (csrc/sh_reconstruct.cu).
        self.factored = bool(factored)
        self.dense_floats = self.layout["shs"][0]                      # means3D, opacities, scales, rotations
        self.slot_floats = -(-(3 * int(P) + 4) // 64) * 64
        self.gather_floats_total = self.world * self.slot_floats if self.factored else 0
        self.n_buffers = int(n_buffers)
        self._ctas_arg = int(ctas) if ctas and ctas > 0 else 0
        self.ctas = self._ctas_arg if self._ctas_arg else (
            self.AUTO_CTAS_FACTORED.get(self.world, 16) if factored else self.AUTO_CTAS.get(self.world, 8))
        self.flag_bytes = -(-int(self._lib.b200s_peer_allreduce_flag_bytes(self.ctas, self.world)) // 256) * 256
        self.buf_bytes = (self.n_floats + self.gather_floats_total) * 4
        self.P, self.sh_coeffs = int(P), int(sh_coeffs)
        # one visibility byte per Gaussian next to every buffer (sparse tail of the collective, see start())
        self.vis_bytes = -(-int(self._lib.b200s_peer_allreduce_vis_bytes(self.P, self.world)) // 256) * 256
        self.slot_bytes = self.buf_bytes + self.vis_bytes
        self.mem = PeerMemory(self.flag_bytes + self.n_buffers * self.slot_bytes, self.device, group, backend)
        self.mem.local.zero_()
        torch.cuda.synchronize(self.device)
        dist.barrier(group=group)                      # nobody raises a flag before every flag array is zero
        # multimem=None (auto): with 4+ ranks on an NVSwitch box the 11 dense floats are summed IN the switch
        # (multimem.ld_reduce) by TWO CTAs — the overlapped step is bound by the SM slots the collective takes from
        # the rasterizer, not by NVLink (measured, ms/step: 4 GPUs 2.00 vs 2.07-2.08 for the peer-load sum; 8 GPUs
        # 2.15 with 2 CTAs, 2.30 with 4, 2.32 for the peer-load sum with 6; one GPU 1.89 — DESIGN.md §7). The switch's summation order is its own: every replica receives identical
        # bits, equal to the rank-order sum to rounding. multimem=False keeps the exact rank-order two-shot sum.
        if multimem is None:
            multimem = self.factored and self.world >= 4 and self.mem.multicast_ptr != 0
            if multimem and not self._ctas_arg:
                self.ctas = 2
        self.multimem = bool(multimem) and self.mem.multicast_ptr != 0
        # factored mode: push the factor slots through the NVSwitch multicast mapping (multimem.st, an exact copy)
        # when the box has one; otherwise every rank pulls the foreign slots with peer loads
        self.gather_push = bool(gather_push) and self.factored and self.mem.multicast_ptr != 0
        self._flat = [self.mem.local[self.flag_bytes + i * self.slot_bytes:
                                     self.flag_bytes + i * self.slot_bytes + self.buf_bytes]
                      .view(torch.float32) for i in range(self.n_buffers)]
        self._cur = -1
        self._epoch = 0
        self._stream = torch.cuda.Stream(device=self.device, priority=int(priority))
        self._done: Optional[torch.cuda.Event] = None
        self._copy_back = None
        self._factors_fresh = False

    # ---- buffers
    def _views(self, i: int) -> Dict[str, torch.Tensor]:
        flat = self._flat[i]
        out = {}
        for name, (off, shape) in self.layout.items():
            n = 1
            for d in shape:
                n *= d
            out[name] = flat[off:off + n].view(shape)
        return out

    def _factor_slot(self, i: int, r: int) -> torch.Tensor:
        off = self.n_floats + r * self.slot_floats
        return self._flat[i][off:off + self.slot_floats]

    def grad_out(self) -> Dict[str, torch.Tensor]:
        """Rotate to the next buffer and return fresh views of it (install as `_C.config["grad_out"]`). In factored
        mode the dict carries "sh_factors" (this rank's gather slot): the backward then writes 3 floats per Gaussian
        there instead of the dL/dsh rows, and the "shs" view is filled by the collective."""
        self._cur = (self._cur + 1) % self.n_buffers
        out = self._views(self._cur)
        if self.factored:
            out["sh_factors"] = self._factor_slot(self._cur, self.rank)
            self._factors_fresh = True
        return out

    # ---- collective
    def start(self, grads: Optional[Dict[str, torch.Tensor]] = None, radii: Optional[torch.Tensor] = None,
              means3D: Optional[torch.Tensor] = None, sh_degree: int = 3, factored: Optional[bool] = None) -> None:
        """Launch the collective for the current buffer. Factored mode (the backward of THIS step wrote its factors
        through grad_out(); every rank must take the same mode) needs `means3D` (the replicated [P,3] positions the
        views were rendered from) and the active `sh_degree` for the local reconstruction of dL/dsh. `radii` (int32[P], this rank's forward output for the
        view the gradients belong to) enables the visibility-sparse SH tail: rows with radii <= 0 are all-zero
        here, so owners skip them — same bits as the dense sum, less NVLink traffic. Only pass it when the
        buffer holds the gradients of exactly that one view (not a local accumulation over several views)."""
        if self._cur < 0:
            self._cur = 0
        views = self._views(self._cur)
        self._copy_back = None
        use_factored = self.factored and self._factors_fresh if factored is None else bool(factored)
        if use_factored and not (self.factored and self._factors_fresh):
            raise RuntimeError("factored collective: this step's backward did not write its factors (grad_out())")
        if use_factored and (means3D is None or means3D.shape != (self.P, 3) or not means3D.is_contiguous()
                             or means3D.dtype != torch.float32 or means3D.device != self.device):
            raise ValueError("factored collective needs means3D: the contiguous float32 [P,3] positions on this device")
        self._factors_fresh = False
        if grads is not None:
            stray = {k: g for k, g in grads.items() if g is not None and k in views
                     and g.data_ptr() != views[k].data_ptr()}
            if stray:
                for k, g in stray.items():
                    if not (use_factored and k == "shs"):          # factored: the rows are rebuilt, not reduced
                        views[k].copy_(g.reshape(views[k].shape))
                self._copy_back = (stray, views)
        a = self._libmod.PeerReduceArgs()
        off = self.flag_bytes + self._cur * self.slot_bytes
        if (radii is not None and self.sh_coeffs == 16 and not self.multimem and self._copy_back is None
                and not use_factored):
            if radii.dtype != torch.int32 or radii.numel() != self.P or not radii.is_cuda or not radii.is_contiguous():
                raise ValueError("radii must be a contiguous int32 CUDA tensor with one entry per Gaussian")
            a.radii = radii.data_ptr()
            a.sparse_rows = self.P
            a.sparse_offset_floats = self.layout["shs"][0]
            a.vis_offset_bytes = self.buf_bytes
            self._radii_keep = radii
        for r in range(self.world):
            a.peer_bufs[r] = self.mem.ptrs[r] + off
            a.peer_flags[r] = self.mem.ptrs[r]
        push = self.gather_push and use_factored
        a.multicast_buf = (self.mem.multicast_ptr + off) if (self.multimem or push) else None
        a.n_floats = self.dense_floats if use_factored else self.n_floats
        if use_factored:
            a.gather_offset_floats, a.gather_floats = self.n_floats, self.slot_floats
        a.rank, a.world = self.rank, self.world
        self._epoch += 1
        a.epoch = self._epoch
        a.ctas = self.ctas
        a.use_multimem = int(self.multimem) | (2 if push else 0)
        ready = torch.cuda.Event()
        ready.record(torch.cuda.current_stream(self.device))
        self._stream.wait_event(ready)
        self._libmod.check(self._lib.b200s_peer_allreduce_sum_f32(a, self._stream.cuda_stream),
                           "b200s_peer_allreduce_sum_f32")
        if use_factored:
            import ctypes as C
            sg = self._libmod.ShGradArgs()
            sg.P, sg.D, sg.M, sg.views = self.P, int(sh_degree), self.sh_coeffs, self.world
            sg.means3D = means3D.data_ptr()
            sg.factors = self._flat[self._cur].data_ptr() + 4 * self.n_floats
            sg.stride_floats = self.slot_floats
            sg.dL_dsh = views["shs"].data_ptr()
            # in-switch dense sum: the step is not a rank-order sum anyway, so the rebuild may use the fused form
            # (half the instructions; identical on every rank); the exact collectives keep the bit-exact rebuild
            sg.fast = 1 if (self.multimem and self.fast_reconstruct) else 0
            self._means_keep = means3D
            self._libmod.check(self._lib.b200s_sh_grad_reconstruct(C.byref(sg), self._stream.cuda_stream),
                               "b200s_sh_grad_reconstruct")
        self._done = torch.cuda.Event()
        self._done.record(self._stream)

    def wait(self) -> None:
        if self._done is not None:
            torch.cuda.current_stream(self.device).wait_event(self._done)
            self._done = None
        if self._copy_back is not None:
            stray, views = self._copy_back
            for k, g in stray.items():
                g.copy_(views[k].reshape(g.shape))
            self._copy_back = None

    def pending(self) -> bool:
        return self._done is not None


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
        # same values as the reference's masked `+=` (gaussian_model.py:551-552) without its boolean-index host sync
        vis = visibility.reshape(-1, 1)
        norm = torch.norm(viewspace_grad[:, :2], dim=-1, keepdim=True)
        self.accum += torch.where(vis, norm, torch.zeros_like(norm))
        self.denom += vis.to(self.denom.dtype)

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
