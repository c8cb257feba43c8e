"""bench.py — the rasterizer hot path on B200, one JSON line (driver contract).

    python bench.py --gpus N --steps K --warmup W                 # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W  # the unmodified reference (oracle/_ref)

Workload (config.workload): BASELINE.json's metric configuration that fits one GPU — C3, the
Mip-NeRF360-shaped training step: 3 M Gaussians, SH degree 3, 1344x896, forward + backward of ONE view per
step per GPU (synthetic seeded Gaussians / cameras). With N > 1 each rank rasterizes its own
view (weak scaling over views) and the 59-float per-Gaussian gradients are SUM-all-reduced over NCCL.

  value            train-step views/s over all ranks, inputs resident in HBM (device-timed, max over ranks)
  e2e              the same step through the drop-in Python API (GaussianRasterizer + autograd), with the
                   per-step host inputs (camera matrices + ground-truth image, pinned) copied H2D and the loss
                   read back D2H inside the timed region — what train.py:86-93 does per iteration
  roofline         the dominant kernel of the step: algorithmic flops (or bytes) / its CUDA-event time, measured
                   live with the library's stage brackets in a separate pass of the same K steps
  cpu_baseline     torch-CPU preprocessing the north_star names (covariance + eval_sh) and the C oracle of the
                   full step, on a bounded sample, on this box's host cores
The reference arm runs the UNMODIFIED reference rasterizer compiled for sm_100 (oracle/_ref, the same-GPU
baseline the north_star's ">= 2x" is quoted against); the reference has no CPU implementation of this path.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

METRIC = "train-step views/s (rasterizer fwd+bwd) at 3M Gaussians, SH3, 1344x896"
UNIT = "views/s"


# ----------------------------------------------------------------------------------------- utilities
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (profiling recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self._stop, self._t = index, [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                                      str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.splitlines()[0].split(",")])
            except Exception:
                pass
            self._stop.wait(0.1)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        reasons = sorted({n for r in self.rows for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


def env_world():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def barrier_sync(world, dev):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)


def timed_region(fn, steps, world, dev, sampler=None, finalize=None):
    """EXACTLY `steps` calls of fn(step) between barrier+synchronize, CUDA events on the current stream,
    MAX over ranks. `finalize` (pipelined loops: last all-reduce wait, last loss read-back) runs INSIDE the region."""
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    barrier_sync(world, dev)
    e0.record()
    for s in range(steps):
        fn(s)
        marks[s].record()
    if finalize is not None:
        finalize()
    e1.record()
    barrier_sync(world, dev)
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    # spread of the individual steps on this rank (median + p10 / p90), from the same events
    edges = [e0] + marks
    per = sorted(edges[i].elapsed_time(edges[i + 1]) for i in range(steps))
    if per:
        q = lambda f: per[min(len(per) - 1, int(round(f * (len(per) - 1))))]
        LAST_STEP_SPREAD.clear()
        LAST_STEP_SPREAD.update({"median_ms": round(q(0.5), 4), "p10_ms": round(q(0.1), 4), "p90_ms": round(q(0.9), 4),
                                 "min_ms": round(per[0], 4), "max_ms": round(per[-1], 4), "n": len(per),
                                 "what": "per-step CUDA-event intervals of the timed region on rank 0"})
    return float(ms.item())


LAST_STEP_SPREAD = {}


def load_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


def load_sm_peaks(sm_max_mhz):
    """FP32-FMA / MUFU / shared-memory / issue-slot peaks of this pool's B200 measured by tools/sm_peaks (a
    register-resident micro-benchmark timed with CUDA events); the newest profiles/r*_sm_peaks.json. Falls back to
    the nominal figures (and says so) when no measurement is committed."""
    files = sorted((ROOT / "profiles").glob("r*_sm_peaks.json"))
    if files:
        d = json.loads(files[-1].read_text())
        return d, f"measured by tools/sm_peaks on this pool's B200 (profiles/{files[-1].name})"
    nominal = {"fp32_tflops": 148 * 128 * 2 * sm_max_mhz * 1e6 / 1e12, "issue_ginst_per_s": 148 * 4 * sm_max_mhz * 1e6 / 1e9}
    return nominal, f"nominal: 148 SM x 128 lanes x 2 x {sm_max_mhz:.0f} MHz (no measured SM peaks committed)"


# ----------------------------------------------------------------------------------------- workload
class Workload:
    """C3: 3 M replicated Gaussians, 8 orbit views, per-view dL/dimage and ground-truth images."""

    def __init__(self, name, dev, rank, world, pin_host=True):
        from neuralgaussiansplatting_b200.synthetic import CONFIGS, make_camera, make_grad_image, make_scene
        self.cfg = cfg = CONFIGS[name]
        self.dev, self.rank, self.world = dev, rank, world
        self.nviews = max(cfg.views, 8)
        sc = make_scene(cfg, seed=0)
        self.scene = {k: v.to(dev) for k, v in sc.items() if k not in ("feature_vector", "normal")}
        self.P = self.scene["means3D"].shape[0]
        self.bg = torch.zeros(3, device=dev)
        self.cams_host = [make_camera(cfg, v) for v in range(self.nviews)]
        self.cams = [make_camera(cfg, v, device=dev) for v in range(self.nviews)]
        self.tan = (math.tan(cfg.fovx * 0.5), math.tan(cfg.fovy * 0.5))
        self.dL = [make_grad_image(cfg, v, device=dev) for v in range(self.nviews)]
        # host-side per-step inputs of the e2e leg: camera block [16 + 16 + 3 floats] and ground truth image
        g = torch.Generator().manual_seed(4000)
        self.gt_host = [torch.rand(3, cfg.height, cfg.width, generator=g) for _ in range(self.nviews)]
        self.cam_host = [torch.cat([c.world_view_transform.reshape(-1), c.full_proj_transform.reshape(-1),
                                    c.camera_center.reshape(-1)]) for c in self.cams_host]
        if pin_host:
            self.gt_host = [t.pin_memory() for t in self.gt_host]
            self.cam_host = [t.pin_memory() for t in self.cam_host]
        self.h2d_bytes = self.gt_host[0].numel() * 4 + self.cam_host[0].numel() * 4
        self.d2h_bytes = 4

    def view_of(self, step):
        return (step * self.world + self.rank) % self.nviews

    def fwd_args(self, v):
        s, c = self.scene, self.cams[v]
        E = torch.Tensor([])
        return (self.bg, s["means3D"], E, s["opacities"], s["scales"], s["rotations"], 1.0, E,
                c.world_view_transform, c.full_proj_transform, self.tan[0], self.tan[1], self.cfg.height,
                self.cfg.width, s["shs"], 3, c.camera_center, False, False)


def stage_table(w: Workload, stage_ms, stage_calls, steps, counts, P_vis, N, hbm_peak, fp32_peak):
    """Per-stage achieved rate against its roofline. Algorithmic bytes/flops: SURVEY.md §8(d) / DESIGN.md."""
    from neuralgaussiansplatting_b200 import _lib
    P, px = w.P, w.cfg.width * w.cfg.height
    T = ((w.cfg.width + 15) // 16) * ((w.cfg.height + 15) // 16)
    P_c = P - P_vis
    passes = -(-(32 + max(T - 1, 1).bit_length()) // 8)
    n_trav, n_contrib, n_trav_bwd = counts
    alg = {
        "preprocess": ("hbm", 311 * P_vis + 20 * P_c),
        "scan": ("hbm", 8 * P),
        "expand": ("hbm", 20 * P + 12 * N),
        "sort": ("hbm", N * (8 + 24 * passes)),
        "ranges": ("hbm", 8 * N + 8 * T),
        "blend_fwd": ("fp32", 14 * n_trav + 9 * n_contrib),
        "blend_bwd": ("fp32", 14 * n_trav_bwd + 69 * n_contrib),
        "preprocess_bwd": ("hbm", 510 * P_vis + 4 * P_c + 236 * P_c),
    }
    # Bytes THIS implementation moves per launch, where they differ from the algorithmic figure above
    # (which, for the binning stages, is the reference's 64-bit CUB sort of all instances): the in-house sort moves
    # u32 tile keys + u32 ids through `tile_passes` onesweep passes and presorts the visible set on depth.
    tile_passes = -(-max(T - 1, 1).bit_length() // 8)
    moved = {
        "scan": 4 * P + 8 * P_vis,
        "expand": 4 * 16 * P_vis + 8 * N,
        "sort": 16 * N * tile_passes + 4 * N,
        "ranges": 4 * N + 8 * T,
    }
    rows = {}
    for i, name in enumerate(_lib.STAGES):
        if stage_calls[i] == 0:
            continue
        ms = stage_ms[i] / steps
        row = {"ms_per_step": round(ms, 5), "launch_brackets_per_step": stage_calls[i] / steps}
        if name in alg and ms > 0:
            bound, units = alg[name]
            if bound == "hbm":
                ach = units / (ms * 1e-3) / 1e9
                row.update(bound="hbm", algorithmic_bytes=int(units), achieved=round(ach, 1), peak=hbm_peak,
                           unit="GB/s", frac=round(ach / hbm_peak, 4))
                if name in moved:
                    row.update(moved_bytes=int(moved[name]),
                               moved_gbps=round(moved[name] / (ms * 1e-3) / 1e9, 1),
                               note="frac is quoted on the reference's algorithmic bytes; moved_bytes is what these "
                                    "kernels actually move (latency / ADU-pipe bound, not HBM)")
            else:
                ach = units / (ms * 1e-3) / 1e12
                row.update(bound="fp32", algorithmic_flops=int(units), achieved=round(ach, 3),
                           peak=round(fp32_peak, 2), unit="TFLOP/s", frac=round(ach / fp32_peak, 4))
        rows[name] = row
    return rows


# ----------------------------------------------------------------------------------------- our arm
def run_ours(args, rank, local_rank, world, dev):
    import ctypes as C
    from neuralgaussiansplatting_b200 import _C, _lib, parallel
    from neuralgaussiansplatting_b200.diff_gaussian_rasterization import (GaussianRasterizationSettings,
                                                                          GaussianRasterizer)
    lib = _lib.load()          # raises if libb200splat.so is missing: there is no fallback path
    _C.config["blend_mode"] = 1 if args.blend_mode == "fast" else 0
    w = Workload(args.workload, dev, rank, world)
    cfg = w.cfg
    reducer, collective_kind = None, "none"
    if world > 1:
        peer_ok = torch.ones(1, device=dev)
        if args.collective == "peer":
            # the collective is one kernel of libb200splat.so over NVLink peer memory; the backward writes its
            # gradients straight into the peer-mapped buffers (no staging copy, no NCCL call in the step)
            try:
                reducer = parallel.PeerGradientAllReduce(w.P, 16, dev, ctas=args.peer_ctas,
                                                         multimem=(True if args.multimem else False if args.exact_collective else None),
                                                         factored=not args.dense_collective,
                                                         gather_push=not args.gather_pull,
                                                         priority=0 if args.compute_priority else -1,
                                                         fast_reconstruct=not args.exact_reconstruct)
            except Exception as e:  # noqa: BLE001 — no peer mappings on this box (no NVLink P2P / IPC): say so, use NCCL
                print(f"[bench] rank {rank}: peer-memory collective unavailable ({e}); falling back to NCCL",
                      file=sys.stderr)
                peer_ok.zero_()
            dist.all_reduce(peer_ok, op=dist.ReduceOp.MIN)
        if args.collective == "peer" and float(peer_ok.item()) > 0:
            collective_kind = (f"b200s_peer_allreduce_sum_f32 ({'multimem/NVLS' if reducer.multimem else 'P2P two-shot'}, "
                               f"{reducer.ctas} CTAs, {reducer.mem.backend} mappings"
                               + (", factored SH gradient: 11 dense floats summed + 3 dL/dRGB floats gathered per Gaussian "
                                  "and view (" + ("multimem.st push" if reducer.gather_push else "peer-load pull")
                                  + "), dL/dsh rebuilt by b200s_sh_grad_reconstruct ("
                                  + ("fused-multiply-add form, identical on every rank, ~1e-7 of the dense sum"
                                     if (reducer.multimem and reducer.fast_reconstruct) else "bit-identical to the dense sum") + ")"
                                  if reducer.factored else "")
                               + (", rasterizer on a high-priority stream" if args.compute_priority else "")
                               + ("" if not args.sparse_collective or reducer.multimem or reducer.factored else
                                  ", dL/dSH rows read only from ranks whose view saw the Gaussian (radii > 0)") + ")")
        else:
            reducer = parallel.FlatGradientAllReduce(w.P, 16, dev)
            collective_kind = "NCCL all_reduce, one call on a flat buffer the backward writes into" + (
                "" if args.collective == "nccl" else " — FALLBACK: peer mappings unavailable on this box")
        _C.config["grad_out"] = reducer.grad_out
    GRAD_KEYS = {"means3D": 3, "opacities": 2, "shs": 5, "scales": 6, "rotations": 7}   # slots of the 8-tuple

    overlap = args.allreduce == "overlap"

    is_peer = isinstance(reducer, parallel.PeerGradientAllReduce)
    sparse = args.sparse_collective and is_peer and not reducer.factored

    def reduce_step(grads, radii=None, means3D=None):
        if reducer is None:
            return
        kw = {"radii": radii} if (sparse and radii is not None) else {}
        if is_peer:
            kw.update(means3D=w.scene["means3D"] if means3D is None else means3D, sh_degree=3)
        if overlap:
            reducer.wait()          # the reduce of the PREVIOUS view ran while this view was rasterized
            reducer.start(grads, **kw)
        else:
            reducer.start(grads, **kw)
            reducer.wait()

    def reduce_finalize():
        if reducer is not None:
            reducer.wait()

    def step_resident(s, collective=True):
        v = w.view_of(s)
        fa = w.fwd_args(v)
        out = _C.rasterize_gaussians(*fa)
        g = _C.rasterize_gaussians_backward(fa[0], fa[1], out[2], fa[2], fa[4], fa[5], fa[6], fa[7], fa[8], fa[9],
                                            fa[10], fa[11], w.dL[v], fa[14], fa[15], fa[16], out[3], out[0], out[4],
                                            out[5], False)
        if collective:
            reduce_step({k: g[i] for k, i in GRAD_KEYS.items()}, radii=out[2])
        return g

    # ---- e2e: the call a user makes (nn.Module + autograd). Every step's host inputs (camera block + ground-truth
    # image, pinned) are copied H2D and its loss is read back D2H inside the timed region; the copies go through
    # the package's view pipeline (pipeline.ViewPrefetcher / DeferredScalar): view k+1 uploads while view k
    # rasterizes and the loss is consumed one step late, as a training loop that logs its loss would do.
    from neuralgaussiansplatting_b200.pipeline import DeferredScalar, ViewPrefetcher
    params = {k: w.scene[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "scales", "rotations", "shs")}
    pf = ViewPrefetcher(dev, {"cam": ((35,), torch.float32), "gt": ((3, cfg.height, cfg.width), torch.float32)})
    losses = DeferredScalar(lag=1)
    state = {"steps": 0, "submitted": 0, "read": 0}

    def e2e_begin(steps):
        state.update(steps=steps, submitted=0, read=0)

    def host_view(s):
        v = w.view_of(s)
        return {"cam": w.cam_host[v], "gt": w.gt_host[v]}

    def step_e2e(s):
        if state["submitted"] == 0:                       # first view of the region: nothing was prefetched
            pf.submit(host_view(s)); state["submitted"] += 1
        buf = pf.acquire()
        if state["submitted"] < state["steps"]:           # upload the next view behind this one's rasterization
            pf.submit(host_view(s + 1)); state["submitted"] += 1
        cam_dev, gt_dev = buf["cam"], buf["gt"]
        settings = GaussianRasterizationSettings(cfg.height, cfg.width, w.tan[0], w.tan[1], w.bg, 1.0,
                                                 cam_dev[0:16].view(4, 4), cam_dev[16:32].view(4, 4), 3,
                                                 cam_dev[32:35], False, False)
        m2d = torch.zeros_like(params["means3D"], requires_grad=True)
        img, radii = GaussianRasterizer(settings)(means3D=params["means3D"], means2D=m2d, shs=params["shs"],
                                                  colors_precomp=None, opacities=params["opacities"],
                                                  scales=params["scales"], rotations=params["rotations"],
                                                  cov3D_precomp=None)
        loss = (img - gt_dev).abs().mean()
        loss.backward()
        reduce_step({k: p.grad for k, p in params.items()}, radii=radii, means3D=params["means3D"].detach())
        if losses.push(loss) is not None:                 # loss of the previous step is now on the host
            state["read"] += 1
        for p in params.values():
            p.grad = None

    def e2e_finalize():
        reduce_finalize()
        pf.release()
        state["read"] += len(losses.drain())
        torch.cuda.current_stream().synchronize()
        assert state["read"] == state["steps"] and state["submitted"] == state["steps"], state

    for s in range(args.warmup):
        step_resident(s)
    reduce_finalize()
    launches0 = lib.b200s_launch_count()
    with ClockSampler(local_rank) as clk:
        ms_total = timed_region(step_resident, args.steps, world, dev, finalize=reduce_finalize)
        spread_resident = dict(LAST_STEP_SPREAD)
    launches = int(lib.b200s_launch_count() - launches0)
    nw = max(args.warmup, 3)
    e2e_begin(nw)
    for s in range(nw):
        step_e2e(s)
    e2e_finalize()
    e2e_begin(args.steps)
    ms_e2e = timed_region(step_e2e, args.steps, world, dev, finalize=e2e_finalize)

    out = {
        "metric": METRIC, "value": round(world * args.steps / (ms_total * 1e-3), 3), "unit": UNIT,
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms_total / args.steps, 4), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{cfg.name}: {w.P} Gaussians SH3, {cfg.width}x{cfg.height}, fwd+bwd, 1 view/step/GPU "
                               f"cycling {w.nviews} orbit views" + (", + SUM all-reduce of 59 floats/Gaussian"
                                                                     if world > 1 else ""),
                   "allreduce": ("none (1 GPU)" if world == 1 else
                                 "overlap: the reduce of view k runs on a side stream while view k+1 is rasterized "
                                 "(north_star); every reduce completes inside the timed region" if overlap else
                                 "blocking: reduce exposed after every view"),
                   "collective": collective_kind,
                   "parallelism": f"view-parallel x{world}, Gaussians replicated",
                   "l2": "inputs (708 MB of Gaussian parameters) exceed the 126 MB L2; no explicit flush",
                   },
        "impl_note": {"blend_mode": BLEND_MODES[_C.config.get("blend_mode", 0)],
                      "what": "this repo's sm_100a kernels (libb200splat.so) through the drop-in _C binding"},
        "step_spread": spread_resident,
        "e2e": {"value": round(world * args.steps / (ms_e2e * 1e-3), 3), "unit": UNIT,
                "ms_per_step": round(ms_e2e / args.steps, 4), "h2d_bytes_per_step": w.h2d_bytes,
                "d2h_bytes_per_step": w.d2h_bytes,
                "api": "GaussianRasterizer(nn.Module)+autograd, L1 loss vs host ground truth; per-step H2D through "
                       "pipeline.ViewPrefetcher (next view uploads behind the current one), loss read one step late"},
        "gpu_launches": launches,
        "clocks": clk.summary(),
    }
    if world > 1:
        cc = collective_check(w, reducer, _C, step_resident, GRAD_KEYS, rank, world, dev)
        out["collective_check"] = cc
        # the driver keeps `config` verbatim and drops nested extras: the verdict on the collective goes there too
        out["config"].update({"collective_ok": cc["all_ranks_ok"],
                              "max_rel_diff_vs_nccl": cc["max_rel_diff_vs_nccl"],
                              "bit_identical_to_fixed_order_sum": cc["bit_identical_to_fixed_order_sum"],
                              "replicas_bit_identical": cc["replicas_identical"]})
    _C.config["grad_out"] = None
    out["full_iteration"] = full_iteration_ours(args, w, _C, dev, reducer=reducer, world=world)
    if rank == 0:
        out.update(roofline_pass(args, w, lib, _C, lambda s: step_resident(s, collective=False)))
        if world == 1 and args.blend_mode == "strict":
            out["fast_mode"] = fast_mode_block(args, w, lib, _C, step_resident, dev)
        if not args.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline(w)
    if not args.no_render_only:
        del w
        torch.cuda.empty_cache()
        ro = render_only(args, rank, world, dev, "ours")
        if rank == 0:
            out["render_only"] = ro
    return out


def fast_mode_block(args, w, lib, _C, step_resident, dev):
    """"perf is reported in both". The same K steps with blend_mode = 1 (ex2.approx falloff in
    both blend kernels), its two blend stage times, and its deviation from strict mode on view 0 (image max-abs,
    n_contrib mismatches, gradient relative L2). The headline `value` stays strict mode."""
    import ctypes as C
    from neuralgaussiansplatting_b200 import _lib
    v = w.view_of(0)
    fa = w.fwd_args(v)

    def fwd_bwd():
        o = _C.forward_with_state(*fa, names=["n_contrib"])
        (R, color, radii, geom, binning, img), st = o
        g = _C.rasterize_gaussians_backward(fa[0], fa[1], radii, fa[2], fa[4], fa[5], fa[6], fa[7], fa[8], fa[9],
                                            fa[10], fa[11], w.dL[v], fa[14], fa[15], fa[16], geom, R, binning, img, False)
        return color, st["n_contrib"], g

    c0, n0, g0 = fwd_bwd()
    try:
        _C.config["blend_mode"] = 1
        c1, n1, g1 = fwd_bwd()
        for s in range(3):
            step_resident(s, collective=False)
        ms = timed_region(lambda s: step_resident(s, collective=False), args.steps, 1, dev)
        saved = w.view_of
        w.view_of = lambda s: v
        try:
            lib.b200s_profile_begin()
            for s in range(args.steps):
                step_resident(s, collective=False)
            st_ms = (C.c_float * len(_lib.STAGES))()
            st_calls = (C.c_int32 * len(_lib.STAGES))()
            _lib.check(lib.b200s_profile_end(st_ms, st_calls), "b200s_profile_end")
        finally:
            w.view_of = saved
    finally:
        _C.config["blend_mode"] = 0

    def rel(a, b):
        return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-300))

    i_f, i_b = _lib.STAGES.index("blend_fwd"), _lib.STAGES.index("blend_bwd")
    names = ("means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations")
    return {"what": BLEND_MODES[1], "value": round(args.steps / (ms * 1e-3), 3), "unit": UNIT,
            "ms_per_step": round(ms / args.steps, 4),
            "blend_fwd_ms": round(st_ms[i_f] / args.steps, 5), "blend_bwd_ms": round(st_ms[i_b] / args.steps, 5),
            "vs_strict": {"image_max_abs": float((c1 - c0).abs().max()),
                          "n_contrib_mismatch_pixels": int((n0 != n1).sum()), "pixels": int(n0.numel()),
                          "grad_rel_l2": {k: round(rel(g1[i], g0[i]), 9) for i, k in enumerate(names)
                                          if g0[i] is not None and k in ("means3D", "opacity", "sh", "scales", "rotations")}}}


BLEND_MODES = {0: "strict (expf, bit-exact n_contrib)", 1: "fast (ex2.approx falloff, forward and backward)"}


def collective_check(w, reducer, _C, step_resident, GRAD_KEYS, rank, world, dev):
    """Correctness of the gradient collective inside the driver-run path (N > 1), after the timed region: one more
    step in deterministic-backward mode (bit-reproducible local gradients) through the collective under test, then
    the same local gradients (a) all-gathered and summed in rank order with torch — must be BIT-identical to what the
    collective delivered — and (b) through one dist.all_reduce (NCCL's own order: equal to rounding)."""
    saved = (_C.config["grad_out"], _C.config["deterministic"])
    s_check = 10_000
    try:
        _C.config["deterministic"] = True
        g = step_resident(s_check, collective=True)
        reducer.wait()
        torch.cuda.synchronize(dev)
        got = {k: g[i].clone() for k, i in GRAD_KEYS.items()}
        _C.config["grad_out"] = None
        g2 = step_resident(s_check, collective=False)
        local = {k: g2[i] for k, i in GRAD_KEYS.items()}
        res = {"max_abs_diff_vs_nccl": 0.0, "max_rel_diff_vs_nccl": 0.0, "bit_identical_to_fixed_order_sum": True,
               "replicas_identical": True}
        for k, t in local.items():
            parts = [torch.empty_like(t) for _ in range(world)]
            dist.all_gather(parts, t.contiguous())
            fixed = parts[0].clone()
            for q in parts[1:]:
                fixed += q
            del parts
            if not torch.equal(fixed, got[k].reshape(fixed.shape)):
                res["bit_identical_to_fixed_order_sum"] = False
                res.setdefault("mismatch", {})[k] = float((fixed - got[k].reshape(fixed.shape)).abs().max())
            nc = t.clone()
            dist.all_reduce(nc, op=dist.ReduceOp.SUM)
            d = (nc - got[k].reshape(nc.shape)).abs().max()
            res["max_abs_diff_vs_nccl"] = max(res["max_abs_diff_vs_nccl"], float(d))
            res["max_rel_diff_vs_nccl"] = max(res["max_rel_diff_vs_nccl"], float(d / nc.abs().max().clamp_min(1e-30)))
            # replicas are compared by bits: two 64-bit wrap-around checksums of the raw words, one of them
            # position-weighted (a float64 sum could hide compensating differences and the sign of zero)
            words = got[k].contiguous().view(torch.int32).reshape(-1).to(torch.int64) & 0xffffffff
            pos = torch.arange(words.numel(), device=dev, dtype=torch.int64)
            mixed = (words * 0x9E3779B1 + pos * 0x85EBCA77) & 0x7fffffffffffffff
            chk = torch.stack([mixed.sum(), words.sum()]).reshape(2)
            del words, pos, mixed
            allc = [torch.empty_like(chk) for _ in range(world)]
            dist.all_gather(allc, chk)
            if not all(torch.equal(allc[0], c) for c in allc[1:]):
                res["replicas_identical"] = False
            del fixed, nc
        # in-switch (multimem.ld_reduce) sums are taken in the switch's order: identical replicas, equal to rounding
        exact_expected = not getattr(reducer, "multimem", False) and hasattr(reducer, "mem")
        ok = res["replicas_identical"] and res["max_rel_diff_vs_nccl"] <= 1e-6 and (
            res["bit_identical_to_fixed_order_sum"] or not exact_expected)
        flag = torch.tensor([1.0 if ok else 0.0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        res["all_ranks_ok"] = bool(flag.item() > 0)
        res["what"] = ("one extra step, deterministic backward; collective result vs rank-order torch sum of an "
                       "all_gather (bit-exact) and vs dist.all_reduce (NCCL order)")
        return res
    finally:
        _C.config["grad_out"], _C.config["deterministic"] = saved



def full_iteration_ours(args, w, _C, dev, reducer=None, world=1):
    """One whole training iteration with this repo's kernels either side of the rasterizer (SURVEY.md §8(f) ranks
    2-3): forward -> fused L1+SSIM loss and dL/dimage (loss.py) -> backward -> fused activation-backward + Adam +
    activation (optimizer.py). Device-resident ground truth; informational, the headline metric stays fwd+bwd.
    N > 1 (BASELINE.json configs[4]): the SYNCHRONOUS view-parallel iteration, every rank one view — the gradient
    collective completes before Adam and nothing is carried into the next iteration, as the reference's loop orders it
    (train.py:86-93,127-128). This is the number the overlapped headline metric does not show."""
    from neuralgaussiansplatting_b200 import parallel
    from neuralgaussiansplatting_b200.loss import fused_l1_ssim
    from neuralgaussiansplatting_b200.optimizer import FusedGaussianAdam
    s, cfg = w.scene, w.cfg
    raw = dict(xyz=s["means3D"].clone(), features=s["shs"].clone(),
               opacity=torch.logit(s["opacities"].clamp(1e-6, 1 - 1e-6)), scaling=torch.log(s["scales"]),
               rotation=s["rotations"].clone())
    opt = FusedGaussianAdam(raw)
    gts = [t.to(dev) for t in w.gt_host]
    E = torch.Tensor([])

    def step(i):
        v = w.view_of(i)
        c, a = w.cams[v], opt.activated
        fa = (w.bg, a["means3D"], E, a["opacities"], a["scales"], a["rotations"], 1.0, E, c.world_view_transform,
              c.full_proj_transform, w.tan[0], w.tan[1], cfg.height, cfg.width, a["shs"], 3, c.camera_center, False,
              False)
        out = _C.rasterize_gaussians(*fa)
        _, dL = fused_l1_ssim(out[1], gts[v], 0.2)
        g = _C.rasterize_gaussians_backward(fa[0], fa[1], out[2], fa[2], fa[4], fa[5], fa[6], fa[7], fa[8], fa[9],
                                            fa[10], fa[11], dL, fa[14], fa[15], fa[16], out[3], out[0], out[4],
                                            out[5], False)
        grads = {"means3D": g[3], "opacities": g[2], "shs": g[5], "scales": g[6], "rotations": g[7]}
        if reducer is not None:
            if is_peer:
                reducer.start(grads, means3D=a["means3D"], sh_degree=3)
            else:
                reducer.start(grads)
            reducer.wait()                       # Adam below consumes the summed gradients: nothing stays in flight
        opt.step(grads)

    is_peer = isinstance(reducer, parallel.PeerGradientAllReduce)
    saved = _C.config["grad_out"]
    try:
        # the backward writes straight into the collective's buffers, as in the headline loop
        _C.config["grad_out"] = reducer.grad_out if reducer is not None else None
        for i in range(max(args.warmup, 3)):
            step(i)
        ms = timed_region(step, args.steps, world, dev)
    finally:
        _C.config["grad_out"] = saved
    res = {"ms_per_step": round(ms / args.steps, 4), "iterations_per_s": round(args.steps / (ms * 1e-3), 2),
           "what": "rasterizer fwd + fused L1/SSIM loss (value+grad) + rasterizer bwd + fused activation/Adam step, "
                   "lambda_dssim 0.2, reference learning rates; all b200splat kernels"}
    if world > 1:
        res.update(views_per_iteration=world, views_per_s=round(world * args.steps / (ms * 1e-3), 2),
                   step_spread=dict(LAST_STEP_SPREAD),
                   replicas_in_sync=bool(parallel.replicas_in_sync(list(raw.values()))))
        res["what"] += ("; synchronous view-parallel iteration: one view per rank, the gradient collective finishes "
                        "before Adam, no overlap across iterations; max over ranks")
    return res


def full_iteration_reference(args, w, ref, kw_of, BK, dev, reducer=None, world=1):
    """The same iteration as the reference runs it: its rasterizer kernels, utils/loss_utils.py's torch expression
    (cuDNN depthwise convs + autograd), torch activations + autograd + torch.optim.Adam(eps=1e-15). N > 1: one view
    per rank and ONE flat ncclAllReduce of the five gradient tensors (the backward accumulates into the flat
    buffer's views) that completes before the optimizer step — the synchronous iteration, as in our arm."""
    import torch.nn.functional as F
    sys.path.insert(0, str(ROOT / "oracle"))
    import loss_ref                      # the reference's loss expression in torch ops, executed on the GPU here
    s = w.scene
    Pm = torch.nn.Parameter
    xyz, fdc, frest = Pm(s["means3D"].clone()), Pm(s["shs"][:, :1].clone()), Pm(s["shs"][:, 1:].clone())
    op = Pm(torch.logit(s["opacities"].clamp(1e-6, 1 - 1e-6)))
    sc, ro = Pm(torch.log(s["scales"])), Pm(s["rotations"].clone())
    adam = torch.optim.Adam([{"params": [xyz], "lr": 0.00016}, {"params": [fdc], "lr": 0.0025},
                             {"params": [frest], "lr": 0.0025 / 20.0}, {"params": [op], "lr": 0.05},
                             {"params": [sc], "lr": 0.005}, {"params": [ro], "lr": 0.001}], lr=0.0, eps=1e-15)
    gts = [t.to(dev) for t in w.gt_host]

    def step(i):
        v = w.view_of(i)
        acts = dict(means3D=xyz, shs=torch.cat((fdc, frest), dim=1), opacities=torch.sigmoid(op), scales=torch.exp(sc),
                    rotations=F.normalize(ro))
        kw = kw_of(v)
        kw.update({k: t.detach() for k, t in acts.items()})
        R, color, radii = ref.forward(sync=False, **kw)
        x = color.detach().requires_grad_(True)
        loss = 0.8 * loss_ref.l1(x, gts[v]) + 0.2 * (1.0 - loss_ref.ssim(x, gts[v]))
        loss.backward()
        gv = None
        if reducer is not None:
            v_ = reducer.grad_out()
            gv = {"means3D": v_["means3D"], "opacity": v_["opacities"], "sh": v_["shs"], "scales": v_["scales"],
                  "rotations": v_["rotations"]}
        g = ref.backward(radii=radii, dL_dpix=x.grad, sync=False, **({"out": gv} if gv is not None else {}),
                         **{k: kw[k] for k in BK})
        if reducer is not None:
            reducer.start({"means3D": g["means3D"], "opacities": g["opacity"], "shs": g["sh"], "scales": g["scales"],
                           "rotations": g["rotations"]})
            reducer.wait()
        torch.autograd.backward([acts["means3D"], acts["shs"], acts["opacities"], acts["scales"], acts["rotations"]],
                                [g["means3D"], g["sh"].reshape(acts["shs"].shape), g["opacity"].reshape(-1, 1),
                                 g["scales"], g["rotations"]])
        adam.step()
        adam.zero_grad(set_to_none=True)

    for i in range(max(args.warmup, 3)):
        step(i)
    ms = timed_region(step, args.steps, world, dev)
    res = {"ms_per_step": round(ms / args.steps, 4), "iterations_per_s": round(args.steps / (ms * 1e-3), 2),
           "what": "reference rasterizer kernels + utils/loss_utils.py torch expression + torch autograd through the "
                   "activations + torch.optim.Adam(eps=1e-15), same learning rates"}
    if world > 1:
        res.update(views_per_iteration=world, views_per_s=round(world * args.steps / (ms * 1e-3), 2),
                   step_spread=dict(LAST_STEP_SPREAD))
        res["what"] += ("; synchronous view-parallel iteration: one view per rank, one flat NCCL all-reduce that "
                        "finishes before Adam; max over ranks")
    return res


def ncu_traffic(stage):
    """(dram__bytes_read.sum + dram__bytes_write.sum, smsp__inst_executed.sum, source) per launch of the stage's
    kernel, from the newest committed `ncu --set full` summary under profiles/ (captured on the C3 workload by
    tools/_round.sh); Nones if absent."""
    import csv
    files = sorted((ROOT / "profiles").glob("r*_ncu_full_summary_C3.csv"))
    if not files:
        return None, None, None
    rows = list(csv.reader(open(files[-1])))
    hdr, units = rows[0], rows[1]
    try:
        ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
    except ValueError:
        return None, None, None
    ii = hdr.index("smsp__inst_executed.sum") if "smsp__inst_executed.sum" in hdr else None
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for r in rows[2:]:
        if f"{stage}_kernel" in r[0]:
            b = float(r[ir]) * scale.get(units[ir], 1.0) + float(r[iw]) * scale.get(units[iw], 1.0)
            inst = int(float(r[ii])) if ii is not None else None
            return int(b), inst, f"profiles/{files[-1].name} (ncu --set full, one launch, view 0)"
    return None, None, None


def roofline_pass(args, w, lib, _C, step_fn):
    """Separate pass of K steps on view_of(0) with the library's per-stage CUDA-event brackets enabled."""
    import ctypes as C
    from neuralgaussiansplatting_b200 import _lib
    dev = w.dev
    hbm_peak, peak_src, sm_max = load_peaks()
    sm_peaks, sm_src = load_sm_peaks(sm_max)
    fp32_peak = float(sm_peaks["fp32_tflops"])
    v = w.view_of(0)
    saved = w.view_of
    w.view_of = lambda s: v
    try:
        for s in range(2):
            step_fn(s)
        torch.cuda.synchronize(dev)
        lib.b200s_profile_begin()
        for s in range(args.steps):
            step_fn(s)
        ms = (C.c_float * len(_lib.STAGES))()
        calls = (C.c_int32 * len(_lib.STAGES))()
        _lib.check(lib.b200s_profile_end(ms, calls), "b200s_profile_end")
    finally:
        w.view_of = saved
    # exact pair counts and visibility of that view
    outp, a, keep = _C._forward_impl(*w.fwd_args(v))
    counts_dev = torch.zeros(3, dtype=torch.int64, device=dev)
    _lib.check(lib.b200s_count_pairs(C.byref(a), counts_dev.data_ptr(), torch.cuda.current_stream().cuda_stream),
               "b200s_count_pairs")
    counts = [int(x) for x in counts_dev.cpu()]
    P_vis, N = int((outp[2] > 0).sum()), int(outp[0])
    rows = stage_table(w, list(ms), list(calls), args.steps, counts, P_vis, N, hbm_peak, fp32_peak)
    top = max((k for k in rows if "frac" in rows[k]), key=lambda k: rows[k]["ms_per_step"])
    r = rows[top]
    traffic, winst, traffic_src = ncu_traffic(top)
    roof = {"kernel": top, "bound": r["bound"], "achieved": r["achieved"], "peak": r["peak"], "unit": r["unit"],
            "frac": r["frac"], "traffic": traffic, "traffic_source": traffic_src, "ms_per_launch": r["ms_per_step"],
            "peak_source": peak_src if r["bound"] == "hbm" else f"FP32 FMA rate {sm_src}",
            "share_of_step": round(r["ms_per_step"] / sum(x["ms_per_step"] for x in rows.values()), 4)}
    if winst and r["bound"] != "hbm":
        # what actually limits the blend kernels: warp-instruction issue (fp32 `achieved` counts only the reference's
        # algorithmic flops; the kernel also issues compares, selects, shuffles and shared-memory loads)
        ach = winst / (r["ms_per_step"] * 1e-3) / 1e9
        roof["issue"] = {"warp_inst_per_launch": winst, "achieved": round(ach, 1), "peak": sm_peaks["issue_ginst_per_s"],
                         "unit": "G warp-inst/s", "frac": round(ach / sm_peaks["issue_ginst_per_s"], 4),
                         "what": "smsp__inst_executed.sum of the committed ncu capture / the live launch time, against "
                                 f"the issue-slot peak {sm_src}"}
        # the same rate against the hardware's 4 issue slots per clock and SM at the nominal maximum clock
        hw_slots = 148 * 4 * 1965.0e6 / 1e9
        roof["issue"]["hw_slot_peak"] = round(hw_slots, 1)
        roof["issue"]["frac_of_hw_slots"] = round(ach / hw_slots, 4)
    return {"roofline": roof, "stages": rows,
            "workload_counts": {"P": w.P, "P_visible": P_vis, "num_rendered": N, "pairs_traversed_fwd": counts[0],
                                "pairs_contributing": counts[1], "pairs_traversed_bwd": counts[2]}}


# ----------------------------------------------------------------------------------------- CPU baseline
def cpu_baseline(w: Workload):
    """Reported baseline, not the target. (1) The CPU column the north_star names: the reference's torch-side
    preprocessing (build_covariance_from_scaling_rotation + eval_sh + clamp, gaussian_renderer/__init__.py:63-81)
    on all P Gaussians on the host cores; (2) the C oracle (oracle/oracle_cpu.c, the one place bench.py executes
    oracle/) for one full forward+backward view of the same workload, OpenMP on all cores."""
    sys.path.insert(0, str(ROOT / "tests"))
    from neuralgaussiansplatting_b200.sh_utils import build_covariance_from_scaling_rotation, eval_sh
    cores = os.cpu_count() or 1
    sc = {k: v.cpu() for k, v in w.scene.items()}
    cam = w.cams_host[0]

    def torch_pre():
        cov = build_covariance_from_scaling_rotation(sc["scales"], 1.0, sc["rotations"])
        d = sc["means3D"] - cam.camera_center[None, :]
        d = d / d.norm(dim=1, keepdim=True)
        col = torch.clamp_min(eval_sh(3, sc["shs"].transpose(1, 2), d) + 0.5, 0.0)
        return cov, col

    res = {}
    for nthreads in (1, cores):
        torch.set_num_threads(nthreads)
        torch_pre()
        best = min(_timeit(torch_pre) for _ in range(3))
        res[f"torch_cpu_preprocess_ms_{nthreads}t"] = round(best * 1e3, 2)
    out = {"unit": UNIT, "cores": cores, "kind": "port",
           "torch_cpu_preprocess": dict(res, P=w.P, what="covariance + eval_sh(deg 3) + clamp, fp32")}
    try:
        import oracle_cpu
        oracle_cpu.set_threads(cores)
        t0 = time.perf_counter()
        o = oracle_cpu.render_scene(w.cfg, sc, cam, w.bg.cpu(), dL_dpix=w.dL[0].cpu(), as_torch=False)
        dt = time.perf_counter() - t0
        out.update(value=round(1.0 / dt, 5), sample=f"1 view of {w.cfg.name} forward+backward through the C oracle "
                   f"({dt:.1f} s, {oracle_cpu.max_threads()} OpenMP threads; binning serial)",
                   oracle_num_rendered=int(o["num_rendered"]))
    except Exception as e:  # the oracle is test infrastructure; its absence must not hide the GPU numbers
        out.update(value=None, sample=f"C oracle unavailable: {e}")
    return out


def _timeit(fn):
    t0 = time.perf_counter()
    fn()
    return time.perf_counter() - t0


# ----------------------------------------------------------------------------------------- reference arm
def run_reference(args, rank, local_rank, world, dev):
    """The UNMODIFIED reference rasterizer (diff-gaussian-rasterization, compiled for sm_100 into oracle/_ref by
    oracle/build_ref.sh) on the same workload, metric and timing rules. Its torch binding's allocations
    (torch::full / nine torch::zeros, rasterize_points.cu:68-69,151-159) are reproduced by tests/refbind.py."""
    sys.path.insert(0, str(ROOT / "tests"))
    import refbind
    from neuralgaussiansplatting_b200 import parallel
    if not refbind.available("ref_dgr"):
        return {"impl": "reference", "unavailable": "oracle/_ref/libref_dgr.so was not built (needs /root/reference)"}
    w = Workload(args.workload, dev, rank, world)
    cfg = w.cfg
    h2d_bytes, d2h_bytes = w.h2d_bytes, w.d2h_bytes
    ref = refbind.RefDGR()
    # N > 1: the strongest plain-NCCL formulation — ONE ncclAllReduce on a flat buffer the reference's backward
    # accumulates into (the binding's nine torch::zeros become zero-fills of views of that buffer: same memset bytes)
    reducer = parallel.FlatGradientAllReduce(w.P, 16, dev) if world > 1 else None
    overlap = args.allreduce == "overlap"      # same communication mode as our arm: the all-reduce is ours in both
    s_ = w.scene

    def grad_views():
        if reducer is None:
            return None
        v = reducer.grad_out()
        return {"means3D": v["means3D"], "opacity": v["opacities"], "sh": v["shs"], "scales": v["scales"],
                "rotations": v["rotations"]}

    def reduce_step(g):
        if reducer is None:
            return
        grads = {"means3D": g["means3D"], "opacities": g["opacity"], "shs": g["sh"], "scales": g["scales"],
                 "rotations": g["rotations"]}
        if overlap:
            reducer.wait()
            reducer.start(grads)
        else:
            reducer.start(grads)
            reducer.wait()

    def reduce_finalize():
        if reducer is not None:
            reducer.wait()

    def kw_of(v, view=None, proj=None, campos=None):
        c = w.cams[v]
        return dict(bg=w.bg, means3D=s_["means3D"], opacities=s_["opacities"],
                    viewmatrix=c.world_view_transform if view is None else view,
                    projmatrix=c.full_proj_transform if proj is None else proj,
                    campos=c.camera_center if campos is None else campos, tan_fovx=w.tan[0], tan_fovy=w.tan[1],
                    W=cfg.width, H=cfg.height, shs=s_["shs"], scales=s_["scales"], rotations=s_["rotations"])

    BK = ("bg", "means3D", "viewmatrix", "projmatrix", "campos", "tan_fovx", "tan_fovy", "shs", "scales", "rotations")

    def fwd_bwd(kw, dL):
        R, color, radii = ref.forward(sync=False, **kw)
        g = ref.backward(radii=radii, dL_dpix=dL, sync=False, out=grad_views(), **{k: kw[k] for k in BK})
        reduce_step(g)
        return color

    def step_resident(s):
        v = w.view_of(s)
        fwd_bwd(kw_of(v), w.dL[v])

    # e2e: the SAME host loop as our arm (pipeline.ViewPrefetcher uploads view k+1 behind view k, the loss is read one
    # step late through DeferredScalar), so the e2e ratio compares rasterizers, not host pipelining.
    from neuralgaussiansplatting_b200.pipeline import DeferredScalar, ViewPrefetcher
    pf = ViewPrefetcher(dev, {"cam": ((35,), torch.float32), "gt": ((3, cfg.height, cfg.width), torch.float32)})
    losses = DeferredScalar(lag=1)
    state = {"steps": 0, "submitted": 0, "read": 0}

    def host_view(s):
        v = w.view_of(s)
        return {"cam": w.cam_host[v], "gt": w.gt_host[v]}

    def step_e2e(s):
        v = w.view_of(s)
        if state["submitted"] == 0:
            pf.submit(host_view(s)); state["submitted"] += 1
        buf = pf.acquire()
        if state["submitted"] < state["steps"]:
            pf.submit(host_view(s + 1)); state["submitted"] += 1
        cam_dev, gt_dev = buf["cam"], buf["gt"]
        kw = kw_of(v, cam_dev[0:16].view(4, 4), cam_dev[16:32].view(4, 4), cam_dev[32:35])
        R, color, radii = ref.forward(sync=False, **kw)
        diff = color - gt_dev
        loss = diff.abs().mean()
        dL = torch.sign(diff) / diff.numel()                 # d(L1 mean)/d(color), what autograd would hand back
        g = ref.backward(radii=radii, dL_dpix=dL, sync=False, out=grad_views(), **{k: kw[k] for k in BK})
        reduce_step(g)
        if losses.push(loss) is not None:
            state["read"] += 1

    def e2e_begin(steps):
        state.update(steps=steps, submitted=0, read=0)

    def e2e_finalize():
        reduce_finalize()
        pf.release()
        state["read"] += len(losses.drain())
        torch.cuda.current_stream().synchronize()
        assert state["read"] == state["steps"] and state["submitted"] == state["steps"], state

    for s in range(args.warmup):
        step_resident(s)
    reduce_finalize()
    with ClockSampler(local_rank) as clk:
        ms_total = timed_region(step_resident, args.steps, world, dev, finalize=reduce_finalize)
        spread_resident = dict(LAST_STEP_SPREAD)
    nw = max(args.warmup, 3)
    e2e_begin(nw)
    for s in range(nw):
        step_e2e(s)
    e2e_finalize()
    e2e_begin(args.steps)
    ms_e2e = timed_region(step_e2e, args.steps, world, dev, finalize=e2e_finalize)
    full_it = full_iteration_reference(args, w, ref, kw_of, BK, dev, reducer=reducer, world=world)
    ref.close()
    value = round(world * args.steps / (ms_total * 1e-3), 3)
    workload_name = (f"{cfg.name}: {w.P} Gaussians SH3, {cfg.width}x{cfg.height}, fwd+bwd, 1 view/step/GPU "
                     f"cycling {w.nviews} orbit views" + (", + SUM all-reduce of 59 floats/Gaussian" if world > 1 else ""))
    ro = None
    if not args.no_render_only:
        del w
        torch.cuda.empty_cache()
        ro = render_only(args, rank, world, dev, "reference")
    return {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms_total / args.steps, 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name,
                   "allreduce": ("none (1 GPU)" if world == 1 else
                                 "overlap: the reduce of view k runs on a side stream while view k+1 is rasterized "
                                 "(north_star); every reduce completes inside the timed region" if overlap else
                                 "blocking: reduce exposed after every view"),
                   "collective": ("none" if world == 1 else
                                  "NCCL all_reduce, one call on a flat buffer the backward accumulates into"),
                   "parallelism": f"view-parallel x{world}, Gaussians replicated",
                   "l2": "inputs (708 MB of Gaussian parameters) exceed the 126 MB L2; no explicit flush",
                   },
        "impl_note": {"blend_mode": "strict (expf; the reference has no other mode)",
                      "what": "unmodified reference diff-gaussian-rasterization kernels compiled for sm_100 "
                              "(oracle/_ref), same GPU, same inputs; the reference has no CPU implementation of this "
                              "path and no multi-GPU code (the all-reduce at N>1 is this repo's NCCL plumbing)"},
        "step_spread": spread_resident,
        "e2e": {"value": round(world * args.steps / (ms_e2e * 1e-3), 3), "unit": UNIT,
                "ms_per_step": round(ms_e2e / args.steps, 4), "h2d_bytes_per_step": h2d_bytes,
                "d2h_bytes_per_step": d2h_bytes,
                "api": "reference kernels behind the same host loop as our arm: pipeline.ViewPrefetcher (next view "
                       "uploads behind the current one), loss read one step late"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": 0, "kind": "reference",
                         "sample": "GPU run of the reference kernels (CUDA-only reference: no host-core path exists); "
                                   f"{args.steps} full steps"},
        "clocks": clk.summary(),
        **({"full_iteration": full_it} if full_it is not None else {}),
        **({"render_only": ro} if ro is not None else {}),
    }


# ----------------------------------------------------------------------------------------- render-only (C4)
def render_only(args, rank, world, dev, impl):
    """The other half of BASELINE.json's metric: render-only views/s, config C4 (6 M Gaussians, SH3, 1920x1080, 64
    orbit views), view-parallel — Gaussians replicated, view v on rank v mod N, no communication (the analogue of
    render.py:24-35's `for view in views: render(view, ...)["render"]` under no_grad). `value` = all 64 views x
    passes / max-over-ranks device time with everything resident; `e2e` adds, per view, the camera block H2D (pinned)
    and the rendered image D2H into pinned host memory (double-buffered, as render.py hands every image to the
    host for torchvision.save_image, :32) inside the timed region. The reference arm loops the unmodified reference
    forward (oracle/_ref) exactly the same way."""
    from neuralgaussiansplatting_b200 import parallel
    from neuralgaussiansplatting_b200.synthetic import CONFIGS, make_camera, make_scene
    cfg = CONFIGS["C4"]
    nviews, passes = 64, max(1, args.render_passes)
    sc = {k: v.to(dev) for k, v in make_scene(cfg, seed=0).items() if k not in ("feature_vector", "normal")}
    mine = parallel.shard_views(nviews, rank, world)
    cams_h = {v: make_camera(cfg, v) for v in mine}
    cams = {v: make_camera(cfg, v, device=dev) for v in mine}
    cam_host = {v: torch.cat([c.world_view_transform.reshape(-1), c.full_proj_transform.reshape(-1),
                              c.camera_center.reshape(-1)]).pin_memory() for v, c in cams_h.items()}
    bg = torch.zeros(3, device=dev)
    tan = (math.tan(cfg.fovx * 0.5), math.tan(cfg.fovy * 0.5))
    E = torch.Tensor([])
    if impl == "ours":
        from neuralgaussiansplatting_b200 import _C

        def render(view, proj, campos):
            return _C.rasterize_gaussians(bg, sc["means3D"], E, sc["opacities"], sc["scales"], sc["rotations"], 1.0, E,
                                          view, proj, tan[0], tan[1], cfg.height, cfg.width, sc["shs"], 3, campos,
                                          False, False)[1]
        closer = None
    else:
        sys.path.insert(0, str(ROOT / "tests"))
        import refbind
        ref = refbind.RefDGR()

        def render(view, proj, campos):
            return ref.forward(sync=False, bg=bg, means3D=sc["means3D"], opacities=sc["opacities"], viewmatrix=view,
                               projmatrix=proj, campos=campos, tan_fovx=tan[0], tan_fovy=tan[1], W=cfg.width,
                               H=cfg.height, shs=sc["shs"], scales=sc["scales"], rotations=sc["rotations"])[1]
        closer = ref.close

    RB = max(1, args.render_batch) if impl == "ours" else 1
    batches = [mine[i:i + RB] for i in range(0, len(mine), RB)]
    img_dev = [torch.empty(RB, 3, cfg.height, cfg.width, device=dev) for _ in range(2)] if impl == "ours" else None
    cnt = {"n": 0}

    def render_views(vs, views_proj_campos, out):
        """ours: ONE batched call (renderer.render_batch's path: shared workspace, one host sync per batch);
        reference: its forward per view, as render.py:31-32 loops."""
        if impl == "ours":
            _C.rasterize_gaussians_batch(bg, sc["means3D"], E, sc["opacities"], sc["scales"], sc["rotations"], 1.0, E,
                                         views_proj_campos, tan[0], tan[1], cfg.height, cfg.width, sc["shs"], 3, False,
                                         out_color=out[:len(vs)])
            return out[:len(vs)]
        return torch.stack([render(*vpc) for vpc in views_proj_campos]) if len(vs) > 1 else render(*views_proj_campos[0])[None]

    def resident_pass(_s):
        for vs in batches:
            i = cnt["n"] & 1
            cnt["n"] += 1
            cams_b = [(cams[v].world_view_transform, cams[v].full_proj_transform, cams[v].camera_center) for v in vs]
            render_views(vs, cams_b, img_dev[i] if img_dev is not None else None)

    img_host = [torch.empty(RB, 3, cfg.height, cfg.width).pin_memory() for _ in range(2)]
    img_ev = [None, None]
    img_keep = [None, None]
    cam_dev = [torch.empty(RB, 35, device=dev) for _ in range(2)]
    cam_host_b = {tuple(vs): torch.stack([cam_host[v] for v in vs]).pin_memory() for vs in batches}
    copy_stream = torch.cuda.Stream(device=dev)

    def e2e_pass(_s):
        for vs in batches:
            i = cnt["n"] & 1
            cnt["n"] += 1
            n = len(vs)
            cd = cam_dev[i]
            cd[:n].copy_(cam_host_b[tuple(vs)], non_blocking=True)
            if img_ev[i] is not None:
                img_ev[i].synchronize()          # the host consumed (here: may overwrite) this pinned buffer
            img = render_views(vs, [(cd[j, 0:16].view(4, 4), cd[j, 16:32].view(4, 4), cd[j, 32:35]) for j in range(n)],
                               img_dev[i] if img_dev is not None else None)
            rendered = torch.cuda.Event()
            rendered.record(torch.cuda.current_stream(dev))
            with torch.cuda.stream(copy_stream):     # batch k goes to the host while batch k+1 is rasterized
                copy_stream.wait_event(rendered)
                img_host[i][:n].copy_(img, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(copy_stream)
            img.record_stream(copy_stream)
            img_ev[i], img_keep[i] = ev, img
        for ev in img_ev:                        # every image of the pass is on the host before the region ends
            if ev is not None:
                torch.cuda.current_stream(dev).wait_event(ev)

    for v in mine[:max(3, min(len(mine), 8))]:                    # warm-up (also sizes the binning workspaces)
        c = cams[v]
        render(c.world_view_transform, c.full_proj_transform, c.camera_center)
    ms = timed_region(resident_pass, passes, world, dev)
    e2e_pass(0)
    ms_e2e = timed_region(e2e_pass, passes, world, dev)
    if closer is not None:
        closer()
    total = nviews * passes
    del sc
    torch.cuda.empty_cache()
    return {"metric": "render-only views/s at 6M Gaussians, SH3, 1920x1080 (C4), view-parallel",
            "value": round(total / (ms * 1e-3), 2), "unit": "views/s", "n_gpus": world, "views": nviews,
            "passes": passes, "scaling": "strong (64 views split over the ranks)",
            "ms_per_view_per_gpu": round(ms / max(len(mine) * passes, 1), 4),
            "e2e": {"value": round(total / (ms_e2e * 1e-3), 2), "unit": "views/s",
                    "ms_per_view_per_gpu": round(ms_e2e / max(len(mine) * passes, 1), 4),
                    "h2d_bytes_per_view": 35 * 4, "d2h_bytes_per_view": 3 * cfg.height * cfg.width * 4,
                    "what": "camera block H2D + rendered image D2H to pinned host memory per view on a copy stream (image k "
                            "leaves while view k+1 is rasterized), double-buffered; "
                            "no gather to rank 0 (every rank keeps / writes its own views, as render.py does per view)"},
            "impl": impl,
            "api": (f"_C.rasterize_gaussians_batch, {RB} views per call: one shared workspace, one host sync per batch"
                    if impl == "ours" else "reference forward per view (render.py:31-32)")}


# ----------------------------------------------------------------------------------------- main
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C3")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--collective", default="peer", choices=["peer", "nccl"],
                    help="N>1, our arm: gradient all-reduce as the library's own NVLink peer-memory kernel, or NCCL")
    ap.add_argument("--peer-ctas", type=int, default=0, help="0 = per-world-size default (16 / 12 / 8 at 2 / 4 / 8)")
    ap.add_argument("--sparse-collective", action="store_true",
                    help="peer collective: use the view's radii to skip all-zero dL/dSH rows (exact; fewer NVLink "
                         "bytes but fewer loads in flight per CTA — slower at 2 GPUs, see DESIGN.md §7)")
    ap.add_argument("--multimem", action="store_true", help="peer collective: NVSwitch multimem.ld_reduce/st")
    ap.add_argument("--exact-reconstruct", action="store_true",
                    help="rebuild dL/dsh with the bit-exact kernel even when the dense part is summed in the switch")
    ap.add_argument("--exact-collective", action="store_true",
                    help="peer collective: always the rank-order two-shot sum over peer loads (bit-identical to a "
                         "fixed-order sum); default: in-switch multimem sum of the dense part with 4+ ranks")
    ap.add_argument("--dense-collective", action="store_true",
                    help="peer collective: reduce the materialised 48-float dL/dsh rows (round-1 behaviour) instead of "
                         "the factored SH gradient")
    ap.set_defaults(compute_priority=True)
    ap.add_argument("--gather-pull", action="store_true",
                    help="factored collective: pull the factor slots with peer loads even when an NVSwitch multicast "
                         "mapping exists (default: multimem.st push)")
    ap.add_argument("--collective-high-priority", dest="compute_priority", action="store_false",
                    help="N>1: round-1 behaviour, the collective's side stream at high priority. Default: the rasterizer "
                         "runs on a high-priority stream and the collective at normal priority, so the collective's "
                         "CTAs only take SM slots the rasterizer leaves free (measured at 2 GPUs: 1.93 vs 2.06 ms/step)")
    ap.add_argument("--no-render-only", action="store_true", help="skip the C4 render-only block")
    ap.add_argument("--render-passes", type=int, default=2, help="timed passes over the 64 C4 views")
    ap.add_argument("--render-batch", type=int, default=8,
                    help="our arm, render-only: views per _C.rasterize_gaussians_batch call (one host sync per batch)")
    ap.add_argument("--blend-mode", default="strict", choices=["strict", "fast"],
                    help="our arm: strict = expf (bit-exact n_contrib, the parity gate); fast = ex2.approx falloff")
    ap.add_argument("--allreduce", default="overlap", choices=["overlap", "blocking"],
                    help="N>1: overlap the gradient all-reduce of view k with the rasterization of view k+1")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    rank, local_rank, world = env_world()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the rasterizer has no CPU path")
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit(f"--gpus {args.gpus} needs torchrun (WORLD_SIZE={world})")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")     # stdout carries exactly one JSON line
        import datetime
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=180))
    if args.compute_priority and world > 1 and args.impl == "ours":
        torch.cuda.set_stream(torch.cuda.Stream(device=dev, priority=-1))
    failed = False
    try:
        res = run_reference(args, rank, local_rank, world, dev) if args.impl == "reference" else \
            run_ours(args, rank, local_rank, world, dev)
        if rank == 0:
            print(json.dumps(res), flush=True)
        if res.get("collective_check", {}).get("all_ranks_ok") is False:
            failed = True
    finally:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
    if failed:
        # the line above is printed for the record; a collective that delivered a wrong sum is not a benchmark result
        raise SystemExit("collective_check failed: the gradient collective did not deliver the expected sum")


if __name__ == "__main__":
    main()
