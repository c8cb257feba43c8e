"""Drop-in for the reference's pybind `_C` modules, on top of the C ABI (include/b200splat.h).

Same function names, positional arity, argument meaning and return tuples as
  diff-gaussian-rasterization/ext.cpp:15-19  (rasterize_gaussians, rasterize_gaussians_backward,
                                              mark_visible; rasterize_points.h:18-67)
  rasterizer2/ext.cpp:15-20                  (+ get_idxmap2, get_idxmap_backward2;
                                              rasterize_points.h:53-79)
so `diff_gaussian_rasterization/__init__.py` and `rasterizer2/__init__.py` read like the
reference's. PyTorch supplies device memory (caching allocator) and the current stream only.
"""
from __future__ import annotations

import ctypes as C
import threading
from typing import Optional, Tuple

import torch

from . import _lib

NUM_CHANNELS = 3
NUM_FEATURES = 64

# Opt-in switches (module-level so callers of the reference API need not change signatures).
#   deterministic: fixed-order gradient reduction (bit-reproducible, no atomics)
#   optimistic   : size the binning workspace from the previous call of the same shape and
#                  enqueue the whole forward without the reference's mid-pipeline host sync
#                  (rasterizer_impl.cu:281); falls back to an exact re-run on overflow.
# grad_out: None, or a callable returning {"means3D","opacities","shs","scales","rotations"} -> preallocated float32
# CUDA tensors the backward writes its gradients into (parallel.PeerGradientAllReduce.grad_out: peer-mapped
# buffers, so the all-reduce kernel sums them in place with no staging copy).
# early_header: on the optimistic path read the header (num_rendered, preprocess error bits) right after the scan and
# let the host return while sort + blend are still running, so the caller's loss / backward launches queue up behind
# them instead of after a full-forward synchronise. Error bits raised by the later stages (look-back watchdogs) are
# then reported by the NEXT forward call of the thread (no extra synchronisation); debug=True keeps the full check.
# blend_mode: 0 = strict (accurate expf, bit-exact n_contrib: the parity gate), 1 = fast (ex2.approx falloff in the forward
# AND the backward blend, consistently; image within 1e-5, n_contrib may differ on a few pixels — tests/test_fast_mode_gpu.py).
config = {"deterministic": False, "optimistic": True, "growth": 1.25, "grad_out": None, "early_header": True,
          "blend_mode": 0}

_tls = threading.local()
_capacity_hint = {}   # (device index, P, W, H) -> last num_rendered


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def _f32c(t: torch.Tensor, device) -> torch.Tensor:
    """contiguous float32 on `device` (the reference calls .contiguous().data<float>())."""
    if t.numel() == 0:
        return t
    if t.dtype != torch.float32:
        raise RuntimeError("expected a float32 tensor")
    if t.device != device:
        raise RuntimeError(f"expected tensor on {device}, got {t.device}")
    t = t.contiguous()
    if t.data_ptr() % 16:
        t = t.clone()
    return t


def _stream(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def _pinned_header() -> Tuple[torch.Tensor, "_lib.Header"]:
    buf = getattr(_tls, "hdr", None)
    if buf is None:
        buf = torch.zeros(C.sizeof(_lib.Header), dtype=torch.uint8).pin_memory()
        _tls.hdr = buf
    return buf, _lib.Header.from_address(buf.data_ptr())


class _DeferredHeaderChecks:
    """Error bits raised by the late stages of a forward (look-back watchdogs of expand / sort / ranges) are not
    known when an early-header forward returns. Every such forward — and every backward, which consumes that state
    — queues an asynchronous copy of the geometry header into a small ring of pinned slots; the slots whose copy
    has completed are checked (without blocking) on entry to every later forward / backward call, and
    `flush_errors()` checks all of them with a synchronisation (render-only callers, end of a run). The copy is
    stream-ordered, so the geometry workspace need not be kept alive."""
    SLOTS = 8

    def __init__(self):
        self.bufs = [torch.zeros(C.sizeof(_lib.Header), dtype=torch.uint8).pin_memory() for _ in range(self.SLOTS)]
        self.events = [None] * self.SLOTS
        self.where = [""] * self.SLOTS
        self.n = 0

    def _check(self, i):
        self.events[i] = None
        hdr = _lib.Header.from_address(self.bufs[i].data_ptr())
        if hdr.error:
            _raise_header_error(hdr.error, f" (raised by the late stages of an earlier {self.where[i]} call)")

    def queue(self, lib, geom_ptr, dev, where):
        i = self.n % self.SLOTS
        self.n += 1
        if self.events[i] is not None:
            self.events[i].synchronize()
            self._check(i)
        _lib.check(lib.b200s_read_header(geom_ptr, self.bufs[i].data_ptr(), _stream(dev)), "b200s_read_header")
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(dev))
        self.events[i], self.where[i] = ev, where

    def poll(self, block=False):
        for i in range(self.SLOTS):
            ev = self.events[i]
            if ev is None:
                continue
            if block:
                ev.synchronize()
            elif not ev.query():
                continue
            self._check(i)


def _deferred() -> "_DeferredHeaderChecks":
    d = getattr(_tls, "deferred", None)
    if d is None:
        d = _tls.deferred = _DeferredHeaderChecks()
    return d


def flush_errors() -> None:
    """Wait for every queued header check of this thread and raise if a device-side watchdog fired."""
    _deferred().poll(block=True)


overflow_count = 0     # optimistic-capacity overflows since import (each one re-runs expand + sort + blend)
sync_count = 0         # host synchronisations issued by the forward paths of this module (tests / tuning)


def _raise_header_error(error: int, where: str = ""):
    raise RuntimeError(f"b200splat device error flags 0x{error:x}{where}"
                       + (" (point filtered although prefiltered is set)" if error & 2 else ""))


def _read_header(lib, geom: torch.Tensor, device) -> "_lib.Header":
    buf, hdr = _pinned_header()
    _lib.check(lib.b200s_read_header(geom.data_ptr(), buf.data_ptr(), _stream(device)), "b200s_read_header")
    torch.cuda.current_stream(device).synchronize()
    global sync_count
    sync_count += 1
    if hdr.error:
        raise RuntimeError(f"b200splat device error flags 0x{hdr.error:x}"
                           + (" (point filtered although prefiltered is set)" if hdr.error & 2 else ""))
    return hdr


def _require_cuda(t: torch.Tensor) -> torch.device:
    if not t.is_cuda:
        raise RuntimeError("b200splat kernels are CUDA-only (sm_100a); got a CPU tensor and there is no CPU path")
    return t.device


def _fill_forward_args(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                       viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W, sh, degree, campos, prefiltered, debug):
    dev = means3D.device
    a = _lib.ForwardArgs()
    P = means3D.size(0)
    M = sh.size(1) if sh.numel() != 0 else 0
    a.P, a.D, a.M, a.W, a.H = P, int(degree), int(M), int(W), int(H)
    a.tan_fovx, a.tan_fovy, a.scale_modifier = float(tan_fovx), float(tan_fovy), float(scale_modifier)
    a.prefiltered, a.debug = int(bool(prefiltered)), int(bool(debug))
    a.blend_mode = int(config["blend_mode"])
    keep = [_f32c(t, dev) for t in (bg, means3D, sh, colors, opacity, scales, rotations, cov3D_precomp,
                                    viewmatrix, projmatrix, campos)]
    (a.background, a.means3D, a.shs, a.colors_precomp, a.opacities, a.scales, a.rotations, a.cov3D_precomp,
     a.viewmatrix, a.projmatrix, a.campos) = [_ptr(t) for t in keep]
    return a, keep


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                        viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                        prefiltered, debug):
    return _forward_impl(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                         viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                         prefiltered, debug)[0]


def _forward_impl(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                  viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                  prefiltered, debug):
    """RasterizeGaussiansCUDA (reference rasterize_points.cu:35-115).

    Returns (num_rendered:int, out_color[3,H,W] f32, radii[P] i32, geomBuffer u8, binningBuffer u8, imgBuffer u8).
    The three buffers are opaque, as in the reference (their layout is private to the library)."""
    global sync_count, overflow_count
    if means3D.ndimension() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    dev = _require_cuda(means3D)
    lib = _lib.load()
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    with torch.cuda.device(dev):
        out_color = torch.empty((NUM_CHANNELS, H, W), dtype=torch.float32, device=dev)
        radii = torch.empty((P,), dtype=torch.int32, device=dev)
        u8 = dict(dtype=torch.uint8, device=dev)
        if P == 0:
            out_color.zero_()
            return (0, out_color, radii, torch.empty(0, **u8), torch.empty(0, **u8), torch.empty(0, **u8)), None, None
        a, keep = _fill_forward_args(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                                     cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W, sh, degree,
                                     campos, prefiltered, debug)
        geom = torch.empty(lib.b200s_geometry_workspace_size(P), **u8)
        img = torch.empty(lib.b200s_image_workspace_size(W, H), **u8)
        a.out_color, a.radii = out_color.data_ptr(), radii.data_ptr()
        a.geom_ws, a.geom_bytes = geom.data_ptr(), geom.numel()
        a.img_ws, a.img_bytes = img.data_ptr(), img.numel()
        stream = _stream(dev)

        def alloc_bin(capacity: int):
            capacity = max(int(capacity), 1)
            nbytes = lib.b200s_binning_workspace_size(capacity, W, H)
            binning = torch.empty(nbytes, **u8)
            a.bin_ws, a.bin_bytes = binning.data_ptr(), nbytes
            a.bin_capacity = lib.b200s_binning_capacity_for_bytes(nbytes)
            return binning

        key = (dev.index, P, W, H)
        hint = _capacity_hint.get(key) if config["optimistic"] and not debug else None
        binning = None
        deferred = _deferred()
        deferred.poll()                                # late-stage error bits of earlier calls (non-blocking)
        if hint is not None and config["early_header"]:
            binning = alloc_bin(int(hint * config["growth"]) + 4096)
            _lib.check(lib.b200s_forward_preprocess(C.byref(a), stream), "b200s_forward_preprocess")
            buf, hdr = _pinned_header()
            _lib.check(lib.b200s_read_header(geom.data_ptr(), buf.data_ptr(), stream), "b200s_read_header")
            ready = torch.cuda.Event()
            ready.record(torch.cuda.current_stream(dev))
            _lib.check(lib.b200s_forward_render(C.byref(a), stream), "b200s_forward_render")
            deferred.queue(lib, geom.data_ptr(), dev, "forward")
            ready.synchronize()                        # scan done: sort + blend of this view are still in flight
            sync_count += 1
            if hdr.error:
                _raise_header_error(hdr.error)
            if int(hdr.num_rendered) > int(a.bin_capacity):      # the device made stage 2 a no-op: redo it exactly
                binning = None
        elif hint is not None:
            binning = alloc_bin(int(hint * config["growth"]) + 4096)
            _lib.check(lib.b200s_forward(C.byref(a), stream), "b200s_forward")
            hdr = _read_header(lib, geom, dev)
            if hdr.overflow:
                binning = None
        else:
            _lib.check(lib.b200s_forward_preprocess(C.byref(a), stream), "b200s_forward_preprocess")
            hdr = _read_header(lib, geom, dev)
        num_rendered = int(hdr.num_rendered)
        if binning is None:
            if hint is not None:
                overflow_count += 1
            binning = alloc_bin(num_rendered)
            _lib.check(lib.b200s_forward_render(C.byref(a), stream), "b200s_forward_render")
            if hint is not None and config["early_header"]:
                deferred.queue(lib, geom.data_ptr(), dev, "forward")
        # capacity hint: running maximum with slow decay, so a loop that alternates small and large views does not
        # overflow (and silently re-run expand + sort + blend) on every large one
        _capacity_hint[key] = max(num_rendered, int(0.98 * _capacity_hint.get(key, 0)))
    return (num_rendered, out_color, radii, geom, binning, img), a, keep


def rasterize_gaussians_batch(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                              cameras, tan_fovx, tan_fovy, image_height, image_width, sh, degree, prefiltered=False,
                              out_color: Optional[torch.Tensor] = None):
    """B forwards of the same Gaussians for B cameras — the view loop of render.py:24-35 — with ONE geometry /
    image / binning workspace shared by all views, the images written straight into `out_color` [B,3,H,W], and ONE
    host synchronisation per batch (b200s_forward_batch): the per-view headers are queued into pinned memory and
    checked together. No gradient state is kept (use rasterize_gaussians per view for training).

    cameras: sequence of (viewmatrix[4,4], projmatrix[4,4], campos[3]) device tensors; tan_fovx / tan_fovy: floats or
    per-view sequences. Returns (num_rendered: list[int], out_color [B,3,H,W], radii int32 [B,P])."""
    global sync_count, overflow_count
    if means3D.ndimension() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    dev = _require_cuda(means3D)
    lib = _lib.load()
    B, P, H, W = len(cameras), means3D.size(0), int(image_height), int(image_width)
    tx = list(tan_fovx) if isinstance(tan_fovx, (list, tuple)) else [float(tan_fovx)] * B
    ty = list(tan_fovy) if isinstance(tan_fovy, (list, tuple)) else [float(tan_fovy)] * B
    with torch.cuda.device(dev):
        if out_color is None:
            out_color = torch.empty((B, NUM_CHANNELS, H, W), dtype=torch.float32, device=dev)
        elif (tuple(out_color.shape) != (B, NUM_CHANNELS, H, W) or out_color.dtype != torch.float32
              or out_color.device != dev or not out_color.is_contiguous()):
            raise RuntimeError(f"out_color must be a contiguous float32 [{B},3,{H},{W}] tensor on {dev}")
        radii = torch.empty((B, P), dtype=torch.int32, device=dev)
        if B == 0:
            return [], out_color, radii
        if P == 0:
            out_color.zero_()
            return [0] * B, out_color, radii
        u8 = dict(dtype=torch.uint8, device=dev)
        key = (dev.index, P, W, H)
        first = 0
        nums = [0] * B
        if _capacity_hint.get(key) is None:
            # no capacity estimate for this shape yet: the first view takes the exact (synchronising) path
            v, pj, cp = cameras[0]
            o = rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                                    cov3D_precomp, v, pj, tx[0], ty[0], H, W, sh, degree, cp, prefiltered, False)
            nums[0] = o[0]
            out_color[0].copy_(o[1]); radii[0].copy_(o[2])
            first = 1
            if B == 1:
                return nums, out_color, radii
        hint = _capacity_hint[key]
        geom = torch.empty(lib.b200s_geometry_workspace_size(P), **u8)
        img = torch.empty(lib.b200s_image_workspace_size(W, H), **u8)
        nbytes = lib.b200s_binning_workspace_size(max(int(hint * config["growth"]) + 4096, 1), W, H)
        binning = torch.empty(nbytes, **u8)
        capacity = lib.b200s_binning_capacity_for_bytes(nbytes)
        n = B - first
        views = (_lib.ForwardArgs * n)()
        keep = []
        for j in range(n):
            b = first + j
            v, pj, cp = cameras[b]
            a, kp = _fill_forward_args(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                                       cov3D_precomp, v, pj, tx[b], ty[b], H, W, sh, degree, cp, prefiltered, False)
            keep.append(kp)
            a.out_color, a.radii = out_color[b].data_ptr(), radii[b].data_ptr()
            a.geom_ws, a.geom_bytes = geom.data_ptr(), geom.numel()
            a.img_ws, a.img_bytes = img.data_ptr(), img.numel()
            a.bin_ws, a.bin_bytes, a.bin_capacity = binning.data_ptr(), nbytes, capacity
            views[j] = a
        hdrs = getattr(_tls, "batch_hdrs", None)
        if hdrs is None or hdrs.numel() < n * C.sizeof(_lib.Header):
            hdrs = _tls.batch_hdrs = torch.zeros(max(n, 16) * C.sizeof(_lib.Header), dtype=torch.uint8).pin_memory()
        stream = _stream(dev)
        _deferred().poll()
        _lib.check(lib.b200s_forward_batch(C.addressof(views), n, hdrs.data_ptr(), stream), "b200s_forward_batch")
        torch.cuda.current_stream(dev).synchronize()           # the ONE synchronisation of the batch
        sync_count += 1
        redo = []
        for j in range(n):
            h = _lib.Header.from_address(hdrs.data_ptr() + j * C.sizeof(_lib.Header))
            if h.error:
                _raise_header_error(h.error, f" (view {first + j} of the batch)")
            nums[first + j] = int(h.num_rendered)
            if h.overflow or int(h.num_rendered) > capacity:
                redo.append(first + j)
        for b in redo:                                          # capacity estimate too small for these views: exact re-run
            overflow_count += 1
            v, pj, cp = cameras[b]
            o = rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                                    cov3D_precomp, v, pj, tx[b], ty[b], H, W, sh, degree, cp, prefiltered, True)
            out_color[b].copy_(o[1]); radii[b].copy_(o[2])
        _capacity_hint[key] = max(max(nums), int(0.98 * _capacity_hint.get(key, 0)))
        del keep
    return nums, out_color, radii


def _backward_impl(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                   projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos, geomBuffer, R, binningBuffer,
                   imageBuffer, debug, want_all: bool):
    dev = _require_cuda(means3D)
    lib = _lib.load()
    P = means3D.size(0)
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    M = sh.size(1) if sh.numel() != 0 else 0
    f32 = dict(dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        has_sh = sh.numel() != 0
        has_sr = scales.numel() != 0
        pre = config["grad_out"]() if (config["grad_out"] is not None and has_sh and has_sr) else {}

        def out(name, shape):
            t = pre.get(name)
            if t is None:
                return torch.empty(shape, **f32)
            if tuple(t.shape) != tuple(shape) or t.dtype != torch.float32 or t.device != dev or not t.is_contiguous():
                raise RuntimeError(f"config['grad_out'][{name!r}] must be a contiguous float32 {tuple(shape)} tensor on {dev}")
            return t

        g_means3D = out("means3D", (P, 3))
        g_means2D = torch.empty((P, 3), **f32)
        g_opacity = out("opacities", (P, 1))
        g_sh = out("shs", (P, M, 3)) if has_sh else torch.zeros((P, M, 3), **f32)
        # factored SH gradient (parallel.PeerGradientAllReduce, factored=True): the backward emits 3 floats per
        # Gaussian + the camera centre instead of the 48-float rows; g_sh is filled by the collective's reconstruction
        sh_factors = pre.get("sh_factors") if has_sh else None
        if sh_factors is not None and (sh_factors.numel() < 3 * P + 4 or sh_factors.dtype != torch.float32
                                       or sh_factors.device != dev or not sh_factors.is_contiguous()):
            raise RuntimeError("config['grad_out']['sh_factors'] must be a contiguous float32 tensor of >= 3*P+4 floats")
        g_scales = out("scales", (P, 3)) if has_sr else torch.zeros((P, 3), **f32)
        g_rot = out("rotations", (P, 4)) if has_sr else torch.zeros((P, 4), **f32)
        g_colors = torch.empty((P, NUM_CHANNELS), **f32) if (want_all or not has_sh) else None
        g_cov3D = torch.empty((P, 6), **f32) if (want_all or not has_sr) else None
        if P == 0:
            return g_means2D, g_colors, g_opacity, g_means3D, g_cov3D, g_sh, g_scales, g_rot
        a = _lib.BackwardArgs()
        a.P, a.D, a.M, a.W, a.H = P, int(degree), int(M), int(W), int(H)
        a.tan_fovx, a.tan_fovy, a.scale_modifier = float(tan_fovx), float(tan_fovy), float(scale_modifier)
        a.debug = int(bool(debug))
        a.deterministic = int(bool(config["deterministic"]))
        a.blend_mode = int(config["blend_mode"])
        keep = [_f32c(t, dev) for t in (background, means3D, sh, colors, scales, rotations, cov3D_precomp, viewmatrix,
                                        projmatrix, campos, dL_dout_color)]
        (a.background, a.means3D, a.shs, a.colors_precomp, a.scales, a.rotations, a.cov3D_precomp, a.viewmatrix,
         a.projmatrix, a.campos, a.dL_dout_color) = [_ptr(t) for t in keep]
        radii_c = radii.contiguous()
        a.radii = radii_c.data_ptr()
        a.geom_ws, a.geom_bytes = geomBuffer.data_ptr(), geomBuffer.numel()
        a.img_ws, a.img_bytes = imageBuffer.data_ptr(), imageBuffer.numel()
        a.bin_ws, a.bin_bytes = _ptr(binningBuffer), binningBuffer.numel()
        a.bin_capacity = lib.b200s_binning_capacity_for_bytes(binningBuffer.numel()) if binningBuffer.numel() else 0
        a.num_rendered = int(R)
        nbytes = lib.b200s_backward_workspace_size(P, int(R) if a.deterministic else 0)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        a.bwd_ws, a.bwd_bytes = ws.data_ptr(), nbytes
        a.dL_dmeans3D, a.dL_dmeans2D, a.dL_dopacity = g_means3D.data_ptr(), g_means2D.data_ptr(), g_opacity.data_ptr()
        a.dL_dsh = g_sh.data_ptr() if (has_sh and sh_factors is None) else None
        a.dL_dsh_factors = sh_factors.data_ptr() if sh_factors is not None else None
        a.dL_dscales = g_scales.data_ptr() if has_sr else None
        a.dL_drotations = g_rot.data_ptr() if has_sr else None
        a.dL_dcolors = g_colors.data_ptr() if g_colors is not None else None
        a.dL_dcov3D = g_cov3D.data_ptr() if g_cov3D is not None else None
        deferred = _deferred()
        deferred.poll()
        _lib.check(lib.b200s_backward(C.byref(a), _stream(dev)), "b200s_backward")
        if config["optimistic"] and config["early_header"] and not debug:
            deferred.queue(lib, geomBuffer.data_ptr(), dev, "forward/backward")   # the state this backward consumed
        del keep
    return g_means2D, g_colors, g_opacity, g_means3D, g_cov3D, g_sh, g_scales, g_rot


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                 cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh, degree,
                                 campos, geomBuffer, R, binningBuffer, imageBuffer, debug):
    """RasterizeGaussiansBackwardCUDA (reference rasterize_points.cu:117-196).

    Returns (dL_dmeans2D[P,3], dL_dcolors[P,3], dL_dopacity[P,1], dL_dmeans3D[P,3], dL_dcov3D[P,6],
             dL_dsh[P,M,3], dL_dscales[P,3], dL_drotations[P,4])."""
    return _backward_impl(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp,
                          viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos, geomBuffer,
                          R, binningBuffer, imageBuffer, debug, want_all=True)


def mark_visible(means3D, viewmatrix, projmatrix):
    """markVisible (reference rasterize_points.cu:198-217): bool[P] near-plane frustum mask."""
    dev = _require_cuda(means3D)
    lib = _lib.load()
    P = means3D.size(0)
    with torch.cuda.device(dev):
        present = torch.zeros((P,), dtype=torch.bool, device=dev)
        if P != 0:
            m, v, pj = _f32c(means3D, dev), _f32c(viewmatrix, dev), _f32c(projmatrix, dev)
            _lib.check(lib.b200s_mark_visible(P, m.data_ptr(), v.data_ptr(), pj.data_ptr(), present.data_ptr(),
                                              _stream(dev)), "b200s_mark_visible")
    return present


def get_idxmap2(background, means3D, normal, scales, rotations, viewmatrix, projmatrix, tan_fovx, tan_fovy,
                image_height, image_width, feature_vector, sh, campos, debug):
    """GetIdxMapCUDA (reference rasterizer2/rasterize_points.cu:183-260).

    Returns (num_rendered:int, out_idx[H,W] i32, out_color[H,W,3], out_depth[H,W], out_feature[H,W,64]).
    `background`, `normal`, `scales`, `rotations`, `sh`, `tan_fov*` are accepted and unused, exactly as the
    reference kernels leave them unused (rasterizer_impl.cu:731-817)."""
    dev = _require_cuda(means3D)
    lib = _lib.load()
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    with torch.cuda.device(dev):
        f32 = dict(dtype=torch.float32, device=dev)
        out_idx = torch.empty((H, W), dtype=torch.int32, device=dev)
        out_color = torch.empty((H, W, NUM_CHANNELS), **f32)
        out_depth = torch.empty((H, W), **f32)
        out_feature = torch.empty((H, W, NUM_FEATURES), **f32)
        radii = torch.empty((max(P, 1),), dtype=torch.int32, device=dev)
        nr = torch.zeros((1,), dtype=torch.int32, device=dev)
        a = _lib.IdxmapArgs()
        a.P, a.W, a.H, a.debug = P, W, H, int(bool(debug))
        keep = [_f32c(t, dev) for t in (means3D, viewmatrix, projmatrix, campos, feature_vector)]
        a.means3D, a.viewmatrix, a.projmatrix, a.campos, a.feature_vector = [_ptr(t) for t in keep]
        a.out_idx, a.out_color, a.out_depth, a.out_feature = (out_idx.data_ptr(), out_color.data_ptr(),
                                                              out_depth.data_ptr(), out_feature.data_ptr())
        a.radii = radii.data_ptr()
        nbytes = lib.b200s_idxmap_workspace_size(P, W, H)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        a.ws, a.ws_bytes = ws.data_ptr(), nbytes
        a.num_rendered_dev = nr.data_ptr()
        _lib.check(lib.b200s_idxmap_forward(C.byref(a), _stream(dev)), "b200s_idxmap_forward")
        num_rendered = int(nr.item())   # the reference returns it as a host int (blocking D2H, :923)
        del keep
    return num_rendered, out_idx, out_color, out_depth, out_feature


def get_idxmap_backward2(means3D, image_height, image_width, campos, out_idx, grad_feature, debug):
    """GetIdxMapBackwardCUDA (reference rasterizer2/rasterize_points.cu:262-295) -> (dL_dfeature_vector[P,64],)"""
    dev = _require_cuda(means3D)
    lib = _lib.load()
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    with torch.cuda.device(dev):
        grad = torch.empty((P, NUM_FEATURES), dtype=torch.float32, device=dev)
        if P != 0:
            idx = out_idx.contiguous()
            gf = _f32c(grad_feature, dev)
            _lib.check(lib.b200s_idxmap_backward(P, W, H, idx.data_ptr(), gf.data_ptr(), grad.data_ptr(), _stream(dev)),
                       "b200s_idxmap_backward")
    return (grad,)


_STATE_DTYPES = {
    "depths": (torch.float32, lambda P, R, T, px: (P,)),
    "means2D": (torch.float32, lambda P, R, T, px: (P, 2)),
    "conic_opacity": (torch.float32, lambda P, R, T, px: (P, 4)),
    "rgb": (torch.float32, lambda P, R, T, px: (P, 3)),
    "tiles_touched": (torch.int32, lambda P, R, T, px: (P,)),
    "point_offsets": (torch.int32, lambda P, R, T, px: (P,)),
    "point_list": (torch.int32, lambda P, R, T, px: (R,)),
    "point_list_keys": (torch.int64, lambda P, R, T, px: (R,)),
    "ranges": (torch.int32, lambda P, R, T, px: (T, 2)),
    "n_contrib": (torch.int32, lambda P, R, T, px: (px,)),
    "final_T": (torch.float32, lambda P, R, T, px: (px,)),
    "clamped": (torch.uint8, lambda P, R, T, px: (P, 3)),
}


def forward_with_state(*args, names=None):
    """Parity-test helper: rasterize_gaussians(*args) plus a dict of the library's private state
    arrays (re-exported in the reference's layouts through b200s_export_state)."""
    out, a, keep = _forward_impl(*args)
    state = {}
    if a is None:
        return out, state
    lib = _lib.load()
    num_rendered, out_color, radii, geom, binning, img = out
    dev = out_color.device
    P, W, H = a.P, a.W, a.H
    T = ((W + 15) // 16) * ((H + 15) // 16)
    with torch.cuda.device(dev):
        for name in (_STATE_DTYPES if names is None else names):
            dtype, shape = _STATE_DTYPES[name]
            t = torch.zeros(shape(P, num_rendered, T, W * H), dtype=dtype, device=dev)
            nbytes = t.numel() * t.element_size()
            rc = lib.b200s_export_state(C.byref(a), _lib.STATE[name], num_rendered, t.data_ptr() if nbytes else None,
                                        nbytes, _stream(dev)) if nbytes else 0
            if rc < 0:
                raise RuntimeError(f"b200s_export_state({name}) failed: {rc}")
            state[name] = t
        torch.cuda.current_stream(dev).synchronize()
    return out, state
