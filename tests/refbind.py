"""TEST INFRASTRUCTURE: ctypes binding of oracle/_ref/libref_dgr.so and libref_r2.so — the UNMODIFIED
reference rasterizers compiled for sm_100 by oracle/build_ref.sh behind the extern "C" shims
oracle/ref_shim_dgr.cu / ref_shim_r2.cu. Used only by tests/, __graft_entry__.smoke() and
`bench.py --impl reference`; never by the product path.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path
from typing import Dict, Optional

import numpy as np
import torch

REF_DIR = Path(__file__).resolve().parent.parent / "oracle" / "_ref"
_libs: Dict[str, C.CDLL] = {}

vp, ci, cf, cll = C.c_void_p, C.c_int, C.c_float, C.c_longlong


def available(name: str = "ref_dgr") -> bool:
    return (REF_DIR / f"lib{name}.so").exists()


def _load(name: str) -> C.CDLL:
    if name in _libs:
        return _libs[name]
    path = REF_DIR / f"lib{name}.so"
    if not path.exists():
        raise FileNotFoundError(f"{path} missing: run oracle/build_ref.sh where /root/reference exists")
    lib = C.CDLL(str(path))
    if name == "ref_dgr":
        lib.ref_dgr_create.restype = vp
        lib.ref_dgr_destroy.argtypes = [vp]
        lib.ref_dgr_forward.restype = ci
        lib.ref_dgr_forward.argtypes = [vp, ci, ci, ci, vp, ci, ci, vp, vp, vp, vp, vp, cf, vp, vp, vp, vp, vp, cf, cf,
                                        ci, vp, vp, ci]
        lib.ref_dgr_backward.restype = ci
        lib.ref_dgr_backward.argtypes = [vp, ci, ci, ci, ci, vp, ci, ci, vp, vp, vp, vp, cf, vp, vp, vp, vp, vp, cf,
                                         cf, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, ci]
        lib.ref_dgr_mark_visible.restype = ci
        lib.ref_dgr_mark_visible.argtypes = [ci, vp, vp, vp, vp]
        lib.ref_dgr_fetch.restype = cll
        lib.ref_dgr_fetch.argtypes = [vp, ci, vp, cll]
    else:
        lib.ref_r2_create.restype = vp
        lib.ref_r2_destroy.argtypes = [vp]
        lib.ref_r2_forward.restype = ci
        lib.ref_r2_forward.argtypes = [vp, ci, ci, ci, vp, ci, ci, vp, vp, vp, vp, cf, vp, vp, vp, vp, cf, cf, ci, vp,
                                       vp, ci]
        lib.ref_r2_backward.restype = ci
        lib.ref_r2_backward.argtypes = [vp, ci, ci, ci, ci, vp, ci, ci, vp, vp, vp, cf, vp, vp, vp, vp, cf, cf, vp,
                                        vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, ci]
        lib.ref_r2_getidxmap.restype = ci
        lib.ref_r2_getidxmap.argtypes = [vp, ci, ci, ci, ci, vp, vp, vp, vp, vp, vp, cf, cf, vp, vp, vp, vp, vp, vp,
                                         vp, vp, ci]
        lib.ref_r2_getidxmap_backward.restype = ci
        lib.ref_r2_getidxmap_backward.argtypes = [ci, ci, ci, vp, vp, vp, vp, vp, ci]
        lib.ref_r2_fetch_idx.restype = cll
        lib.ref_r2_fetch_idx.argtypes = [vp, ci, vp, cll]
    _libs[name] = lib
    return lib


def _p(t: Optional[torch.Tensor]):
    return None if t is None or t.numel() == 0 else t.data_ptr()


DGR_STATE = {
    "depths": (0, np.float32), "clamped": (1, np.uint8), "means2D": (2, np.float32), "cov3D": (3, np.float32),
    "conic_opacity": (4, np.float32), "rgb": (5, np.float32), "tiles_touched": (6, np.uint32),
    "point_offsets": (7, np.uint32), "point_list": (8, np.uint32), "point_list_keys": (9, np.uint64),
    "ranges": (10, np.uint32), "n_contrib": (11, np.uint32), "accum_alpha": (12, np.float32),
    "point_list_keys_unsorted": (13, np.uint64), "point_list_unsorted": (14, np.uint32),
}


class RefDGR:
    """The reference diff-gaussian-rasterization (CudaRasterizer::Rasterizer) on raw device pointers.
    NB: the reference launches on the legacy default stream; callers synchronise around it."""

    def __init__(self):
        self.lib = _load("ref_dgr")
        self.h = self.lib.ref_dgr_create()
        self.R = 0

    def close(self):
        if self.h:
            self.lib.ref_dgr_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def forward(self, *, bg, means3D, opacities, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, W, H,
                shs=None, colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None, degree=3,
                scale_modifier=1.0, prefiltered=False, debug=False, sync=True):
        dev = means3D.device
        P = means3D.shape[0]
        M = shs.shape[1] if shs is not None and shs.numel() else 0
        out = torch.zeros(3, H, W, dtype=torch.float32, device=dev)
        radii = torch.zeros(P, dtype=torch.int32, device=dev)
        if sync:   # the reference launches on the legacy default stream (= torch's default stream)
            torch.cuda.synchronize(dev)
        R = self.lib.ref_dgr_forward(self.h, P, degree, M, _p(bg), W, H, _p(means3D), _p(shs), _p(colors_precomp),
                                     _p(opacities), _p(scales), scale_modifier, _p(rotations), _p(cov3D_precomp),
                                     _p(viewmatrix), _p(projmatrix), _p(campos), tan_fovx, tan_fovy,
                                     int(prefiltered), out.data_ptr(), radii.data_ptr(), int(debug))
        if sync:
            torch.cuda.synchronize(dev)
        if R < 0:
            raise RuntimeError("reference forward failed")
        self.R, self.P, self.W, self.H, self.M, self.D = R, P, W, H, M, degree
        return R, out, radii

    def backward(self, *, bg, means3D, radii, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, dL_dpix,
                 shs=None, colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None,
                 scale_modifier=1.0, debug=False, sync=True, out=None):
        """`out`: optional {name: preallocated tensor} (e.g. views of one flat buffer for a single-call all-reduce);
        those are zero-filled in place instead of allocated with torch.zeros (same memset bytes as the binding)."""
        dev = means3D.device
        P, M = self.P, self.M

        def z(name, *s):
            t = out.get(name) if out else None
            if t is None:
                return torch.zeros(*s, dtype=torch.float32, device=dev)
            assert t.numel() == int(np.prod(s)) and t.dtype == torch.float32 and t.is_contiguous()
            return t.zero_().view(*s)

        g = dict(means2D=z("means2D", P, 3), conic=z("conic", P, 2, 2), opacity=z("opacity", P, 1),
                 colors=z("colors", P, 3), means3D=z("means3D", P, 3), cov3D=z("cov3D", P, 6), sh=z("sh", P, M, 3),
                 scales=z("scales", P, 3), rotations=z("rotations", P, 4))
        if sync:
            torch.cuda.synchronize(dev)
        rc = self.lib.ref_dgr_backward(self.h, P, self.D, M, self.R, _p(bg), self.W, self.H, _p(means3D), _p(shs),
                                       _p(colors_precomp), _p(scales), scale_modifier, _p(rotations),
                                       _p(cov3D_precomp), _p(viewmatrix), _p(projmatrix), _p(campos), tan_fovx,
                                       tan_fovy, radii.data_ptr(), dL_dpix.data_ptr(), g["means2D"].data_ptr(),
                                       g["conic"].data_ptr(), g["opacity"].data_ptr(), g["colors"].data_ptr(),
                                       g["means3D"].data_ptr(), g["cov3D"].data_ptr(), _p(g["sh"]),
                                       g["scales"].data_ptr(), g["rotations"].data_ptr(), int(debug))
        if sync:
            torch.cuda.synchronize(dev)
        if rc != 0:
            raise RuntimeError("reference backward failed")
        return g

    def mark_visible(self, means3D, viewmatrix, projmatrix):
        P = means3D.shape[0]
        present = torch.zeros(P, dtype=torch.bool, device=means3D.device)
        torch.cuda.synchronize()
        self.lib.ref_dgr_mark_visible(P, _p(means3D), _p(viewmatrix), _p(projmatrix), _p(present))
        torch.cuda.synchronize()
        return present

    def fetch(self, name: str) -> np.ndarray:
        which, dt = DGR_STATE[name]
        T = ((self.W + 15) // 16) * ((self.H + 15) // 16)
        n = {"depths": self.P, "clamped": 3 * self.P, "means2D": 2 * self.P, "cov3D": 6 * self.P,
             "conic_opacity": 4 * self.P, "rgb": 3 * self.P, "tiles_touched": self.P, "point_offsets": self.P,
             "point_list": self.R, "point_list_keys": self.R, "ranges": 2 * T, "n_contrib": self.W * self.H,
             "accum_alpha": self.W * self.H, "point_list_keys_unsorted": self.R, "point_list_unsorted": self.R}[name]
        buf = np.zeros(n, dtype=dt)
        if n == 0:
            return buf
        got = self.lib.ref_dgr_fetch(self.h, which, buf.ctypes.data, buf.nbytes)
        if got != buf.nbytes:
            raise RuntimeError(f"ref fetch {name}: got {got}, want {buf.nbytes}")
        return buf


class RefR2:
    """The reference rasterizer2: idxmap pipeline (getidxmap / getidxmap_backward) and stripped alpha path."""

    def __init__(self):
        self.lib = _load("ref_r2")
        self.h = self.lib.ref_r2_create()

    def close(self):
        if self.h:
            self.lib.ref_r2_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def forward(self, *, bg, means3D, opacities, scales, rotations, viewmatrix, projmatrix, campos, tan_fovx,
                tan_fovy, W, H, shs, degree=3, scale_modifier=1.0, prefiltered=False, debug=False):
        """rasterizer2's stripped alpha path, `_C.rasterize_gaussians` with 17 arguments (ext.cpp:16,
        rasterize_points.cu:27-100): no colors_precomp / cov3D_precomp inputs."""
        dev = means3D.device
        P, M = means3D.shape[0], shs.shape[1]
        out = torch.zeros(3, H, W, dtype=torch.float32, device=dev)
        radii = torch.zeros(P, dtype=torch.int32, device=dev)
        torch.cuda.synchronize(dev)
        R = self.lib.ref_r2_forward(self.h, P, degree, M, _p(bg), W, H, _p(means3D), _p(shs), _p(opacities),
                                    _p(scales), scale_modifier, _p(rotations), _p(viewmatrix), _p(projmatrix),
                                    _p(campos), tan_fovx, tan_fovy, int(prefiltered), out.data_ptr(),
                                    radii.data_ptr(), int(debug))
        torch.cuda.synchronize(dev)
        if R < 0:
            raise RuntimeError("reference rasterizer2 forward failed")
        self.R, self.P, self.W, self.H, self.M, self.D = R, P, W, H, M, degree
        return R, out, radii

    def backward(self, *, bg, means3D, radii, shs, scales, rotations, viewmatrix, projmatrix, campos, tan_fovx,
                 tan_fovy, dL_dpix, scale_modifier=1.0, debug=False):
        """`_C.rasterize_gaussians_backward` of rasterizer2 (ext.cpp:17, rasterize_points.cu:105-180): returns the
        binding's 6-tuple (dL_dmeans2D, dL_dopacity, dL_dmeans3D, dL_dsh, dL_dscales, dL_drotations) (:177)."""
        dev = means3D.device
        P, M = self.P, self.M
        z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=dev)  # noqa: E731
        g2d, gcon, gop, gcol, g3d, gcov, gsh, gs, gr = (z(P, 3), z(P, 2, 2), z(P, 1), z(P, 3), z(P, 3), z(P, 6),
                                                         z(P, M, 3), z(P, 3), z(P, 4))
        torch.cuda.synchronize(dev)
        rc = self.lib.ref_r2_backward(self.h, P, self.D, M, self.R, _p(bg), self.W, self.H, _p(means3D), _p(shs),
                                      _p(scales), scale_modifier, _p(rotations), _p(viewmatrix), _p(projmatrix),
                                      _p(campos), tan_fovx, tan_fovy, radii.data_ptr(), dL_dpix.data_ptr(),
                                      g2d.data_ptr(), gcon.data_ptr(), gop.data_ptr(), gcol.data_ptr(), g3d.data_ptr(),
                                      gcov.data_ptr(), gsh.data_ptr(), gs.data_ptr(), gr.data_ptr(), int(debug))
        torch.cuda.synchronize(dev)
        if rc != 0:
            raise RuntimeError("reference rasterizer2 backward failed")
        return g2d, gop, g3d, gsh, gs, gr

    def getidxmap(self, *, means3D, normal, scales, rotations, viewmatrix, projmatrix, tan_fovx, tan_fovy,
                  feature_vector, shs, campos, W, H, debug=False):
        dev = means3D.device
        P = means3D.shape[0]
        M = shs.shape[1] if shs is not None and shs.numel() else 0
        idx = torch.zeros(H, W, dtype=torch.int32, device=dev)
        col = torch.zeros(H, W, 3, dtype=torch.float32, device=dev)
        dep = torch.zeros(H, W, dtype=torch.float32, device=dev)
        feat = torch.zeros(H, W, 64, dtype=torch.float32, device=dev)
        radii = torch.zeros(P, dtype=torch.int32, device=dev)
        torch.cuda.synchronize(dev)
        R = self.lib.ref_r2_getidxmap(self.h, P, M, W, H, _p(means3D), _p(normal), _p(scales), _p(rotations),
                                      _p(viewmatrix), _p(projmatrix), tan_fovx, tan_fovy, _p(feature_vector), _p(shs),
                                      _p(campos), idx.data_ptr(), col.data_ptr(), dep.data_ptr(), feat.data_ptr(),
                                      radii.data_ptr(), int(debug))
        torch.cuda.synchronize(dev)
        if R < 0:
            raise RuntimeError("reference getidxmap failed")
        return R, idx, col, dep, feat, radii

    def getidxmap_backward(self, *, means3D, campos, out_idx, grad_feature, W, H):
        P = means3D.shape[0]
        g = torch.zeros(P, 64, dtype=torch.float32, device=means3D.device)
        torch.cuda.synchronize()
        self.lib.ref_r2_getidxmap_backward(P, W, H, _p(means3D), _p(campos), out_idx.data_ptr(),
                                           grad_feature.data_ptr(), g.data_ptr(), 0)
        torch.cuda.synchronize()
        return g
