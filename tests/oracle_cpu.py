"""TEST INFRASTRUCTURE: ctypes binding of oracle/_build/liboracle_cpu.so (the plain-C restatement of the
reference hot path, oracle/oracle_cpu.c). Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
reference legs may import this."""
from __future__ import annotations

import ctypes as C
import math
import subprocess
from pathlib import Path
from typing import Dict, Optional

import numpy as np
import torch

ORACLE_DIR = Path(__file__).resolve().parent.parent / "oracle"
LIB = ORACLE_DIR / "_build" / "liboracle_cpu.so"
_lib = None


def available() -> bool:
    try:
        load()
        return True
    except Exception:
        return False


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    src = ORACLE_DIR / "oracle_cpu.c"
    if not LIB.exists() or (src.exists() and src.stat().st_mtime > LIB.stat().st_mtime):
        subprocess.run(["bash", str(ORACLE_DIR / "build_cpu.sh")], check=True, capture_output=True)
    lib = C.CDLL(str(LIB))
    lib.oc_scan.restype = C.c_int64
    lib.oc_idxmap_forward.restype = C.c_int64
    lib.oc_higher_msb.restype = C.c_uint32
    lib.oc_max_threads.restype = C.c_int
    _lib = lib
    return lib


def _np(t, dtype=np.float32) -> Optional[np.ndarray]:
    if t is None:
        return None
    if isinstance(t, torch.Tensor):
        if t.numel() == 0:
            return None
        t = t.detach().cpu().numpy()
    return np.ascontiguousarray(t, dtype=dtype)


def _p(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def set_threads(n: int) -> None:
    load().oc_set_threads(C.c_int(int(n)))


def max_threads() -> int:
    return int(load().oc_max_threads())


def rasterize(*, means3D, opacities, viewmatrix, projmatrix, campos, bg, W, H, tan_fovx, tan_fovy, shs=None,
              colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None, degree=3, scale_modifier=1.0,
              dL_dpix=None, stop_after: Optional[str] = None) -> Dict[str, np.ndarray]:
    """Full pipeline (forward, optionally backward) of the CPU oracle. Returns numpy arrays in the reference's
    layouts. stop_after in {"preprocess", "bin"} truncates the pipeline (used for timing / big cases)."""
    lib = load()
    m = _np(means3D)
    P = m.shape[0]
    sh = _np(shs); cp = _np(colors_precomp); sc = _np(scales); rot = _np(rotations); c3p = _np(cov3D_precomp)
    op = _np(opacities).reshape(-1)
    view, proj, cam, bgc = _np(viewmatrix).reshape(-1), _np(projmatrix).reshape(-1), _np(campos), _np(bg)
    M = sh.shape[1] if sh is not None else 0
    W, H = int(W), int(H)
    o: Dict[str, np.ndarray] = {}
    o["radii"] = np.zeros(P, np.int32)
    o["means2D"] = np.zeros((P, 2), np.float32)
    o["depths"] = np.zeros(P, np.float32)
    o["cov3D"] = np.zeros((P, 6), np.float32)
    o["conic_opacity"] = np.zeros((P, 4), np.float32)
    o["rgb"] = np.zeros((P, 3), np.float32)
    o["clamped"] = np.zeros((P, 3), np.uint8)
    o["tiles_touched"] = np.zeros(P, np.uint32)
    o["rects"] = np.zeros((P, 4), np.int32)
    f = C.c_float
    lib.oc_preprocess(C.c_int(P), C.c_int(degree), C.c_int(M), C.c_int(W), C.c_int(H), f(tan_fovx), f(tan_fovy),
                      f(scale_modifier), _p(m), _p(sh), _p(cp), _p(op), _p(sc), _p(rot), _p(c3p), _p(view), _p(proj),
                      _p(cam), _p(o["radii"]), _p(o["means2D"]), _p(o["depths"]), _p(o["cov3D"]),
                      _p(o["conic_opacity"]), _p(o["rgb"]), _p(o["clamped"]), _p(o["tiles_touched"]), _p(o["rects"]))
    if stop_after == "preprocess":
        return o
    o["point_offsets"] = np.zeros(P, np.uint32)
    R = int(lib.oc_scan(C.c_int(P), _p(o["tiles_touched"]), _p(o["point_offsets"])))
    o["num_rendered"] = R
    T = ((W + 15) // 16) * ((H + 15) // 16)
    o["point_list_keys"] = np.zeros(max(R, 1), np.uint64)[:R]
    o["point_list"] = np.zeros(max(R, 1), np.uint32)[:R]
    keys = np.zeros(max(R, 1), np.uint64)
    vals = np.zeros(max(R, 1), np.uint32)
    o["ranges"] = np.zeros((T, 2), np.uint32)
    rc = lib.oc_bin(C.c_int(P), C.c_int(W), C.c_int(H), _p(o["radii"]), _p(o["depths"]), _p(o["rects"]),
                    _p(o["point_offsets"]), C.c_int64(R), _p(keys), _p(vals), _p(o["ranges"]))
    assert rc == 0
    o["point_list_keys"], o["point_list"] = keys[:R], vals[:R]
    if stop_after == "bin":
        return o
    o["final_T"] = np.zeros((H, W), np.float32)
    o["n_contrib"] = np.zeros((H, W), np.uint32)
    o["color"] = np.zeros((3, H, W), np.float32)
    counts = np.zeros(2, np.uint64)
    lib.oc_blend_forward(C.c_int(W), C.c_int(H), _p(o["ranges"]), _p(vals), _p(o["means2D"]), _p(o["conic_opacity"]),
                         _p(o["rgb"]), _p(bgc), _p(o["final_T"]), _p(o["n_contrib"]), _p(o["color"]), _p(counts))
    o["pairs_traversed"], o["pairs_contributing"] = int(counts[0]), int(counts[1])
    if dL_dpix is None:
        return o
    dl = _np(dL_dpix)
    g = dict(means2D=np.zeros((P, 3), np.float32), conic=np.zeros((P, 4), np.float32),
             opacity=np.zeros((P, 1), np.float32), colors=np.zeros((P, 3), np.float32),
             means3D=np.zeros((P, 3), np.float32), cov3D=np.zeros((P, 6), np.float32),
             sh=np.zeros((P, M, 3), np.float32), scales=np.zeros((P, 3), np.float32),
             rotations=np.zeros((P, 4), np.float32))
    lib.oc_blend_backward(C.c_int(P), C.c_int(W), C.c_int(H), _p(o["ranges"]), _p(vals), _p(o["means2D"]),
                          _p(o["conic_opacity"]), _p(o["rgb"]), _p(bgc), _p(o["final_T"]), _p(o["n_contrib"]), _p(dl),
                          _p(g["means2D"]), _p(g["conic"]), _p(g["opacity"]), _p(g["colors"]))
    cov_in = c3p if c3p is not None else o["cov3D"]
    lib.oc_preprocess_backward(C.c_int(P), C.c_int(degree), C.c_int(M), C.c_int(W), C.c_int(H), f(tan_fovx),
                               f(tan_fovy), f(scale_modifier), _p(m), _p(o["radii"]), _p(sh), _p(o["clamped"]),
                               _p(sc), _p(rot), _p(cov_in), _p(view), _p(proj), _p(cam), _p(g["means2D"]),
                               _p(g["conic"]), _p(g["colors"]), _p(g["means3D"]), _p(g["cov3D"]),
                               _p(g["sh"]) if M else None, _p(g["scales"]), _p(g["rotations"]))
    o["grads"] = g
    return o


def render_scene(cfg, sc, cam, bg, dL_dpix=None, degree=3, colors_precomp=None, cov3D_precomp=None,
                 scale_modifier=1.0, as_torch=True, stop_after=None):
    """Convenience: run on a synthetic scene dict + SyntheticCamera (CPU tensors)."""
    use_sh, use_sr = colors_precomp is None, cov3D_precomp is None
    o = rasterize(means3D=sc["means3D"], opacities=sc["opacities"], viewmatrix=cam.world_view_transform,
                  projmatrix=cam.full_proj_transform, campos=cam.camera_center, bg=bg, W=cam.image_width,
                  H=cam.image_height, tan_fovx=math.tan(cam.FoVx * 0.5), tan_fovy=math.tan(cam.FoVy * 0.5),
                  shs=sc["shs"] if use_sh else None, colors_precomp=colors_precomp,
                  scales=sc["scales"] if use_sr else None, rotations=sc["rotations"] if use_sr else None,
                  cov3D_precomp=cov3D_precomp, degree=degree, scale_modifier=scale_modifier, dL_dpix=dL_dpix,
                  stop_after=stop_after)
    if as_torch:
        conv = lambda v: torch.from_numpy(v.astype(np.int64) if v.dtype in (np.uint32, np.uint64) else v)  # noqa: E731
        o = {k: ({kk: conv(vv) for kk, vv in v.items()} if isinstance(v, dict) else
                 (conv(v) if isinstance(v, np.ndarray) else v)) for k, v in o.items()}
    return o


def mark_visible(means3D, viewmatrix) -> np.ndarray:
    m = _np(means3D)
    out = np.zeros(m.shape[0], np.uint8)
    load().oc_mark_visible(C.c_int(m.shape[0]), _p(m), _p(_np(viewmatrix).reshape(-1)), _p(out))
    return out.astype(bool)


def idxmap(*, means3D, viewmatrix, projmatrix, campos, feature_vector, W, H, grad_feature=None):
    lib = load()
    m, fv = _np(means3D), _np(feature_vector)
    P = m.shape[0]
    o = dict(idx=np.zeros((H, W), np.int32), color=np.zeros((H, W, 3), np.float32),
             depth=np.zeros((H, W), np.float32), feature=np.zeros((H, W, 64), np.float32),
             radii=np.zeros(P, np.int32))
    o["num_rendered"] = int(lib.oc_idxmap_forward(
        C.c_int(P), C.c_int(W), C.c_int(H), _p(m), _p(_np(viewmatrix).reshape(-1)), _p(_np(projmatrix).reshape(-1)),
        _p(_np(campos)), _p(fv), _p(o["idx"]), _p(o["color"]), _p(o["depth"]), _p(o["feature"]), _p(o["radii"])))
    if grad_feature is not None:
        o["dL_dfeature_vector"] = np.zeros((P, 64), np.float32)
        lib.oc_idxmap_backward(C.c_int(P), C.c_int(W), C.c_int(H), _p(o["idx"]), _p(_np(grad_feature)),
                               _p(o["dL_dfeature_vector"]))
    return o
