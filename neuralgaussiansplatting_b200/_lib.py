"""ctypes binding of libb200splat.so — the C ABI declared in include/b200splat.h.

There is no CPU or PyTorch fallback: if the shared library is missing it is built in-tree
with nvcc (see build.py); if that is impossible the import fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

from . import build as _build

c_f32p = C.c_void_p  # device pointers are passed as integers


class Header(C.Structure):
    _fields_ = [
        ("num_rendered", C.c_uint32),
        ("overflow", C.c_uint32),
        ("error", C.c_uint32),
        ("num_visible", C.c_uint32),
        ("reserved", C.c_uint32 * 12),
    ]


class ForwardArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("D", C.c_int32), ("M", C.c_int32), ("W", C.c_int32), ("H", C.c_int32),
        ("tan_fovx", C.c_float), ("tan_fovy", C.c_float), ("scale_modifier", C.c_float),
        ("prefiltered", C.c_int32), ("debug", C.c_int32),
        ("background", C.c_void_p), ("means3D", C.c_void_p), ("shs", C.c_void_p),
        ("colors_precomp", C.c_void_p), ("opacities", C.c_void_p), ("scales", C.c_void_p),
        ("rotations", C.c_void_p), ("cov3D_precomp", C.c_void_p), ("viewmatrix", C.c_void_p),
        ("projmatrix", C.c_void_p), ("campos", C.c_void_p),
        ("out_color", C.c_void_p), ("radii", C.c_void_p),
        ("geom_ws", C.c_void_p), ("geom_bytes", C.c_size_t),
        ("img_ws", C.c_void_p), ("img_bytes", C.c_size_t),
        ("bin_ws", C.c_void_p), ("bin_bytes", C.c_size_t),
        ("bin_capacity", C.c_int64),
        ("blend_mode", C.c_int32), ("reserved", C.c_int32),
    ]


class BackwardArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("D", C.c_int32), ("M", C.c_int32), ("W", C.c_int32), ("H", C.c_int32),
        ("tan_fovx", C.c_float), ("tan_fovy", C.c_float), ("scale_modifier", C.c_float),
        ("debug", C.c_int32), ("deterministic", C.c_int32),
        ("background", C.c_void_p), ("means3D", C.c_void_p), ("shs", C.c_void_p),
        ("colors_precomp", C.c_void_p), ("opacities", C.c_void_p), ("scales", C.c_void_p),
        ("rotations", C.c_void_p), ("cov3D_precomp", C.c_void_p), ("viewmatrix", C.c_void_p),
        ("projmatrix", C.c_void_p), ("campos", C.c_void_p), ("radii", C.c_void_p),
        ("dL_dout_color", C.c_void_p),
        ("geom_ws", C.c_void_p), ("geom_bytes", C.c_size_t),
        ("img_ws", C.c_void_p), ("img_bytes", C.c_size_t),
        ("bin_ws", C.c_void_p), ("bin_bytes", C.c_size_t),
        ("bin_capacity", C.c_int64),
        ("bwd_ws", C.c_void_p), ("bwd_bytes", C.c_size_t),
        ("num_rendered", C.c_int64),
        ("dL_dmeans3D", C.c_void_p), ("dL_dmeans2D", C.c_void_p), ("dL_dopacity", C.c_void_p),
        ("dL_dsh", C.c_void_p), ("dL_dscales", C.c_void_p), ("dL_drotations", C.c_void_p),
        ("dL_dcolors", C.c_void_p), ("dL_dcov3D", C.c_void_p), ("dL_dconic", C.c_void_p),
        ("blend_mode", C.c_int32), ("reserved", C.c_int32),
        ("dL_dsh_factors", C.c_void_p),
    ]


class IdxmapArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("W", C.c_int32), ("H", C.c_int32), ("debug", C.c_int32),
        ("means3D", C.c_void_p), ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p),
        ("campos", C.c_void_p), ("feature_vector", C.c_void_p),
        ("out_idx", C.c_void_p), ("out_color", C.c_void_p), ("out_depth", C.c_void_p),
        ("out_feature", C.c_void_p), ("radii", C.c_void_p),
        ("ws", C.c_void_p), ("ws_bytes", C.c_size_t),
        ("num_rendered_dev", C.c_void_p),
    ]


MAX_PEERS = 16
ABI_VERSION = 3


class PeerReduceArgs(C.Structure):
    _fields_ = [
        ("peer_bufs", C.c_void_p * MAX_PEERS), ("peer_flags", C.c_void_p * MAX_PEERS),
        ("multicast_buf", C.c_void_p), ("n_floats", C.c_uint64),
        ("rank", C.c_int32), ("world", C.c_int32), ("epoch", C.c_uint32), ("ctas", C.c_int32),
        ("use_multimem", C.c_int32), ("sparse_rows", C.c_int32),
        ("radii", C.c_void_p), ("sparse_offset_floats", C.c_uint64), ("vis_offset_bytes", C.c_uint64),
        ("gather_offset_floats", C.c_uint64), ("gather_floats", C.c_uint64),
    ]


class ShGradArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("D", C.c_int32), ("M", C.c_int32), ("views", C.c_int32),
        ("means3D", C.c_void_p), ("factors", C.c_void_p), ("stride_floats", C.c_uint64), ("dL_dsh", C.c_void_p),
        ("fast", C.c_int32), ("reserved_", C.c_int32),
    ]


ADAM_MAX_TENSORS = 8
ADAM_MODES = {"identity": 0, "sh": 1, "exp": 2, "sigmoid": 3, "normalize": 4}


class AdamTensor(C.Structure):
    _fields_ = [
        ("param", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p), ("grad", C.c_void_p),
        ("activated", C.c_void_p), ("n", C.c_int64), ("neg_step", C.c_float), ("neg_step_rest", C.c_float),
        ("mode", C.c_int32), ("reserved", C.c_int32),
    ]


class AdamArgs(C.Structure):
    _fields_ = [
        ("t", AdamTensor * ADAM_MAX_TENSORS), ("count", C.c_int32), ("beta1", C.c_float), ("beta2", C.c_float),
        ("one_minus_beta1", C.c_float), ("one_minus_beta2", C.c_float), ("eps", C.c_float),
        ("bias_correction2_sqrt", C.c_float), ("reserved", C.c_int32),
    ]


class LossArgs(C.Structure):
    _fields_ = [
        ("C", C.c_int32), ("H", C.c_int32), ("W", C.c_int32), ("lambda_dssim", C.c_float), ("grad_scale", C.c_float),
        ("reserved", C.c_int32), ("image", C.c_void_p), ("gt", C.c_void_p), ("loss_out", C.c_void_p),
        ("dL_dimage", C.c_void_p), ("ws", C.c_void_p), ("ws_bytes", C.c_size_t),
    ]


STATE = {
    "depths": 0, "means2D": 1, "conic_opacity": 2, "rgb": 3, "tiles_touched": 4, "point_offsets": 5,
    "point_list": 6, "point_list_keys": 7, "ranges": 8, "n_contrib": 9, "final_T": 10, "clamped": 11,
}

STAGES = ("preprocess", "scan", "expand", "sort", "ranges", "blend_fwd", "blend_bwd", "preprocess_bwd", "idxmap",
          "idxmap_bwd", "misc")

ERRORS = {
    -1: "B200S_E_BADARG (NULL or inconsistent arguments)",
    -2: "B200S_E_WORKSPACE (workspace smaller than required)",
    -3: "B200S_E_OVERFLOW (num_rendered exceeds the binning capacity)",
    -4: "B200S_E_UNSUPPORTED",
    -5: "B200S_E_INTERNAL (device watchdog)",
}

# every symbol include/b200splat.h declares: (name, restype, argtypes)
SYMBOLS = [
    ("b200s_abi_version", C.c_int, []),
    ("b200s_build_info", C.c_char_p, []),
    ("b200s_sizeof", C.c_size_t, [C.c_int32]),
    ("b200s_geometry_workspace_size", C.c_size_t, [C.c_int32]),
    ("b200s_image_workspace_size", C.c_size_t, [C.c_int32, C.c_int32]),
    ("b200s_binning_workspace_size", C.c_size_t, [C.c_int64, C.c_int32, C.c_int32]),
    ("b200s_binning_capacity_for_bytes", C.c_int64, [C.c_size_t]),
    ("b200s_backward_workspace_size", C.c_size_t, [C.c_int32, C.c_int64]),
    ("b200s_forward_preprocess", C.c_int, [C.POINTER(ForwardArgs), C.c_void_p]),
    ("b200s_forward_render", C.c_int, [C.POINTER(ForwardArgs), C.c_void_p]),
    ("b200s_forward", C.c_int, [C.POINTER(ForwardArgs), C.c_void_p]),
    ("b200s_read_header", C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    ("b200s_forward_batch", C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]),
    ("b200s_backward", C.c_int, [C.POINTER(BackwardArgs), C.c_void_p]),
    ("b200s_mark_visible", C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    ("b200s_idxmap_workspace_size", C.c_size_t, [C.c_int32, C.c_int32, C.c_int32]),
    ("b200s_idxmap_forward", C.c_int, [C.POINTER(IdxmapArgs), C.c_void_p]),
    ("b200s_idxmap_forward_static", C.c_int, [C.POINTER(IdxmapArgs), C.c_void_p]),
    ("b200s_idxmap_backward", C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    ("b200s_profile_begin", C.c_int, []),
    ("b200s_profile_end", C.c_int, [C.c_void_p, C.c_void_p]),
    ("b200s_launch_count", C.c_uint64, []),
    ("b200s_count_pairs", C.c_int, [C.POINTER(ForwardArgs), C.c_void_p, C.c_void_p]),
    ("b200s_peer_allreduce_flag_bytes", C.c_size_t, [C.c_int32, C.c_int32]),
    ("b200s_peer_allreduce_rows_per_rank", C.c_int64, [C.c_int32, C.c_int32]),
    ("b200s_peer_allreduce_vis_bytes", C.c_size_t, [C.c_int32, C.c_int32]),
    ("b200s_peer_allreduce_sum_f32", C.c_int, [C.POINTER(PeerReduceArgs), C.c_void_p]),
    ("b200s_sh_grad_reconstruct", C.c_int, [C.POINTER(ShGradArgs), C.c_void_p]),
    ("b200s_adam_step", C.c_int, [C.POINTER(AdamArgs), C.c_void_p]),
    ("b200s_loss_workspace_size", C.c_size_t, [C.c_int32, C.c_int32, C.c_int32]),
    ("b200s_l1_ssim_loss", C.c_int, [C.POINTER(LossArgs), C.c_void_p]),
    ("b200s_export_state", C.c_int64, [C.POINTER(ForwardArgs), C.c_int32, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p]),
]

_lib = None


def lib_path() -> Path:
    # B200S_LIB selects another in-tree build of the same sources (tuning variants, python build.py --tag NAME -D...)
    alt = os.environ.get("B200S_LIB")
    return Path(alt).resolve() if alt else _build.LIB_PATH


def load() -> C.CDLL:
    """Load (building first if needed) the shared library and bind every declared symbol."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not path.exists() or os.environ.get("B200S_REBUILD"):
        # several ranks of a torchrun job may arrive here together: one builds, the others wait for the lock and
        # then find the library in place
        import fcntl
        _build.PKG_DIR.joinpath("_obj").mkdir(exist_ok=True)
        with open(_build.PKG_DIR / "_obj" / ".build.lock", "w") as lock:
            fcntl.flock(lock, fcntl.LOCK_EX)
            try:
                if not path.exists() or os.environ.get("B200S_REBUILD"):
                    _build.build(force=bool(os.environ.get("B200S_REBUILD")))
            finally:
                fcntl.flock(lock, fcntl.LOCK_UN)
    if not path.exists():
        raise RuntimeError(f"{path} is missing and could not be built; b200splat has no fallback path")
    if path == _build.LIB_PATH:
        # an existing library is used as it is (the GPU box receives it prebuilt); say so when the sources are newer
        try:
            newest = max(p.stat().st_mtime for p in [*(_build.CSRC / s for s in _build.SOURCES), *_build.HEADERS])
            if newest > path.stat().st_mtime + 1.0:
                import sys
                print(f"[b200splat] {path.name} is older than csrc/: rebuild with "
                      f"`python -m neuralgaussiansplatting_b200.build` or B200S_REBUILD=1", file=sys.stderr)
        except OSError:
            pass
    lib = C.CDLL(str(path))
    for name, restype, argtypes in SYMBOLS:
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = restype
        fn.argtypes = argtypes
    if lib.b200s_abi_version() != ABI_VERSION:
        raise RuntimeError("libb200splat.so ABI version mismatch")
    for which, struct in enumerate((Header, ForwardArgs, BackwardArgs, IdxmapArgs, PeerReduceArgs, AdamArgs, LossArgs,
                                   ShGradArgs)):
        if lib.b200s_sizeof(which) != C.sizeof(struct):
            raise RuntimeError(f"ctypes layout of {struct.__name__} does not match the library")
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    """Non-zero status -> RuntimeError (the reference binding throws std::runtime_error)."""
    if rc == 0:
        return
    if rc < 0:
        raise RuntimeError(f"{what}: {ERRORS.get(rc, rc)}")
    raise RuntimeError(f"{what}: CUDA error {rc}")
