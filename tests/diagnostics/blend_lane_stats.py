"""Lane-mapping statistics of the blend kernels for one synthetic view. Not a test: a diagnostic
that runs the CPU oracle's forward (tests/ may use oracle/) and replays the per-tile traversal in C
(blend_lane_stats.c) to count instances read, tile-level survivors, (instance, 8x4 block) visits, contributing pairs,
and the serial visits half-warp (4x4) / quarter-warp (4x2) survivor lists would need.
Usage: python tests/diagnostics/blend_lane_stats.py [C3] > profiles/r02_blend_lane_stats.txt"""
import ctypes as C
import subprocess
import sys
import tempfile
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

import numpy as np
import torch

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import oracle_cpu as oc  # noqa: E402
from neuralgaussiansplatting_b200 import synthetic as syn  # noqa: E402

NAMES = ["bwd instances read (pos < tile_last)", "bwd tile-level survivors", "bwd (instance, 8x4 block) visits",
         "  of which with a contributing pixel", "contributing (pixel, Gaussian) pairs", "bwd half-warp 4x4 lists, max per 32-chunk",
         "bwd half-warp 4x4 lists, whole-list max", "bwd quarter-warp 4x2 lists, max per 32-chunk", "bwd quarter-warp lists, whole-list max",
         "sum of (instance, 4x4 block) hits", "sum of (instance, 4x2 block) hits", "fwd (instance, 8x4 block) visits",
         "fwd half-warp 4x4 lists, max per 32-chunk", "fwd contributing pairs", "fwd instances read (until the tile is done)"]

if __name__ == "__main__":
    name = sys.argv[1] if len(sys.argv) > 1 else "C3"
    so = Path(tempfile.gettempdir()) / "libblend_lane_stats.so"
    subprocess.run(["gcc", "-O2", "-shared", "-fPIC", "-o", str(so), str(HERE / "blend_lane_stats.c"), "-lm"], check=True)
    lib = C.CDLL(str(so))
    cfg = syn.CONFIGS[name]
    sc = syn.make_scene(cfg)
    cam = syn.make_camera(cfg, 0)
    o = oc.render_scene(cfg, sc, cam, torch.zeros(3), as_torch=False)
    arrs = [np.ascontiguousarray(o[k]) for k in ("ranges", "point_list", "means2D", "conic_opacity", "n_contrib")]
    W, H = cam.image_width, cam.image_height
    NT = 8
    outs = [np.zeros(16, np.float64) for _ in range(NT)]

    def run(i):
        lib.blend_stats(C.c_int(W), C.c_int(H), *[a.ctypes.data_as(C.c_void_p) for a in arrs],
                        outs[i].ctypes.data_as(C.c_void_p), C.c_int(NT), C.c_int(i))

    with ThreadPoolExecutor(NT) as ex:
        list(ex.map(run, range(NT)))
    tot = sum(outs)
    print(f"{name} view 0: num_rendered {o['num_rendered']}, pairs traversed by the reference {o['pairs_traversed']}, "
          f"contributing {o['pairs_contributing']}")
    for n_, v in zip(NAMES, tot):
        print(f"{n_:52s} {v / 1e6:10.2f} M")
    print(f"lanes with a contributing pixel per 8x4 visit: {tot[4] / tot[2]:.1f} of 32")
