"""Per-stage CUDA-event times of the forward+backward on one config/view (library stage brackets)."""
import argparse
import ctypes as C
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import helpers as Hh  # noqa: E402
from neuralgaussiansplatting_b200 import _C, _lib  # noqa: E402
from neuralgaussiansplatting_b200.synthetic import make_grad_image  # noqa: E402

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfg", default="C3")
    ap.add_argument("--view", type=int, default=0)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--det", action="store_true")
    ap.add_argument("--fwd-only", action="store_true")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    lib = _lib.load()
    cfg, sc, cam = Hh.scene_on(a.cfg, dev, view=a.view)
    sc = {k: v for k, v in sc.items() if k not in ("feature_vector", "normal")}
    bg = torch.zeros(3, device=dev)
    dL = make_grad_image(cfg, a.view, device=dev)
    args = Hh.fwd_args(cfg, sc, cam, bg)
    _C.config["deterministic"] = a.det

    def step():
        out = _C.rasterize_gaussians(*args)
        if not a.fwd_only:
            Hh.ours_backward(args, out, dL)

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        step()
    e1.record(); torch.cuda.synchronize()
    total = e0.elapsed_time(e1) / a.steps
    lib.b200s_profile_begin()
    for _ in range(a.steps):
        step()
    ms = (C.c_float * len(_lib.STAGES))(); calls = (C.c_int32 * len(_lib.STAGES))()
    lib.b200s_profile_end(ms, calls)
    row = {n: round(ms[i] / a.steps, 4) for i, n in enumerate(_lib.STAGES) if calls[i]}
    print(json.dumps({"cfg": a.cfg, "ms_per_step": round(total, 4), "stages": row, "sum": round(sum(row.values()), 4)}))
