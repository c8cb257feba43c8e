"""Measured deviation of blend_mode=1 from strict mode (and from nothing else): the numbers the tolerances in
tests/test_fast_mode_gpu.py are set from.  python tools/fast_mode_dev.py [T1 C1 C2]"""
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import helpers as Hh  # noqa: E402
from neuralgaussiansplatting_b200 import _C  # noqa: E402
from neuralgaussiansplatting_b200.synthetic import make_grad_image  # noqa: E402


def run(cfg_name, dev, mode):
    cfg, sc, cam = Hh.scene_on(cfg_name, dev)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    dL = make_grad_image(cfg, 0, device=dev)
    try:
        _C.config["blend_mode"] = mode
        args, out, state = Hh.ours_forward(cfg, sc, cam, bg, names=["n_contrib", "final_T"])
        g = Hh.ours_backward(args, out, dL)
    finally:
        _C.config["blend_mode"] = 0
    return out, state, g


def main():
    dev = torch.device("cuda:0")
    for name in (sys.argv[1:] or ["T1", "C1", "C2"]):
        o0, s0, g0 = run(name, dev, 0)
        o1, s1, g1 = run(name, dev, 1)
        d = (o0[1] - o1[1]).abs()
        rec = {"cfg": name, "img_max": float(d.max()), "img_mean": float(d.mean()),
               "img_frac_gt_1e-5": float((d > 1e-5).float().mean()), "img_frac_gt_1e-4": float((d > 1e-4).float().mean()),
               "n_contrib_frac": float((s0["n_contrib"] != s1["n_contrib"]).float().mean()),
               "final_T_max": float((s0["final_T"] - s1["final_T"]).abs().max()),
               "grad_rel_l2": {k: Hh.rel_l2(g1[k], g0[k]) for k in ("means2D", "opacity", "means3D", "sh", "scales", "rotations")}}
        print(json.dumps(rec))


if __name__ == "__main__":
    main()
