"""How far apart are two runs of the SAME implementation? Float atomics make both the reference's and this repo's
backward order-dependent; this prints, per gradient, rel-L2 between repeated runs (ours / ours, ref / ref), between
the two implementations, and of each against this repo's deterministic mode — the numbers the parity tolerance in
tests/test_parity_gpu.py has to be read against."""
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import helpers as Hh  # noqa: E402
import refbind  # noqa: E402
from neuralgaussiansplatting_b200 import _C  # noqa: E402
from neuralgaussiansplatting_b200.synthetic import make_grad_image  # noqa: E402

NAMES = ["means2D", "opacity", "means3D", "sh", "scales", "rotations"]


def run(cfg_name, view, reps=4):
    dev = torch.device("cuda:0")
    cfg, sc, cam = Hh.scene_on(cfg_name, dev, view=view)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    dL = make_grad_image(cfg, view, device=dev)
    ref = refbind.RefDGR()
    args, out, _ = Hh.ours_forward(cfg, sc, cam, bg)
    kw, (R, rcolor, rradii) = Hh.ref_forward(ref, cfg, sc, cam, bg)
    _C.config["deterministic"] = False
    ours = [Hh.ours_backward(args, out, dL) for _ in range(reps)]
    refs = [Hh.ref_backward(ref, kw, rradii, dL) for _ in range(reps)]
    _C.config["deterministic"] = True
    det = Hh.ours_backward(args, out, dL)
    _C.config["deterministic"] = False
    row = {"cfg": cfg_name, "view": view}
    for k in NAMES:
        f = lambda a, b: Hh.rel_l2(a[k].reshape(b[k].shape), b[k])  # noqa: E731
        row[k] = {"ours_vs_ref_max": max(f(o, r) for o in ours for r in refs),
                  "ours_vs_ours_max": max(f(ours[i], ours[0]) for i in range(1, reps)),
                  "ref_vs_ref_max": max(f(refs[i], refs[0]) for i in range(1, reps)),
                  "ours_vs_det": f(ours[0], det), "ref_vs_det": f(refs[0], det)}
    ref.close()
    return row


if __name__ == "__main__":
    cases = [("T0", 0), ("T1", 0), ("T1", 1), ("T1", 3), ("T1", 6), ("C1", 0), ("C2", 0)]
    for c, v in cases:
        r = run(c, v)
        print(json.dumps({k: ({kk: float(f"{vv:.3g}") for kk, vv in val.items()} if isinstance(val, dict) else val)
                          for k, val in r.items()}))
