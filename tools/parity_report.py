"""Run the product and the reference-compiled oracle on the same seeded inputs, print a parity table and
A/B timings. Diagnostic tool for the GPU box (writes gpurun_out/parity_<cfg>.json)."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

import helpers as Hh  # noqa: E402
import refbind  # noqa: E402
from neuralgaussiansplatting_b200 import _C  # noqa: E402
from neuralgaussiansplatting_b200.synthetic import make_grad_image  # noqa: E402


def time_cuda(fn, warmup=3, iters=10):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        ts.append((time.perf_counter() - t0) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def run(cfg_name, P=None, timing=True, det=False, view=0):
    dev = torch.device("cuda:0")
    cfg, sc, cam = Hh.scene_on(cfg_name, dev, P=P, view=view)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    dL = make_grad_image(cfg, view, device=dev)
    res = {"cfg": cfg_name, "P": sc["means3D"].shape[0], "W": cfg.width, "H": cfg.height}
    args, out, state = Hh.ours_forward(cfg, sc, cam, bg)
    ref = refbind.RefDGR()
    kw, (R, rcolor, rradii) = Hh.ref_forward(ref, cfg, sc, cam, bg)
    res["num_rendered_ref"] = R
    res["visible"] = int((rradii > 0).sum())
    res["state"] = Hh.compare_forward_state(state, ref, out[2], rradii)
    res["image_maxabs"] = float((out[1] - rcolor).abs().max())
    _C.config["deterministic"] = det
    g_o = Hh.ours_backward(args, out, dL)
    g_r = Hh.ref_backward(ref, kw, rradii, dL)
    res["grads_rel_l2"] = {k: Hh.rel_l2(g_o[k].reshape(g_r[k].shape), g_r[k]) for k in g_o if g_o[k] is not None}
    res["grads_maxabs"] = {k: float((g_o[k].reshape(g_r[k].shape) - g_r[k]).abs().max()) for k in g_o
                           if g_o[k] is not None}
    if timing:
        res["ms_ours_fwd"] = time_cuda(lambda: _C.rasterize_gaussians(*args))
        o2 = _C.rasterize_gaussians(*args)
        res["ms_ours_fwd_bwd"] = time_cuda(lambda: Hh.ours_backward(args, _C.rasterize_gaussians(*args), dL))
        res["ms_ref_fwd"] = time_cuda(lambda: ref.forward(**kw))
        res["ms_ref_fwd_bwd"] = time_cuda(lambda: Hh.ref_backward(ref, kw, ref.forward(**kw)[2], dL))
        del o2
    ref.close()
    return res


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfgs", default="T0,T1,C1")
    ap.add_argument("--no-timing", action="store_true")
    ap.add_argument("--det", action="store_true")
    a = ap.parse_args()
    os.makedirs(ROOT / "gpurun_out", exist_ok=True)
    for c in a.cfgs.split(","):
        r = run(c, timing=not a.no_timing, det=a.det)
        print(json.dumps(r, indent=1))
        (ROOT / "gpurun_out" / f"parity_{c}{'_det' if a.det else ''}.json").write_text(json.dumps(r, indent=1))
