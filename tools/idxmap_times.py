"""rasterizer2 idxmap path (K11..K16) at BASELINE.json's full sizes: CUDA-event time of this repo's forward and
feature-gradient backward, the reference rasterizer2 (oracle/_ref, compiled from /root/reference) on the same inputs
when it travelled with the repo, and bit-parity of the index map between the two.

    python tools/idxmap_times.py [--cfg C3 C4] [--steps 20]
"""
import argparse
import ctypes as C
import json
import math
import sys
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import helpers as Hh  # noqa: E402
import refbind  # noqa: E402
from neuralgaussiansplatting_b200 import _C, _lib  # noqa: E402
from neuralgaussiansplatting_b200.synthetic import make_grad_featuremap  # noqa: E402


def timed(fn, steps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfg", nargs="+", default=["C3", "C4"])
    ap.add_argument("--steps", type=int, default=20)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    lib = _lib.load()
    for name in a.cfg:
        cfg, sc, cam = Hh.scene_on(name, dev)
        tx, ty = math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5)
        bg = torch.zeros(3, device=dev)
        gf = make_grad_featuremap(cfg, 0, device=dev)

        def fwd():
            return _C.get_idxmap2(bg, sc["means3D"], sc["normal"], sc["scales"], sc["rotations"], cam.world_view_transform,
                                  cam.full_proj_transform, tx, ty, cfg.height, cfg.width, sc["feature_vector"], sc["shs"],
                                  cam.camera_center, False)

        n, idx, col, dep, feat = fwd()

        def bwd():
            return _C.get_idxmap_backward2(sc["means3D"], cfg.height, cfg.width, cam.camera_center, idx, gf, False)

        t_f, t_b = timed(fwd, a.steps), timed(bwd, a.steps)
        lib.b200s_profile_begin()
        for _ in range(a.steps):
            fwd(); bwd()
        ms = (C.c_float * len(_lib.STAGES))(); calls = (C.c_int32 * len(_lib.STAGES))()
        lib.b200s_profile_end(ms, calls)
        stages = {s: round(ms[i] / a.steps, 4) for i, s in enumerate(_lib.STAGES) if calls[i]}
        px, P = cfg.width * cfg.height, cfg.P
        covered = int((idx >= 0).sum())
        # bytes the forward must move: means (12 B/point), the 8-byte z-buffer written + read once, the four maps
        # (4 + 12 + 4 + 256 B/pixel) and the 39 learned channels gathered per covered pixel
        alg_f = 12 * P + 4 * P + 16 * px + 276 * px + 156 * covered
        alg_b = 4 * px + 256 * covered + 256 * P + 156 * covered
        rec = {"cfg": name, "P": P, "pixels": px, "covered_pixels": covered, "pairs_the_reference_sorts": int(n),
               "ours": {"forward_ms": round(t_f, 4), "backward_ms": round(t_b, 4), "stages_ms": stages,
                        "forward_alg_bytes": alg_f, "forward_gbps": round(alg_f / (t_f * 1e-3) / 1e9, 1),
                        "backward_alg_bytes": alg_b, "backward_gbps": round(alg_b / (t_b * 1e-3) / 1e9, 1)}}
        if refbind.available("ref_r2"):
            r2 = refbind.RefR2()
            kw = dict(means3D=sc["means3D"], normal=sc["normal"], scales=sc["scales"], rotations=sc["rotations"],
                      viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform, tan_fovx=tx, tan_fovy=ty,
                      feature_vector=sc["feature_vector"], shs=sc["shs"], campos=cam.camera_center, W=cfg.width,
                      H=cfg.height)
            R, ridx, rcol, rdep, rfeat, rrad = r2.getidxmap(**kw)       # warm-up + parity
            best = 1e9
            for _ in range(5):                                          # the shim synchronises on both sides
                torch.cuda.synchronize(); t0 = time.perf_counter()
                r2.getidxmap(**kw)
                best = min(best, time.perf_counter() - t0)
            bb = 1e9
            for _ in range(5):
                torch.cuda.synchronize(); t0 = time.perf_counter()
                r2.getidxmap_backward(means3D=sc["means3D"], campos=cam.camera_center, out_idx=ridx, grad_feature=gf,
                                      W=cfg.width, H=cfg.height)
                bb = min(bb, time.perf_counter() - t0)
            rec["reference"] = {"forward_ms": round(best * 1e3, 3), "backward_ms": round(bb * 1e3, 3),
                                "what": "oracle/_ref rasterizer2 getidxmap, host wall clock around a synchronised call "
                                        "(includes its output allocations), best of 5"}
            rec["parity"] = {"num_rendered_equal": int(n) == int(R), "idx_bit_exact": bool(torch.equal(idx, ridx)),
                             "depth_bit_exact": bool(torch.equal(dep.view(torch.int32), rdep.view(torch.int32))),
                             "learned_channels_equal": bool(torch.equal(feat[..., 25:], rfeat[..., 25:])),
                             "encoding_max_abs": float((feat[..., :25] - rfeat[..., :25]).abs().max())}
            rec["speedup_forward"] = round(best * 1e3 / t_f, 1)
            r2.close()
            del ridx, rcol, rdep, rfeat, rrad
        print(json.dumps(rec))
        del sc, idx, col, dep, feat, gf
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
