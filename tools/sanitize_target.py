"""The workload compute-sanitizer runs over (tools/sanitize.sh): every kernel family once on the tiny T0/T1 scenes —
forward + backward (atomic and deterministic accumulation, strict and fast blend), batched forward, idxmap forward +
backward, the factored SH-gradient reconstruction."""
import math
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import helpers as Hh  # noqa: E402
from neuralgaussiansplatting_b200 import _C  # noqa: E402
from neuralgaussiansplatting_b200.synthetic import make_grad_featuremap, make_grad_image  # noqa: E402

dev = torch.device("cuda:0")
for name in ("T0", "T1"):
    cfg, sc, cam = Hh.scene_on(name, dev)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    dL = make_grad_image(cfg, 0, device=dev)
    args = Hh.fwd_args(cfg, sc, cam, bg)
    for det in (False, True):
        for mode in (0, 1):
            _C.config["deterministic"], _C.config["blend_mode"] = det, mode
            out = _C.rasterize_gaussians(*args)
            g = Hh.ours_backward(args, out, dL)
            assert all(torch.isfinite(v).all() for v in g.values() if v is not None)
    _C.config["deterministic"], _C.config["blend_mode"] = False, 0
    tx, ty = math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5)
    n, idx, col, dep, feat = _C.get_idxmap2(bg, sc["means3D"], sc["normal"], sc["scales"], sc["rotations"],
                                            cam.world_view_transform, cam.full_proj_transform, tx, ty, cfg.height,
                                            cfg.width, sc["feature_vector"], sc["shs"], cam.camera_center, False)
    gf = make_grad_featuremap(cfg, 0, device=dev)
    _C.get_idxmap_backward2(sc["means3D"], cfg.height, cfg.width, cam.camera_center, idx, gf, False)
    torch.cuda.synchronize()
    print(name, "ok", int(out[0]), int(n))
print("sanitize target ok")
