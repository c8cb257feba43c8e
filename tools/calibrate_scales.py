"""Bisection of SceneConfig.log_scale_mu so that num_rendered / visible Gaussians = 6.0 at camera 0
(SURVEY.md §8d), using the CPU oracle's preprocess (tools-only; not part of the product)."""
import dataclasses
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import oracle_cpu  # noqa: E402
from neuralgaussiansplatting_b200.synthetic import CONFIGS, make_camera, make_scene  # noqa: E402

if __name__ == "__main__":
    for name in sys.argv[1:] or ["C1", "C3", "C4"]:
        cfg0 = CONFIGS[name]
        cam = make_camera(cfg0, 0)
        lo, hi = -7.0, -2.0
        for _ in range(12):
            mu = 0.5 * (lo + hi)
            cfg = dataclasses.replace(cfg0, log_scale_mu=mu)
            sc = make_scene(cfg, 0, P=min(cfg.P, 300_000))
            o = oracle_cpu.render_scene(cfg, sc, cam, torch.zeros(3), stop_after="preprocess", as_torch=False)
            vis = int((o["radii"] > 0).sum()); n = int(o["tiles_touched"].astype(np.int64).sum())
            if n / max(vis, 1) > 6.0: hi = mu
            else: lo = mu
        print(f"{name}: log_scale_mu={mu:.3f} N/P_vis={n / max(vis, 1):.3f} visible fraction={vis / sc['means3D'].shape[0]:.3f}")
