"""Render-only throughput, view-parallel (BASELINE.json config C4: 6 M Gaussians, 1920x1080, 64 orbit views; the analogue
of render.py:24-35's loop over a camera list). Gaussians are replicated, view v belongs to rank v mod N
(parallel.shard_views), there is no communication unless --gather collects the images on rank 0.

    python tools/render_views.py                                   # 1 GPU
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/render_views.py

Timing: CUDA events around the whole shard after a warm-up pass, max over ranks; views/s = all views / that time.
Prints one JSON line on rank 0. (Not part of bench.py's contract: the headline metric is the training step.)"""
import argparse
import json
import math
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from neuralgaussiansplatting_b200 import _C, parallel  # noqa: E402
from neuralgaussiansplatting_b200.synthetic import CONFIGS, make_camera, make_scene  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfg", default="C4")
    ap.add_argument("--views", type=int, default=64)
    ap.add_argument("--repeats", type=int, default=3, help="timed passes over the rank's shard")
    ap.add_argument("--gather", action="store_true", help="also collect every image on rank 0 inside the timed region")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank, local = int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cfg = CONFIGS[a.cfg]
    sc = {k: v.to(dev) for k, v in make_scene(cfg, seed=0).items() if k not in ("feature_vector", "normal")}
    mine = parallel.shard_views(a.views, rank, world)
    cams = {v: make_camera(cfg, v, device=dev) for v in mine}
    bg = torch.zeros(3, device=dev)
    E = torch.Tensor([])
    tan = (math.tan(cfg.fovx * 0.5), math.tan(cfg.fovy * 0.5))

    def render(v):
        c = cams[v]
        return _C.rasterize_gaussians(bg, sc["means3D"], E, sc["opacities"], sc["scales"], sc["rotations"], 1.0, E,
                                      c.world_view_transform, c.full_proj_transform, tan[0], tan[1], cfg.height,
                                      cfg.width, sc["shs"], 3, c.camera_center, False, False)[1]

    def one_pass():
        imgs = [render(v) for v in mine]
        if a.gather and world > 1:
            # equal-sized shards are required by gather; pad the short ranks
            per = -(-a.views // world)
            stack = torch.zeros(per, 3, cfg.height, cfg.width, device=dev)
            if imgs:
                stack[:len(imgs)] = torch.stack(imgs)
            out = [torch.empty_like(stack) for _ in range(world)] if rank == 0 else None
            dist.gather(stack, out, dst=0)
        return imgs

    one_pass()                                                    # warm-up (also sizes the binning workspaces)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.repeats):
        one_pass()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_views = a.views * a.repeats
    rep = {"metric": f"render-only views/s, {a.cfg}: {cfg.P} Gaussians SH3 {cfg.width}x{cfg.height}", "n_gpus": world,
           "views": a.views, "repeats": a.repeats, "gather_to_rank0": bool(a.gather and world > 1),
           "value": round(total_views / (float(ms.item()) * 1e-3), 2), "unit": "views/s",
           "ms_per_view_per_gpu": round(float(ms.item()) / max(len(mine) * a.repeats, 1), 4)}
    if rank == 0:
        print(json.dumps(rep))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
