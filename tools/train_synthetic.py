"""Fits a perturbed copy of a synthetic scene to renders of the original with ViewParallelTrainer — every kernel of
the iteration is this repo's. Single process, or `torchrun --nproc-per-node N tools/train_synthetic.py` (one view per
rank per iteration, gradients summed by the peer-memory all-reduce, replicas checked bit-identical)."""
import argparse
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from neuralgaussiansplatting_b200.synthetic import CONFIGS, make_camera, make_scene  # noqa: E402
from neuralgaussiansplatting_b200.trainer import ViewParallelTrainer  # noqa: E402


def raw_from_scene(sc, dev, noise=0.0, seed=0):
    g = torch.Generator().manual_seed(seed)
    n = lambda t, s: t + noise * s * torch.randn(t.shape, generator=g)  # noqa: E731
    return dict(xyz=n(sc["means3D"], 0.02).to(dev).contiguous(), features=n(sc["shs"], 0.3).to(dev).contiguous(),
                opacity=torch.logit(sc["opacities"].clamp(1e-4, 1 - 1e-4)).to(dev).contiguous(),
                scaling=n(torch.log(sc["scales"]), 0.1).to(dev).contiguous(),
                rotation=n(sc["rotations"], 0.05).to(dev).contiguous())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfg", default="T1")
    ap.add_argument("--iters", type=int, default=60)
    ap.add_argument("--views", type=int, default=8)
    ap.add_argument("--densify-at", type=int, default=0, help="iteration after which to run densify_and_prune once")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank, local = int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cfg = CONFIGS[a.cfg]
    sc = make_scene(cfg, seed=0)
    cams = [make_camera(cfg, v, device=dev) for v in range(a.views)]
    target = ViewParallelTrainer(raw_from_scene(sc, dev), collective="none")
    gts = [target.render(c).clone() for c in cams]
    tr = ViewParallelTrainer(raw_from_scene(sc, dev, noise=1.0, seed=7))          # same perturbation on every rank
    losses, densified = [], None
    for it in range(a.iters):
        v = (it * world + rank) % a.views
        losses.append(tr.step(cams[v], gts[v]))
        if a.densify_at and it + 1 == a.densify_at:
            p0 = tr.P
            p1 = tr.densify_and_prune(max_grad=2e-6, min_opacity=0.005, extent=4.0)
            densified = (p0, p1)
    torch.cuda.synchronize()
    vals = [float(x[0]) for x in losses]
    rep = {"world": world, "collective": tr.collective, "P": tr.P, "iters": a.iters,
           "loss_first5": round(sum(vals[:5]) / 5, 5), "loss_last5": round(sum(vals[-5:]) / 5, 5),
           "replicas_in_sync": tr.replicas_in_sync(), "densified": densified}
    if rank == 0:
        print(json.dumps(rep))
    if densified is None:       # a densification step re-opens the loss (clones double their parents' density at first)
        assert rep["loss_last5"] < rep["loss_first5"], rep
    else:
        assert densified[1] != densified[0], rep
    assert rep["replicas_in_sync"], rep
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
