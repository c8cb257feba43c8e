"""Times the two fused kernels either side of the rasterizer against the reference's
torch formulation on the same GPU: L1+SSIM loss forward+backward at the C3 image size, and activation-backward + Adam
+ activation over P Gaussians. CUDA events, warm-up 5, 20 timed iterations each."""
import argparse
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "oracle"))
import loss_ref  # noqa: E402  (only its torch formulation, run on the GPU here as the reference arm)
from neuralgaussiansplatting_b200 import loss as L  # noqa: E402
from neuralgaussiansplatting_b200 import optimizer as opt  # noqa: E402


def timeit(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--P", type=int, default=3_000_000)
    ap.add_argument("--H", type=int, default=896)
    ap.add_argument("--W", type=int, default=1344)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(0)
    gt = torch.rand(3, a.H, a.W, device=dev, generator=g)
    img = (gt + 0.1 * torch.randn(3, a.H, a.W, device=dev, generator=g)).clamp(0, 1)
    rep = {}

    def torch_loss():
        x = img.clone().requires_grad_(True)
        loss = 0.8 * loss_ref.l1(x, gt) + 0.2 * (1.0 - loss_ref.ssim(x, gt))
        loss.backward()
        return x.grad

    rep["loss_torch_ms"] = round(timeit(torch_loss), 4)
    rep["loss_fused_ms"] = round(timeit(lambda: L.fused_l1_ssim(img, gt, 0.2)), 4)
    gref = torch_loss()
    _, gf = L.fused_l1_ssim(img, gt, 0.2)
    rep["loss_grad_rel_l2_vs_torch_gpu"] = float((gf - gref).norm() / gref.norm())

    P = a.P
    raw = dict(xyz=torch.randn(P, 3, device=dev, generator=g), features=torch.randn(P, 16, 3, device=dev, generator=g),
               opacity=torch.randn(P, 1, device=dev, generator=g), scaling=torch.randn(P, 3, device=dev, generator=g) - 4,
               rotation=torch.randn(P, 4, device=dev, generator=g))
    grads = dict(means3D=torch.randn(P, 3, device=dev, generator=g), shs=torch.randn(P, 16, 3, device=dev, generator=g),
                 opacities=torch.randn(P, 1, device=dev, generator=g), scales=torch.randn(P, 3, device=dev, generator=g),
                 rotations=torch.randn(P, 4, device=dev, generator=g))
    ours = opt.FusedGaussianAdam({k: v.clone() for k, v in raw.items()})
    rep["adam_fused_ms"] = round(timeit(lambda: ours.step(grads)), 4)
    bytes_per_step = P * 59 * 4 * 7 + P * (1 + 3 + 4) * 4     # grad + p,m,v read + p,m,v write + 8 activated floats written
    rep["adam_fused_GBps"] = round(bytes_per_step / rep["adam_fused_ms"] / 1e6, 1)
    rep["adam_bytes_per_step"] = bytes_per_step

    # the reference's formulation on the GPU: separate f_dc / f_rest parameters, activations, autograd, torch Adam
    Pm = torch.nn.Parameter
    xyz, fdc, frest = Pm(raw["xyz"].clone()), Pm(raw["features"][:, :1].clone()), Pm(raw["features"][:, 1:].clone())
    op, sc, ro = Pm(raw["opacity"].clone()), Pm(raw["scaling"].clone()), Pm(raw["rotation"].clone())
    lrs = opt.DEFAULT_LRS
    adam = torch.optim.Adam([{"params": [xyz], "lr": lrs["xyz"]}, {"params": [fdc], "lr": lrs["f_dc"]},
                             {"params": [frest], "lr": lrs["f_rest"]}, {"params": [op], "lr": lrs["opacity"]},
                             {"params": [sc], "lr": lrs["scaling"]}, {"params": [ro], "lr": lrs["rotation"]}],
                            lr=0.0, eps=1e-15)

    def torch_step():
        acts = [xyz, torch.cat((fdc, frest), dim=1), torch.sigmoid(op), torch.exp(sc), torch.nn.functional.normalize(ro)]
        torch.autograd.backward(acts, [grads["means3D"], grads["shs"], grads["opacities"], grads["scales"],
                                       grads["rotations"]])
        adam.step()
        adam.zero_grad(set_to_none=True)

    rep["adam_torch_ms"] = round(timeit(torch_step), 4)
    print(json.dumps(rep))
