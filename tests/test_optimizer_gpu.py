"""b200s_adam_step (fused activation-backward + Adam + activation) against the reference's own optimizer path
(oracle/optimizer_ref.py = torch autograd + torch.optim.Adam on the CPU) on the same seeded inputs.
Tolerances (fp32, fused multiply-adds vs torch's separate kernels): parameters 1e-6, moments 1e-5 relative L2."""
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "oracle"))
import optimizer_ref  # noqa: E402
from neuralgaussiansplatting_b200 import optimizer as opt  # noqa: E402
from test_optimizer_cpu import _grads, _raw  # noqa: E402

pytestmark = pytest.mark.gpu
PARAM_TOL, MOMENT_TOL = 1e-6, 1e-5


def rel_l2(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


@pytest.mark.parametrize("P", [1, 7, 10_007])
def test_fused_adam_matches_reference_optimizer(P):
    dev = torch.device("cuda:0")
    raw = _raw(P)
    ref = optimizer_ref.ReferenceGaussianOptimizer(**raw, lrs=opt.DEFAULT_LRS)
    ours = opt.FusedGaussianAdam({k: v.to(dev).contiguous() for k, v in raw.items()})
    for k, v in ref.activated().items():                                    # activations at construction
        assert rel_l2(ours.activated[k].cpu(), v.detach()) < 1e-6, k
    for it in range(6):
        grads = _grads(P, 10 + it)
        if P > 4:
            for v in grads.values():
                v[it % 3::3] = 0.0                                           # rows of culled Gaussians are exact zeros
        if it == 3:                                                          # update_learning_rate reschedules xyz
            ref.set_lr("xyz", 7.3e-5); ours.set_lr("xyz", 7.3e-5)
        ref.step(grads)
        ours.step({k: v.to(dev) for k, v in grads.items()})
        torch.cuda.synchronize()
        new, mom = ref.raw(), ref.moments()
        for k in new:
            assert rel_l2(ours.raw[k].cpu(), new[k]) < PARAM_TOL, (it, k)
            assert rel_l2(ours.exp_avg[k].cpu(), mom[k]["exp_avg"]) < MOMENT_TOL, (it, k)
            assert rel_l2(ours.exp_avg_sq[k].cpu(), mom[k]["exp_avg_sq"]) < MOMENT_TOL, (it, k)
        for k, v in ref.activated().items():
            assert rel_l2(ours.activated[k].cpu(), v.detach()) < PARAM_TOL, (it, k)
    # reference-layout export (checkpoint interchange)
    rt, rs = ours.reference_tensors(), ours.reference_state()
    assert rt["_features_dc"].shape == (P, 1, 3) and rt["_features_rest"].shape == (P, 15, 3)
    assert rs["f_rest"]["exp_avg"].shape == (P, 15, 3) and rs["xyz"]["step"] == 6


def test_fused_adam_reads_gradients_in_place_and_keeps_aliases():
    dev = torch.device("cuda:0")
    raw = {k: v.to(dev).contiguous() for k, v in _raw(257).items()}
    ours = opt.FusedGaussianAdam(raw)
    assert ours.activated["means3D"].data_ptr() == raw["xyz"].data_ptr()     # identity activations are not copied
    assert ours.activated["shs"].data_ptr() == raw["features"].data_ptr()
    before = raw["opacity"].clone()
    g = {k: v.to(dev) for k, v in _grads(257, 3).items()}
    ours.step(g)
    torch.cuda.synchronize()
    assert not torch.equal(raw["opacity"], before)                           # updated in place
    assert torch.allclose(ours.activated["opacities"], torch.sigmoid(raw["opacity"]), atol=1e-6)
    assert torch.allclose(ours.activated["rotations"].norm(dim=1), torch.ones(257, device=dev), atol=1e-5)
