"""CPU checks around the fused optimizer: the torch-CPU oracle against Adam's closed form,
and the C-ABI argument checks of b200s_adam_step (no device needed: rejected before any launch)."""
import ctypes as C
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "oracle"))
import optimizer_ref  # noqa: E402
from neuralgaussiansplatting_b200 import _lib  # noqa: E402
from neuralgaussiansplatting_b200 import optimizer as opt  # noqa: E402


def _raw(P, seed=0):
    g = torch.Generator().manual_seed(seed)
    return dict(xyz=torch.randn(P, 3, generator=g), features=torch.randn(P, 16, 3, generator=g) * 0.3,
                opacity=torch.randn(P, 1, generator=g) * 2, scaling=torch.randn(P, 3, generator=g) * 0.6 - 4.0,
                rotation=torch.randn(P, 4, generator=g))


def _grads(P, seed):
    g = torch.Generator().manual_seed(seed)
    return dict(means3D=torch.randn(P, 3, generator=g) * 1e-3, shs=torch.randn(P, 16, 3, generator=g) * 1e-4,
                opacities=torch.randn(P, 1, generator=g) * 1e-3, scales=torch.randn(P, 3, generator=g) * 1e-2,
                rotations=torch.randn(P, 4, generator=g) * 1e-4)


def test_oracle_first_step_is_lr_times_sign_of_chained_gradient():
    # step 1 of Adam: m_hat = g, v_hat = g^2  ->  p1 = p0 - lr * g / (|g| + eps)
    P = 64
    raw, grads = _raw(P), _grads(P, 1)
    ref = optimizer_ref.ReferenceGaussianOptimizer(**raw, lrs=opt.DEFAULT_LRS)
    ref.step(grads)
    new = ref.raw()
    assert torch.allclose(new["xyz"], raw["xyz"] - opt.DEFAULT_LRS["xyz"] * torch.sign(grads["means3D"]), atol=1e-7)
    s = torch.sigmoid(raw["opacity"])
    chained = grads["opacities"] * s * (1 - s)                              # sigmoid backward
    assert torch.allclose(new["opacity"], raw["opacity"] - opt.DEFAULT_LRS["opacity"] * torch.sign(chained), atol=1e-6)
    f = new["features"] - raw["features"]
    assert torch.allclose(f[:, :1].abs(), torch.full_like(f[:, :1], opt.DEFAULT_LRS["f_dc"]), atol=1e-7)
    assert torch.allclose(f[:, 1:].abs(), torch.full_like(f[:, 1:], opt.DEFAULT_LRS["f_rest"]), atol=1e-7)


def test_oracle_zero_gradient_rows_do_not_move():
    P = 32
    raw, grads = _raw(P), _grads(P, 2)
    for v in grads.values():
        v[::2] = 0.0                                                         # culled Gaussians: exact zeros
    ref = optimizer_ref.ReferenceGaussianOptimizer(**raw, lrs=opt.DEFAULT_LRS)
    ref.step(grads)
    new = ref.raw()
    for k in raw:
        assert torch.equal(new[k][::2], raw[k][::2]), k                     # 0 / (0 + 1e-15) == 0


def test_fused_adam_rejects_cpu_tensors():
    with pytest.raises(RuntimeError, match="CUDA"):
        opt.FusedGaussianAdam(_raw(8))


def test_adam_abi_argument_checks():
    lib = _lib.load()
    assert lib.b200s_adam_step(None, None) == -1
    a = _lib.AdamArgs()
    a.count, a.bias_correction2_sqrt = 9, 1.0
    assert lib.b200s_adam_step(C.byref(a), None) == -1                       # more than B200S_ADAM_MAX_TENSORS
    a.count = 1
    a.t[0].n, a.t[0].mode = 48, 1
    assert lib.b200s_adam_step(C.byref(a), None) == -1                       # NULL tensors
    a.t[0].param = a.t[0].exp_avg = a.t[0].exp_avg_sq = a.t[0].grad = 4096
    a.t[0].n = 50
    assert lib.b200s_adam_step(C.byref(a), None) == -1                       # SH rows are 48 floats
    a.t[0].n, a.t[0].mode = 6, 4
    assert lib.b200s_adam_step(C.byref(a), None) == -1                       # quaternion rows are 4 floats
    a.t[0].mode, a.t[0].grad = 0, 4100
    assert lib.b200s_adam_step(C.byref(a), None) == -1                       # 16-byte alignment
    a.t[0].grad, a.t[0].mode = 4096, 7
    assert lib.b200s_adam_step(C.byref(a), None) == -1                       # unknown activation
    a.t[0].mode, a.bias_correction2_sqrt = 0, 0.0
    assert lib.b200s_adam_step(C.byref(a), None) == -1
    a.bias_correction2_sqrt, a.count = 1.0, 0
    assert lib.b200s_adam_step(C.byref(a), None) == 0                        # nothing to do, no launch
