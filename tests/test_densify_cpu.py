"""densify.py against the reference's density control driven through a real torch.optim.Adam (oracle/densify_ref.py):
same seeded state, statistics and noise -> bit-identical parameters, moments, statistics and extras."""
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "oracle"))
import densify_ref  # noqa: E402
from neuralgaussiansplatting_b200 import densify as D  # noqa: E402


def _setup(P=400, seed=0, extent=5.0):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)  # noqa: E731
    raw = dict(xyz=r(P, 3), features=r(P, 16, 3) * 0.3, opacity=r(P, 1) * 3.0,
               scaling=r(P, 3) * 1.2 + torch.log(torch.tensor(0.01 * extent)), rotation=r(P, 4))
    normals = r(P, 3)
    ref = densify_ref.ReferenceDensifier(**raw, normals=normals)
    for it in range(3):                                       # populate Adam state
        grads = {"xyz": r(P, 3), "f_dc": r(P, 1, 3), "f_rest": r(P, 15, 3), "opacity": r(P, 1), "scaling": r(P, 3),
                 "rotation": r(P, 4)}
        ref.warm(grads)
    accum = torch.rand(P, 1, generator=g) * 0.001
    denom = torch.randint(0, 4, (P, 1), generator=g).float()   # zeros -> NaN gradients -> treated as 0
    ref.accum, ref.denom = accum.clone(), denom.clone()
    ref.max_radii2D = torch.rand(P, generator=g) * 40
    st = D.GaussianState({k: v.clone() for k, v in ref.params().items()},
                         {k: v.clone() for k, v in ref.moments("exp_avg").items()},
                         {k: v.clone() for k, v in ref.moments("exp_avg_sq").items()},
                         accum.clone(), denom.clone(), ref.max_radii2D.clone(), {"normals": normals.clone()})
    return ref, st, g


def _same(ref, st):
    assert st.P == ref.t["xyz"].shape[0]
    for k, v in ref.params().items():
        assert torch.equal(st.params[k], v), k
    for key, mine in (("exp_avg", st.exp_avg), ("exp_avg_sq", st.exp_avg_sq)):
        for k, v in ref.moments(key).items():
            assert torch.equal(mine[k], v), (key, k)
    assert torch.equal(st.extras["normals"], ref.normals)
    assert torch.equal(st.grad_accum, ref.accum) and torch.equal(st.denom, ref.denom)
    assert torch.equal(st.max_radii2D, ref.max_radii2D)


@pytest.mark.parametrize("max_screen_size", [None, 20])
@pytest.mark.parametrize("seed", [0, 1])
def test_densify_and_prune_is_bit_identical_to_the_reference_flow(seed, max_screen_size):
    extent, thr = 5.0, 0.0002
    ref, st, g = _setup(seed=seed, extent=extent)
    P0 = st.P
    noise = {}

    def draws(k):
        noise["u"] = torch.randn(2 * k, 3, generator=g)
        return noise["u"]

    ref.densify_and_prune(thr, 0.005, extent, max_screen_size, draws)
    out = D.densify_and_prune(st, thr, 0.005, extent, max_screen_size, unit_normals=noise["u"])
    _same(ref, out)
    assert out.P != P0 and noise["u"].shape[0] > 0            # the case exercises clone, split and prune
    assert float(out.grad_accum.abs().sum()) == 0.0 and float(out.max_radii2D.abs().sum()) == 0.0


def test_clone_split_prune_pieces_and_reset_opacity():
    ref, st, g = _setup(P=300, seed=3)
    grads = st.grad_accum / st.denom
    grads[grads.isnan()] = 0.0
    ref.clone(grads, 0.0002, 5.0)
    st = D.clone_small(st, grads, 0.0002, 5.0, 0.01)
    _same(ref, st)
    k = ref.split_count(grads, 0.0002, 5.0)
    u = torch.randn(2 * k, 3, generator=g)
    ref.split(grads, 0.0002, 5.0, u)
    st = D.split_large(st, grads, 0.0002, 5.0, 0.01, 2, unit_normals=u)
    _same(ref, st)
    mask = torch.rand(st.P, generator=g) < 0.3
    ref._prune(mask)
    st = D.prune(st, mask)
    _same(ref, st)
    ref.reset_opacity()
    st = D.reset_opacity(st)
    _same(ref, st)
    assert float(torch.sigmoid(st.params["opacity"]).max()) <= 0.01 + 1e-6
    assert float(st.exp_avg["opacity"].abs().sum()) == 0.0 and float(st.exp_avg["xyz"].abs().sum()) > 0.0


def test_split_children_follow_the_parent_frame_and_generator_is_reproducible():
    ref, st, g = _setup(P=200, seed=5)
    grads = torch.full((st.P, 1), 1.0)
    a = D.split_large(st, grads, 0.5, 5.0, 0.01, 2, generator=torch.Generator().manual_seed(9))
    b = D.split_large(st, grads, 0.5, 5.0, 0.01, 2, generator=torch.Generator().manual_seed(9))
    assert torch.equal(a.params["xyz"], b.params["xyz"])      # identically seeded replicas take identical decisions
    sel = torch.exp(st.params["scaling"]).max(dim=1).values > 0.05
    K = int(sel.sum())
    assert a.P == st.P - K + 2 * K
    R = D.rotation_matrices(st.params["rotation"][:4])
    assert torch.allclose(R @ R.transpose(1, 2), torch.eye(3).expand(4, 3, 3), atol=1e-5)
    kids = a.params["scaling"][-2 * K:]
    assert torch.allclose(torch.exp(kids), torch.exp(st.params["scaling"][sel]).repeat(2, 1) / 1.6, rtol=1e-5)
