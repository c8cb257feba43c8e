"""chkpnt*.pth interchange (checkpoint.py) with a REAL reference-style optimizer: a 7-group torch.optim.Adam
(scene/gaussian_model.py:257-265) is stepped, captured into the reference's 14-tuple, converted to this repo's layout
and back, loaded into a fresh Adam with load_state_dict (what GaussianModel.restore does, :93-111), and both optimizers
must then take an identical next step."""
import torch
from torch import nn

from neuralgaussiansplatting_b200 import checkpoint as ck

LRS = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 1.25e-4, "features": 2.5e-3, "opacity": 0.05, "scaling": 5e-3,
       "rotation": 1e-3}
SHAPES = {"xyz": (3,), "f_dc": (1, 3), "f_rest": (15, 3), "features": (64,), "opacity": (1,), "scaling": (3,),
          "rotation": (4,)}


def _reference_model(P, seed):
    g = torch.Generator().manual_seed(seed)
    t = {k: nn.Parameter(torch.randn(P, *s, generator=g)) for k, s in SHAPES.items()}
    opt = torch.optim.Adam([{"params": [t[k]], "lr": LRS[k], "name": k} for k in ck.GROUP_ORDER], lr=0.0, eps=1e-15)
    return t, opt, g


def _step(t, opt, g, skip=("features",)):
    for k, p in t.items():
        p.grad = None if k in skip else torch.randn(p.shape, generator=g)
    opt.step()


def _capture(t, opt, P):
    return (3, t["xyz"], nn.Parameter(torch.zeros(P, 3)), t["f_dc"], t["f_rest"], t["features"], t["scaling"],
            t["rotation"], t["opacity"], torch.arange(P).float(), torch.ones(P, 1), torch.full((P, 1), 2.0),
            opt.state_dict(), 1.7)


def test_round_trip_through_this_repos_layout_restores_an_identical_optimizer(tmp_path):
    P = 37
    t, opt, g = _reference_model(P, 0)
    for _ in range(3):
        _step(t, opt, g)
    path = str(tmp_path / "chkpnt3.pth")
    ck.save_checkpoint(path, _capture(t, opt, P), 3)
    model_args, it = ck.load_checkpoint(path)
    assert it == 3 and len(model_args) == 14
    mine = ck.from_reference_tuple(model_args)
    assert mine["step"] == 3 and mine["raw"]["features"].shape == (P, 16, 3) and mine["spatial_lr_scale"] == 1.7
    assert torch.equal(mine["raw"]["features"][:, :1], t["f_dc"].detach())
    assert torch.equal(mine["exp_avg"]["features"][:, 1:], opt.state[t["f_rest"]]["exp_avg"])
    assert abs(mine["lrs"]["f_rest"] - LRS["f_rest"]) < 1e-12 and torch.equal(mine["max_radii2D"], torch.arange(P).float())

    back = ck.to_reference_tuple(mine["raw"], mine["exp_avg"], mine["exp_avg_sq"], mine["step"], mine["lrs"],
                                 mine["grad_accum"], mine["denom"], mine["max_radii2D"], mine["normals"],
                                 mine["feature_vector"], mine["active_sh_degree"], mine["spatial_lr_scale"])
    assert len(back) == 14 and back[0] == 3 and back[13] == 1.7
    # what GaussianModel.restore does: rebuild the groups from the tuple's tensors, then load_state_dict
    t2 = {"xyz": back[1], "f_dc": back[3], "f_rest": back[4], "features": back[5], "opacity": back[8], "scaling": back[6],
          "rotation": back[7]}                                           # same iteration order as `t` (gradient draws)
    opt2 = torch.optim.Adam([{"params": [t2[k]], "lr": 0.123, "name": k} for k in ck.GROUP_ORDER], lr=0.0, eps=1e-15)
    opt2.load_state_dict(back[12])
    assert [g_["name"] for g_ in opt2.param_groups] == list(ck.GROUP_ORDER)
    assert abs(opt2.param_groups[0]["lr"] - LRS["xyz"]) < 1e-12
    gen_a, gen_b = torch.Generator().manual_seed(99), torch.Generator().manual_seed(99)
    _step(t, opt, gen_a)
    _step(t2, opt2, gen_b)
    for k in t:
        assert torch.equal(t[k].detach(), t2[k].detach()), k            # identical next step
    assert torch.equal(back[10], torch.ones(P, 1)) and torch.equal(back[11], torch.full((P, 1), 2.0))


def test_fresh_run_has_no_optimizer_state_and_defaults():
    P = 5
    raw = {"xyz": torch.zeros(P, 3), "features": torch.zeros(P, 16, 3), "opacity": torch.zeros(P, 1),
           "scaling": torch.zeros(P, 3), "rotation": torch.ones(P, 4)}
    z = {k: torch.zeros_like(v) for k, v in raw.items()}
    tup = ck.to_reference_tuple(raw, z, z, 0, LRS, torch.zeros(P, 1), torch.zeros(P, 1))
    assert tup[12]["state"] == {} and len(tup[12]["param_groups"]) == 7
    assert tup[2].shape == (P, 3) and tup[5].shape == (P, 64) and tup[9].shape == (P,)
    mine = ck.from_reference_tuple(tup)
    assert mine["step"] == 0 and float(mine["exp_avg"]["features"].abs().sum()) == 0.0
