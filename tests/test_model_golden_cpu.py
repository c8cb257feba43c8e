"""densify.py / ply_io.py / checkpoint.py against golden files written by the REFERENCE's own
scene/gaussian_model.py (tests/golden/make_model_golden.py imports it from /root/reference in the build container;
here only the committed files are read): densify_and_prune / reset_opacity outputs, point_cloud.ply bytes, and the
`(capture(), iteration)` checkpoint pickle."""
from pathlib import Path

import numpy as np
import pytest
import torch

from neuralgaussiansplatting_b200 import checkpoint as ck
from neuralgaussiansplatting_b200 import densify as D
from neuralgaussiansplatting_b200 import ply_io

GOLD = Path(__file__).resolve().parent / "golden"
T = lambda a: torch.from_numpy(np.ascontiguousarray(a))  # noqa: E731


def _state(z, pre):
    feats = lambda k: torch.cat((T(z[f"{pre}{k}f_dc"]), T(z[f"{pre}{k}f_rest"])), dim=1)  # noqa: E731
    params = {"xyz": T(z[f"{pre}xyz"]), "features": feats(""), "opacity": T(z[f"{pre}opacity"]),
              "scaling": T(z[f"{pre}scaling"]), "rotation": T(z[f"{pre}rotation"])}
    mom = {}
    for key in ("m_", "v_"):
        mom[key] = {k: T(z[f"{pre}{key}{k}"]) for k in ("xyz", "opacity", "scaling", "rotation")}
        mom[key]["features"] = feats(key)
    return D.GaussianState(params, mom["m_"], mom["v_"], T(z[f"{pre}grad_accum"]), T(z[f"{pre}denom"]),
                           T(z[f"{pre}max_radii2D"]), {"normals": T(z[f"{pre}normals"]),
                                                       "feature_vector": T(z[f"{pre}feature_vector"])})


def _assert_same(st, z, pre):
    want = _state(z, pre)
    assert st.P == want.P
    for k in D.PARAMS:
        assert torch.equal(st.params[k], want.params[k]), ("param", k)
        assert torch.equal(st.exp_avg[k], want.exp_avg[k]), ("exp_avg", k)
        assert torch.equal(st.exp_avg_sq[k], want.exp_avg_sq[k]), ("exp_avg_sq", k)
    for k in ("normals", "feature_vector"):
        assert torch.equal(st.extras[k], want.extras[k]), k
    assert torch.equal(st.grad_accum, want.grad_accum) and torch.equal(st.denom, want.denom)
    assert torch.equal(st.max_radii2D, want.max_radii2D)


@pytest.mark.parametrize("tag", ["a_", "b_"])
def test_densify_and_prune_and_reset_opacity_match_the_reference_model(tag):
    """GaussianModel.densify_and_prune (gaussian_model.py:534-548, incl. clone :517-533, split :490-515, the Adam
    state surgery :407-460) and reset_opacity (:331-334) of the reference, bit for bit."""
    z = np.load(GOLD / "ref_densify.npz")
    st = _state(z, f"{tag}in_")
    mss = float(z[f"{tag}max_screen_size"])
    out = D.densify_and_prune(st, float(z["max_grad"]), float(z["min_opacity"]), float(z["extent"]),
                              None if mss < 0 else mss, float(z["percent_dense"]),
                              unit_normals=T(z[f"{tag}unit_normals"]))
    assert out.P != st.P
    _assert_same(out, z, f"{tag}out_")
    _assert_same(D.reset_opacity(out), z, f"{tag}reset_")


def test_ply_written_by_the_reference_is_read_and_rewritten_byte_for_byte(tmp_path):
    """GaussianModel.save_ply (:311-329): property order, channel-major SH, raw opacity / scale."""
    z = np.load(GOLD / "ref_model_small.npz")
    got = ply_io.load_ply(str(GOLD / "ref_point_cloud.ply"))
    feats = np.concatenate((z["f_dc"], z["f_rest"]), axis=1)
    for k, want in (("xyz", z["xyz"]), ("normals", z["normals"]), ("features", feats), ("opacity", z["opacity"]),
                    ("scaling", z["scaling"]), ("rotation", z["rotation"]), ("feature_vector", z["feature_vector"])):
        assert np.array_equal(got[k], want), k
    path = tmp_path / "point_cloud.ply"
    ply_io.save_ply(str(path), z["xyz"], feats, z["opacity"], z["scaling"], z["rotation"], normals=z["normals"],
                    feature_vector=z["feature_vector"])
    assert path.read_bytes() == (GOLD / "ref_point_cloud.ply").read_bytes()


def test_checkpoint_written_by_the_reference_converts_both_ways():
    """torch.save((gaussians.capture(), iteration)) (train.py:132; capture :75-91) read into this repo's layout and
    converted back: same 14-tuple, same optimizer groups and moments."""
    z = np.load(GOLD / "ref_model_small.npz")
    model_args, iteration = ck.load_checkpoint(str(GOLD / "ref_chkpnt.pth"))
    assert iteration == 4242 and len(model_args) == 14
    d = ck.from_reference_tuple(model_args)
    feats = np.concatenate((z["f_dc"], z["f_rest"]), axis=1)
    assert np.array_equal(d["raw"]["xyz"].numpy(), z["xyz"]) and np.array_equal(d["raw"]["features"].numpy(), feats)
    for k in ("opacity", "scaling", "rotation"):
        assert np.array_equal(d["raw"][k].numpy(), z[k]), k
    for key, pre in (("exp_avg", "m_"), ("exp_avg_sq", "v_")):
        for k in ("xyz", "opacity", "scaling", "rotation"):
            assert np.array_equal(d[key][k].numpy(), z[f"{pre}{k}"]), (key, k)
        assert np.array_equal(d[key]["features"].numpy(), np.concatenate((z[f"{pre}f_dc"], z[f"{pre}f_rest"]), axis=1))
    assert d["step"] == 3 and d["active_sh_degree"] == 3 and abs(d["spatial_lr_scale"] - 1.7) < 1e-12
    assert np.array_equal(d["feature_vector"].numpy(), z["feature_vector"])
    assert np.array_equal(d["max_radii2D"].numpy(), z["max_radii2D"])
    # back to the reference's tuple: identical tensors, identical optimizer serialisation
    back = ck.to_reference_tuple(d["raw"], d["exp_avg"], d["exp_avg_sq"], d["step"], d["lrs"], d["grad_accum"],
                                 d["denom"], d["max_radii2D"], d["normals"], d["feature_vector"],
                                 d["active_sh_degree"], d["spatial_lr_scale"])
    assert back[0] == model_args[0] and back[13] == model_args[13]
    for i in (1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11):
        assert torch.equal(back[i].detach(), model_args[i].detach()), i
    ref_opt, my_opt = model_args[12], back[12]
    ref_groups = [g for g in ref_opt["param_groups"]]
    assert [g["name"] for g in my_opt["param_groups"]] == [g["name"] for g in ref_groups]
    for a, b in zip(my_opt["param_groups"], ref_groups):
        assert a["params"] == b["params"] and abs(a["lr"] - b["lr"]) < 1e-15 and a["eps"] == b["eps"]
        assert tuple(a["betas"]) == tuple(b["betas"])
    assert set(my_opt["state"]) == set(ref_opt["state"])
    for i, s in ref_opt["state"].items():
        assert float(my_opt["state"][i]["step"]) == float(s["step"])
        assert torch.equal(my_opt["state"][i]["exp_avg"], s["exp_avg"])
        assert torch.equal(my_opt["state"][i]["exp_avg_sq"], s["exp_avg_sq"])
