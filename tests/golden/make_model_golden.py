"""Golden files for SURVEY.md §8(f4) produced by the REFERENCE's own scene/gaussian_model.py (run in the build
container, where /root/reference exists; the files travel, the reference does not):

    tests/golden/ref_densify.npz        inputs + outputs of GaussianModel.densify_and_prune (:534-548) and reset_opacity (:331-334)
    tests/golden/ref_point_cloud.ply    GaussianModel.save_ply (:311-329)
    tests/golden/ref_chkpnt.pth         torch.save((gaussians.capture(), iteration)) as train.py:132 writes it (:75-91)

    python tests/golden/make_model_golden.py

The reference module cannot be imported as is in this container: it imports `simple_knn._C` (CUDA extension, not
built), `tinycudann` and `plyfile` (not installed) at module level and hard-codes device="cuda". This script
  * registers empty stand-in modules for simple_knn._C / tinycudann (neither is touched by the methods used here);
  * registers a minimal `plyfile` stand-in with the two entry points save_ply / load_ply use — PlyElement.describe +
    PlyData.write / PlyData.read for ONE binary-little-endian `vertex` element of float32 properties, the framing
    plyfile itself writes for such an element (header lines `property float <name>`); column order and values come
    from the reference's code;
  * executes the sources of scene/gaussian_model.py and utils/general_utils.py with the literal device="cuda" /
    device='cuda' replaced by "cpu" (the only edit; there is no GPU in the container).
Everything else — tensor layouts, Adam state surgery, split sampling (`torch.normal` on the global CPU generator),
pruning — is the reference's code running unmodified.
"""
from __future__ import annotations

import sys
import types
from pathlib import Path

import numpy as np
import torch

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent


# ------------------------------------------------------------------ stand-ins for modules missing here
class _PlyElement:
    def __init__(self, data, name):
        self.data, self.name = data, name
        self.properties = [types.SimpleNamespace(name=n) for n in data.dtype.names]

    @staticmethod
    def describe(data, name):
        return _PlyElement(data, name)

    def __getitem__(self, key):
        return self.data[key]


class _PlyData:
    def __init__(self, elements):
        self.elements = list(elements)

    def write(self, path):
        (el,) = self.elements
        hdr = ["ply", "format binary_little_endian 1.0", f"element {el.name} {len(el.data)}"]
        hdr += [f"property float {n}" for n in el.data.dtype.names] + ["end_header"]
        with open(path, "wb") as f:
            f.write(("\n".join(hdr) + "\n").encode("ascii"))
            f.write(el.data.astype(el.data.dtype.newbyteorder("<")).tobytes())

    @staticmethod
    def read(path):
        with open(path, "rb") as f:
            names, count = [], 0
            while True:
                line = f.readline().decode("ascii").strip()
                if line.startswith("element vertex"):
                    count = int(line.split()[-1])
                elif line.startswith("property float"):
                    names.append(line.split()[-1])
                elif line == "end_header":
                    break
            data = np.frombuffer(f.read(), dtype=[(n, "<f4") for n in names], count=count)
        return _PlyData([_PlyElement(data, "vertex")])


def import_reference_gaussian_model():
    for name in ("simple_knn", "simple_knn._C", "tinycudann"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["simple_knn._C"].distCUDA2 = None
    ply = types.ModuleType("plyfile")
    ply.PlyData, ply.PlyElement = _PlyData, _PlyElement
    sys.modules.setdefault("plyfile", ply)
    sys.path.insert(0, str(REF))

    def load_patched(rel, modname):
        src = (REF / rel).read_text().replace('device="cuda"', 'device="cpu"').replace("device='cuda'", "device='cpu'")
        mod = types.ModuleType(modname)
        mod.__file__ = str(REF / rel)
        sys.modules[modname] = mod
        exec(compile(src, mod.__file__, "exec"), mod.__dict__)
        return mod

    import utils  # noqa: F401  (the reference's package; its general_utils hard-codes the device too: :65,83,102)
    load_patched("utils/general_utils.py", "utils.general_utils")
    return load_patched("scene/gaussian_model.py", "ref_gaussian_model")


class Opt:
    """arguments/__init__.py:73-89 (OptimizationParams defaults)."""
    position_lr_init, position_lr_final, position_lr_delay_mult, position_lr_max_steps = 0.00016, 0.0000016, 0.01, 30_000
    feature_lr, opacity_lr, scaling_lr, rotation_lr, percent_dense = 0.0025, 0.05, 0.005, 0.001, 0.01


def seeded_model(mod, P, seed, extent):
    from torch import nn
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)  # noqa: E731
    m = mod.GaussianModel(3)
    m.active_sh_degree = 3
    Pm = lambda t: nn.Parameter(t.requires_grad_(True))  # noqa: E731
    m._xyz, m._normals = Pm(r(P, 3)), Pm(r(P, 3))
    m._features_dc, m._features_rest, m._features = Pm(r(P, 1, 3) * 0.3), Pm(r(P, 15, 3) * 0.1), Pm(r(P, 64))
    m._scaling = Pm(r(P, 3) * 1.2 + torch.log(torch.tensor(0.01 * extent)))
    m._rotation, m._opacity = Pm(r(P, 4)), Pm(r(P, 1) * 3.0)
    m.max_radii2D = torch.zeros(P)
    m.spatial_lr_scale = 1.7
    m.training_setup(Opt)
    for _ in range(3):                                             # non-trivial Adam moments (the 64-d group gets none)
        for p in (m._xyz, m._features_dc, m._features_rest, m._opacity, m._scaling, m._rotation):
            p.grad = r(*p.shape)
        m.optimizer.step()
        m.optimizer.zero_grad(set_to_none=True)
    m.xyz_gradient_accum = torch.rand(P, 1, generator=g) * 0.001
    m.denom = torch.randint(0, 4, (P, 1), generator=g).float()     # zeros -> NaN gradients -> treated as 0
    m.max_radii2D = torch.rand(P, generator=g) * 40
    return m


def snapshot(m, prefix):
    d = {}
    for name, t in (("xyz", m._xyz), ("normals", m._normals), ("f_dc", m._features_dc), ("f_rest", m._features_rest),
                    ("feature_vector", m._features), ("scaling", m._scaling), ("rotation", m._rotation),
                    ("opacity", m._opacity), ("max_radii2D", m.max_radii2D), ("grad_accum", m.xyz_gradient_accum),
                    ("denom", m.denom)):
        d[f"{prefix}{name}"] = t.detach().numpy().copy()
    for grp in m.optimizer.param_groups:
        st = m.optimizer.state.get(grp["params"][0])
        if st is not None and "exp_avg" in st:
            d[f"{prefix}m_{grp['name']}"] = st["exp_avg"].numpy().copy()
            d[f"{prefix}v_{grp['name']}"] = st["exp_avg_sq"].numpy().copy()
    return d


def main():
    mod = import_reference_gaussian_model()
    extent, thr, min_opacity = 5.0, 0.0002, 0.005
    out = {"extent": extent, "max_grad": thr, "min_opacity": min_opacity, "percent_dense": Opt.percent_dense}
    for tag, seed, max_screen in (("a_", 0, None), ("b_", 1, 20)):
        m = seeded_model(mod, 300, seed, extent)
        out.update(snapshot(m, f"{tag}in_"))
        grads = m.xyz_gradient_accum / m.denom
        grads[grads.isnan()] = 0.0
        # the split's draws: same global-generator state the reference's torch.normal will consume
        n_clone = int(torch.logical_and(torch.norm(grads, dim=-1) >= thr,
                                        torch.max(m.get_scaling, dim=1).values <= Opt.percent_dense * extent).sum())
        pad = torch.cat((grads.squeeze(), torch.zeros(n_clone)))
        scal = torch.cat((m.get_scaling, m.get_scaling[torch.logical_and(
            torch.norm(grads, dim=-1) >= thr, torch.max(m.get_scaling, dim=1).values <= Opt.percent_dense * extent)]))
        n_split = int(torch.logical_and(pad >= thr, torch.max(scal, dim=1).values > Opt.percent_dense * extent).sum())
        torch.manual_seed(4321 + seed)
        unit = torch.randn(2 * n_split, 3)
        torch.manual_seed(4321 + seed)
        m.densify_and_prune(thr, min_opacity, extent, max_screen)
        out[f"{tag}unit_normals"] = unit.numpy()
        out[f"{tag}max_screen_size"] = -1.0 if max_screen is None else float(max_screen)
        out.update(snapshot(m, f"{tag}out_"))
        m.reset_opacity()
        out.update(snapshot(m, f"{tag}reset_"))
        print(tag, "P", 300, "->", m.get_xyz.shape[0], "clones", n_clone, "splits", n_split)
    np.savez_compressed(OUT / "ref_densify.npz", **out)

    # ---- PLY and checkpoint of a small model (after one densification, so the state is not pristine)
    m = seeded_model(mod, 120, 7, extent)
    m.save_ply(str(OUT / "ref_point_cloud.ply"))
    torch.save((m.capture(), 4242), str(OUT / "ref_chkpnt.pth"))
    np.savez_compressed(OUT / "ref_model_small.npz", **snapshot(m, ""))
    # the reference reading its own file back (load_ply :336-391): must equal what it wrote
    m2 = mod.GaussianModel(3)
    m2.load_ply(str(OUT / "ref_point_cloud.ply"))
    assert torch.equal(m2._xyz, m._xyz) and torch.equal(m2._features_rest, m._features_rest)
    assert torch.equal(m2._features, m._features) and torch.equal(m2._rotation, m._rotation)
    print("wrote", [p.name for p in OUT.glob("ref_*") if p.stat().st_mtime > 0])


if __name__ == "__main__":
    main()
