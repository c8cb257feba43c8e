"""Generates the golden vectors under tests/golden/ by running the UNMODIFIED reference rasterizers
(oracle/_ref, compiled from /root/reference by oracle/build_ref.sh for sm_100) on a B200:

    gpurun -- 'python tests/golden/make_golden.py --out gpurun_out/golden'   # then copy *.npz to tests/golden/

The reference ships no golden vectors of its own, so these outputs of the reference itself are
what pins both the CPU oracle (tests/test_oracle_cpu.py) and the CUDA path (tests/test_golden_gpu.py).
Inputs are stored alongside the outputs (T0) or re-derived from seeds and guarded by a checksum (T1).
"""
from __future__ import annotations

import argparse
import hashlib
import math
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

import helpers as Hh  # noqa: E402
import refbind  # noqa: E402
from neuralgaussiansplatting_b200.synthetic import make_grad_featuremap, make_grad_image  # noqa: E402


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def input_checksum(sc, cam) -> str:
    h = hashlib.sha256()
    for k in sorted(sc):
        h.update(np.ascontiguousarray(sc[k].cpu().numpy()).tobytes())
    for t in (cam.world_view_transform, cam.full_proj_transform, cam.camera_center):
        h.update(np.ascontiguousarray(t.cpu().numpy()).tobytes())
    return h.hexdigest()


def make(cfg_name: str, out: Path, store_inputs: bool, variant: str = "sh_sr"):
    dev = torch.device("cuda:0")
    cfg, sc, cam = Hh.scene_on(cfg_name, dev)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    dL = make_grad_image(cfg, 0, device=dev)
    ref = refbind.RefDGR()
    kwv = {}
    if variant == "precomp":
        cov, col = Hh.cpu_covariance_and_colors({k: v.cpu() for k, v in sc.items()}, cam.to("cpu"))
        cam.to(dev)
        kwv = dict(colors_precomp=col.to(dev), cov3D_precomp=cov.to(dev))
    kw, (R, color, radii) = Hh.ref_forward(ref, cfg, sc, cam, bg, **kwv)
    g = Hh.ref_backward(ref, kw, radii, dL)
    d = dict(cfg=cfg_name, variant=variant, num_rendered=R, bg=bg.cpu().numpy(),
             input_sha=input_checksum(sc, cam),
             radii=radii.cpu().numpy(), tiles_touched=ref.fetch("tiles_touched"),
             ranges=ref.fetch("ranges"), n_contrib=ref.fetch("n_contrib"),
             final_T=ref.fetch("accum_alpha"), color=color.cpu().numpy(),
             keys_sha=sha(ref.fetch("point_list_keys")), list_sha=sha(ref.fetch("point_list")),
             depths=ref.fetch("depths"), means2D=ref.fetch("means2D"), conic_opacity=ref.fetch("conic_opacity"),
             rgb=ref.fetch("rgb"), clamped=ref.fetch("clamped"))
    vis = d["radii"] > 0
    for name in ("depths",):
        d[name] = np.where(vis, d[name], 0).astype(np.float32)
    d["means2D"] = np.where(vis[:, None], d["means2D"].reshape(-1, 2), 0).astype(np.float32)
    d["conic_opacity"] = np.where(vis[:, None], d["conic_opacity"].reshape(-1, 4), 0).astype(np.float32)
    d["rgb"] = np.where(vis[:, None], d["rgb"].reshape(-1, 3), 0).astype(np.float32)
    d["clamped"] = np.where(vis[:, None], d["clamped"].reshape(-1, 3), 0).astype(np.uint8)
    for k in ("means2D", "opacity", "means3D", "scales", "rotations", "colors", "cov3D", "conic"):
        d["g_" + k] = g[k].cpu().numpy()
    if store_inputs:
        d["point_list_keys"] = ref.fetch("point_list_keys")
        d["point_list"] = ref.fetch("point_list")
        d["g_sh"] = g["sh"].cpu().numpy()
        for k, v in sc.items():
            d["in_" + k] = v.cpu().numpy()
        d["in_viewmatrix"] = cam.world_view_transform.cpu().numpy()
        d["in_projmatrix"] = cam.full_proj_transform.cpu().numpy()
        d["in_campos"] = cam.camera_center.cpu().numpy()
        d["in_tanfov"] = np.array([math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5)], np.float64)
        d["in_dL_dpix"] = dL.cpu().numpy()
        for k, v in kwv.items():
            d["in_" + k] = v.cpu().numpy()
    else:
        d["g_sh_rowsum"] = g["sh"].sum(dim=(1, 2)).cpu().numpy()
    ref.close()

    if variant == "sh_sr":
        # idxmap extra channels from the reference rasterizer2
        r2 = refbind.RefR2()
        R2, idx, col, dep, feat, rad2 = r2.getidxmap(
            means3D=sc["means3D"], normal=sc["normal"], scales=sc["scales"], rotations=sc["rotations"],
            viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform,
            tan_fovx=math.tan(cam.FoVx * 0.5), tan_fovy=math.tan(cam.FoVy * 0.5),
            feature_vector=sc["feature_vector"], shs=sc["shs"], campos=cam.camera_center, W=cfg.width, H=cfg.height)
        d["idx_num_rendered"] = R2
        d["idx_idx"] = idx.cpu().numpy()
        d["idx_depth"] = dep.cpu().numpy()
        d["idx_color"] = col.cpu().numpy()
        d["idx_radii"] = rad2.cpu().numpy()
        d["idx_feature_head"] = feat[..., :25].cpu().numpy() if store_inputs else None
        d["idx_feature_sha"] = sha(feat.cpu().numpy())
        if not store_inputs:
            d.pop("idx_feature_head")
        r2.close()
    out.mkdir(parents=True, exist_ok=True)
    path = out / f"ref_{cfg_name}_{variant}.npz"
    np.savez_compressed(path, **d)
    print("wrote", path, path.stat().st_size, "bytes; num_rendered", R)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=str(ROOT / "gpurun_out" / "golden"))
    a = ap.parse_args()
    out = Path(a.out)
    make("T0", out, store_inputs=True)
    make("T0", out, store_inputs=True, variant="precomp")
    make("T1", out, store_inputs=False)
