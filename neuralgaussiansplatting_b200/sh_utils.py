"""Device-agnostic torch helpers the `render()` caller needs for the reference's
`convert_SHs_python` / `compute_cov3D_python` pipeline toggles (arguments/__init__.py:66-67).

They restate utils/sh_utils.py:57-112 (eval_sh) and utils/general_utils.py:64-110 +
scene/gaussian_model.py:29-33 (covariance from scaling/rotation); the reference versions
hard-code device="cuda" (general_utils.py:65,83,102), these take the device from the input.
"""
from __future__ import annotations

import torch

C0 = 0.28209479177387814
C1 = 0.4886025119029199
C2 = (1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396)
C3 = (-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
      1.445305721320277, -0.5900435899266435)


def eval_sh(deg: int, sh: torch.Tensor, dirs: torch.Tensor) -> torch.Tensor:
    """sh: [..., C, (deg+1)^2], dirs: [..., 3] unit vectors -> [..., C]"""
    assert 0 <= deg <= 3
    assert sh.shape[-1] >= (deg + 1) ** 2
    out = C0 * sh[..., 0]
    if deg > 0:
        x, y, z = dirs[..., 0:1], dirs[..., 1:2], dirs[..., 2:3]
        out = out - C1 * y * sh[..., 1] + C1 * z * sh[..., 2] - C1 * x * sh[..., 3]
        if deg > 1:
            xx, yy, zz = x * x, y * y, z * z
            xy, yz, xz = x * y, y * z, x * z
            out = (out + C2[0] * xy * sh[..., 4] + C2[1] * yz * sh[..., 5] +
                   C2[2] * (2.0 * zz - xx - yy) * sh[..., 6] + C2[3] * xz * sh[..., 7] +
                   C2[4] * (xx - yy) * sh[..., 8])
            if deg > 2:
                out = (out + C3[0] * y * (3 * xx - yy) * sh[..., 9] + C3[1] * xy * z * sh[..., 10] +
                       C3[2] * y * (4 * zz - xx - yy) * sh[..., 11] +
                       C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[..., 12] +
                       C3[4] * x * (4 * zz - xx - yy) * sh[..., 13] + C3[5] * z * (xx - yy) * sh[..., 14] +
                       C3[6] * x * (xx - 3 * yy) * sh[..., 15])
    return out


def build_rotation(r: torch.Tensor) -> torch.Tensor:
    q = r / r.norm(dim=1, keepdim=True)
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
        2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
        2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], dim=1)
    return R.view(-1, 3, 3)


def build_covariance_from_scaling_rotation(scaling: torch.Tensor, scaling_modifier: float,
                                           rotation: torch.Tensor) -> torch.Tensor:
    """[P,6] upper triangle of R S S^T R^T."""
    L = build_rotation(rotation) * (scaling_modifier * scaling)[:, None, :]
    cov = L @ L.transpose(1, 2)
    return torch.stack([cov[:, 0, 0], cov[:, 0, 1], cov[:, 0, 2], cov[:, 1, 1], cov[:, 1, 2], cov[:, 2, 2]], dim=1)
