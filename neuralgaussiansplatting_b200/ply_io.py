"""point_cloud.ply in the reference's exact on-disk layout (data format next to the path).

Reference: GaussianModel.construct_list_of_attributes / save_ply / load_ply (scene/gaussian_model.py:293-329, :336-390).
One binary little-endian `vertex` element, every property `float`, in this order:
    x y z | nx ny nz | f_dc_0..2 | f_rest_0..44 | features_0..63 | opacity | scale_0..2 | rot_0..3
f_dc / f_rest are CHANNEL-major on disk ([P,K,3] transposed to [P,3,K] and flattened, gaussian_model.py:316-317), the
stored opacity / scale values are the RAW (pre-activation) parameters. The reference goes through the `plyfile`
package; this module writes and parses the same bytes with numpy only, so files interchange in both directions.
In memory the SH coefficients use this repo's concatenated layout `features[P,16,3]` (coefficient 0 = f_dc).
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional

import numpy as np

SH_COEFFS = 16
NUM_FEATURES = 64


def attribute_names(num_features: int = NUM_FEATURES, sh_coeffs: int = SH_COEFFS) -> List[str]:
    """construct_list_of_attributes (gaussian_model.py:293-309)."""
    names = ["x", "y", "z", "nx", "ny", "nz"]
    names += [f"f_dc_{i}" for i in range(3)]
    names += [f"f_rest_{i}" for i in range(3 * (sh_coeffs - 1))]
    names += [f"features_{i}" for i in range(num_features)]
    names += ["opacity"]
    names += [f"scale_{i}" for i in range(3)]
    names += [f"rot_{i}" for i in range(4)]
    return names


def _np(t) -> np.ndarray:
    if hasattr(t, "detach"):
        t = t.detach().cpu().numpy()
    return np.asarray(t, dtype=np.float32)


def save_ply(path: str, xyz, features, opacity, scaling, rotation, normals=None, feature_vector=None) -> None:
    """save_ply (gaussian_model.py:311-329). features: [P,16,3]; opacity [P,1] and scaling [P,3] raw; feature_vector
    [P,64] (the fork's per-point descriptor) and normals [P,3] default to zeros."""
    xyz = _np(xyz)
    P = xyz.shape[0]
    f = _np(features).reshape(P, SH_COEFFS if P == 0 else -1, 3)
    K = f.shape[1]
    normals = np.zeros((P, 3), np.float32) if normals is None else _np(normals).reshape(P, 3)
    fv = np.zeros((P, NUM_FEATURES), np.float32) if feature_vector is None else _np(feature_vector).reshape(P, NUM_FEATURES if P == 0 else -1)
    f_dc = f[:, :1].transpose(0, 2, 1).reshape(P, 3)
    f_rest = f[:, 1:].transpose(0, 2, 1).reshape(P, 3 * (K - 1))
    table = np.concatenate((xyz.reshape(P, 3), normals, f_dc, f_rest, fv, _np(opacity).reshape(P, 1),
                            _np(scaling).reshape(P, 3), _np(rotation).reshape(P, 4)), axis=1).astype("<f4")
    names = attribute_names(fv.shape[1], K)
    assert table.shape[1] == len(names)
    header = ["ply", "format binary_little_endian 1.0", f"element vertex {P}"]
    header += [f"property float {n}" for n in names]
    header += ["end_header"]
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)                      # mkdir_p(os.path.dirname(path)), :312
    with open(path, "wb") as fh:
        fh.write(("\n".join(header) + "\n").encode("ascii"))
        fh.write(np.ascontiguousarray(table).tobytes())


_PLY_TYPES = {"float": "f4", "float32": "f4", "double": "f8", "float64": "f8", "uchar": "u1", "uint8": "u1",
              "char": "i1", "int8": "i1", "short": "i2", "int16": "i2", "ushort": "u2", "uint16": "u2",
              "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4"}


def read_vertex_table(path: str) -> np.ndarray:
    """The first element of a PLY file as a numpy structured array (binary little/big endian or ascii)."""
    with open(path, "rb") as fh:
        if fh.readline().strip() != b"ply":
            raise ValueError(f"{path}: not a PLY file")
        fmt, count, props, in_first, seen_element = None, None, [], False, False
        while True:
            line = fh.readline()
            if not line:
                raise ValueError(f"{path}: unterminated PLY header")
            tok = line.decode("ascii").split()
            if not tok or tok[0] == "comment" or tok[0] == "obj_info":
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                in_first = not seen_element
                if in_first:
                    count = int(tok[2])
                seen_element = True
            elif tok[0] == "property" and in_first:
                if tok[1] == "list":
                    raise ValueError("list properties are not supported in the vertex element")
                props.append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if count is None:
            raise ValueError(f"{path}: no element in header")
        if fmt == "ascii":
            out = np.empty(count, dtype=[(n, t) for n, t in props])
            if count == 0:
                return out
            rows = np.loadtxt(fh, max_rows=count, ndmin=2)
            for i, (n, _) in enumerate(props):
                out[n] = rows[:, i]
            return out
        order = "<" if fmt == "binary_little_endian" else ">"
        dt = np.dtype([(n, order + t) for n, t in props])
        return np.frombuffer(fh.read(count * dt.itemsize), dtype=dt, count=count)


def load_ply(path: str, max_sh_degree: int = 3) -> Dict[str, np.ndarray]:
    """load_ply (gaussian_model.py:336-390) -> float32 arrays {xyz[P,3], normals[P,3], features[P,16,3], opacity[P,1],
    scaling[P,3], rotation[P,4], feature_vector[P,F]} (raw parameters; F may be 0 for files of the original 3DGS)."""
    v = read_vertex_table(path)
    names = v.dtype.names
    col = lambda n: np.asarray(v[n], dtype=np.float32)  # noqa: E731
    P = v.shape[0]
    xyz = np.stack((col("x"), col("y"), col("z")), axis=1)
    normals = np.stack((col("nx"), col("ny"), col("nz")), axis=1) if "nx" in names else np.zeros_like(xyz)

    def numbered(prefix: str) -> List[str]:
        return sorted((n for n in names if n.startswith(prefix)), key=lambda n: int(n.split("_")[-1]))

    K = (max_sh_degree + 1) ** 2
    rest_names = numbered("f_rest_")
    if len(rest_names) != 3 * K - 3:                       # the reference asserts the same (:357)
        raise ValueError(f"{path}: {len(rest_names)} f_rest_* properties, expected {3 * K - 3}")
    f_dc = np.stack((col("f_dc_0"), col("f_dc_1"), col("f_dc_2")), axis=1).reshape(P, 3, 1)
    f_rest = np.stack([col(n) for n in rest_names], axis=1).reshape(P, 3, K - 1) if K > 1 else np.zeros((P, 3, 0), np.float32)
    features = np.concatenate((f_dc, f_rest), axis=2).transpose(0, 2, 1)          # [P,K,3]
    scaling = np.stack([col(n) for n in numbered("scale_")], axis=1)
    rotation = np.stack([col(n) for n in numbered("rot_")], axis=1)               # the reference matches "rot" (:370)
    fv_names = numbered("features_")
    fv = np.stack([col(n) for n in fv_names], axis=1) if fv_names else np.zeros((P, 0), np.float32)
    return {"xyz": xyz, "normals": normals, "features": np.ascontiguousarray(features, dtype=np.float32),
            "opacity": col("opacity")[:, None], "scaling": scaling, "rotation": rotation, "feature_vector": fv}
