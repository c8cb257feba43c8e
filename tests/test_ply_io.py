"""point_cloud.ply layout (ply_io.py) against the reference's save_ply / load_ply (scene/gaussian_model.py:293-390):
property order, channel-major SH on disk, raw parameters, byte-level header, round trips, original-3DGS files."""
import numpy as np
import pytest

from neuralgaussiansplatting_b200 import ply_io


def _model(P, seed=0):
    r = np.random.default_rng(seed)
    return dict(xyz=r.normal(size=(P, 3)).astype(np.float32), features=r.normal(size=(P, 16, 3)).astype(np.float32),
                opacity=r.normal(size=(P, 1)).astype(np.float32), scaling=r.normal(size=(P, 3)).astype(np.float32),
                rotation=r.normal(size=(P, 4)).astype(np.float32), normals=r.normal(size=(P, 3)).astype(np.float32),
                feature_vector=r.normal(size=(P, 64)).astype(np.float32))


def test_attribute_order_is_the_reference_order():
    n = ply_io.attribute_names()
    assert n[:9] == ["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"]
    assert n[9] == "f_rest_0" and n[53] == "f_rest_44" and n[54] == "features_0" and n[117] == "features_63"
    assert n[118:] == ["opacity", "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3"]
    assert len(n) == 3 + 3 + 3 + 45 + 64 + 1 + 3 + 4


@pytest.mark.parametrize("P", [0, 1, 257])
def test_round_trip_is_bit_exact(tmp_path, P):
    m = _model(P)
    path = str(tmp_path / "sub" / "point_cloud.ply")          # save_ply creates the directory (mkdir_p)
    ply_io.save_ply(path, **m)
    back = ply_io.load_ply(path)
    for k, v in m.items():
        assert back[k].dtype == np.float32 and back[k].shape == v.shape, k
        assert np.array_equal(back[k], v), k


def test_header_and_record_bytes(tmp_path):
    m = _model(3, 1)
    path = str(tmp_path / "pc.ply")
    ply_io.save_ply(path, **m)
    raw = open(path, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    lines = head.decode().splitlines()
    assert lines[:3] == ["ply", "format binary_little_endian 1.0", "element vertex 3"]
    assert lines[3:] == [f"property float {n}" for n in ply_io.attribute_names()]
    rec = np.frombuffer(body, dtype="<f4").reshape(3, 126)
    assert np.array_equal(rec[:, 0:3], m["xyz"]) and np.array_equal(rec[:, 3:6], m["normals"])
    # SH is channel-major on disk: f_dc_c = features[:,0,c]; f_rest_(c*15 + k) = features[:,1+k,c]
    assert np.array_equal(rec[:, 6:9], m["features"][:, 0, :])
    for c in range(3):
        for k in (0, 7, 14):
            assert np.array_equal(rec[:, 9 + c * 15 + k], m["features"][:, 1 + k, c])
    assert np.array_equal(rec[:, 118], m["opacity"][:, 0]) and np.array_equal(rec[:, 122:126], m["rotation"])


def test_reads_original_3dgs_files_without_feature_vector_and_ascii(tmp_path):
    # the original 3DGS writes no features_* columns; load_ply then returns an empty [P,0] feature_vector
    m = _model(4, 2)
    names = [n for n in ply_io.attribute_names() if not n.startswith("features_")]
    f = m["features"]
    table = np.concatenate((m["xyz"], m["normals"], f[:, 0, :], f[:, 1:].transpose(0, 2, 1).reshape(4, 45), m["opacity"],
                            m["scaling"], m["rotation"]), axis=1)
    path = str(tmp_path / "ascii.ply")
    with open(path, "w") as fh:
        fh.write("ply\nformat ascii 1.0\ncomment made by a test\nelement vertex 4\n")
        fh.write("".join(f"property float {n}\n" for n in names) + "end_header\n")
        for row in table:
            fh.write(" ".join(repr(float(x)) for x in row) + "\n")
    back = ply_io.load_ply(path)
    assert back["feature_vector"].shape == (4, 0)
    assert np.array_equal(back["features"], f) and np.array_equal(back["scaling"], m["scaling"])
    bad = str(tmp_path / "bad.ply")
    open(bad, "w").write("ply\nformat ascii 1.0\nelement vertex 0\nproperty float x\nproperty float y\nproperty float z\n"
                         "property float opacity\nproperty float f_dc_0\nproperty float f_dc_1\nproperty float f_dc_2\nend_header\n")
    with pytest.raises(ValueError, match="f_rest"):
        ply_io.load_ply(bad)
