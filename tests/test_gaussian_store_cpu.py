"""PLY store (splatloc_b200/gaussian_store.py) against the layout of the reference's
GaussianModel.save_ply / load_ply (gaussian_model.py:327-377, 397-475): attribute order, channel-major
SH coefficients, packed little-endian float32 records, exact round trip, foreign PLY flavours."""
import struct

import numpy as np
import pytest
import torch

from splatloc_b200 import gaussian_store as GS


def _model(P, deg, seed=0):
    g = torch.Generator().manual_seed(seed)
    K = (deg + 1) ** 2 - 1
    r = lambda *s: torch.randn(*s, generator=g)
    return dict(xyz=r(P, 3), features_dc=r(P, 1, 3), features_rest=r(P, K, 3), opacity=r(P, 1), scaling=r(P, 3),
                rotation=r(P, 4), marker=r(P, 1), kp_score=r(P, 1))


@pytest.mark.parametrize("deg", [0, 1, 3])
def test_round_trip_is_exact(tmp_path, deg):
    m = _model(257, deg)
    path = str(tmp_path / "point_cloud.ply")
    GS.save_ply(path, **m)
    back = GS.load_ply(path, deg)
    for k, v in m.items():
        assert back[k].shape == v.shape and torch.equal(back[k], v), k


def test_file_layout(tmp_path):
    m = _model(3, 1)
    path = str(tmp_path / "p.ply")
    GS.save_ply(path, **m)
    raw = open(path, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    lines = head.decode().split("\n")
    names = GS.attribute_names(3, 9)
    assert lines[:3] == ["ply", "format binary_little_endian 1.0", "element vertex 3"]
    assert lines[3:3 + len(names)] == [f"property float {n}" for n in names]
    assert names == ["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"] + [f"f_rest_{i}" for i in range(9)] + \
        ["opacity", "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3", "marker", "kp_score"]
    assert len(body) == 3 * len(names) * 4
    row0 = struct.unpack("<" + "f" * len(names), body[: len(names) * 4])
    assert row0[:3] == tuple(m["xyz"][0].tolist()) and row0[3:6] == (0.0, 0.0, 0.0)
    # SH coefficients are channel-major: f_rest_{c * K + k} = features_rest[:, k, c]
    K = 3
    for c in range(3):
        for k in range(K):
            assert row0[9 + c * K + k] == float(m["features_rest"][0, k, c])
    assert row0[6:9] == tuple(m["features_dc"][0, 0].tolist())


def test_reads_ascii_big_endian_and_mixed_types(tmp_path):
    names = GS.attribute_names(3, 0)
    vals = np.arange(2 * len(names), dtype=np.float64).reshape(2, len(names)) / 8.0
    p_ascii = tmp_path / "a.ply"
    with open(p_ascii, "w") as f:
        f.write("ply\nformat ascii 1.0\ncomment made by hand\nelement vertex 2\n")
        f.write("".join(f"property {'double' if i % 2 else 'float'} {n}\n" for i, n in enumerate(names)))
        f.write("element face 0\nproperty list uchar int vertex_indices\nend_header\n")
        f.write("\n".join(" ".join(repr(float(x)) for x in row) for row in vals) + "\n")
    p_be = tmp_path / "b.ply"
    with open(p_be, "wb") as f:
        f.write(("ply\nformat binary_big_endian 1.0\nelement vertex 2\n" + "".join(f"property float {n}\n" for n in names) +
                 "end_header\n").encode())
        f.write(vals.astype(">f4").tobytes())
    a, b = GS.load_ply(str(p_ascii), 0), GS.load_ply(str(p_be), 0)
    for k in a:
        assert torch.equal(a[k], b[k]), k
    assert a["features_rest"].shape == (2, 0, 3) and a["features_dc"].shape == (2, 1, 3)
    assert torch.equal(a["xyz"], torch.tensor(vals[:, :3], dtype=torch.float))
    assert float(a["kp_score"][1, 0]) == float(vals[1, -1])


def test_wrong_sh_degree_is_rejected_like_the_reference(tmp_path):
    path = str(tmp_path / "p.ply")
    GS.save_ply(path, **_model(4, 1))
    with pytest.raises(AssertionError, match="9 not eq 45"):
        GS.load_ply(path, 3)
