"""Gaussian map <-> PLY file, in the layout SplatLoc writes and reads (first part).

Host-side mirror of `GaussianModel.save_ply` / `load_ply`
(/root/reference/gaussian_splatting/scene/gaussian_model.py:327-377, 397-475): one `vertex` element
of float32 properties

    x y z  nx ny nz  f_dc_0..2  f_rest_0..(3K-1)  opacity  scale_0..2  rot_0..3  marker  kp_score

with the SH coefficients stored channel-major (`features.transpose(1, 2).flatten(1)`), raw
(pre-activation) opacity / scale / rotation, zero normals.  The reference goes through the `plyfile`
package, which is not in this image; the writer below emits the same file that package produces for
such an element (header lines `property float <name>`, `format binary_little_endian 1.0`, packed
little-endian records) and the reader accepts any PLY whose first element carries these properties
(binary little/big endian or ascii, any scalar property types).  Parity with `plyfile` itself is
unpinned; the format is the published PLY specification and the tests fix the byte layout.
"""
from __future__ import annotations

import numpy as np
import torch

_PLY_TYPES = {"char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2",
              "ushort": "u2", "uint16": "u2", "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4",
              "float": "f4", "float32": "f4", "double": "f8", "float64": "f8"}


def attribute_names(n_dc: int, n_rest: int, n_scale: int = 3, n_rot: int = 4) -> list[str]:
    """`construct_list_of_attributes`, gaussian_model.py:327-343."""
    names = ["x", "y", "z", "nx", "ny", "nz"]
    names += [f"f_dc_{i}" for i in range(n_dc)]
    names += [f"f_rest_{i}" for i in range(n_rest)]
    names.append("opacity")
    names += [f"scale_{i}" for i in range(n_scale)]
    names += [f"rot_{i}" for i in range(n_rot)]
    names += ["marker", "kp_score"]
    return names


def _np(t) -> np.ndarray:
    return t.detach().cpu().numpy() if torch.is_tensor(t) else np.asarray(t)


def save_ply(path: str, xyz, features_dc, features_rest, opacity, scaling, rotation, marker, kp_score) -> None:
    """Arguments as the model holds them: xyz (P,3), features_dc (P,1,3), features_rest (P,K,3),
    opacity (P,1), scaling (P,3), rotation (P,4), marker (P,1), kp_score (P,1) -- raw parameters."""
    xyz = _np(xyz).astype(np.float32)
    P = xyz.shape[0]
    f_dc = np.ascontiguousarray(np.transpose(_np(features_dc), (0, 2, 1))).reshape(P, -1)
    f_rest = np.ascontiguousarray(np.transpose(_np(features_rest), (0, 2, 1))).reshape(P, -1)
    cols = [xyz, np.zeros_like(xyz), f_dc, f_rest, _np(opacity).reshape(P, -1), _np(scaling).reshape(P, -1),
            _np(rotation).reshape(P, -1), _np(marker).reshape(P, -1), _np(kp_score).reshape(P, -1)]
    table = np.concatenate([c.astype(np.float32) for c in cols], axis=1)
    names = attribute_names(f_dc.shape[1], f_rest.shape[1], cols[5].shape[1], cols[6].shape[1])
    assert table.shape[1] == len(names)
    header = ["ply", "format binary_little_endian 1.0", f"element vertex {P}"]
    header += [f"property float {n}" for n in names]
    header.append("end_header")
    with open(path, "wb") as f:
        f.write(("\n".join(header) + "\n").encode("ascii"))
        f.write(np.ascontiguousarray(table.astype("<f4")).tobytes())


def read_vertex_table(path: str) -> dict[str, np.ndarray]:
    """All scalar properties of the file's first element, by name."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError(f"{path}: not a PLY file")
        fmt, count, props, element = None, None, [], 0     # element: 0 before, 1 inside, 2 after the first one
        while True:
            line = f.readline()
            if not line:
                raise ValueError(f"{path}: header has no end_header")
            tok = line.decode("ascii", "replace").split()
            if not tok or tok[0] in ("comment", "obj_info"):
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                element += 1
                if element == 1:
                    count = int(tok[2])
            elif tok[0] == "property" and element == 1:
                if tok[1] == "list":
                    raise ValueError(f"{path}: list property in the vertex element")
                props.append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if fmt is None or count is None:
            raise ValueError(f"{path}: incomplete PLY header")
        if fmt == "ascii":
            rows = np.loadtxt(f, dtype=np.float64, max_rows=count, ndmin=2)
            return {n: rows[:, i].astype(t) for i, (n, t) in enumerate(props)}
        order = "<" if fmt == "binary_little_endian" else ">"
        dt = np.dtype([(n, order + t) for n, t in props])
        rec = np.frombuffer(f.read(count * dt.itemsize), dtype=dt, count=count)
        return {n: np.ascontiguousarray(rec[n]) for n, _ in props}


def load_ply(path: str, max_sh_degree: int, device="cpu") -> dict[str, torch.Tensor]:
    """`load_ply`, gaussian_model.py:397-475: returns the raw parameter tensors in the model's shapes
    (features_dc (P,1,3), features_rest (P,K,3) with K = (max_sh_degree + 1)^2 - 1)."""
    v = read_vertex_table(path)
    P = v["x"].shape[0]

    def numbered(prefix):
        names = sorted((n for n in v if n.startswith(prefix)), key=lambda n: int(n.split("_")[-1]))
        return np.stack([v[n] for n in names], axis=1) if names else np.zeros((P, 0))

    rest = numbered("f_rest_")
    expected = 3 * (max_sh_degree + 1) ** 2 - 3
    assert rest.shape[1] == expected, "{} not eq {}".format(rest.shape[1], expected)
    dc = np.stack([v["f_dc_0"], v["f_dc_1"], v["f_dc_2"]], axis=1).reshape(P, 3, 1)
    rest = rest.reshape(P, 3, (max_sh_degree + 1) ** 2 - 1)

    def t(a):
        return torch.tensor(np.asarray(a).astype(np.float64), dtype=torch.float, device=device)   # (native byte order)

    return {"xyz": t(np.stack([v["x"], v["y"], v["z"]], axis=1)),
            "features_dc": t(dc).transpose(1, 2).contiguous(),
            "features_rest": t(rest).transpose(1, 2).contiguous(),
            "opacity": t(v["opacity"][:, None]), "scaling": t(numbered("scale_")), "rotation": t(numbered("rot")),
            "marker": t(v["marker"][:, None]), "kp_score": t(v["kp_score"][:, None])}
