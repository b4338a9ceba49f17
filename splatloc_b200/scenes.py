"""Seeded synthetic scenes, cameras and upstream gradients (no datasets offline).

Shapes follow SURVEY.md section 8(d): a Replica room_0-sized box
(`configs/replica_nerf/room_0.yaml:6` bounds) with most Gaussians on the six
faces, SplatLoc's 2 cm primitive prior (`train_gaussians.py:226`), 4-channel
precomputed colours `[rgb, kp_score]` as `gaussian_renderer/__init__.py:85-102`
builds them, Replica / 12-Scenes intrinsics
(`configs/replica_nerf/base_config.yaml:16-29`, `configs/scenes12/base_config.yaml:16-29`)
and the projection of `getProjectionMatrix2`
(`gaussian_splatting/utils/graphics_utils.py:72-93`, znear 0.01, zfar 100).
Camera tensors use the reference convention: `viewmatrix = W2C^T`,
`projmatrix = (P W2C)^T`, `campos = inverse(viewmatrix)[3, :3]`
(`utils/camera_utils.py:129-139`).

Everything is generated on the CPU with a `torch.Generator` so that the same
seed gives the same bits on every machine; callers move tensors to the GPU.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch

ROOM_LO = (-1.0, -1.3, -1.7)
ROOM_HI = (7.0, 3.7, 1.4)
SH_C0 = 0.28209479177387814

REPLICA_INTRINSICS = dict(W=640, H=480, fx=320.0, fy=320.0, cx=319.5, cy=239.5)
SCENES12_INTRINSICS = dict(W=640, H=480, fx=572.0, fy=572.0, cx=320.0, cy=240.0)


@dataclass
class Camera:
    W: int
    H: int
    tanfovx: float
    tanfovy: float
    viewmatrix: torch.Tensor   # (4,4) float32, transposed world-to-camera
    projmatrix: torch.Tensor   # (4,4) float32, transposed full projection
    campos: torch.Tensor       # (3,)

    def to(self, device):
        return Camera(self.W, self.H, self.tanfovx, self.tanfovy, self.viewmatrix.to(device),
                      self.projmatrix.to(device), self.campos.to(device))


def _gen(seed: int) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed(int(seed))
    return g


def room_points(P: int, g: torch.Generator, lo=ROOM_LO, hi=ROOM_HI, face_frac=0.85, jitter=0.01):
    """(P,3) positions and (P,) face-normal axis (-1 for interior clutter)."""
    lo_t, hi_t = torch.tensor(lo, dtype=torch.float64), torch.tensor(hi, dtype=torch.float64)
    ext = hi_t - lo_t
    xyz = lo_t + ext * torch.rand((P, 3), generator=g, dtype=torch.float64)
    n_face = int(round(P * face_frac))
    # choose a face with probability proportional to its area
    areas = torch.stack([ext[1] * ext[2], ext[0] * ext[2], ext[0] * ext[1]]).repeat_interleave(2)
    face = torch.multinomial(areas / areas.sum(), n_face, replacement=True, generator=g)
    axis = face // 2
    side = face % 2
    wall = torch.where(side == 0, lo_t[axis], hi_t[axis])
    wall = wall + jitter * torch.randn(n_face, generator=g, dtype=torch.float64)
    xyz[torch.arange(n_face), axis] = wall
    normal_axis = torch.full((P,), -1, dtype=torch.long)
    normal_axis[:n_face] = axis
    perm = torch.randperm(P, generator=g)
    return xyz[perm].float(), normal_axis[perm]


def make_scene(P: int, seed: int = 0, channels: int = 4, sh_degree: int | None = None, scale: float = 1.0):
    """Gaussian parameters as CPU float32 tensors.

    channels=4 -> `colors_precomp` (P,4) = [rgb, kp_score] (SplatLoc's mode);
    sh_degree=k -> `shs` (P,(k+1)^2,3) instead, rendered as 3 channels.
    `scale` stretches the room (config 4 uses 1.5)."""
    g = _gen(seed)
    lo = tuple(v * scale for v in ROOM_LO)
    hi = tuple(v * scale for v in ROOM_HI)
    xyz, normal_axis = room_points(P, g, lo, hi)
    log_s = math.log(0.02) + 0.5 * torch.randn((P, 3), generator=g)
    flat = normal_axis >= 0
    idx = torch.nonzero(flat).squeeze(1)
    log_s[idx, normal_axis[idx]] += math.log(0.3)
    scales = torch.exp(log_s)
    q = torch.randn((P, 4), generator=g)
    rotations = q / q.norm(dim=1, keepdim=True)
    opacities = torch.sigmoid(1.0 + 2.0 * torch.randn((P, 1), generator=g))
    rgb = torch.rand((P, 3), generator=g)
    out = dict(means3D=xyz, scales=scales, rotations=rotations, opacities=opacities)
    if sh_degree is None:
        if channels == 4:
            kp = torch.sigmoid(-3.0 + 2.0 * torch.randn((P, 1), generator=g))
            out["colors_precomp"] = torch.cat([rgb, kp], dim=1)
        else:
            out["colors_precomp"] = rgb
    else:
        M = (sh_degree + 1) ** 2
        shs = 0.1 * torch.randn((P, M, 3), generator=g)
        shs[:, 0, :] = (rgb - 0.5) / SH_C0
        out["shs"] = shs
    return {k: v.contiguous() for k, v in out.items()}


def projection_matrix(znear, zfar, cx, cy, fx, fy, W, H) -> torch.Tensor:
    """Off-centre perspective matrix with w = z_cam and NDC z in [0, 1]."""
    left = znear / fx * (((2 * cx - W) / W - 1.0) * W / 2.0)
    right = znear / fx * (((2 * cx - W) / W + 1.0) * W / 2.0)
    top = znear / fy * (((2 * cy - H) / H + 1.0) * H / 2.0)
    bottom = znear / fy * (((2 * cy - H) / H - 1.0) * H / 2.0)
    Pm = torch.zeros(4, 4)
    Pm[0, 0] = 2.0 * znear / (right - left)
    Pm[1, 1] = 2.0 * znear / (top - bottom)
    Pm[0, 2] = (right + left) / (right - left)
    Pm[1, 2] = (top + bottom) / (top - bottom)
    Pm[3, 2] = 1.0
    Pm[2, 2] = zfar / (zfar - znear)
    Pm[2, 3] = -(zfar * znear) / (zfar - znear)
    return Pm


def camera_from_w2c(w2c: torch.Tensor, W, H, fx, fy, cx, cy, znear=0.01, zfar=100.0) -> Camera:
    w2c = w2c.float()
    view = w2c.transpose(0, 1).contiguous()
    proj = projection_matrix(znear, zfar, cx, cy, fx, fy, W, H).transpose(0, 1)
    full = (view.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0).contiguous()
    campos = view.inverse()[3, :3].contiguous()
    return Camera(W, H, W / (2.0 * fx), H / (2.0 * fy), view, full, campos)


def look_at_w2c(eye: torch.Tensor, target: torch.Tensor, roll: float = 0.0) -> torch.Tensor:
    """World-to-camera with +z forward, +x right, +y down (the reference's convention)."""
    eye, target = eye.double(), target.double()
    fwd = target - eye
    fwd = fwd / fwd.norm()
    up_w = torch.tensor([0.0, 0.0, 1.0], dtype=torch.float64)
    if abs(float(fwd @ up_w)) > 0.99:
        up_w = torch.tensor([0.0, 1.0, 0.0], dtype=torch.float64)
    right = torch.linalg.cross(fwd, up_w)
    right = right / right.norm()
    down = torch.linalg.cross(fwd, right)
    c, s = math.cos(roll), math.sin(roll)
    right, down = c * right + s * down, -s * right + c * down
    Rm = torch.stack([right, down, fwd])  # rows: camera axes in world coordinates
    w2c = torch.eye(4, dtype=torch.float64)
    w2c[:3, :3] = Rm
    w2c[:3, 3] = -Rm @ eye
    return w2c


def make_camera(seed: int = 0, intrinsics: dict | None = None, scale: float = 1.0, size: tuple | None = None) -> Camera:
    """Eye uniform in the inner 60 % of the room, looking at a random wall point, |roll| <= 10 deg.
    `size=(W,H)` rescales the intrinsics to another resolution."""
    k = dict(REPLICA_INTRINSICS if intrinsics is None else intrinsics)
    if size is not None:
        sx, sy = size[0] / k["W"], size[1] / k["H"]
        k = dict(W=size[0], H=size[1], fx=k["fx"] * sx, fy=k["fy"] * sy,
                 cx=(k["cx"] + 0.5) * sx - 0.5, cy=(k["cy"] + 0.5) * sy - 0.5)
    g = _gen(1000003 + seed)
    lo = torch.tensor(ROOM_LO, dtype=torch.float64) * scale
    hi = torch.tensor(ROOM_HI, dtype=torch.float64) * scale
    mid, half = (lo + hi) / 2, (hi - lo) / 2
    eye = mid + 0.6 * half * (2 * torch.rand(3, generator=g, dtype=torch.float64) - 1)
    target = lo + (hi - lo) * torch.rand(3, generator=g, dtype=torch.float64)
    axis = int(torch.randint(0, 3, (1,), generator=g))
    side = int(torch.randint(0, 2, (1,), generator=g))
    target[axis] = hi[axis] if side else lo[axis]
    roll = math.radians(10.0) * (2 * float(torch.rand(1, generator=g)) - 1)
    return camera_from_w2c(look_at_w2c(eye, target, roll), **k)


def make_pixel_grads(seed: int, C: int, H: int, W: int, normalised: bool = True):
    """Upstream dL/dcolor (C,H,W), dL/ddepth (1,H,W), dL/dalpha (1,H,W) ~ N(0,1) (/ N)."""
    g = _gen(2000003 + seed)
    s = 1.0 / (H * W) if normalised else 1.0
    return (s * torch.randn((C, H, W), generator=g), s * torch.randn((1, H, W), generator=g),
            s * torch.randn((1, H, W), generator=g))


def make_knn_points(P: int, seed: int = 0, dup_frac: float = 0.001) -> torch.Tensor:
    """Config 5 point cloud: 70 % on the room faces, 30 % in the volume, 0.1 % exact duplicates."""
    g = _gen(3000003 + seed)
    xyz, _ = room_points(P, g, face_frac=0.7)
    n_dup = int(P * dup_frac)
    if n_dup > 0:
        src = torch.randint(0, P, (n_dup,), generator=g)
        dst = torch.randint(0, P, (n_dup,), generator=g)
        xyz[dst] = xyz[src]
    return xyz.contiguous()
