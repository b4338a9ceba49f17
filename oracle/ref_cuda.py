"""TEST INFRASTRUCTURE -- ctypes front end of the UNMODIFIED reference CUDA code.

`oracle/_ref/libref_raster_c{3,4}.so` and `libref_knn.so` are the reference's own
translation units (forward.cu, backward.cu, rasterizer_impl.cu, simple_knn.cu)
compiled in place from /root/reference for sm_100a by `oracle/build_ref.sh`,
behind the thin C shim `oracle/ref_shim.cu`.  They are the golden for the GPU
parity tests, the source of `tests/golden/`, and bench.py's reference arm.
Only `tests/`, `__graft_entry__.smoke()` and `bench.py` may import this module;
nothing in `splatloc_b200/` does.

The functions mirror the reference's pybind entry points
(diff-gaussian-rasterization/rasterize_points.cu:35-234, simple-knn/spatial.cu:15-25)
including their allocation pattern (outputs pre-zeroed with torch.full / zeros).
The reference launches on the legacy default stream, so callers must stay on
torch's default stream.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF = os.path.join(_HERE, "_ref")
ALLOC_FN = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_size_t)
_P = C.c_void_p

_libs: dict[str, C.CDLL] = {}


def available(name: str = "libref_raster_c4.so") -> bool:
    return os.path.exists(os.path.join(_REF, name))


def _load(name: str) -> C.CDLL:
    if name not in _libs:
        path = os.path.join(_REF, name)
        if not os.path.exists(path):
            raise RuntimeError(f"{path} missing: run oracle/build_ref.sh where /root/reference exists")
        lib = C.CDLL(path)
        if "raster" in name:
            lib.ref_num_channels.restype = C.c_int
            lib.ref_forward.restype = C.c_int
            lib.ref_forward.argtypes = ([ALLOC_FN, _P] * 3 + [C.c_int] * 3 + [_P, C.c_int, C.c_int] +
                                        [_P] * 5 + [C.c_float] + [_P] * 6 + [C.c_float, C.c_float, C.c_int] +
                                        [_P] * 4 + [C.c_int])
            lib.ref_backward.restype = C.c_int
            lib.ref_backward.argtypes = ([C.c_int] * 4 + [_P, C.c_int, C.c_int] + [_P] * 4 + [C.c_float] +
                                         [_P] * 5 + [C.c_float, C.c_float] + [_P] * 19 + [C.c_int])
            lib.ref_mark_visible.restype = None
            lib.ref_mark_visible.argtypes = [C.c_int, _P, _P, _P, _P]
        else:
            lib.ref_knn.restype = None
            lib.ref_knn.argtypes = [C.c_int, _P, _P]
        _libs[name] = lib
    return _libs[name]


def _raster_lib(channels: int) -> C.CDLL:
    lib = _load(f"libref_raster_c{channels}.so")
    assert lib.ref_num_channels() == channels
    return lib


def _ptr(t):
    return None if t is None or t.numel() == 0 else t.data_ptr()


class _Chunk:
    def __init__(self, device):
        self.tensor = torch.empty(0, dtype=torch.uint8, device=device)

        def _alloc(_ctx, n):
            self.tensor = torch.empty(int(n), dtype=torch.uint8, device=self.tensor.device)
            return self.tensor.data_ptr()

        self.cb = ALLOC_FN(_alloc)


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                        cov3D_precomp, tile_mask, viewmatrix, projmatrix, tan_fovx, tan_fovy,
                        image_height, image_width, sh, degree, campos, prefiltered, debug, channels=None):
    """Reference RasterizeGaussiansCUDA; `channels` picks the NUM_CHANNELS build."""
    if channels is None:
        channels = int(colors.shape[1]) if colors.numel() else 3
    lib = _raster_lib(channels)
    dev = means3D.device
    P, H, W = means3D.shape[0], image_height, image_width
    out_color = torch.full((channels, H, W), 0.0, device=dev)
    out_depth = torch.full((1, H, W), 0.0, device=dev)
    out_alpha = torch.full((1, H, W), 0.0, device=dev)
    radii = torch.full((P,), 0, dtype=torch.int32, device=dev)
    geom, binning, img = _Chunk(dev), _Chunk(dev), _Chunk(dev)
    rendered = 0
    if P != 0:
        M = sh.shape[1] if sh.numel() else 0
        tm = tile_mask.contiguous() if tile_mask is not None and tile_mask.numel() else None
        rendered = lib.ref_forward(
            geom.cb, None, binning.cb, None, img.cb, None, P, degree, M,
            _ptr(background.contiguous()), W, H, _ptr(means3D.contiguous()), _ptr(sh.contiguous()),
            _ptr(colors.contiguous()), _ptr(opacity.contiguous()), _ptr(scales.contiguous()),
            float(scale_modifier), _ptr(rotations.contiguous()), _ptr(cov3D_precomp.contiguous()), _ptr(tm),
            _ptr(viewmatrix.contiguous()), _ptr(projmatrix.contiguous()), _ptr(campos.contiguous()),
            float(tan_fovx), float(tan_fovy), int(prefiltered),
            _ptr(out_color), _ptr(out_depth), _ptr(out_alpha), _ptr(radii), int(debug))
        if rendered < 0:
            raise RuntimeError("reference forward raised")
    return rendered, out_color, out_depth, out_alpha, radii, geom.tensor, binning.tensor, img.tensor


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                 cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy,
                                 dL_dout_color, dL_dout_depth, dL_dout_alpha, sh, degree, campos,
                                 geomBuffer, R, binningBuffer, imageBuffer, out_alpha, debug):
    """Reference RasterizeGaussiansBackwardCUDA."""
    channels = dL_dout_color.shape[0]
    lib = _raster_lib(channels)
    dev = means3D.device
    P = means3D.shape[0]
    H, W = dL_dout_color.shape[1], dL_dout_color.shape[2]
    M = sh.shape[1] if sh.numel() else 0
    z = lambda *s: torch.zeros(s, device=dev)  # noqa: E731
    g_means3D, g_means2D, g_abs2D = z(P, 3), z(P, 3), z(P, 3)
    g_colors, g_depths, g_conic = z(P, channels), z(P, 1), z(P, 2, 2)
    g_opacity, g_cov3D, g_sh = z(P, 1), z(P, 6), z(P, M, 3)
    g_scales, g_rot = z(P, 3), z(P, 4)
    if P != 0:
        rc = lib.ref_backward(
            P, degree, M, R, _ptr(background.contiguous()), W, H, _ptr(means3D.contiguous()),
            _ptr(sh.contiguous()), _ptr(colors.contiguous()), _ptr(scales.contiguous()), float(scale_modifier),
            _ptr(rotations.contiguous()), _ptr(cov3D_precomp.contiguous()), _ptr(viewmatrix.contiguous()),
            _ptr(projmatrix.contiguous()), _ptr(campos.contiguous()), float(tan_fovx), float(tan_fovy),
            _ptr(radii.contiguous()), _ptr(geomBuffer), _ptr(binningBuffer), _ptr(imageBuffer),
            _ptr(out_alpha.contiguous()), _ptr(dL_dout_color.contiguous()), _ptr(dL_dout_depth.contiguous()),
            _ptr(dL_dout_alpha.contiguous()), g_means2D.data_ptr(), g_abs2D.data_ptr(), g_conic.data_ptr(),
            g_opacity.data_ptr(), g_colors.data_ptr(), g_depths.data_ptr(), g_means3D.data_ptr(),
            g_cov3D.data_ptr(), _ptr(g_sh), g_scales.data_ptr(), g_rot.data_ptr(), int(debug))
        if rc < 0:
            raise RuntimeError("reference backward raised")
    return g_abs2D, g_means2D, g_colors, g_opacity, g_means3D, g_cov3D, g_sh, g_scales, g_rot


def mark_visible(means3D, viewmatrix, projmatrix):
    lib = _raster_lib(4)
    P = means3D.shape[0]
    present = torch.empty((P,), dtype=torch.bool, device=means3D.device)
    if P != 0:
        lib.ref_mark_visible(P, _ptr(means3D.contiguous()), _ptr(viewmatrix.contiguous()),
                             _ptr(projmatrix.contiguous()), present.data_ptr())
    return present


def distCUDA2(points):
    lib = _load("libref_knn.so")
    P = points.shape[0]
    means = torch.full((P,), 0.0, device=points.device)
    lib.ref_knn(P, _ptr(points.contiguous()), means.data_ptr())
    return means
