"""`diff_gauss`-compatible front end of the B200 rasterizer.

Keeps the public surface SplatLoc's `gaussian_renderer.render` imports from the
fork (submodules/diff-gaussian-rasterization/diff_gauss/__init__.py):
`GaussianRasterizationSettings` (:188-200, same 12 fields in the same order),
`GaussianRasterizer` (:203-258), `rasterize_gaussians` (:41-64), `mark_visible`
(:261-269), `BLOCK_X/BLOCK_Y` (:17-18), the `means2D.absgrad` side channel (:183)
and the scratch-buffer interpreters (:399-502).  The rasterizer returns the
fork's 4-tuple `(color, depth, alpha, radii)`.

The 4D-Gaussian helpers of the fork (`compute_cov_4d`, `compute_sh_4d`,
`fused_preprocess_4d*`) are not used by SplatLoc and are not provided.
"""
from __future__ import annotations

from typing import NamedTuple

import torch
from torch import nn

from .. import _C

BLOCK_X = 16
BLOCK_Y = 16

_DUMP_FW = "snapshot_fw.dump"
_DUMP_BW = "snapshot_bw.dump"


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def _host_copy(args):
    return tuple(a.detach().cpu().clone() if isinstance(a, torch.Tensor) else a for a in args)


def _call_with_snapshot(fn, args, debug, dump_path, what):
    """In debug mode keep a CPU copy of the inputs and save it if the call raises
    (the fork's snapshot_{fw,bw}.dump behaviour, :108-115,159-166)."""
    if not debug:
        return fn(*args)
    saved = _host_copy(args)
    try:
        return fn(*args)
    except Exception:
        torch.save(saved, dump_path)
        print(f"\nAn error occured in {what}. Inputs written to {dump_path}.")
        raise


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                cov3Ds_precomp, tile_mask, raster_settings, theta=None, rho=None, want_n_touched=False,
                max_pairs=None, status_out=None):
        rs = raster_settings
        fw_args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier,
                   cov3Ds_precomp, tile_mask, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
                   rs.image_height, rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
        res = _call_with_snapshot(_C.rasterize_gaussians, fw_args + (want_n_touched, max_pairs), rs.debug, _DUMP_FW,
                                  "forward")
        (num_rendered, color, depth, alpha, radii, geom, binning, img) = res[:8]
        if max_pairs is not None and status_out is not None:
            status_out.append(res[-1])     # device int32: 1 if the frame did not fit `max_pairs`

        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.pose = (theta is not None, rho is not None)     # which of the two placeholders were given
        ctx.want_pose = ctx.pose[0] or ctx.pose[1]
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
                              geom, binning, img, alpha, means2D)
        ctx.mark_non_differentiable(radii)
        if want_n_touched:
            ctx.mark_non_differentiable(res[8])
            return color, depth, alpha, radii, res[8]
        return color, depth, alpha, radii

    @staticmethod
    def backward(ctx, grad_color, grad_depth, grad_alpha, _grad_radii, _grad_n_touched=None):
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning, img,
         alpha, means2D) = ctx.saved_tensors
        bw_args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier,
                   cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
                   grad_color, grad_depth, grad_alpha, sh, rs.sh_degree, rs.campos,
                   geom, ctx.num_rendered, binning, img, alpha, rs.debug)
        res = _call_with_snapshot(_C.rasterize_gaussians_backward, bw_args + (ctx.want_pose,), rs.debug, _DUMP_BW,
                                  "backward")
        (g_abs2D, g_means2D, g_colors, g_opac, g_means3D, g_cov3D, g_sh, g_scales, g_rot) = res[:9]
        g_theta = g_rho = None
        if ctx.want_pose:   # extension: camera-pose gradient, (rho, theta) order in the C ABI
            g_rho = res[9][:3] if ctx.pose[1] else None       # an input that was None must get None back
            g_theta = res[9][3:] if ctx.pose[0] else None

        # densification statistics read |dL/dmean2D| from this attribute
        means2D.absgrad = g_abs2D
        # one gradient per forward input; tile_mask and the settings have none
        return (g_means3D, g_means2D, g_sh, g_colors, g_opac, g_scales, g_rot, g_cov3D, None, None,
                g_theta, g_rho, None, None, None)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                        cov3Ds_precomp, tile_mask, raster_settings, theta=None, rho=None, want_n_touched=False,
                        max_pairs=None, status_out=None):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, tile_mask, raster_settings, theta, rho, want_n_touched,
                                     max_pairs, status_out)


def mark_visible(positions: torch.Tensor, viewmatrix: torch.Tensor, projmatrix: torch.Tensor):
    with torch.no_grad():
        return _C.mark_visible(positions, viewmatrix, projmatrix)


class GaussianRasterizer(nn.Module):
    """The fork's module.  Two opt-in extensions the reference does not have (SURVEY.md section 0,
    D1/D2) leave the default call unchanged:
      * `theta=` / `rho=` (3-vectors, MonoGS's `cam_rot_delta` / `cam_trans_delta`): placeholders whose
        values are taken as zero; after backward their `.grad` holds dL/dtheta, dL/drho for the left
        perturbation W2C' = Exp(rho, theta) W2C of `viewmatrix` (and of `projmatrix`, `campos`);
      * `return_n_touched=True` at construction: forward returns a 5th tensor, n_touched (P,) int32.
    A third option changes how, not what: `max_pairs=N` (sl_forward_nosync) sizes the binning scratch
    for N (tile, Gaussian) pairs up front instead of reading the pair count back from the device, so
    the call never blocks the host and a forward + backward step can be captured in a CUDA graph
    (`splatloc_b200.graphs`).  After each forward `self.overflow` is a (1,) int32 device tensor: 1 if
    the frame needed more than N pairs (or a tile more than 4096), in which case the outputs are
    background-only and the gradients zero -- check it when convenient and redo that frame."""

    def __init__(self, raster_settings: GaussianRasterizationSettings, return_n_touched: bool = False,
                 max_pairs: int | None = None):
        super().__init__()
        self.raster_settings = raster_settings
        self.return_n_touched = return_n_touched
        self.max_pairs = max_pairs
        self.overflow = None

    def markVisible(self, positions):
        rs = self.raster_settings
        return mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None, tile_mask=None, *, theta=None, rho=None):
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        has_sr = scales is not None or rotations is not None
        if (cov3D_precomp is None and (scales is None or rotations is None)) or \
                (cov3D_precomp is not None and has_sr):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')

        # absent optional inputs travel as empty tensors (-> NULL pointers in the C ABI)
        none = torch.Tensor([])
        shs = none if shs is None else shs
        colors_precomp = none if colors_precomp is None else colors_precomp
        scales = none if scales is None else scales
        rotations = none if rotations is None else rotations
        cov3D_precomp = none if cov3D_precomp is None else cov3D_precomp
        tile_mask = none.bool() if tile_mask is None else tile_mask
        if tile_mask.dtype != torch.bool:
            tile_mask = tile_mask.bool()

        status = []
        out = rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                  cov3D_precomp, tile_mask, self.raster_settings, theta, rho,
                                  self.return_n_touched, self.max_pairs, status)
        self.overflow = status[0] if status else None
        return out


# ------------------------------------------------------------------ several cameras per call
class _RasterizeGaussiansBatch(torch.autograd.Function):
    """V cameras of one image size in one call (sl_forward_batch / sl_backward_batch).  No counterpart
    in the reference, whose callers loop over the cameras of a window (train_gaussians.py:195-229)."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                tile_mask, raster_settings, theta, rho, want_n_touched, max_pairs, status_out, pose_only):
        rs = raster_settings
        res = _C.rasterize_gaussians_batch(
            rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
            tile_mask, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, sh,
            rs.sh_degree, rs.campos, rs.prefiltered, rs.debug, want_n_touched, max_pairs)
        (num_rendered, color, depth, alpha, radii, geom, binning, img) = res[:8]
        if max_pairs is not None and status_out is not None:
            status_out.append(res[-1])
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.pose = (theta is not None, rho is not None)
        ctx.pose_only = pose_only
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
                              geom, binning, img, alpha, means2D)
        ctx.mark_non_differentiable(radii)
        if want_n_touched:
            ctx.mark_non_differentiable(res[8])
            return color, depth, alpha, radii, res[8]
        return color, depth, alpha, radii

    @staticmethod
    def backward(ctx, grad_color, grad_depth, grad_alpha, _grad_radii, _grad_n_touched=None):
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning, img,
         alpha, means2D) = ctx.saved_tensors
        want_pose = ctx.pose[0] or ctx.pose[1]
        res = _C.rasterize_gaussians_backward_batch(
            rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
            rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_color, grad_depth, grad_alpha, sh,
            rs.sh_degree, rs.campos, geom, ctx.num_rendered, binning, img, alpha, rs.debug, want_pose, ctx.pose_only)
        (g_abs2D, g_means2D, g_colors, g_opac, g_means3D, g_cov3D, g_sh, g_scales, g_rot) = res[:9]
        g_theta = g_rho = None
        if want_pose:
            g_rho = res[9][:, :3] if ctx.pose[1] else None
            g_theta = res[9][:, 3:] if ctx.pose[0] else None
        means2D.absgrad = g_abs2D
        if ctx.pose_only:
            g_means2D = g_colors = g_opac = None
        elif sh.numel() != 0:
            g_colors = None       # per-view SH colour gradients are an intermediate, not an input's gradient
        return (g_means3D, g_means2D, g_sh, g_colors, g_opac, g_scales, g_rot, g_cov3D, None, None,
                g_theta, g_rho, None, None, None, None)


class BatchedGaussianRasterizer(nn.Module):
    """`GaussianRasterizer` for V cameras at once.  `raster_settings.viewmatrix` / `.projmatrix` are
    (V, 4, 4) and `.campos` (V, 3) (same conventions per camera); image size, field of view, background
    and SH degree are shared.  forward() returns color (V, C, H, W), depth (V, 1, H, W), alpha (V, 1, H, W),
    radii (V, P) [, n_touched (V, P)]; each view equals what `GaussianRasterizer` returns for that camera,
    bit for bit.  After backward the Gaussians' `.grad` hold the SUM over the views (one store per value,
    no atomics across views), `means2D` -- a (V, P, 3) placeholder -- holds the per-view screen-space
    gradients and `means2D.absgrad` their absolute counterparts, and `theta` / `rho` (V, 3) the per-view
    camera-pose gradients.  `pose_only=True`: the map is fixed, only theta / rho receive gradients."""

    def __init__(self, raster_settings: GaussianRasterizationSettings, return_n_touched: bool = False,
                 max_pairs: int | None = None, pose_only: bool = False):
        super().__init__()
        self.raster_settings = raster_settings
        self.return_n_touched = return_n_touched
        self.max_pairs = max_pairs
        self.pose_only = pose_only
        self.overflow = None

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None, tile_mask=None, *, theta=None, rho=None):
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        has_sr = scales is not None or rotations is not None
        if (cov3D_precomp is None and (scales is None or rotations is None)) or \
                (cov3D_precomp is not None and has_sr):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        if self.pose_only and theta is None and rho is None:
            raise Exception('pose_only needs theta and/or rho')
        none = torch.Tensor([])
        shs = none if shs is None else shs
        colors_precomp = none if colors_precomp is None else colors_precomp
        scales = none if scales is None else scales
        rotations = none if rotations is None else rotations
        cov3D_precomp = none if cov3D_precomp is None else cov3D_precomp
        tile_mask = none.bool() if tile_mask is None else tile_mask.bool()
        status = []
        out = _RasterizeGaussiansBatch.apply(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                             cov3D_precomp, tile_mask, self.raster_settings, theta, rho,
                                             self.return_n_touched, self.max_pairs, status, self.pose_only)
        self.overflow = status[0] if status else None
        return out


# ------------------------------------------------------------------ scratch-buffer views
def _aligned_fields(buf, spec):
    """Walk a chunk laid out as consecutive 128-byte-aligned arrays
    (csrc/state.cuh; same rule as rasterizer_impl.h:22-31)."""
    base = buf.data_ptr()
    cur = 0
    out = {}
    for name, dtype, count, shape in spec:
        cur += (-(base + cur)) % 128
        nbytes = count * torch.empty(0, dtype=dtype).element_size()
        view = buf[cur:cur + nbytes].view(dtype)
        out[name] = view.view(*shape) if shape else view
        cur += nbytes
    return out


def interpret_geomBuffer(geomBuffer: torch.Tensor, N: int):
    """Typed views of the geometry chunk for N Gaussians."""
    return _aligned_fields(geomBuffer, [
        ("depths", torch.float32, N, None),
        ("clamped", torch.bool, 3 * N, (N, 3)),
        ("internal_radii", torch.int32, N, None),
        ("means2D", torch.float32, 2 * N, (N, 2)),
        ("cov3D", torch.float32, 6 * N, (N, 6)),
        ("conic_opacity", torch.float32, 4 * N, (N, 4)),
        ("rgb", torch.float32, 3 * N, (N, 3)),
        ("tiles_touched", torch.int32, N, None),
        ("point_offsets", torch.int32, N, None),
    ])


def interpret_binningBuffer(binningBuffer: torch.Tensor, N: int):
    """Typed views of the binning chunk for N (tile, Gaussian) pairs."""
    d = _aligned_fields(binningBuffer, [
        ("point_list", torch.int32, N, None),
        ("point_list_unsorted", torch.int32, N, None),
        ("point_list_keys", torch.int64, N, None),
        ("point_list_keys_unsorted", torch.int64, N, None),
    ])
    halves = d["point_list_keys"].view(torch.int32).view(N, 2)  # little endian: depth bits, tile id
    d["depths"] = d["point_list_keys"].view(torch.float32).view(N, 2)[:, 0]
    d["tile_ids"] = halves[:, 1]
    return d


def interpret_imgBuffer(imgBuffer: torch.Tensor, N: int, M: int):
    """Typed views of the image chunk for N pixels and M tiles."""
    return _aligned_fields(imgBuffer, [
        ("n_contrib", torch.int32, N, None),
        ("ranges", torch.int32, 2 * M, (M, 2)),
    ])
