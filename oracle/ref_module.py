"""TEST INFRASTRUCTURE -- autograd front end over the unmodified reference CUDA library.

The reference's own Python package (`diff_gauss/__init__.py`) imports a pybind
module that is not built here (its two torch-binding translation units take
minutes each and add nothing to the arithmetic); this module restates its
autograd wiring (`_RasterizeGaussians`, :67-185; `GaussianRasterizer.forward`,
:219-258) on top of `oracle/ref_cuda.py` so that bench.py's reference arm goes
through the same kind of API call as the product arm.
"""
from typing import NamedTuple

import torch
from torch import nn

from . import ref_cuda as _C


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


class _Rasterize(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors, opacities, scales, rotations, cov3D, tile_mask, rs):
        R, color, depth, alpha, radii, geom, binning, img = _C.rasterize_gaussians(
            rs.bg, means3D, colors, opacities, scales, rotations, rs.scale_modifier, cov3D, tile_mask,
            rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, sh,
            rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
        ctx.rs, ctx.R = rs, R
        ctx.save_for_backward(colors, means3D, scales, rotations, cov3D, radii, sh, geom, binning, img, alpha, means2D)
        ctx.mark_non_differentiable(radii)
        return color, depth, alpha, radii

    @staticmethod
    def backward(ctx, g_color, g_depth, g_alpha, _):
        rs = ctx.rs
        colors, means3D, scales, rotations, cov3D, radii, sh, geom, binning, img, alpha, means2D = ctx.saved_tensors
        g_abs, g_m2, g_col, g_op, g_m3, g_cov, g_sh, g_sc, g_rot = _C.rasterize_gaussians_backward(
            rs.bg, means3D, radii, colors, scales, rotations, rs.scale_modifier, cov3D, rs.viewmatrix,
            rs.projmatrix, rs.tanfovx, rs.tanfovy, g_color, g_depth, g_alpha, sh, rs.sh_degree, rs.campos,
            geom, ctx.R, binning, img, alpha, rs.debug)
        means2D.absgrad = g_abs
        return g_m3, g_m2, g_sh, g_col, g_op, g_sc, g_rot, g_cov, None, None


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None, tile_mask=None):
        e = torch.Tensor([])
        return _Rasterize.apply(
            means3D, means2D, e if shs is None else shs, e if colors_precomp is None else colors_precomp,
            opacities, e if scales is None else scales, e if rotations is None else rotations,
            e if cov3D_precomp is None else cov3D_precomp, e.bool() if tile_mask is None else tile_mask.bool(),
            self.raster_settings)
