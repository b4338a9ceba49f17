"""`render()` of SplatLoc with its pre- and post-processing folded into the rasterizer.

Mirrors `gaussian_splatting/gaussian_renderer/__init__.py:12-141` of the reference: same signature,
same returned dict (`render`, `kp_prob`, `viewspace_points`, `visibility_filter`, `radii`, `depth`,
`opacity`).  The reference builds the rasterizer's inputs with PyTorch ops on every call --

    opacity  = sigmoid(pc._opacity)                    scene/gaussian_model.py:103-105
    scales   = exp(pc._scaling) (.repeat(1, 3))        :75-77, gaussian_renderer/__init__.py:73-76
    rotation = normalize(pc._rotation)                 :79-81
    colours  = clamp_min(eval_sh(deg, f, dir) + 0.5, 0)   gaussian_renderer/__init__.py:85-90, utils/sh_utils.py:55-118
    features = cat(colours, pc._kp_score)              :98-102

-- four (P, .) tensors plus their autograd nodes per view.  In the configuration SplatLoc trains with
(SH degree 0, `convert_SHs_python`, covariance from scale / rotation: configs/*/base_config.yaml:88-97)
all of that runs inside the kernels here (`sl_forward_render` / `sl_backward_render`): the raw
parameters go in, gradients of the raw parameters come out.  Other configurations (higher SH degree,
`compute_cov3D_python`, `mask`, `override_color`) take the same steps as the reference, as PyTorch ops
in front of this library's `GaussianRasterizer`.

To use it from the reference tree, replace the body of its `render` by a call to this one.
"""
from __future__ import annotations

import math

import torch

from . import _C
from .diff_gauss import GaussianRasterizationSettings, GaussianRasterizer

_EMPTY = torch.Tensor([])

C0 = 0.28209479177387814
C1 = 0.4886025119029199
C2 = (1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396)
C3 = (-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
      1.445305721320277, -0.5900435899266435)


def eval_sh(deg: int, sh: torch.Tensor, dirs: torch.Tensor) -> torch.Tensor:
    """Real spherical harmonics up to degree 3: sh (..., C, (deg+1)^2), dirs (..., 3) unit -> (..., C).
    Same basis and term order as the reference's `eval_sh` (utils/sh_utils.py:55-118)."""
    assert 0 <= deg <= 3 and sh.shape[-1] >= (deg + 1) ** 2
    result = C0 * sh[..., 0]
    if deg > 0:
        x, y, z = dirs[..., 0:1], dirs[..., 1:2], dirs[..., 2:3]
        result = result - C1 * y * sh[..., 1] + C1 * z * sh[..., 2] - C1 * x * sh[..., 3]
        if deg > 1:
            xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
            result = (result + C2[0] * xy * sh[..., 4] + C2[1] * yz * sh[..., 5] + C2[2] * (2.0 * zz - xx - yy) * sh[..., 6]
                      + C2[3] * xz * sh[..., 7] + C2[4] * (xx - yy) * sh[..., 8])
            if deg > 2:
                result = (result + C3[0] * y * (3 * xx - yy) * sh[..., 9] + C3[1] * xy * z * sh[..., 10]
                          + C3[2] * y * (4 * zz - xx - yy) * sh[..., 11] + C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[..., 12]
                          + C3[4] * x * (4 * zz - xx - yy) * sh[..., 13] + C3[5] * z * (xx - yy) * sh[..., 14]
                          + C3[6] * x * (xx - 3 * yy) * sh[..., 15])
    return result


class _FusedRender(torch.autograd.Function):
    """One view through sl_forward_render / sl_backward_render."""

    @staticmethod
    def forward(ctx, xyz, screenspace_points, features_dc, kp_score, raw_opacity, raw_scaling, raw_rotation,
                bg, scale_modifier, viewmatrix, projmatrix, campos, tanfovx, tanfovy, H, W):
        res = _C.render_forward(bg, xyz, features_dc, kp_score, raw_opacity, raw_scaling, raw_rotation, scale_modifier,
                                None, viewmatrix[None], projmatrix[None], tanfovx, tanfovy, H, W, campos[None])
        R, color, depth, alpha, radii, geom, binning, img = res
        ctx.meta = (bg, scale_modifier, viewmatrix, projmatrix, campos, tanfovx, tanfovy, R)
        ctx.save_for_backward(xyz, features_dc, kp_score, raw_opacity, raw_scaling, raw_rotation, radii, geom, binning,
                              img, alpha, screenspace_points)
        ctx.mark_non_differentiable(radii)
        return color[0], depth[0], alpha[0], radii[0]

    @staticmethod
    def backward(ctx, g_color, g_depth, g_alpha, _g_radii):
        bg, scale_modifier, viewmatrix, projmatrix, campos, tanfovx, tanfovy, R = ctx.meta
        (xyz, features_dc, kp_score, raw_opacity, raw_scaling, raw_rotation, radii, geom, binning, img, alpha,
         screenspace_points) = ctx.saved_tensors
        (g_abs2D, g_means2D, g_xyz, g_fdc, g_kp, g_opac, g_scale, g_rot) = _C.render_backward(
            bg, xyz, radii, features_dc, kp_score, raw_opacity, raw_scaling, raw_rotation, scale_modifier,
            viewmatrix[None], projmatrix[None], tanfovx, tanfovy, g_color[None], g_depth[None], g_alpha[None],
            campos[None], geom, R, binning, img, alpha)
        screenspace_points.absgrad = g_abs2D[0]     # densification statistics read |dL/dmean2D| from here
        return (g_xyz, g_means2D[0], g_fdc.view(features_dc.shape), g_kp.view(kp_score.shape) if kp_score.numel() else None,
                g_opac.view(raw_opacity.shape), g_scale.view(raw_scaling.shape), g_rot, None, None, None, None, None,
                None, None, None, None)


def _can_fuse(pc, pipe, override_color, mask) -> bool:
    return (mask is None and override_color is None and getattr(pipe, "convert_SHs_python", False)
            and not getattr(pipe, "compute_cov3D_python", False) and pc.active_sh_degree == 0
            and pc._features_dc.numel() == 3 * pc._xyz.shape[0])


def render(viewpoint_camera, pc, pipe, bg_color: torch.Tensor, scaling_modifier=1.0, override_color=None, mask=None,
           fuse: bool = True):
    """Render the scene (reference: gaussian_renderer/__init__.py:12-141).  `pc` is a GaussianModel (any
    object with `_xyz`, `_features_dc`, `_features_rest`, `_kp_score`, `_opacity`, `_scaling`, `_rotation`,
    `active_sh_degree`, `max_sh_degree` and the reference's getters).  `fuse=False` (not in the
    reference) keeps the pre-processing as PyTorch ops even where the kernels could do it."""
    if pc._xyz.shape[0] == 0:
        return None
    xyz = pc._xyz
    screenspace_points = torch.zeros_like(xyz, dtype=xyz.dtype, requires_grad=True, device=xyz.device) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    tanfovx = math.tan(viewpoint_camera.FoVx * 0.5)
    tanfovy = math.tan(viewpoint_camera.FoVy * 0.5)
    H, W = int(viewpoint_camera.image_height), int(viewpoint_camera.image_width)

    if fuse and _can_fuse(pc, pipe, override_color, mask):
        rendered_image, depth, opacity, radii = _FusedRender.apply(
            xyz, screenspace_points, pc._features_dc, pc._kp_score, pc._opacity, pc._scaling, pc._rotation, bg_color,
            float(scaling_modifier), viewpoint_camera.world_view_transform, viewpoint_camera.full_proj_transform,
            viewpoint_camera.camera_center, tanfovx, tanfovy, H, W)
    else:
        rs = GaussianRasterizationSettings(
            image_height=H, image_width=W, tanfovx=tanfovx, tanfovy=tanfovy, bg=bg_color, scale_modifier=scaling_modifier,
            viewmatrix=viewpoint_camera.world_view_transform, projmatrix=viewpoint_camera.full_proj_transform,
            sh_degree=pc.active_sh_degree, campos=viewpoint_camera.camera_center, prefiltered=False, debug=False)
        rasterizer = GaussianRasterizer(raster_settings=rs)
        opacity_in = torch.sigmoid(pc._opacity)
        scales = rotations = cov3D_precomp = None
        if getattr(pipe, "compute_cov3D_python", False):
            cov3D_precomp = pc.get_covariance(scaling_modifier)
        else:
            scales = torch.exp(pc._scaling)
            if scales.shape[-1] == 1:
                scales = scales.repeat(1, 3)
            rotations = torch.nn.functional.normalize(pc._rotation)
        shs = colors_precomp = None
        if override_color is not None:
            colors_precomp = override_color
        elif getattr(pipe, "convert_SHs_python", False):
            feats = torch.cat((pc._features_dc, pc._features_rest), dim=1)
            shs_view = feats.transpose(1, 2).view(-1, 3, (pc.max_sh_degree + 1) ** 2)
            dir_pp = xyz - viewpoint_camera.camera_center.repeat(feats.shape[0], 1)
            dir_pp = dir_pp / dir_pp.norm(dim=1, keepdim=True)
            colors_precomp = torch.clamp_min(eval_sh(pc.active_sh_degree, shs_view, dir_pp) + 0.5, 0.0)
        else:
            shs = torch.cat((pc._features_dc, pc._features_rest), dim=1)
        features = torch.cat((colors_precomp, pc._kp_score), dim=1) if colors_precomp is not None else None
        sel = (lambda t: t) if mask is None else (lambda t: None if t is None else t[mask])
        rendered_image, depth, opacity, radii = rasterizer(
            means3D=sel(xyz), means2D=sel(screenspace_points), shs=sel(shs), colors_precomp=sel(features),
            opacities=sel(opacity_in), scales=sel(scales), rotations=sel(rotations), cov3D_precomp=sel(cov3D_precomp))
    return {
        "render": rendered_image[:3, :, :],
        "kp_prob": rendered_image[-1, :, :],
        "viewspace_points": screenspace_points,
        "visibility_filter": radii > 0,
        "radii": radii,
        "depth": depth,
        "opacity": opacity,
    }
