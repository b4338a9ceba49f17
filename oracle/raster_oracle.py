"""TEST INFRASTRUCTURE -- CPU restatement (PyTorch, float64 by default) of the reference
rasterizer's algorithm.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
reference-fallback legs may import this; the product path never does.

Parity status: PINNED against outputs of the reference's own CUDA code, run on a B200 from
oracle/_ref and committed as tests/golden/*.npz by tests/golden/make_golden.py (the reference
ships no tests, golden vectors or CPU implementation of this path -- SURVEY.md 8(c)).  Because
this restatement computes in float64 and the reference in float32 with hard thresholds
(alpha < 1/255, T < 1e-4, power > 0, frustum planes, det <= 0, ceil of the radius), a small
budgeted fraction of elements may land on the other side of a decision; tests state the budget.
The explicit backward below is additionally checked against autograd through the forward.

Each function cites the reference code it follows (paths relative to
submodules/diff-gaussian-rasterization/cuda_rasterizer/).
"""
from __future__ import annotations

import math
import time

import torch

TILE = 16
SH_C0 = 0.28209479177387814
SH_C1 = 0.4886025119029199
SH_C2 = (1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396)
SH_C3 = (-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154,
         -0.4570457994644658, 1.445305721320277, -0.5900435899266435)


# ----------------------------------------------------------------------------- per-Gaussian forward
def _xform43(p, m):
    """auxiliary.h:62-70 -- m is the flat 16-vector of the (transposed) 4x4."""
    x, y, z = p.unbind(-1)
    return torch.stack([m[0] * x + m[4] * y + m[8] * z + m[12],
                        m[1] * x + m[5] * y + m[9] * z + m[13],
                        m[2] * x + m[6] * y + m[10] * z + m[14]], -1)


def _xform44(p, m):
    """auxiliary.h:72-81."""
    x, y, z = p.unbind(-1)
    return torch.stack([m[0] * x + m[4] * y + m[8] * z + m[12],
                        m[1] * x + m[5] * y + m[9] * z + m[13],
                        m[2] * x + m[6] * y + m[10] * z + m[14],
                        m[3] * x + m[7] * y + m[11] * z + m[15]], -1)


def cov3d(scales, rotations, mod=1.0):
    """forward.cu:394-428 (computeCov3D): Sigma = (S R)^T (S R) in GLM's column-major notation,
    quaternion used as given (not normalised).  Returns (P,6)."""
    r, x, y, z = rotations.unbind(-1)
    # GLM fills columns: R_glm[col][row]; as a conventional row-major matrix this is Rm[row][col].
    Rm = torch.stack([
        torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y + r * z), 2 * (x * z - r * y)], -1),
        torch.stack([2 * (x * y - r * z), 1 - 2 * (x * x + z * z), 2 * (y * z + r * x)], -1),
        torch.stack([2 * (x * z + r * y), 2 * (y * z - r * x), 1 - 2 * (x * x + y * y)], -1)], -2)
    S = torch.diag_embed(mod * scales)
    M = S @ Rm                       # GLM "S * R"
    Sigma = M.transpose(-1, -2) @ M
    return torch.stack([Sigma[:, 0, 0], Sigma[:, 0, 1], Sigma[:, 0, 2], Sigma[:, 1, 1], Sigma[:, 1, 2],
                        Sigma[:, 2, 2]], -1)


def cov2d(means3D, cov3D, view, focal_x, focal_y, tan_fovx, tan_fovy):
    """forward.cu:347-389 (computeCov2D): EWA projection + 0.3 px low-pass.  Returns (a, b, c)."""
    t = _xform43(means3D, view)
    limx, limy = 1.3 * tan_fovx, 1.3 * tan_fovy
    tz = t[:, 2]
    tx = torch.clamp(t[:, 0] / tz, -limx, limx) * tz
    ty = torch.clamp(t[:, 1] / tz, -limy, limy) * tz
    zero = torch.zeros_like(tz)
    # conventional Jacobian rows
    J = torch.stack([torch.stack([focal_x / tz, zero, -(focal_x * tx) / (tz * tz)], -1),
                     torch.stack([zero, focal_y / tz, -(focal_y * ty) / (tz * tz)], -1)], -2)   # (P,2,3)
    Wr = torch.stack([torch.stack([view[0], view[4], view[8]]),
                      torch.stack([view[1], view[5], view[9]]),
                      torch.stack([view[2], view[6], view[10]])])                             # world->cam rotation
    Vrk = torch.stack([torch.stack([cov3D[:, 0], cov3D[:, 1], cov3D[:, 2]], -1),
                       torch.stack([cov3D[:, 1], cov3D[:, 3], cov3D[:, 4]], -1),
                       torch.stack([cov3D[:, 2], cov3D[:, 4], cov3D[:, 5]], -1)], -2)
    A = J @ Wr                                # (P,2,3)
    cov = A @ Vrk @ A.transpose(-1, -2)       # (P,2,2)
    return cov[:, 0, 0] + 0.3, cov[:, 0, 1], cov[:, 1, 1] + 0.3


def sh_basis(deg, d):
    """forward.cu:20-71 basis functions for unit directions d (P,3) -> (P,(deg+1)^2)."""
    x, y, z = d.unbind(-1)
    b = [torch.full_like(x, SH_C0)]
    if deg > 0:
        b += [-SH_C1 * y, SH_C1 * z, -SH_C1 * x]
    if deg > 1:
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        b += [SH_C2[0] * xy, SH_C2[1] * yz, SH_C2[2] * (2 * zz - xx - yy), SH_C2[3] * xz, SH_C2[4] * (xx - yy)]
    if deg > 2:
        b += [SH_C3[0] * y * (3 * xx - yy), SH_C3[1] * xy * z, SH_C3[2] * y * (4 * zz - xx - yy),
              SH_C3[3] * z * (2 * zz - 3 * xx - 3 * yy), SH_C3[4] * x * (4 * zz - xx - yy),
              SH_C3[5] * z * (xx - yy), SH_C3[6] * x * (xx - 3 * yy)]
    return torch.stack(b, -1)


def sh_to_rgb(deg, shs, means3D, campos):
    """forward.cu:20-71 (computeColorFromSH): returns (rgb clamped at 0, clamped flags)."""
    d = means3D - campos
    d = d / d.norm(dim=-1, keepdim=True)
    b = sh_basis(deg, d)
    res = (b.unsqueeze(-1) * shs[:, : b.shape[1], :]).sum(1) + 0.5
    return torch.clamp_min(res, 0.0), res < 0


def get_rect(xy, radius, grid_x, grid_y):
    """auxiliary.h:50-60 (getRect): truncation toward zero, clamped to the grid."""
    r = radius.to(xy.dtype)
    def lo(v, g):
        return torch.clamp(torch.trunc((v - r - 0.5) / TILE), 0, g).long()
    def hi(v, g):
        return torch.clamp(torch.trunc((v + r + TILE - 1 + 0.5) / TILE), 0, g).long()
    return lo(xy[:, 0], grid_x), lo(xy[:, 1], grid_y), hi(xy[:, 0], grid_x), hi(xy[:, 1], grid_y)


def preprocess(means3D, opacities, view, proj, W, H, tan_fovx, tan_fovy, scales=None, rotations=None,
               cov3D_precomp=None, scale_modifier=1.0, prefiltered=False, tile_mask=None):
    """forward.cu:694-839 (preprocessCUDA) + auxiliary.h:148-168 (in_frustum).  All inputs float64.
    Returns a dict; entries of culled Gaussians are unspecified except radii / tiles_touched."""
    P = means3D.shape[0]
    view, proj = view.reshape(-1), proj.reshape(-1)
    focal_y, focal_x = H / (2.0 * tan_fovy), W / (2.0 * tan_fovx)
    grid_x, grid_y = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE

    p_view = _xform43(means3D, view)
    p_hom = _xform44(means3D, proj)
    p_w = 1.0 / (p_hom[:, 3] + 0.0000001)
    p_proj = p_hom[:, :3] * p_w[:, None]
    alive = torch.ones(P, dtype=torch.bool)
    if not prefiltered:
        alive = ((p_proj[:, 2] > -1.01) & (p_proj[:, 2] < 1.01) & (p_proj[:, 0] > -1.2) & (p_proj[:, 0] < 1.2) &
                 (p_proj[:, 1] > -1.2) & (p_proj[:, 1] < 1.2))
    alive = alive & ~(opacities.reshape(-1) < 0.0001)

    c3 = cov3D_precomp if cov3D_precomp is not None else cov3d(scales, rotations, scale_modifier)
    a, b, c = cov2d(means3D, c3, view, focal_x, focal_y, tan_fovx, tan_fovy)
    det = a * c - b * b
    radii = torch.zeros(P, dtype=torch.int32)
    bad = alive & ((det <= 0) | (a <= 0) | (c <= 0))
    det_safe = torch.where(det == 0, torch.ones_like(det), det)
    conic = torch.stack([c / det_safe, -b / det_safe, a / det_safe], -1)
    mid = 0.5 * (a + c)
    root = torch.sqrt(torch.clamp_min(mid * mid - det, 0.1))
    l1, l2 = mid + root, mid - root
    bad2 = alive & ~bad & ((l1 <= 0.01) | (l2 <= 0.01) | (l1 < l2) | (l1 / l2 > 10000.0))
    radii[bad | bad2] = -1
    ok = alive & ~bad & ~bad2
    my_radius = torch.ceil(3.0 * torch.sqrt(torch.maximum(l1, l2)))
    xy = torch.stack([((p_proj[:, 0] + 1.0) * W - 1.0) * 0.5, ((p_proj[:, 1] + 1.0) * H - 1.0) * 0.5], -1)
    rad_i = torch.where(ok, my_radius, torch.zeros_like(my_radius))
    xy_ok = torch.where(ok[:, None], xy, torch.zeros_like(xy))     # culled entries may hold inf / nan
    x0, y0, x1, y1 = get_rect(xy_ok, rad_i, grid_x, grid_y)
    touched = (y1 - y0) * (x1 - x0)
    touched = torch.where(ok, touched, torch.zeros_like(touched))
    if tile_mask is not None:
        tm = tile_mask.reshape(grid_y, grid_x).to(torch.int64)
        ii = torch.zeros((grid_y + 1, grid_x + 1), dtype=torch.int64)
        ii[1:, 1:] = tm.cumsum(0).cumsum(1)                      # integral image of the mask
        cnt = ii[y1, x1] - ii[y0, x1] - ii[y1, x0] + ii[y0, x0]
        touched = torch.where(touched > 0, cnt, touched)
    vis = ok & (touched > 0)
    radii[vis] = my_radius[vis].to(torch.int32)
    return dict(radii=radii, tiles_touched=touched.to(torch.int32), depths=p_view[:, 2], means2D=xy,
                depths_f32=depth_f32(means3D, view),
                cov3D=c3, conic_opacity=torch.cat([conic, opacities.reshape(-1, 1)], 1),
                rect=(x0, y0, x1, y1), visible=vis, p_view=p_view)


# ----------------------------------------------------------------------------- binning
def f32_bits(x):
    """Bit pattern of float32(x) as a non-negative int64 (the low word of the sort key)."""
    return x.to(torch.float32).view(torch.int32).to(torch.int64) & 0xFFFFFFFF


def depth_f32(means3D, view):
    """View-space z with the reference's float32 rounding sequence (auxiliary.h:62-70 as nvcc contracts
    it: fadd(fma(z, m10, fma(x, m2, y * m6)), m14)), emulated in float64 where each product is exact.
    The depth's bit pattern is half of every sort key, so it is the one float the oracle reproduces
    bit for bit rather than in float64."""
    f32 = lambda t: t.to(torch.float32).to(torch.float64)  # noqa: E731
    x, y, z = (f32(c) for c in means3D.unbind(-1))
    m2, m6, m10, m14 = (f32(view[i]) for i in (2, 6, 10, 14))
    t = f32(y * m6)
    t = f32(x * m2 + t)
    t = f32(z * m10 + t)
    return f32(t + m14)


def build_pairs(pre, grid_x, grid_y, tile_mask=None):
    """rasterizer_impl.cu:70-132 (duplicateWithKeys) + :518 (prefix sum) + :546-551 (sort) + :137-159
    (identifyTileRanges).  Emission order is Gaussian-major, row-major over the rectangle; the sort is
    stable on (tile << 32 | depth bits).  Returns unsorted/sorted keys and ids, offsets and ranges."""
    x0, y0, x1, y1 = pre["rect"]
    vis = pre["radii"] > 0
    touched = pre["tiles_touched"].to(torch.int64)
    offsets = touched.cumsum(0)
    idx = torch.nonzero(vis).squeeze(1)
    w = (x1 - x0)[idx]
    h = (y1 - y0)[idx]
    n = w * h
    gid = torch.repeat_interleave(idx, n)
    start = torch.cumsum(n, 0) - n
    k = torch.arange(int(n.sum())) - torch.repeat_interleave(start, n)
    wk = torch.repeat_interleave(w, n)
    ty = torch.repeat_interleave(y0[idx], n) + k // wk
    tx = torch.repeat_interleave(x0[idx], n) + k % wk
    tile = ty * grid_x + tx
    if tile_mask is not None:
        keep = tile_mask.reshape(-1).bool()[tile]
        gid, tile = gid[keep], tile[keep]
    keys = (tile << 32) | f32_bits(pre["depths_f32"])[gid]
    order = torch.sort(keys, stable=True).indices
    keys_sorted, point_list = keys[order], gid[order]
    n_tiles = grid_x * grid_y
    ranges = torch.zeros((n_tiles, 2), dtype=torch.int32)
    if keys_sorted.numel():
        t_sorted = keys_sorted >> 32
        first = torch.searchsorted(t_sorted, torch.arange(n_tiles))
        last = torch.searchsorted(t_sorted, torch.arange(n_tiles), right=True)
        nonempty = last > first
        ranges[nonempty, 0] = first[nonempty].to(torch.int32)
        ranges[nonempty, 1] = last[nonempty].to(torch.int32)
    return dict(keys_unsorted=keys, vals_unsorted=gid.to(torch.int32), keys_sorted=keys_sorted,
                point_list=point_list.to(torch.int32), point_offsets=offsets.to(torch.int32), ranges=ranges)


# ----------------------------------------------------------------------------- blending
def _tile_pixels(tile, grid_x, W, H, dtype):
    ty, tx = divmod(tile, grid_x)
    ys = torch.arange(ty * TILE, min(ty * TILE + TILE, H))
    xs = torch.arange(tx * TILE, min(tx * TILE + TILE, W))
    py, px = torch.meshgrid(ys, xs, indexing="ij")
    return px.reshape(-1), py.reshape(-1)


def _tile_alpha(px, py, xy, co):
    """forward.cu:921-931: power, the power > 0 / alpha < 1/255 skips, alpha = min(0.99, o exp(power))."""
    dx = xy[None, :, 0] - px[:, None]
    dy = xy[None, :, 1] - py[:, None]
    power = -0.5 * (co[None, :, 0] * dx * dx + co[None, :, 2] * dy * dy) - co[None, :, 1] * dx * dy
    G = torch.exp(power)
    raw = co[None, :, 3] * G
    # value: min(0.99, o G).  Derivative: the reference's backward differentiates alpha = o G and
    # ignores the clamp (backward.cu:1131 vs :1184, dL_dG = o * dL_dalpha with no mask), i.e. a
    # straight-through clamp; written this way autograd through this function agrees with it.
    alpha = raw + (torch.clamp_max(raw, 0.99) - raw).detach()
    keep = ~(power > 0) & ~(alpha < 1.0 / 255.0)
    return dx, dy, G, alpha, keep


def blend_forward(pre, pairs, features, bg, W, H):
    """forward.cu:844-963 (renderCUDA): front-to-back compositing per tile; a pixel stops after the
    Gaussian that takes T below 1e-4 (which is still blended).  Returns color (C,H,W), depth (H,W),
    alpha (H,W), n_contrib (H,W) and, for the backward, nothing else (it re-walks the lists)."""
    dtype = features.dtype
    C = features.shape[1]
    grid_x = (W + TILE - 1) // TILE
    color = torch.zeros((C, H, W), dtype=dtype)
    depth = torch.zeros((H, W), dtype=dtype)
    alpha_out = torch.zeros((H, W), dtype=dtype)
    n_contrib = torch.zeros((H, W), dtype=torch.int32)
    n_touched = torch.zeros(features.shape[0], dtype=torch.int64)
    xy_all, co_all, dep_all = pre["means2D"], pre["conic_opacity"], pre["depths"]
    pl = pairs["point_list"].long()
    for tile, (lo, hi) in enumerate(pairs["ranges"].tolist()):
        px, py = _tile_pixels(tile, grid_x, W, H, dtype)
        npix = px.numel()
        T = torch.ones(npix, dtype=dtype)
        acc = torch.zeros((npix, C), dtype=dtype)
        D = torch.zeros(npix, dtype=dtype)
        last = torch.zeros(npix, dtype=torch.int64)
        done = torch.zeros(npix, dtype=torch.bool)
        for s in range(lo, hi, 256):
            if bool(done.all()):
                break
            ids = pl[s:min(s + 256, hi)]
            _, _, _, alpha, keep = _tile_alpha(px.to(dtype), py.to(dtype), xy_all[ids], co_all[ids])
            a_eff = torch.where(keep, alpha, torch.zeros_like(alpha))
            T_after = T[:, None] * torch.cumprod(1 - a_eff, 1)
            T_before = torch.cat([T[:, None], T_after[:, :-1]], 1)
            inc = keep & ~done[:, None] & ~(T_before < 0.0001)
            w = torch.where(inc, alpha * T_before, torch.zeros_like(alpha))
            # extension (no reference counterpart): pixels where the Gaussian is blended and more than
            # half of the light still passes behind it
            n_touched.index_add_(0, ids, (inc & (T_after.detach() > 0.5)).sum(0))
            acc = acc + w @ features[ids]
            D = D + w @ dep_all[ids]
            pos = torch.arange(s - lo + 1, s - lo + 1 + ids.numel())
            last = torch.maximum(last, (inc.long() * pos[None, :]).max(1).values)
            # transmittance after the last included entry of this chunk
            T_inc = torch.where(inc, T_after, torch.full_like(T_after, float("inf"))).min(1).values
            T = torch.where(torch.isinf(T_inc), T, T_inc)
            done = done | (T < 0.0001)
        color[:, py, px] = (acc + T[:, None] * bg[None, :]).transpose(0, 1)
        depth[py, px] = D
        alpha_out[py, px] = 1 - T
        n_contrib[py, px] = last.to(torch.int32)
    pre["n_touched"] = n_touched.to(torch.int32)
    return color, depth, alpha_out, n_contrib


def blend_backward(pre, pairs, features, bg, W, H, out_alpha, n_contrib, dL_dpix, dL_ddepth_pix, dL_dalpha_pix):
    """backward.cu:1018-1209 (renderCUDA): gradients of the compositing w.r.t. the per-Gaussian 2D
    quantities, in closed form of the reference's back-to-front recurrences:
        accum_rec_i = (sum_{j>i} c_j w_j) / T_after_i          (:1146-1156, w_j = alpha_j T_before_j)
        accum_red_i likewise with depths (:1161-1165), accum_rea_i with 1 (:1167-1168)
        dL_dalpha_i = T_before_i [ sum_{ch<3}(c_i - rec_i) g_ch + (d_i - red_i) g_d + (1 - rea_i) g_a ]
                      - T_final / (1 - alpha_i) * sum_ch bg_ch g_ch       (:1171-1181; channel mask :1153)
    and the per-pair outputs of :1186-1206.  Returns dict of (P,...) tensors incl. absgrad."""
    dtype = features.dtype
    P, C = features.shape
    grid_x = (W + TILE - 1) // TILE
    g = dict(mean2D=torch.zeros((P, 2), dtype=dtype), absmean2D=torch.zeros((P, 2), dtype=dtype),
             conic=torch.zeros((P, 3), dtype=dtype), opacity=torch.zeros(P, dtype=dtype),
             color=torch.zeros((P, C), dtype=dtype), depth=torch.zeros(P, dtype=dtype))
    xy_all, co_all, dep_all = pre["means2D"], pre["conic_opacity"], pre["depths"]
    pl = pairs["point_list"].long()
    chan_w = torch.tensor([1.0 if ch <= 2 else 0.0 for ch in range(C)], dtype=dtype)
    for tile, (lo, hi) in enumerate(pairs["ranges"].tolist()):
        px, py = _tile_pixels(tile, grid_x, W, H, dtype)
        last = n_contrib[py, px].long()
        L = int(last.max()) if last.numel() else 0
        if L == 0:
            continue
        ids = pl[lo:lo + L]
        dx, dy, G, alpha, keep = _tile_alpha(px.to(dtype), py.to(dtype), xy_all[ids], co_all[ids])
        pos = torch.arange(1, L + 1)
        inc = keep & (pos[None, :] <= last[:, None])
        a_eff = torch.where(inc, alpha, torch.zeros_like(alpha))
        T_after = torch.cumprod(1 - a_eff, 1)
        T_before = torch.cat([torch.ones_like(T_after[:, :1]), T_after[:, :-1]], 1)
        T_final = 1 - out_alpha[py, px]
        w = a_eff * T_before
        gc = dL_dpix[:, py, px].transpose(0, 1)            # (npix, C)
        gd = dL_ddepth_pix[py, px]
        ga = dL_dalpha_pix[py, px]
        f = features[ids]                                   # (L, C)
        dep = dep_all[ids]

        def behind(v):   # sum_{j>i} v_j, along the list
            return torch.flip(torch.cumsum(torch.flip(v, [1]), 1), [1]) - v
        rec = torch.stack([behind(w * f[None, :, ch]) for ch in range(C)], -1) / T_after[..., None]
        red = behind(w * dep[None, :]) / T_after
        rea = behind(w) / T_after
        dL_dalpha = (((f[None, :, :] - rec) * (gc * chan_w[None, :])[:, None, :]).sum(-1)
                     + (dep[None, :] - red) * gd[:, None] + (1 - rea) * ga[:, None]) * T_before
        bg_dot = (gc * bg[None, :]).sum(1)
        dL_dalpha = dL_dalpha + (-T_final[:, None] / (1 - alpha)) * bg_dot[:, None]
        dL_dalpha = torch.where(inc, dL_dalpha, torch.zeros_like(dL_dalpha))
        co = co_all[ids]
        dL_dG = co[None, :, 3] * dL_dalpha
        gdx, gdy = G * dx, G * dy
        dG_dx = -gdx * co[None, :, 0] - gdy * co[None, :, 1]
        dG_dy = -gdy * co[None, :, 2] - gdx * co[None, :, 1]
        mx = dL_dG * dG_dx * (0.5 * W)
        my = dL_dG * dG_dy * (0.5 * H)
        g["mean2D"].index_add_(0, ids, torch.stack([mx.sum(0), my.sum(0)], -1))
        g["absmean2D"].index_add_(0, ids, torch.stack([mx.abs().sum(0), my.abs().sum(0)], -1))
        g["conic"].index_add_(0, ids, torch.stack([(-0.5 * gdx * dx * dL_dG).sum(0), (-0.5 * gdx * dy * dL_dG).sum(0),
                                                  (-0.5 * gdy * dy * dL_dG).sum(0)], -1))
        g["opacity"].index_add_(0, ids, (G * dL_dalpha).sum(0))
        g["color"].index_add_(0, ids, w.transpose(0, 1) @ gc)
        g["depth"].index_add_(0, ids, (w * gd[:, None]).sum(0))
    return g


# ----------------------------------------------------------------------------- full pipeline
def _f64(t):
    return None if t is None else t.detach().to(torch.float64).cpu()


def forward(means3D, opacities, view, proj, campos, W, H, tan_fovx, tan_fovy, bg, colors_precomp=None, shs=None,
            sh_degree=0, scales=None, rotations=None, cov3D_precomp=None, tile_mask=None, scale_modifier=1.0,
            prefiltered=False):
    """rasterizer_impl.cu:435-582 (Rasterizer::forward) end to end in float64."""
    m3, op = _f64(means3D), _f64(opacities)
    view64, proj64 = _f64(view).reshape(-1), _f64(proj).reshape(-1)
    pre = preprocess(m3, op, view64, proj64, W, H, tan_fovx, tan_fovy, _f64(scales), _f64(rotations),
                     _f64(cov3D_precomp), scale_modifier, prefiltered, tile_mask)
    grid_x, grid_y = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
    if colors_precomp is not None:
        feats = _f64(colors_precomp)
    else:
        feats, clamped = sh_to_rgb(sh_degree, _f64(shs), m3, _f64(campos))
        pre["clamped"] = clamped
    pre["features"] = feats
    pairs = build_pairs(pre, grid_x, grid_y, tile_mask)
    bg64 = _f64(bg).reshape(-1)[: feats.shape[1]]
    color, depth, alpha, n_contrib = blend_forward(pre, pairs, feats, bg64, W, H)
    return dict(pre=pre, pairs=pairs, color=color, depth=depth, alpha=alpha, n_contrib=n_contrib, bg=bg64,
                n_touched=pre["n_touched"], R=int(pairs["keys_sorted"].numel()))


def perturbed_camera(view0, proj0, tau):
    """viewmatrix / projmatrix / campos (reference convention: transposed, utils/camera_utils.py:129-139) of
    the camera W2C' = Exp(rho, theta) W2C, to first order in tau = (rho, theta) -- exact for the gradient at
    tau = 0, which is where MonoGS-style pose refinement evaluates it.  Extension; no reference counterpart."""
    rho, th = tau[:3], tau[3:]
    z = torch.zeros((), dtype=tau.dtype)
    skew = torch.stack([torch.stack([z, -th[2], th[1]]), torch.stack([th[2], z, -th[0]]), torch.stack([-th[1], th[0], z])])
    M = torch.eye(4, dtype=tau.dtype)
    M = torch.cat([torch.cat([torch.eye(3, dtype=tau.dtype) + skew, rho.reshape(3, 1)], 1), M[3:]], 0)
    w2c0 = view0.reshape(4, 4).transpose(0, 1)
    w2c = M @ w2c0
    view = w2c.transpose(0, 1)
    proj_t = torch.linalg.inv(view0.reshape(4, 4)) @ proj0.reshape(4, 4)       # P^T
    proj = view @ proj_t
    campos = torch.linalg.inv(view)[3, :3]
    return view.reshape(-1), proj.reshape(-1), campos


def backward(fw, means3D, view, proj, campos, W, H, tan_fovx, tan_fovy, dL_dcolor, dL_ddepth, dL_dalpha,
             shs=None, sh_degree=0, scales=None, rotations=None, cov3D_precomp=None, scale_modifier=1.0, pose=False):
    """rasterizer_impl.cu:586-692 (Rasterizer::backward): blend backward (explicit, above), then the
    per-Gaussian chain of backward.cu:534-664 (computeCov2DCUDA) and :949-1015 (preprocessCUDA) obtained
    by reverse-mode differentiation of the float64 forward expressions -- which is what those two
    kernels implement by hand -- restricted, like them, to Gaussians with radii > 0."""
    pre, pairs = fw["pre"], fw["pairs"]
    feats = pre["features"]
    gb = blend_backward(pre, pairs, feats, fw["bg"], W, H, fw["alpha"], fw["n_contrib"], _f64(dL_dcolor),
                        _f64(dL_ddepth).reshape(H, W), _f64(dL_dalpha).reshape(H, W))
    vis = pre["radii"] > 0
    vi = torch.nonzero(vis).squeeze(1)
    m3 = _f64(means3D).clone().requires_grad_(True)
    view64, proj64 = _f64(view).reshape(-1), _f64(proj).reshape(-1)
    campos64 = _f64(campos)
    tau = None
    if pose:
        tau = torch.zeros(6, dtype=torch.float64, requires_grad=True)
        view64, proj64, campos64 = perturbed_camera(view64, proj64, tau)
    sc = rot = c3p = sh = None
    # the chain is evaluated on the visible subset only (culled Gaussians may have singular
    # covariances); autograd scatters zeros to the rest, like the kernels' `radii > 0` guard
    m3v = m3[vi]
    if cov3D_precomp is not None:
        c3p = _f64(cov3D_precomp).clone().requires_grad_(True)
        c3 = c3p[vi]
    else:
        sc = _f64(scales).clone().requires_grad_(True)
        rot = _f64(rotations).clone().requires_grad_(True)
        c3 = cov3d(sc[vi], rot[vi], scale_modifier)
    c3.retain_grad()
    focal_y, focal_x = H / (2.0 * tan_fovy), W / (2.0 * tan_fovx)
    a, b, c = cov2d(m3v, c3, view64, focal_x, focal_y, tan_fovx, tan_fovy)
    det = a * c - b * b
    conic = torch.stack([c / det, -b / det, a / det], -1)
    p_hom = _xform44(m3v, proj64)
    p_w = 1.0 / (p_hom[:, 3] + 0.0000001)
    # dL/dmean2D is per NDC unit (the 0.5 W / 0.5 H factors are already inside it, backward.cu:1188-1189)
    ndc = p_hom[:, :2] * p_w[:, None]
    depth = _xform43(m3v, view64)[:, 2]
    # The blend backward stores HALF of dL/d(conic_xy) (backward.cu:1201: -0.5 * gdx * d.y * dL_dG is the
    # share of one of the two symmetric off-diagonal entries); computeCov2DCUDA's formulas account for
    # both entries (the factors 2 at backward.cu:593-595), hence the 2 here.
    gconic = gb["conic"][vi] * torch.tensor([1.0, 2.0, 1.0], dtype=torch.float64)
    loss = (conic * gconic).sum() + (ndc * gb["mean2D"][vi]).sum() + (depth * gb["depth"][vi]).sum()
    g_sh = None
    if shs is not None:
        sh = _f64(shs).clone().requires_grad_(True)
        rgb, _ = sh_to_rgb(sh_degree, sh[vi], m3v, campos64)
        loss = loss + (rgb * gb["color"][vi]).sum()
    loss.backward()
    P = m3.shape[0]
    z = lambda t, shape: torch.zeros(shape, dtype=torch.float64) if t is None or t.grad is None else t.grad  # noqa: E731
    if shs is not None:
        g_sh = z(sh, sh.shape)
    g_cov3D = torch.zeros((P, 6), dtype=torch.float64)
    if c3.grad is not None:
        g_cov3D[vi] = c3.grad
    return dict(g_means2D=torch.cat([gb["mean2D"], torch.zeros((P, 1), dtype=torch.float64)], 1),
                g_abs2D=torch.cat([gb["absmean2D"], torch.zeros((P, 1), dtype=torch.float64)], 1),
                g_colors=gb["color"], g_opacity=gb["opacity"].reshape(P, 1), g_means3D=z(m3, (P, 3)),
                g_cov3D=g_cov3D, g_sh=g_sh if g_sh is not None else torch.zeros((P, 0, 3), dtype=torch.float64),
                g_scales=z(sc, (P, 3)), g_rot=z(rot, (P, 4)), g_conic=gb["conic"], g_depth=gb["depth"],
                g_tau=None if tau is None else (tau.grad if tau.grad is not None else torch.zeros(6, dtype=torch.float64)))


def mark_visible(means3D, view, proj):
    """rasterizer_impl.cu:54-66 (checkFrustum)."""
    p_hom = _xform44(_f64(means3D), _f64(proj).reshape(-1))
    p = p_hom[:, :3] / (p_hom[:, 3:4] + 0.0000001)
    return ((p[:, 2] > -1.01) & (p[:, 2] < 1.01) & (p[:, 0] > -1.2) & (p[:, 0] < 1.2) & (p[:, 1] > -1.2) & (p[:, 1] < 1.2))


# ----------------------------------------------------------------------------- bench.py's CPU leg
def timed_sample(budget_s=20.0, P=500_000, seed=0):
    """Time fwd+bwd of this float64 port on the host cores on a bounded sample of bench.py's workload
    (the config-2 scene and camera 0): a centred crop of the 640x480 view, rendered through a tile
    mask that keeps only the crop's tiles.  The crop is grown until one pass costs ~budget_s/3."""
    import os
    from splatloc_b200 import scenes
    sc = scenes.make_scene(P, seed=seed, channels=4)
    cam = scenes.make_camera(seed=0)
    W, H = cam.W, cam.H
    gx, gy = (W + 15) // 16, (H + 15) // 16
    gcol, gdep, galp = scenes.make_pixel_grads(0, 4, H, W)
    bg = torch.zeros(4)

    def run(tw, th):
        mask = torch.zeros((gy, gx), dtype=torch.bool)
        y0, x0 = (gy - th) // 2, (gx - tw) // 2
        mask[y0:y0 + th, x0:x0 + tw] = True
        t0 = time.perf_counter()
        fw = forward(sc["means3D"], sc["opacities"], cam.viewmatrix, cam.projmatrix, cam.campos, W, H, cam.tanfovx,
                     cam.tanfovy, bg, colors_precomp=sc["colors_precomp"], scales=sc["scales"],
                     rotations=sc["rotations"], tile_mask=mask)
        backward(fw, sc["means3D"], cam.viewmatrix, cam.projmatrix, cam.campos, W, H, cam.tanfovx, cam.tanfovy,
                 gcol, gdep, galp, scales=sc["scales"], rotations=sc["rotations"])
        return time.perf_counter() - t0, int(mask.sum()) * 256

    tw, th = 4, 3
    dt, pix = run(tw, th)            # warm-up + size probe
    while dt < budget_s / 3 and (tw < gx or th < gy):
        tw, th = min(gx, tw * 2), min(gy, th * 2)
        dt, pix = run(tw, th)
    return {"value": round(pix / dt / 1e6, 6), "unit": "Mpix/s", "cores": torch.get_num_threads(),
            "host_cpus": os.cpu_count(), "kind": "port",
            "sample": f"float64 PyTorch port, {P} Gaussians (all preprocessed), centred {tw * 16}x{th * 16} px crop of "
                      f"the 640x480 view via tile mask, 1 fwd+bwd pass, {dt:.1f} s"}
