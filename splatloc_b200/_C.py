"""Host-side operator layer: torch tensors in, C-ABI calls out.

Mirrors the functions the reference registers in its pybind module `diff_gauss._C`
for this path (submodules/diff-gaussian-rasterization/ext.cpp:15-27 and
rasterize_points.h:18-67,105-108) and simple-knn's `simple_knn._C.distCUDA2`
(submodules/simple-knn/ext.cpp:15-17, spatial.h:14): same names, positional
argument order, return tuples and error behaviour.  PyTorch is used for device
memory and the current stream only; all computation is in libsplatloc_b200.so.
"""
from __future__ import annotations

import ctypes as C
import threading

import torch

from . import _capi


def _ptr(t: torch.Tensor | None):
    """Device address, or NULL for an absent (None / empty) optional tensor."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def _f32(t: torch.Tensor, name: str) -> torch.Tensor:
    if t.numel() == 0:
        return t
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (the rasterizer has no CPU path)")
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32, got {t.dtype}")
    return t.contiguous()


def _stream(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


class _Chunk:
    """A growable byte tensor handed to the library as an allocation callback
    (the reference's resizeFunctional, rasterize_points.cu:27-33).  The ctypes callback object is
    created once and reused: `take()` hands the tensor of the finished call to the caller."""

    def __init__(self):
        self.device = None
        self.tensor = None

        def _alloc(_ctx, nbytes):
            self.tensor = torch.empty(int(nbytes), dtype=torch.uint8, device=self.device)
            return self.tensor.data_ptr()

        self.callback = _capi.ALLOC_FN(_alloc)

    def take(self):
        t, self.tensor = self.tensor, None
        return t if t is not None else torch.empty(0, dtype=torch.uint8, device=self.device)


_tls = threading.local()


def _chunks(device):
    """This thread's three allocation callbacks (geometry, binning, image), bound to `device`."""
    c = getattr(_tls, "chunks", None)
    if c is None:
        c = _tls.chunks = (_Chunk(), _Chunk(), _Chunk())
    for x in c:
        x.device = device
        x.tensor = None
    return c


class _on_device:
    """`torch.cuda.device(dev)` only when `dev` is not already current (the common case costs nothing)."""

    def __init__(self, dev):
        cur = torch.cuda.current_device()
        self.ctx = None if dev.index is None or dev.index == cur else torch.cuda.device(dev)

    def __enter__(self):
        if self.ctx is not None:
            self.ctx.__enter__()

    def __exit__(self, *exc):
        if self.ctx is not None:
            self.ctx.__exit__(*exc)


def _pad_bg(bg: torch.Tensor, n_ch: int) -> torch.Tensor:
    """SplatLoc hands a 3-float background to its 4-channel build, which then reads
    one float past the end (train_gaussians.py:70 vs forward.cu:960).  Here missing
    channels are defined as 0 instead of out-of-bounds memory."""
    bg = bg.reshape(-1)
    if bg.numel() >= n_ch:
        return bg
    return torch.cat([bg, bg.new_zeros(n_ch - bg.numel())])


def _channels(colors: torch.Tensor) -> int:
    if colors.numel() == 0:
        return 3  # SH colours are RGB
    if colors.dim() != 2 or colors.shape[1] not in (3, 4):
        raise RuntimeError(f"colors_precomp must have shape (num_points, 3 or 4), got {tuple(colors.shape)}")
    return int(colors.shape[1])


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                        cov3D_precomp, tile_mask, viewmatrix, projmatrix, tan_fovx, tan_fovy,
                        image_height, image_width, sh, degree, campos, prefiltered, debug, want_n_touched=False,
                        max_pairs=None):
    """-> (num_rendered, color, depth, alpha, radii, geomBuffer, binningBuffer, imgBuffer)
    [+ n_touched (P,) int32 when `want_n_touched` -- an extension, see sl_forward_ext]
    [+ status (1,) int32 on the device when `max_pairs` is given -- sl_forward_nosync: no host
       synchronisation (CUDA-graph capturable), `num_rendered` is then `max_pairs`, and status becomes 1
       if the frame did not fit (outputs are background-only in that case)]."""
    if means3D.dim() != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    lib = _capi.load()
    if not means3D.is_cuda:
        raise RuntimeError("means3D must be a CUDA tensor (the rasterizer has no CPU path)")
    dev = means3D.device
    P, H, W = int(means3D.shape[0]), int(image_height), int(image_width)

    means3D = _f32(means3D, "means3D")
    colors = _f32(colors, "colors_precomp")
    opacity = _f32(opacity, "opacities")
    scales = _f32(scales, "scales")
    rotations = _f32(rotations, "rotations")
    cov3D_precomp = _f32(cov3D_precomp, "cov3D_precomp")
    sh = _f32(sh, "shs")
    background = _f32(background, "bg")
    viewmatrix = _f32(viewmatrix, "viewmatrix")
    projmatrix = _f32(projmatrix, "projmatrix")
    campos = _f32(campos, "campos")
    n_ch = _channels(colors)
    background = _pad_bg(background, n_ch)
    if tile_mask is not None and tile_mask.numel() != 0:
        if not tile_mask.is_cuda:
            raise RuntimeError("tile_mask must be a CUDA tensor")
        tile_mask = tile_mask.to(torch.bool).contiguous()
        tiles = ((H + 15) // 16) * ((W + 15) // 16)
        if tile_mask.numel() != tiles:
            raise RuntimeError(f"tile_mask must have {tiles} elements, got {tile_mask.numel()}")
    M = int(sh.shape[1]) if sh.numel() != 0 else 0

    with _on_device(dev):
        # colour, depth and alpha are views of one allocation (one device -> host copy moves them all)
        images = torch.empty((n_ch + 2, H, W), dtype=torch.float32, device=dev)
        out_color, out_depth, out_alpha = images[:n_ch], images[n_ch:n_ch + 1], images[n_ch + 1:]
        radii = torch.empty((P,), dtype=torch.int32, device=dev)
        geom, binning, img = _chunks(dev)
        n_touched = torch.zeros((P,), dtype=torch.int32, device=dev) if want_n_touched else None
        common = (geom.callback, None, binning.callback, None, img.callback, None,
                  P, int(degree), M, n_ch,
                  _ptr(background), W, H,
                  _ptr(means3D), _ptr(sh), _ptr(colors),
                  _ptr(opacity), _ptr(scales), float(scale_modifier),
                  _ptr(rotations), _ptr(cov3D_precomp), _ptr(tile_mask),
                  _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos),
                  float(tan_fovx), float(tan_fovy), int(bool(prefiltered)),
                  out_color.data_ptr(), out_depth.data_ptr(), out_alpha.data_ptr(), _ptr(radii), _ptr(n_touched))
        status = None
        if max_pairs is None:
            rc = lib.sl_forward_ext(*common, int(bool(debug)), _stream(dev))
        else:
            status = torch.empty(1, dtype=torch.int32, device=dev)
            rc = lib.sl_forward_nosync(*common, int(max_pairs), status.data_ptr(), _stream(dev))
        geom_t, binning_t, img_t = geom.take(), binning.take(), img.take()
        rendered = _capi.check(rc)
    out = (rendered, out_color, out_depth, out_alpha, radii, geom_t, binning_t, img_t)
    if want_n_touched:
        out = out + (n_touched,)
    if status is not None:
        out = out + (status,)
    return out


def trainable_grad_numel(P: int, n_ch: int, M: int) -> int:
    """Length of the leading span of the packed gradient block that holds the gradients of
    means3D, opacity, scales, rotations and colours / SH (the blocks are 16-byte aligned)."""
    return sum((P * w + 3) // 4 * 4 for w in (3, 1, 3, 4, n_ch, 3 * M))


def _grad_block(P: int, widths: list[int], device) -> list[torch.Tensor]:
    """One zero-filled allocation (one memset) carved into per-gradient blocks.  The base tensor
    (`block._base`) carries `trainable_numel`: the length of its leading span that holds the
    gradients of the trainable inputs (means3D, opacity, scales, rotations, colours / SH)."""
    sizes = [(P * w + 3) // 4 * 4 for w in widths]  # keep every block 16-byte aligned
    flat = torch.zeros(sum(sizes), dtype=torch.float32, device=device)
    out, off = [], 0
    for w, s in zip(widths, sizes):
        out.append(flat[off:off + P * w])
        off += s
    return out


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                 cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy,
                                 dL_dout_color, dL_dout_depth, dL_dout_alpha, sh, degree, campos,
                                 geomBuffer, R, binningBuffer, imageBuffer, out_alpha, debug, want_pose=False):
    """-> (dL_dabsmeans2D, dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D,
           dL_dsh, dL_dscales, dL_drotations)
    [+ dL_dtau (6,) = (dL/drho, dL/dtheta) as a 10th element when `want_pose` -- see sl_backward_ext]."""
    lib = _capi.load()
    dev = means3D.device
    P = int(means3D.shape[0])
    H, W = int(dL_dout_color.shape[1]), int(dL_dout_color.shape[2])
    n_ch = int(dL_dout_color.shape[0])
    M = int(sh.shape[1]) if sh.numel() != 0 else 0

    means3D = _f32(means3D, "means3D")
    colors = _f32(colors, "colors_precomp")
    scales = _f32(scales, "scales")
    rotations = _f32(rotations, "rotations")
    cov3D_precomp = _f32(cov3D_precomp, "cov3D_precomp")
    sh = _f32(sh, "shs")
    dL_dout_color = _f32(dL_dout_color, "dL_dout_color")
    dL_dout_depth = _f32(dL_dout_depth, "dL_dout_depth")
    dL_dout_alpha = _f32(dL_dout_alpha, "dL_dout_alpha")
    out_alpha = _f32(out_alpha, "out_alpha")
    background = _pad_bg(_f32(background, "bg"), n_ch)
    viewmatrix = _f32(viewmatrix, "viewmatrix")
    projmatrix = _f32(projmatrix, "projmatrix")
    campos = _f32(campos, "campos")
    radii = radii.contiguous()

    with _on_device(dev):
        # what an optimiser consumes first (one contiguous span for a multi-GPU all-reduce), then the
        # per-view densification statistics, then the intermediates of the chain
        (g_mean3D, g_opac, g_scale, g_rot, g_color, g_sh, g_mean2D, g_abs2D, g_depth, g_conic,
         g_cov3D) = _grad_block(P, [3, 1, 3, 4, n_ch, 3 * M, 3, 3, 1, 4, 6], dev)
        flat = g_mean3D._base
        flat.trainable_numel = (flat.numel() if cov3D_precomp.numel() != 0
                                else g_mean2D.storage_offset() - flat.storage_offset())
        g_tau = torch.zeros(6, dtype=torch.float32, device=dev) if want_pose else None
        if P != 0:
            rc = lib.sl_backward_ext(
                P, int(degree), M, int(R), n_ch,
                _ptr(background), W, H,
                _ptr(means3D), _ptr(sh), _ptr(colors),
                _ptr(scales), float(scale_modifier), _ptr(rotations),
                _ptr(cov3D_precomp), _ptr(viewmatrix), _ptr(projmatrix),
                _ptr(campos), float(tan_fovx), float(tan_fovy), _ptr(radii),
                _ptr(geomBuffer), _ptr(binningBuffer), _ptr(imageBuffer),
                _ptr(out_alpha), _ptr(dL_dout_color), _ptr(dL_dout_depth), _ptr(dL_dout_alpha),
                g_mean2D.data_ptr(), g_abs2D.data_ptr(), g_conic.data_ptr(), g_opac.data_ptr(),
                g_color.data_ptr(), g_depth.data_ptr(), g_mean3D.data_ptr(), g_cov3D.data_ptr(),
                _ptr(g_sh), g_scale.data_ptr(), g_rot.data_ptr(), _ptr(g_tau),
                int(bool(debug)), _stream(dev))
            _capi.check(rc)
    out = (g_abs2D.view(P, 3), g_mean2D.view(P, 3), g_color.view(P, n_ch), g_opac.view(P, 1),
           g_mean3D.view(P, 3), g_cov3D.view(P, 6), g_sh.view(P, M, 3), g_scale.view(P, 3),
           g_rot.view(P, 4))
    return out + (g_tau,) if want_pose else out


def rasterize_gaussians_batch(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                              cov3D_precomp, tile_mask, viewmatrices, projmatrices, tan_fovx, tan_fovy,
                              image_height, image_width, sh, degree, campos, prefiltered, debug,
                              want_n_touched=False, max_pairs=None):
    """Several cameras in one call (sl_forward_batch; the reference renders one camera per call and its
    callers loop, train_gaussians.py:195-229).  Arguments as `rasterize_gaussians`, except
    `viewmatrices`, `projmatrices` (V, 4, 4) and `campos` (V, 3): V cameras of one image size.
    -> (num_rendered over all views, color (V, C, H, W), depth (V, 1, H, W), alpha (V, 1, H, W),
        radii (V, P), geomBuffer, binningBuffer, imgBuffer) [+ n_touched (V, P)] [+ status (1,)]
    Each view's results are bit-identical to the single-view call's."""
    if means3D.dim() != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if viewmatrices.dim() != 3 or projmatrices.shape != viewmatrices.shape or tuple(viewmatrices.shape[1:]) != (4, 4):
        raise RuntimeError("viewmatrices and projmatrices must have dimensions (num_views, 4, 4)")
    lib = _capi.load()
    if not means3D.is_cuda:
        raise RuntimeError("means3D must be a CUDA tensor (the rasterizer has no CPU path)")
    dev = means3D.device
    V, P, H, W = int(viewmatrices.shape[0]), int(means3D.shape[0]), int(image_height), int(image_width)
    if V < 1 or campos.numel() != 3 * V:
        raise RuntimeError(f"campos must have dimensions ({V}, 3)")

    means3D = _f32(means3D, "means3D")
    colors = _f32(colors, "colors_precomp")
    opacity = _f32(opacity, "opacities")
    scales = _f32(scales, "scales")
    rotations = _f32(rotations, "rotations")
    cov3D_precomp = _f32(cov3D_precomp, "cov3D_precomp")
    sh = _f32(sh, "shs")
    viewmatrices = _f32(viewmatrices, "viewmatrices")
    projmatrices = _f32(projmatrices, "projmatrices")
    campos = _f32(campos, "campos")
    n_ch = _channels(colors)
    background = _pad_bg(_f32(background, "bg"), n_ch)
    if tile_mask is not None and tile_mask.numel() != 0:
        if not tile_mask.is_cuda:
            raise RuntimeError("tile_mask must be a CUDA tensor")
        tile_mask = tile_mask.to(torch.bool).contiguous()
        tiles = ((H + 15) // 16) * ((W + 15) // 16)
        if tile_mask.numel() != tiles:
            raise RuntimeError(f"tile_mask must have {tiles} elements, got {tile_mask.numel()}")
    M = int(sh.shape[1]) if sh.numel() != 0 else 0

    with _on_device(dev):
        # one allocation [colour of all views | depth | alpha]: one device -> host copy moves them all
        N = H * W
        images = torch.empty(V * (n_ch + 2) * N, dtype=torch.float32, device=dev)
        color = images[:V * n_ch * N].view(V, n_ch, H, W)
        depth = images[V * n_ch * N:V * (n_ch + 1) * N].view(V, 1, H, W)
        alpha = images[V * (n_ch + 1) * N:].view(V, 1, H, W)
        radii = torch.empty((V, P), dtype=torch.int32, device=dev)
        geom, binning, img = _chunks(dev)
        n_touched = torch.zeros((V, P), dtype=torch.int32, device=dev) if want_n_touched else None
        status = torch.empty(1, dtype=torch.int32, device=dev) if max_pairs is not None else None
        rc = lib.sl_forward_batch(
            geom.callback, None, binning.callback, None, img.callback, None,
            V, P, int(degree), M, n_ch, _ptr(background), W, H,
            _ptr(means3D), _ptr(sh), _ptr(colors), _ptr(opacity), _ptr(scales), float(scale_modifier),
            _ptr(rotations), _ptr(cov3D_precomp), _ptr(tile_mask),
            _ptr(viewmatrices), _ptr(projmatrices), _ptr(campos),
            float(tan_fovx), float(tan_fovy), int(bool(prefiltered)),
            color.data_ptr(), depth.data_ptr(), alpha.data_ptr(), _ptr(radii), _ptr(n_touched),
            -1 if max_pairs is None else int(max_pairs), _ptr(status), int(bool(debug)), _stream(dev))
        geom_t, binning_t, img_t = geom.take(), binning.take(), img.take()
        rendered = _capi.check(rc)
    out = (rendered, color, depth, alpha, radii, geom_t, binning_t, img_t)
    if want_n_touched:
        out = out + (n_touched,)
    if status is not None:
        out = out + (status,)
    return out


def _carve(flat: torch.Tensor, counts: list[int]) -> list[torch.Tensor]:
    """Consecutive 16-byte-aligned slices of `flat` with the given element counts."""
    out, off = [], 0
    for n in counts:
        out.append(flat[off:off + n])
        off += (n + 3) // 4 * 4
    return out


def rasterize_gaussians_backward_batch(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                       cov3D_precomp, viewmatrices, projmatrices, tan_fovx, tan_fovy,
                                       dL_dout_color, dL_dout_depth, dL_dout_alpha, sh, degree, campos,
                                       geomBuffer, R, binningBuffer, imageBuffer, out_alpha, debug,
                                       want_pose=False, pose_only=False, exchange_multicast: int | None = None):
    """Backward of `rasterize_gaussians_batch` (sl_backward_batch).  Upstream gradients are (V, C, H, W),
    (V, 1, H, W), (V, 1, H, W).
    -> (dL_dabsmeans2D (V, P, 3), dL_dmeans2D (V, P, 3), dL_dcolors, dL_dopacity (P, 1), dL_dmeans3D (P, 3),
        dL_dcov3D (P, 6), dL_dsh (P, M, 3), dL_dscales (P, 3), dL_drotations (P, 4)) [+ dL_dtau (V, 6)]
    The per-Gaussian gradients are summed over the views; dL_dcolors is (P, C) (summed) with precomputed
    colours and (V, P, 3) with SH.  `pose_only`: only dL_dtau is computed (a fixed map); the per-Gaussian
    gradients come back as None.  `exchange_multicast`: multicast address of a multi-GPU exchange buffer
    (splatloc_b200.exchange): the summed gradients are also pushed into every GPU's buffer from the last
    kernel (sl_backward_batch_push)."""
    lib = _capi.load()
    dev = means3D.device
    P = int(means3D.shape[0])
    V, n_ch, H, W = (int(x) for x in dL_dout_color.shape)
    M = int(sh.shape[1]) if sh.numel() != 0 else 0
    with_sh = M != 0

    means3D = _f32(means3D, "means3D")
    colors = _f32(colors, "colors_precomp")
    scales = _f32(scales, "scales")
    rotations = _f32(rotations, "rotations")
    cov3D_precomp = _f32(cov3D_precomp, "cov3D_precomp")
    sh = _f32(sh, "shs")
    dL_dout_color = _f32(dL_dout_color, "dL_dout_color")
    dL_dout_depth = _f32(dL_dout_depth, "dL_dout_depth")
    dL_dout_alpha = _f32(dL_dout_alpha, "dL_dout_alpha")
    out_alpha = _f32(out_alpha, "out_alpha")
    background = _pad_bg(_f32(background, "bg"), n_ch)
    viewmatrices = _f32(viewmatrices, "viewmatrices")
    projmatrices = _f32(projmatrices, "projmatrices")
    campos = _f32(campos, "campos")
    radii = radii.contiguous()
    want_pose = want_pose or pose_only

    with _on_device(dev):
        # summed over the views, what an optimiser consumes first (one contiguous span for a multi-GPU
        # all-reduce); then the per-view statistics and chain intermediates
        n_color_sum = 0 if with_sh else P * n_ch
        counts = [P * 3, P, P * 3, P * 4, n_color_sum, P * 3 * M, P * 6,
                  V * P * 3, V * P * 3, V * P, V * P * 4, V * P * 3 if with_sh else 0]
        if pose_only:       # nothing per Gaussian leaves the last stage
            counts[0] = counts[2] = counts[3] = counts[5] = counts[6] = 0
        flat = torch.zeros(sum((n + 3) // 4 * 4 for n in counts), dtype=torch.float32, device=dev)
        (g_mean3D, g_opac, g_scale, g_rot, g_color_sum, g_sh, g_cov3D, g_mean2D, g_abs2D, g_depth, g_conic,
         g_color_view) = _carve(flat, counts)
        flat.trainable_numel = (flat.numel() if cov3D_precomp.numel() != 0
                                else g_cov3D.storage_offset() - flat.storage_offset())
        g_color = g_color_view if with_sh else g_color_sum
        g_tau = torch.zeros((V, 6), dtype=torch.float32, device=dev) if want_pose else None
        if P != 0:
            push = () if exchange_multicast is None else (int(exchange_multicast),)
            rc = (lib.sl_backward_batch if exchange_multicast is None else lib.sl_backward_batch_push)(
                V, P, int(degree), M, int(R), n_ch, _ptr(background), W, H,
                _ptr(means3D), _ptr(sh), _ptr(colors), _ptr(scales), float(scale_modifier), _ptr(rotations),
                _ptr(cov3D_precomp), _ptr(viewmatrices), _ptr(projmatrices), _ptr(campos),
                float(tan_fovx), float(tan_fovy), _ptr(radii),
                _ptr(geomBuffer), _ptr(binningBuffer), _ptr(imageBuffer),
                _ptr(out_alpha), _ptr(dL_dout_color), _ptr(dL_dout_depth), _ptr(dL_dout_alpha),
                g_mean2D.data_ptr(), g_abs2D.data_ptr(), g_conic.data_ptr(), g_opac.data_ptr(),
                g_color.data_ptr(), g_depth.data_ptr(), _ptr(g_mean3D), _ptr(g_cov3D),
                _ptr(g_sh), _ptr(g_scale), _ptr(g_rot), _ptr(g_tau), *push,
                int(bool(debug)), _stream(dev))
            _capi.check(rc)
    colors_out = g_color.view(V, P, 3) if with_sh else g_color.view(P, n_ch)
    if pose_only:
        out = (g_abs2D.view(V, P, 3), g_mean2D.view(V, P, 3), colors_out, g_opac.view(P, 1), None, None, None, None, None)
    else:
        out = (g_abs2D.view(V, P, 3), g_mean2D.view(V, P, 3), colors_out, g_opac.view(P, 1), g_mean3D.view(P, 3),
               g_cov3D.view(P, 6), g_sh.view(P, M, 3), g_scale.view(P, 3), g_rot.view(P, 4))
    return out + (g_tau,) if want_pose else out


def render_forward(background, means3D, features_dc, kp_score, raw_opacity, raw_scaling, raw_rotation,
                   scale_modifier, tile_mask, viewmatrices, projmatrices, tan_fovx, tan_fovy, image_height,
                   image_width, campos, debug=False, max_pairs=None):
    """render()'s pre-processing + rasterizer forward in one call (sl_forward_render): raw optimiser
    parameters in (gaussian_model.py:75-110 getters and gaussian_renderer/__init__.py:85-102 are applied
    inside the kernels).  `features_dc` (P, 3) or (P, 1, 3); `kp_score` (P, 1) or an empty tensor;
    `raw_scaling` (P, 3) or (P, 1); cameras stacked as for `rasterize_gaussians_batch`.
    -> (num_rendered, color (V, C, H, W), depth (V, 1, H, W), alpha (V, 1, H, W), radii (V, P),
        geomBuffer, binningBuffer, imgBuffer) [+ status]"""
    lib = _capi.load()
    if not means3D.is_cuda:
        raise RuntimeError("means3D must be a CUDA tensor (the rasterizer has no CPU path)")
    dev = means3D.device
    V, P, H, W = int(viewmatrices.shape[0]), int(means3D.shape[0]), int(image_height), int(image_width)
    means3D = _f32(means3D, "means3D")
    features_dc = _f32(features_dc, "features_dc")
    kp_score = _f32(kp_score, "kp_score")
    raw_opacity = _f32(raw_opacity, "raw_opacity")
    raw_scaling = _f32(raw_scaling, "raw_scaling")
    raw_rotation = _f32(raw_rotation, "raw_rotation")
    viewmatrices, projmatrices, campos = _f32(viewmatrices, "viewmatrices"), _f32(projmatrices, "projmatrices"), _f32(campos, "campos")
    if P and features_dc.numel() != 3 * P:
        raise RuntimeError("features_dc must have 3 values per Gaussian (SH degree 0)")
    scale_dim = int(raw_scaling.numel() // P) if P else 3
    n_ch = 4 if kp_score.numel() != 0 else 3
    background = _pad_bg(_f32(background, "bg"), n_ch)
    if tile_mask is not None and tile_mask.numel() != 0:
        tile_mask = tile_mask.to(torch.bool).contiguous()
    with _on_device(dev):
        N = H * W
        images = torch.empty(V * (n_ch + 2) * N, dtype=torch.float32, device=dev)
        color = images[:V * n_ch * N].view(V, n_ch, H, W)
        depth = images[V * n_ch * N:V * (n_ch + 1) * N].view(V, 1, H, W)
        alpha = images[V * (n_ch + 1) * N:].view(V, 1, H, W)
        radii = torch.empty((V, P), dtype=torch.int32, device=dev)
        geom, binning, img = _chunks(dev)
        status = torch.empty(1, dtype=torch.int32, device=dev) if max_pairs is not None else None
        rc = lib.sl_forward_render(
            geom.callback, None, binning.callback, None, img.callback, None,
            V, P, _ptr(background), W, H, _ptr(means3D), _ptr(features_dc), _ptr(kp_score),
            _ptr(raw_opacity), _ptr(raw_scaling), scale_dim, float(scale_modifier), _ptr(raw_rotation), _ptr(tile_mask),
            _ptr(viewmatrices), _ptr(projmatrices), _ptr(campos), float(tan_fovx), float(tan_fovy),
            color.data_ptr(), depth.data_ptr(), alpha.data_ptr(), _ptr(radii),
            -1 if max_pairs is None else int(max_pairs), _ptr(status), int(bool(debug)), _stream(dev))
        geom_t, binning_t, img_t = geom.take(), binning.take(), img.take()
        rendered = _capi.check(rc)
    out = (rendered, color, depth, alpha, radii, geom_t, binning_t, img_t)
    return out + (status,) if status is not None else out


def render_backward(background, means3D, radii, features_dc, kp_score, raw_opacity, raw_scaling, raw_rotation,
                    scale_modifier, viewmatrices, projmatrices, tan_fovx, tan_fovy, dL_dout_color, dL_dout_depth,
                    dL_dout_alpha, campos, geomBuffer, R, binningBuffer, imageBuffer, out_alpha, debug=False):
    """Backward of `render_forward` (sl_backward_render): gradients of the RAW parameters.
    -> (dL_dabsmeans2D (V, P, 3), dL_dmeans2D (V, P, 3), dL_dmeans3D (P, 3), dL_dfeatures_dc (P, 3),
        dL_dkp_score (P, 1), dL_draw_opacity (P, 1), dL_draw_scaling (P, scale_dim), dL_draw_rotation (P, 4))"""
    lib = _capi.load()
    dev = means3D.device
    P = int(means3D.shape[0])
    V, n_ch, H, W = (int(x) for x in dL_dout_color.shape)
    means3D = _f32(means3D, "means3D")
    features_dc, kp_score = _f32(features_dc, "features_dc"), _f32(kp_score, "kp_score")
    raw_opacity, raw_scaling, raw_rotation = _f32(raw_opacity, "raw_opacity"), _f32(raw_scaling, "raw_scaling"), _f32(raw_rotation, "raw_rotation")
    dL_dout_color, dL_dout_depth, dL_dout_alpha = (_f32(dL_dout_color, "dL_dout_color"), _f32(dL_dout_depth, "dL_dout_depth"),
                                                   _f32(dL_dout_alpha, "dL_dout_alpha"))
    out_alpha = _f32(out_alpha, "out_alpha")
    background = _pad_bg(_f32(background, "bg"), n_ch)
    viewmatrices, projmatrices, campos = _f32(viewmatrices, "viewmatrices"), _f32(projmatrices, "projmatrices"), _f32(campos, "campos")
    scale_dim = int(raw_scaling.numel() // P) if P else 3
    with_kp = kp_score.numel() != 0
    with _on_device(dev):
        # results first (one contiguous span for a multi-GPU all-reduce), then per-view statistics, then scratch
        counts = [P * 3, P, P * scale_dim, P * 4, P * 3, P if with_kp else 0,
                  V * P * 3, V * P * 3, V * P, V * P * 4, P * n_ch, P * 6]
        flat = torch.zeros(sum((n + 3) // 4 * 4 for n in counts), dtype=torch.float32, device=dev)
        (g_mean3D, g_opac, g_scale, g_rot, g_fdc, g_kp, g_mean2D, g_abs2D, g_depth, g_conic, g_color, g_cov3D) = _carve(flat, counts)
        flat.trainable_numel = g_mean2D.storage_offset() - flat.storage_offset()
        if P != 0:
            rc = lib.sl_backward_render(
                V, P, int(R), _ptr(background), W, H, _ptr(means3D), _ptr(features_dc), _ptr(kp_score),
                _ptr(raw_opacity), _ptr(raw_scaling), scale_dim, float(scale_modifier), _ptr(raw_rotation),
                _ptr(viewmatrices), _ptr(projmatrices), _ptr(campos), float(tan_fovx), float(tan_fovy), _ptr(radii.contiguous()),
                _ptr(geomBuffer), _ptr(binningBuffer), _ptr(imageBuffer),
                _ptr(out_alpha), _ptr(dL_dout_color), _ptr(dL_dout_depth), _ptr(dL_dout_alpha),
                g_mean2D.data_ptr(), g_abs2D.data_ptr(), g_conic.data_ptr(), g_color.data_ptr(), g_depth.data_ptr(),
                g_cov3D.data_ptr(), g_mean3D.data_ptr(), g_fdc.data_ptr(), _ptr(g_kp), g_opac.data_ptr(),
                g_scale.data_ptr(), g_rot.data_ptr(), int(bool(debug)), _stream(dev))
            _capi.check(rc)
    return (g_abs2D.view(V, P, 3), g_mean2D.view(V, P, 3), g_mean3D.view(P, 3), g_fdc.view(P, 3),
            g_kp.view(P, 1) if with_kp else g_kp, g_opac.view(P, 1), g_scale.view(P, scale_dim), g_rot.view(P, 4))


def mark_visible(means3D, viewmatrix, projmatrix):
    """-> bool tensor (P,): inside the padded view frustum."""
    lib = _capi.load()
    if not means3D.is_cuda:
        raise RuntimeError("means3D must be a CUDA tensor (the rasterizer has no CPU path)")
    dev = means3D.device
    P = int(means3D.shape[0])
    present = torch.empty((P,), dtype=torch.bool, device=dev)
    if P != 0:
        means3D = _f32(means3D, "means3D")
        viewmatrix = _f32(viewmatrix, "viewmatrix")
        projmatrix = _f32(projmatrix, "projmatrix")
        with torch.cuda.device(dev):
            _capi.check(lib.sl_mark_visible(P, _ptr(means3D), _ptr(viewmatrix), _ptr(projmatrix),
                                            present.data_ptr(), _stream(dev)))
    return present


def distCUDA2(points):
    """-> float tensor (P,): mean squared distance to the 3 nearest neighbours."""
    lib = _capi.load()
    if not points.is_cuda:
        raise RuntimeError("points must be a CUDA tensor (distCUDA2 has no CPU path)")
    dev = points.device
    P = int(points.shape[0])
    means = torch.zeros((P,), dtype=torch.float32, device=dev)
    if P != 0:
        points = _f32(points, "points")
        with torch.cuda.device(dev):
            nbytes = int(lib.sl_knn_scratch_bytes(P))
            scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            _capi.check(lib.sl_knn(P, _ptr(points), means.data_ptr(), scratch.data_ptr(), nbytes, _stream(dev)))
    return means
