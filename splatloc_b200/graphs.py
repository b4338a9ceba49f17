"""CUDA-graph replay of a forward + backward step through the rasterizer, one or several views.

The reference's forward blocks the host once per call to read the pair count back
(`cuda_rasterizer/rasterizer_impl.cu:522`), so its training step cannot be captured; between that
read, the Python glue and the autograd engine a step spends more host time than the GPU needs for
its kernels.  The no-read-back forward (sl_forward_nosync / sl_forward_batch with a pair capacity)
has no such read, so the whole step -- forward, optional device -> host copy of the images, backward
-- is captured once and replayed with a single launch (there is no reference counterpart to cite).

    step = GraphedViewStep(params, H, W, tanfovx, tanfovy, bg, max_pairs=2_000_000)
    for view in views:
        step.set_camera(view.world_view_transform, view.full_proj_transform, view.camera_center)
        step.upstream.copy_(dL_dimages)          # (C + 2, H, W): d loss / d (colour | depth | alpha)
        step.replay()
        ... params[k].grad, step.color, step.depth, step.alpha, step.radii, step.overflow ...

`n_views=V` renders V cameras of the same size per replay with the batch entry points (the mapping
loop's window of views, `train_gaussians.py:195-229`): the static inputs become `viewmatrices`
(V, 4, 4), `projmatrices`, `campos` (V, 3) and `upstream_color` (V, C, H, W), `upstream_depth`,
`upstream_alpha` (V, 1, H, W); the images get a leading view dimension and `params[k].grad` is the
sum over the V views.

`step.overflow` is a (1,) int32 device tensor: 1 if the frame had more than `max_pairs` pairs (then
the images are background-only and the gradients zero); read it when convenient and redo that frame
with a plain `GaussianRasterizer`.
"""
from __future__ import annotations

import torch

from . import _C, _capi

_EMPTY = torch.Tensor([])


class GraphedViewStep:
    def __init__(self, params: dict, H: int, W: int, tanfovx: float, tanfovy: float, bg: torch.Tensor,
                 max_pairs: int, sh_degree: int = 0, scale_modifier: float = 1.0, host_images: torch.Tensor | None = None,
                 warmup: int = 3, n_views: int = 1, exchange_multicast: int | None = None,
                 exchange=None, exchange_buffer: int = 0, consume=None, exchange_mode: str = "backward",
                 exchange_push_ctas: int = 0):
        """`params`: tensors `means3D`, `opacities`, `scales`, `rotations` and `colors_precomp` (P, C) or
        `shs` (P, M, 3); after every replay their `.grad` is this step's gradient (a view of the step's
        own packed block, `self.grad_block` -- overwritten by the next replay, not accumulated).
        `max_pairs`: (tile, Gaussian) pair capacity of one replay, all views together.
        `host_images`: optional pinned tensor of V * (C + 2) * H * W floats (for V = 1: (C + 2, H, W));
        the images are copied into it inside the graph, overlapped with the backward, laid out
        [colour of all views | depth | alpha], and are complete when the replay is.
        `exchange_multicast`: multicast address of a multi-GPU exchange buffer (splatloc_b200.exchange);
        every replay then adds this step's summed gradients into all GPUs' buffers from its last kernel.
        `exchange` (an `NvlsGradientExchange`) with `exchange_buffer` b: the same push into buffer b, with the
        split-phase barrier inside the captured step -- forward | wait for every GPU's rows of the PREVIOUS
        step | `consume(exchange.rows(1 - b))` if given, then those rows are zeroed (beside the backward) |
        backward + push | arrive.  Alternate two such steps (b = 0, 1); see splatloc_b200/exchange.py.
        `exchange_mode="next_step"`: the rows leave in a kernel of their own at the start of the NEXT replay
        (of the step built with the other `exchange_buffer`), on a second stream beside its forward:
        [zero rows(b) | push the other step's rows | arrive] beside [forward | backward], then wait and
        `consume(exchange.rows(1 - b))`: the sums of the previous step.  `exchange_push_ctas`: CTAs of that
        push kernel (0: the library's default, few enough to leave the forward its SMs; < 0: no push at all,
        to measure the protocol alone)."""
        self.params = params
        self.exchange, self.exchange_buffer, self.consume = exchange, int(exchange_buffer), consume
        if exchange_mode not in ("backward", "next_step"):
            raise ValueError("exchange_mode: 'backward' or 'next_step'")
        self.exchange_mode, self.exchange_push_ctas = exchange_mode, int(exchange_push_ctas)
        if exchange is not None and exchange_mode == "backward":
            exchange_multicast = exchange.multicast_ptr(self.exchange_buffer)
        self.exchange_multicast = exchange_multicast
        dev = params["means3D"].device
        self.H, self.W, self.V = int(H), int(W), int(n_views)
        self.max_pairs = int(max_pairs)
        self.tanfovx, self.tanfovy = float(tanfovx), float(tanfovy)
        self.bg, self.sh_degree, self.scale_modifier = bg, int(sh_degree), float(scale_modifier)
        colors = params.get("colors_precomp")
        self.C = C = int(colors.shape[1]) if colors is not None else 3
        V, N = self.V, self.H * self.W
        # static inputs of the captured step
        self.viewmatrices = torch.zeros(V, 4, 4, device=dev)
        self.projmatrices = torch.zeros(V, 4, 4, device=dev)
        self.cam_positions = torch.zeros(V, 3, device=dev)
        self.upstream_flat = torch.zeros(V * (C + 2) * N, device=dev)
        self.upstream_color = self.upstream_flat[:V * C * N].view(V, C, self.H, self.W)
        self.upstream_depth = self.upstream_flat[V * C * N:V * (C + 1) * N].view(V, 1, self.H, self.W)
        self.upstream_alpha = self.upstream_flat[V * (C + 1) * N:].view(V, 1, self.H, self.W)
        if V == 1:      # single-view names: one camera, one packed (C + 2, H, W) gradient block
            self.viewmatrix, self.projmatrix, self.campos = self.viewmatrices[0], self.projmatrices[0], self.cam_positions[0]
            self.upstream = self.upstream_flat.view(C + 2, self.H, self.W)
        self.host_images = host_images
        if host_images is not None and host_images.numel() != V * (C + 2) * N:
            raise ValueError(f"host_images must hold {V * (C + 2) * N} floats")
        self._side = torch.cuda.Stream(device=dev) if host_images is not None else None
        self._zero_side = torch.cuda.Stream(device=dev) if exchange is not None else None
        self.graph = None
        self.color = self.depth = self.alpha = self.radii = self.overflow = None
        self.grad_block = None           # the packed gradient block of the captured backward (flat float32)
        self.means2D_grad = self.means2D_absgrad = None

        with torch.cuda.device(dev):
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                for _ in range(max(1, warmup)):
                    self._step()
            torch.cuda.current_stream().wait_stream(s)
            torch.cuda.synchronize()
            lib = _capi.load()
            n0 = lib.sl_launch_count()
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self._step()
            # every launch of the captured pass went through the library's launch counter
            self.kernels_per_replay = int(lib.sl_launch_count() - n0)
        self._bind_grads()

    def _step(self):
        p, V = self.params, self.V
        deferred = self.exchange if self.exchange_mode == "next_step" else None
        if deferred is not None:
            self._zero_side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self._zero_side):
                deferred.zero(self.exchange_buffer)              # read at the end of the previous replay
                if self.exchange_push_ctas >= 0:                 # the previous step's rows, from the other step's arrays
                    deferred.push_source(1 - self.exchange_buffer, self.exchange_push_ctas)
                deferred.arrive()
        shs = p.get("shs")
        colors = p.get("colors_precomp")
        shs = _EMPTY if shs is None else shs.detach()
        colors = _EMPTY if colors is None else colors.detach()
        means3D, scales, rot = p["means3D"].detach(), p["scales"].detach(), p["rotations"].detach()
        fw = _C.rasterize_gaussians_batch(
            self.bg, means3D, colors, p["opacities"].detach(), scales, rot, self.scale_modifier, _EMPTY, _EMPTY.bool(),
            self.viewmatrices, self.projmatrices, self.tanfovx, self.tanfovy, self.H, self.W, shs, self.sh_degree,
            self.cam_positions, False, False, False, self.max_pairs)
        R, color, depth, alpha, radii, geom, binning, img, status = fw
        if self.host_images is not None:
            main = torch.cuda.current_stream()
            self._side.wait_stream(main)
            with torch.cuda.stream(self._side):
                self.host_images.view(-1).copy_(color._base, non_blocking=True)
        xch = self.exchange if self.exchange_mode == "backward" else None
        if xch is not None:
            other = 1 - self.exchange_buffer
            xch.wait_previous()                      # every GPU's rows of the previous step are in rows(other)
            if self.consume is not None:
                self.consume(xch.rows(other))
            self._zero_side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self._zero_side):
                xch.zero(other)                      # nobody pushes into it before having seen this step's arrive
        bw = _C.rasterize_gaussians_backward_batch(
            self.bg, means3D, radii, colors, scales, rot, self.scale_modifier, _EMPTY, self.viewmatrices,
            self.projmatrices, self.tanfovx, self.tanfovy, self.upstream_color, self.upstream_depth, self.upstream_alpha,
            shs, self.sh_degree, self.cam_positions, geom, R, binning, img, alpha, False,
            exchange_multicast=self.exchange_multicast)
        if self.host_images is not None:
            torch.cuda.current_stream().wait_stream(self._side)
        if xch is not None:
            torch.cuda.current_stream().wait_stream(self._zero_side)
            xch.arrive()
        if deferred is not None:
            torch.cuda.current_stream().wait_stream(self._zero_side)
            deferred.wait_current()                  # every GPU's rows of the previous step have landed
            if self.consume is not None:
                self.consume(deferred.rows(1 - self.exchange_buffer))
        self._fw, self._bw = fw, bw
        if V == 1:
            self.color, self.depth, self.alpha, self.radii = color[0], depth[0], alpha[0], radii[0]
        else:
            self.color, self.depth, self.alpha, self.radii = color, depth, alpha, radii
        self.overflow = status

    def _bind_grads(self):
        """Point `.grad` of the parameters at the captured backward's outputs (views of one block)."""
        (g_abs2D, g_means2D, g_colors, g_opac, g_means3D, g_cov3D, g_sh, g_scales, g_rot) = self._bw[:9]
        self.grad_block = g_means3D._base
        self.means2D_grad = g_means2D[0] if self.V == 1 else g_means2D
        self.means2D_absgrad = g_abs2D[0] if self.V == 1 else g_abs2D
        p = self.params
        pairs = [("means3D", g_means3D), ("opacities", g_opac), ("scales", g_scales), ("rotations", g_rot),
                 ("colors_precomp", g_colors), ("shs", g_sh)]
        for name, g in pairs:
            t = p.get(name)
            if t is not None and t.requires_grad:
                t.grad = g.view(t.shape)
        if self.exchange is not None and self.exchange_mode == "next_step":
            if p.get("colors_precomp") is None:
                raise ValueError("the exchange rows hold colors_precomp gradients (no SH)")
            self.bind_exchange_source()

    def bind_exchange_source(self) -> None:
        """`exchange_mode="next_step"`: make this step's gradient arrays the ones the OTHER step's replay pushes
        into buffer `exchange_buffer` (done at construction; call again after another pair of steps has been
        built on the same exchange)."""
        (_, _, g_colors, g_opac, g_means3D, _, _, g_scales, g_rot) = self._bw[:9]
        self.exchange.bind_source(self.exchange_buffer, self._fw[4], g_means3D, g_opac, g_scales, g_rot, g_colors)

    def set_camera(self, viewmatrix: torch.Tensor, projmatrix: torch.Tensor, campos: torch.Tensor, view: int = 0) -> None:
        """Copy a camera (reference convention: transposed world-to-view, full projection, centre) into
        the static inputs of view `view`; host or device tensors."""
        self.viewmatrices[view].copy_(viewmatrix, non_blocking=True)
        self.projmatrices[view].copy_(projmatrix, non_blocking=True)
        self.cam_positions[view].copy_(campos, non_blocking=True)

    def replay(self) -> None:
        self.graph.replay()

    def trainable_grads(self) -> torch.Tensor:
        """The gradients of means3D, opacity, scales, rotations and colours / SH of the last replay as ONE
        flat tensor (the leading span of the step's packed block): what a multi-GPU step all-reduces."""
        return self.grad_block[:self.grad_block.trainable_numel]


class GraphedRefinementIteration:
    """One iteration of SplatLoc's colour-refinement loop (train_gaussians.py:270-295) -- render, the
    (1 - lambda) L1 + lambda (1 - SSIM) loss, backward, Adam -- captured once and replayed with one launch.

    Everything between the optimiser's raw parameters and their update runs inside the library: render()'s
    pre-processing and the rasterizer (sl_forward_render without read-back, sl_backward_render), the loss and
    its gradient (sl_ssim_forward / sl_ssim_backward), the Adam step of all groups with the step count on the
    device (sl_adam_step_counted).  No reference counterpart: the reference's iteration is ~150 launches and
    one host synchronisation.

        it = GraphedRefinementIteration(opt, H, W, tanfovx, tanfovy, bg, max_pairs=2_000_000, lambda_dssim=0.2)
        for view in views:
            it.set_view(view.world_view_transform, view.full_proj_transform, view.camera_center, view.original_image)
            it.replay()
        ... it.loss (device scalar of the last replay), it.image (3, H, W), it.radii, it.overflow ...

    `opt` is a `GaussianStoreOptimizer` with groups xyz, f_dc, opacity, scaling, rotation (and kp_score, which
    this loss does not train); its tensors are updated in place by every replay."""

    GROUPS = ("xyz", "f_dc", "opacity", "scaling", "rotation")

    def __init__(self, opt, H: int, W: int, tanfovx: float, tanfovy: float, bg: torch.Tensor, max_pairs: int,
                 lambda_dssim: float = 0.2, scale_modifier: float = 1.0, warmup: int = 2):
        from .ssim_loss import _WINDOW
        self.opt, self.H, self.W = opt, int(H), int(W)
        self.tanfovx, self.tanfovy, self.bg = float(tanfovx), float(tanfovy), bg
        self.max_pairs, self.lam, self.scale_modifier = int(max_pairs), float(lambda_dssim), float(scale_modifier)
        self._window = _WINDOW
        dev = opt.params["xyz"].device
        for n in self.GROUPS:
            if n not in opt.params:
                raise ValueError(f"the optimiser has no group {n}")
        self.viewmatrix = torch.zeros(1, 4, 4, device=dev)
        self.projmatrix = torch.zeros(1, 4, 4, device=dev)
        self.campos = torch.zeros(1, 3, device=dev)
        self.gt_image = torch.zeros(3, self.H, self.W, device=dev)
        self.loss = self.image = self.radii = self.overflow = None
        lib = _capi.load()
        with torch.cuda.device(dev):
            self._ws = torch.zeros(int(lib.sl_ssim_workspace_bytes(self.W, self.H, 3)), dtype=torch.uint8, device=dev)
            if not hasattr(opt, "device_step"):
                opt.device_step = torch.zeros(1, dtype=torch.int32, device=dev)
            # warm-up runs would move the parameters: they are put back before the capture
            saved = {n: (opt.params[n].detach().clone(), opt.exp_avg[n].clone(), opt.exp_avg_sq[n].clone()) for n in self.GROUPS}
            step0 = opt.device_step.clone()
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                for _ in range(max(1, warmup)):
                    self._iteration()
            torch.cuda.current_stream().wait_stream(s)
            torch.cuda.synchronize()
            n0 = lib.sl_launch_count()
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self._iteration()
            self.kernels_per_replay = int(lib.sl_launch_count() - n0)
            with torch.no_grad():
                for n, (p, m, v) in saved.items():
                    opt.params[n].copy_(p)
                    opt.exp_avg[n].copy_(m)
                    opt.exp_avg_sq[n].copy_(v)
                opt.device_step.copy_(step0)
            torch.cuda.synchronize()

    def _iteration(self):
        lib, o = _capi.load(), self.opt
        p = {n: o.params[n].detach() for n in o.params}
        kp = p.get("kp_score", _EMPTY)
        dev = self.gt_image.device
        fw = _C.render_forward(self.bg, p["xyz"], p["f_dc"], kp, p["opacity"], p["scaling"], p["rotation"], self.scale_modifier,
                               None, self.viewmatrix, self.projmatrix, self.tanfovx, self.tanfovy, self.H, self.W, self.campos,
                               False, self.max_pairs)
        R, color, depth, alpha, radii, geom, binning, img, status = fw
        C, n = int(color.shape[1]), float(3 * self.H * self.W)
        image = color[0, :3]
        maps = torch.empty((3, 3, self.H, self.W), device=dev)
        out = torch.empty(2, device=dev)
        g_color = torch.zeros_like(color)             # the key-point channel, depth and alpha are not in this loss
        g_rgb = g_color[0, :3]
        stream = torch.cuda.current_stream(dev).cuda_stream
        _capi.check(lib.sl_ssim_forward(self.W, self.H, 3, image.data_ptr(), self.gt_image.data_ptr(), self._window,
                                        maps.data_ptr(), out.data_ptr(), self._ws.data_ptr(), stream))
        _capi.check(lib.sl_ssim_backward(self.W, self.H, 3, image.data_ptr(), self.gt_image.data_ptr(), self._window,
                                         maps.data_ptr(), None, -self.lam / n, (1.0 - self.lam) / n, g_rgb.data_ptr(), stream))
        bw = _C.render_backward(self.bg, p["xyz"], radii, p["f_dc"], kp, p["opacity"], p["scaling"], p["rotation"],
                                self.scale_modifier, self.viewmatrix, self.projmatrix, self.tanfovx, self.tanfovy, g_color,
                                torch.zeros_like(depth), torch.zeros_like(alpha), self.campos, geom, R, binning, img, alpha)
        (_, g_means2D, g_xyz, g_fdc, _, g_opac, g_scale, g_rot) = bw
        grads = {"xyz": g_xyz.view(o.params["xyz"].shape), "f_dc": g_fdc.view(o.params["f_dc"].shape),
                 "opacity": g_opac.view(o.params["opacity"].shape), "scaling": g_scale.view(o.params["scaling"].shape),
                 "rotation": g_rot.view(o.params["rotation"].shape)}
        o.step_counted({k: v.contiguous() for k, v in grads.items()}, names=list(self.GROUPS))
        self.loss = (1.0 - self.lam) * out[1] + self.lam * (1.0 - out[0])
        self.image, self.radii, self.overflow, self.means2D_grad = image, radii[0], status, g_means2D[0]
        self._keep = (fw, bw, maps, out, g_color)

    def set_view(self, viewmatrix, projmatrix, campos, gt_image) -> None:
        """Copy a camera (reference convention: transposed world-to-view, full projection, centre) and its
        ground-truth image (3, H, W) into the static inputs; host or device tensors."""
        self.viewmatrix[0].copy_(viewmatrix, non_blocking=True)
        self.projmatrix[0].copy_(projmatrix, non_blocking=True)
        self.campos[0].copy_(campos, non_blocking=True)
        self.gt_image.copy_(gt_image, non_blocking=True)

    def replay(self) -> None:
        self.graph.replay()
