"""CUDA-graph replay of one view's forward + backward through the rasterizer.

The reference's forward blocks the host once per call to read the pair count back
(`cuda_rasterizer/rasterizer_impl.cu:522`), so its training step cannot be captured; between that
read, the Python glue and the autograd engine a step spends more host time than the GPU needs for
its kernels.  `GaussianRasterizer(..., max_pairs=N)` (sl_forward_nosync) has no such read, so the
whole step -- forward, optional device -> host copy of the images, backward -- is captured once and
replayed with a single launch (there is no reference counterpart to cite).

    step = GraphedViewStep(params, H, W, tanfovx, tanfovy, bg, max_pairs=2_000_000)
    for view in views:
        step.set_camera(view.world_view_transform, view.full_proj_transform, view.camera_center)
        step.upstream.copy_(dL_dimages)          # (C + 2, H, W): d loss / d (colour | depth | alpha)
        step.replay()
        ... params[k].grad, step.color, step.depth, step.alpha, step.radii, step.overflow ...

`step.overflow` is a (1,) int32 device tensor: 1 if the frame had more than `max_pairs` pairs (then
the images are background-only and the gradients zero); read it when convenient and redo that frame
with a plain `GaussianRasterizer`.
"""
from __future__ import annotations

import torch

from .diff_gauss import GaussianRasterizationSettings, GaussianRasterizer


class GraphedViewStep:
    def __init__(self, params: dict, H: int, W: int, tanfovx: float, tanfovy: float, bg: torch.Tensor,
                 max_pairs: int, sh_degree: int = 0, scale_modifier: float = 1.0, host_images: torch.Tensor | None = None,
                 warmup: int = 3):
        """`params`: leaf tensors `means3D`, `opacities`, `scales`, `rotations` and `colors_precomp` (P, C)
        or `shs` (P, M, 3); those that require grad receive `.grad` on every replay.
        `host_images`: optional pinned (C + 2, H, W) tensor; the images are copied into it inside the
        graph, overlapped with the backward, and are complete when the replay is."""
        self.params = params
        dev = params["means3D"].device
        self.H, self.W = int(H), int(W)
        self.max_pairs = int(max_pairs)
        colors = params.get("colors_precomp")
        self.C = int(colors.shape[1]) if colors is not None else 3
        # static inputs of the captured step
        self.viewmatrix = torch.zeros(4, 4, device=dev)
        self.projmatrix = torch.zeros(4, 4, device=dev)
        self.campos = torch.zeros(3, device=dev)
        self.upstream = torch.zeros(self.C + 2, self.H, self.W, device=dev)
        self.means2D = torch.zeros_like(params["means3D"], requires_grad=True)
        self.host_images = host_images
        self._settings = GaussianRasterizationSettings(self.H, self.W, tanfovx, tanfovy, bg, scale_modifier,
                                                       self.viewmatrix, self.projmatrix, sh_degree, self.campos, False, False)
        self._side = torch.cuda.Stream(device=dev) if host_images is not None else None
        self.graph = None
        self.color = self.depth = self.alpha = self.radii = self.overflow = None

        with torch.cuda.device(dev):
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                for _ in range(max(1, warmup)):
                    self._clear_grads()
                    self._step()
            torch.cuda.current_stream().wait_stream(s)
            torch.cuda.synchronize()
            self._clear_grads()
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self.color, self.depth, self.alpha, self.radii, self.overflow = self._step()

    def _clear_grads(self):
        for t in list(self.params.values()) + [self.means2D]:
            if t is not None and t.requires_grad:
                t.grad = None

    def _step(self):
        p = self.params
        rast = GaussianRasterizer(self._settings, max_pairs=self.max_pairs)
        color, depth, alpha, radii = rast(means3D=p["means3D"], means2D=self.means2D, opacities=p["opacities"],
                                          shs=p.get("shs"), colors_precomp=p.get("colors_precomp"),
                                          scales=p["scales"], rotations=p["rotations"])
        if self.host_images is not None:
            main = torch.cuda.current_stream()
            self._side.wait_stream(main)
            with torch.cuda.stream(self._side):
                C = self.C
                self.host_images[:C].copy_(color.detach(), non_blocking=True)
                self.host_images[C:C + 1].copy_(depth.detach(), non_blocking=True)
                self.host_images[C + 1:].copy_(alpha.detach(), non_blocking=True)
        C = self.C
        torch.autograd.backward([color, depth, alpha], [self.upstream[:C], self.upstream[C:C + 1], self.upstream[C + 1:]])
        if self.host_images is not None:
            torch.cuda.current_stream().wait_stream(self._side)
        return color, depth, alpha, radii, rast.overflow

    def set_camera(self, viewmatrix: torch.Tensor, projmatrix: torch.Tensor, campos: torch.Tensor) -> None:
        """Copy a camera (reference convention: transposed world-to-view, full projection, centre) into
        the step's static inputs; host or device tensors."""
        self.viewmatrix.copy_(viewmatrix, non_blocking=True)
        self.projmatrix.copy_(projmatrix, non_blocking=True)
        self.campos.copy_(campos, non_blocking=True)

    def replay(self) -> None:
        self.graph.replay()

    def trainable_grads(self) -> torch.Tensor:
        """The gradients of means3D, opacity, scales, rotations and colours / SH of the last replay as ONE
        flat tensor (they are the leading blocks of a single allocation): what a multi-GPU step
        all-reduces.  Requires all of them to be leaves of this step."""
        from ._C import trainable_grad_numel
        g = self.params["means3D"].grad
        shs = self.params.get("shs")
        M = int(shs.shape[1]) if shs is not None else 0
        n = trainable_grad_numel(int(g.shape[0]), self.C if shs is None else 3, M)
        return g.as_strided((n,), (1,), g.storage_offset())
