"""Optimiser step and densification over the Gaussian store, on the GPU.

Mirrors what `GaussianModel` does between two rasterizer calls in the reference
(gaussian_splatting/scene/gaussian_model.py): `training_setup` / `optimizer.step()` (:250-300,
torch.optim.Adam with one learning rate per parameter group, eps 1e-15), `add_densification_stats`
(:677-679) with the `max_radii2D` update of train_gaussians.py:241-244, and `densify_and_prune`
(:650-675) with `densify_and_clone`, `densify_and_split`, `prune_points` and the optimizer-state
surgery of `cat_tensors_to_optimizer` / `_prune_optimizer` (:493-551).  The reference does these with
dozens of PyTorch indexing, `torch.cat` and dict operations; here each is one or two kernels of
libsplatloc_b200.so (csrc/optimizer.cu) over the store's arrays.  Same group names, same row order of
the densified store, same moments (kept rows keep theirs, new rows start at zero).

    opt = GaussianStoreOptimizer(params, lrs)           # params: the reference's eight groups, name -> tensor
    opt.step(grads)                                     # grads: name -> tensor (missing / None: group skipped)
    opt.add_densification_stats(means2D_grad, radii)    # (V, P, 3) / (V, P) or one view
    opt.densify_and_prune(0.0002, 0.7, 1.0, 20)         # params, moments, statistics replaced; returns new P
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _capi

GROUPS = ("xyz", "f_dc", "f_rest", "opacity", "marker", "kp_score", "scaling", "rotation")


def _stream(dev) -> int:
    return torch.cuda.current_stream(dev).cuda_stream


class GaussianStoreOptimizer:
    def __init__(self, params: dict, lrs: dict, betas=(0.9, 0.999), eps: float = 1e-15, percent_dense: float = 0.01,
                 primitive_reg: bool = False):
        """`params`: float32 CUDA tensors (P, ...) under the reference's group names (GROUPS; `f_rest` may
        have zero width); they are updated in place by step() and REPLACED by densify_and_prune().
        `primitive_reg`: key primitives (marker > 0.005) keep their position gradient at zero
        (train_gaussians.py:232-234) and are never pruned (gaussian_model.py:671-673)."""
        self.lib = _capi.load()
        self.params = {}
        for n in GROUPS:
            if n in params:
                t = params[n]
                if not t.is_cuda or t.dtype != torch.float32:
                    raise RuntimeError(f"{n} must be a float32 CUDA tensor (the store has no CPU path)")
                self.params[n] = t.detach().contiguous()
        for need in ("xyz", "opacity", "scaling", "rotation"):
            if need not in self.params:
                raise RuntimeError(f"parameter group {need} is required")
        self.P = int(self.params["xyz"].shape[0])
        self.dev = self.params["xyz"].device
        self.lrs = {n: float(lrs.get(n, 0.0)) for n in self.params}
        self.betas, self.eps = (float(betas[0]), float(betas[1])), float(eps)
        self.percent_dense, self.primitive_reg = float(percent_dense), bool(primitive_reg)
        self.exp_avg = {n: torch.zeros_like(t) for n, t in self.params.items()}
        self.exp_avg_sq = {n: torch.zeros_like(t) for n, t in self.params.items()}
        self.steps = {n: 0 for n in self.params}            # torch keeps one step count per parameter
        self._reset_stats()

    def _reset_stats(self):
        self.xyz_gradient_accum = torch.zeros((self.P, 1), device=self.dev)
        self.denom = torch.zeros((self.P, 1), device=self.dev)
        self.max_radii2D = torch.zeros((self.P,), device=self.dev)

    def _width(self, n) -> int:
        return int(self.params[n].numel() // self.P) if self.P else 0

    # ------------------------------------------------------------------ optimizer.step()
    def step(self, grads: dict) -> None:
        """One Adam step over every group that has a gradient, in one kernel launch per distinct step
        count (groups normally share it)."""
        by_step = {}
        for n, p in self.params.items():
            g = grads.get(n)
            if g is None or p.numel() == 0:
                continue
            if g.shape != p.shape or not g.is_cuda or g.dtype != torch.float32:
                raise RuntimeError(f"gradient of {n} must be a float32 CUDA tensor of shape {tuple(p.shape)}")
            self.steps[n] += 1
            by_step.setdefault(self.steps[n], []).append((n, g.contiguous()))
        # The reference forms its key-primitive mask BEFORE optimizer.step() (train_gaussians.py:232-234), and
        # the marker is itself a trained group of the same launch: freeze by a snapshot, not by the live tensor.
        marker = self.params.get("marker") if self.primitive_reg else None
        if marker is not None:
            marker = marker.clone()
        with torch.cuda.device(self.dev):
            for step, items in by_step.items():
                arr = (_capi.AdamGroup * len(items))()
                for k, (n, g) in enumerate(items):
                    arr[k] = _capi.AdamGroup(self.params[n].data_ptr(), g.data_ptr(), self.exp_avg[n].data_ptr(),
                                             self.exp_avg_sq[n].data_ptr(), self._width(n), self.lrs[n],
                                             1 if (n == "xyz" and marker is not None) else 0)
                _capi.check(self.lib.sl_adam_step(self.P, len(items), arr, step, self.betas[0], self.betas[1], self.eps,
                                                  None if marker is None else marker.data_ptr(), 0.005, _stream(self.dev)))

    def step_counted(self, grads: dict, names: list | None = None) -> None:
        """`step()` with the step count on the device (`self.device_step`, one int32 shared by the groups named
        -- all groups that have a gradient by default): nothing in the call depends on a host-side number that
        changes from step to step, so it can be captured in a CUDA graph and replayed.  The host-side counts
        are not advanced; do not mix with `step()` on the same groups."""
        if not hasattr(self, "device_step"):
            self.device_step = torch.zeros(1, dtype=torch.int32, device=self.dev)
        items = []
        for n in (names if names is not None else list(self.params)):
            g = grads.get(n)
            p = self.params[n]
            if g is None or p.numel() == 0:
                continue
            if g.shape != p.shape or not g.is_cuda or g.dtype != torch.float32 or not g.is_contiguous():
                raise RuntimeError(f"gradient of {n} must be a contiguous float32 CUDA tensor of shape {tuple(p.shape)}")
            items.append((n, g))
        marker = self.params.get("marker") if self.primitive_reg else None
        if marker is not None:
            marker = marker.clone()
        with torch.cuda.device(self.dev):
            arr = (_capi.AdamGroup * len(items))()
            for k, (n, g) in enumerate(items):
                arr[k] = _capi.AdamGroup(self.params[n].data_ptr(), g.data_ptr(), self.exp_avg[n].data_ptr(),
                                         self.exp_avg_sq[n].data_ptr(), self._width(n), self.lrs[n],
                                         1 if (n == "xyz" and marker is not None) else 0)
            _capi.check(self.lib.sl_adam_step_counted(self.P, len(items), arr, self.device_step.data_ptr(), self.betas[0],
                                                      self.betas[1], self.eps, None if marker is None else marker.data_ptr(),
                                                      0.005, _stream(self.dev)))

    # ------------------------------------------------------------------ add_densification_stats
    def add_densification_stats(self, means2D_grad: torch.Tensor, radii: torch.Tensor) -> None:
        """`means2D_grad` (V, P, 3) or (P, 3): the screen-space gradients of the window's views
        (`viewspace_points.grad`); `radii` (V, P) or (P,) int32: visible where > 0."""
        if means2D_grad.dim() == 2:
            means2D_grad, radii = means2D_grad[None], radii[None]
        V = int(radii.shape[0])
        means2D_grad, radii = means2D_grad.contiguous(), radii.to(torch.int32).contiguous()
        with torch.cuda.device(self.dev):
            _capi.check(self.lib.sl_densify_stats(self.P, V, means2D_grad.data_ptr(), radii.data_ptr(),
                                                  self.xyz_gradient_accum.data_ptr(), self.denom.data_ptr(),
                                                  self.max_radii2D.data_ptr(), _stream(self.dev)))

    # ------------------------------------------------------------------ densify_and_prune
    def densify_and_prune(self, max_grad: float, min_opacity: float, extent: float, max_screen_size=None,
                          noise: torch.Tensor | None = None) -> int:
        """Clone, split (N = 2) and prune in one compaction.  `noise` (2 * n_split, 3) replaces the
        standard normal draws of the split (tests replay the reference's); by default torch.randn.
        Returns the new number of Gaussians; `params`, the moments and the statistics are new tensors."""
        P, dev = self.P, self.dev
        if P == 0:
            return 0
        marker = self.params.get("marker")
        with torch.cuda.device(dev):
            nbytes = int(self.lib.sl_densify_scratch_bytes(P))
            scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            totals = torch.zeros(4, dtype=torch.int32, device=dev)
            _capi.check(self.lib.sl_densify_plan(
                P, self.xyz_gradient_accum.data_ptr(), self.denom.data_ptr(), self.params["scaling"].data_ptr(),
                self._width("scaling"), self.params["opacity"].data_ptr(), None if marker is None else marker.data_ptr(),
                float(max_grad), self.percent_dense, float(extent), float(min_opacity), 1 if max_screen_size else 0,
                1 if (self.primitive_reg and marker is not None) else 0, scratch.data_ptr(), nbytes, totals.data_ptr(),
                _stream(dev)))
            n_keep, n_clone, n_child, n_split = (int(x) for x in totals.tolist())      # the one host synchronisation
            new_P = n_keep + n_clone + 2 * n_child
            if noise is None:
                noise = torch.randn((2 * n_split, 3), device=dev)
            elif noise.shape[0] != 2 * n_split:
                raise RuntimeError(f"noise must have {2 * n_split} rows, got {noise.shape[0]}")
            noise = noise.to(dev, torch.float32).contiguous()
            names = [n for n in self.params]
            new_params = {n: torch.empty((new_P,) + tuple(self.params[n].shape[1:]), device=dev) for n in names}
            new_m = {n: torch.empty_like(new_params[n]) for n in names}
            new_v = {n: torch.empty_like(new_params[n]) for n in names}
            arr = (_capi.DensifyGroup * len(names))()
            for k, n in enumerate(names):
                arr[k] = _capi.DensifyGroup(self.params[n].data_ptr(), new_params[n].data_ptr(), self.exp_avg[n].data_ptr(),
                                            new_m[n].data_ptr(), self.exp_avg_sq[n].data_ptr(), new_v[n].data_ptr(),
                                            self._width(n))
            if new_P:
                _capi.check(self.lib.sl_densify_apply(P, len(names), arr, names.index("xyz"), names.index("scaling"),
                                                      names.index("rotation"), scratch.data_ptr(), totals.data_ptr(),
                                                      noise.data_ptr() if noise.numel() else scratch.data_ptr(), _stream(dev)))
        self.params, self.exp_avg, self.exp_avg_sq, self.P = new_params, new_m, new_v, new_P
        self.last_densify = dict(kept=n_keep, clones=n_clone, split_children_kept=n_child, split=n_split)
        self._reset_stats()
        return new_P
