"""Photometric loss of SplatLoc's colour-refinement loop, fused with its gradient (SURVEY.md 8(f) row 2,
SSIM variant).

Host-side mirror of the reference's
    ssim(img1, img2, window_size=11, size_average=True)     gaussian_splatting/utils/loss_utils.py:61-102
    l1_loss(network_output, gt)                             gaussian_splatting/utils/loss_utils.py
(same names, arguments and return value: a scalar tensor to call `.backward()` on) plus `l1_ssim_loss`,
which evaluates what the refinement loop adds up (train_gaussians.py:283-286),
    (1 - lambda_dssim) * l1_loss(image, gt) + lambda_dssim * (1 - ssim(image, gt)),
with ONE forward and ONE backward kernel instead of five depth-wise convolutions, ~15 elementwise
kernels and their autograd mirrors.  All arithmetic is in libsplatloc_b200.so (`sl_ssim_forward`,
`sl_ssim_backward`); there is no PyTorch fallback.  Gradients flow to the first image only (the second
is the ground truth).
"""
from __future__ import annotations

import ctypes
import math

import torch

from . import _capi
from ._C import _on_device, _stream


def gaussian_window(window_size: int = 11, sigma: float = 1.5) -> torch.Tensor:
    """The reference's 1-D window (loss_utils.py:40-46): exp in double, stored and normalised in float32."""
    g = torch.tensor([math.exp(-((x - window_size // 2) ** 2) / float(2 * sigma ** 2)) for x in range(window_size)],
                     dtype=torch.float32)
    return g / g.sum()


_WINDOW = (ctypes.c_float * 11)(*[float(v) for v in gaussian_window()])


_workspaces: dict = {}


def _workspace(dev: torch.device, nbytes: int) -> torch.Tensor:
    """One scratch block per (device, stream), zeroed once (the kernel leaves its ticket zero): calls on one
    stream are ordered."""
    key = (dev.index, torch.cuda.current_stream(dev).cuda_stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = _workspaces[key] = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
    return ws


def _as_f32(t: torch.Tensor, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (the SSIM loss has no CPU path)")
    return t.detach().to(torch.float32).contiguous()


def _planes(img: torch.Tensor) -> tuple[int, int, int]:
    if img.dim() not in (3, 4):
        raise RuntimeError(f"image must be (C, H, W) or (B, C, H, W), got {tuple(img.shape)}")
    H, W = int(img.shape[-2]), int(img.shape[-1])
    return int(img.numel() // (H * W)), H, W        # the window is depth-wise: batch and channels are all planes


class _L1Ssim(torch.autograd.Function):
    """-> (mean ssim, mean |img1 - img2|, weight_ssim * mean ssim + weight_l1 * mean L1 + offset)."""

    @staticmethod
    def forward(ctx, img1, img2, weight_ssim, weight_l1, offset):
        if img1.shape != img2.shape:
            raise RuntimeError(f"image shapes differ: {tuple(img1.shape)} vs {tuple(img2.shape)}")
        lib = _capi.load()
        dev = img1.device
        C, H, W = _planes(img1)
        a, b = _as_f32(img1, "img1"), _as_f32(img2, "img2")
        with _on_device(dev):
            maps = torch.empty((3, C, H, W), dtype=torch.float32, device=dev)
            out = torch.empty(2, dtype=torch.float32, device=dev)
            ws = _workspace(dev, int(lib.sl_ssim_workspace_bytes(W, H, C)))
            _capi.check(lib.sl_ssim_forward(W, H, C, a.data_ptr(), b.data_ptr(), _WINDOW, maps.data_ptr(), out.data_ptr(),
                                            ws.data_ptr(), _stream(dev)))
        ctx.save_for_backward(a, b, maps)
        ctx.weights = (float(weight_ssim), float(weight_l1))
        ctx.shape = tuple(img1.shape)
        return out[0] * weight_ssim + out[1] * weight_l1 + offset

    @staticmethod
    def backward(ctx, grad):
        a, b, maps = ctx.saved_tensors
        lib = _capi.load()
        dev = a.device
        C, H, W = _planes(a)
        n = float(C * H * W)
        up = grad.detach().to(torch.float32).reshape(1).contiguous()
        with _on_device(dev):
            g = torch.empty_like(a)
            _capi.check(lib.sl_ssim_backward(W, H, C, a.data_ptr(), b.data_ptr(), _WINDOW, maps.data_ptr(), up.data_ptr(),
                                             ctx.weights[0] / n, ctx.weights[1] / n, g.data_ptr(), _stream(dev)))
        return g.view(ctx.shape), None, None, None, None


def ssim(img1: torch.Tensor, img2: torch.Tensor, window_size: int = 11, size_average: bool = True) -> torch.Tensor:
    """loss_utils.py:61-70 (the 11-tap window and the mean over all elements are what SplatLoc uses)."""
    if window_size != 11 or not size_average:
        raise NotImplementedError("ssim: window_size 11 and size_average=True (SplatLoc's call) are built")
    return _L1Ssim.apply(img1, img2, 1.0, 0.0, 0.0)


def l1_loss(network_output: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    """loss_utils.py: torch.abs(network_output - gt).mean()."""
    return _L1Ssim.apply(network_output, gt, 0.0, 1.0, 0.0)


def l1_ssim_loss(image: torch.Tensor, gt_image: torch.Tensor, lambda_dssim: float) -> torch.Tensor:
    """train_gaussians.py:285: (1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image, gt)), one kernel each way."""
    lam = float(lambda_dssim)
    return _L1Ssim.apply(image, gt_image, -lam, 1.0 - lam, lam)
