"""Per-view mapping loss of SplatLoc's training loop, fused with its gradient seed.

Host-side mirror of the reference's
    get_loss_mapping(config, image, depth, viewpoint, opacity, initialization=False)   utils/utils.py:55-59
    get_loss_marker(config, pred, gt)                                                  train_gaussians.py:38-42
(same names, arguments and return value: a scalar tensor to call `.backward()` on) plus
`fused_mapping_loss`, which evaluates both in ONE kernel over the rasterizer's (4, H, W) output --
what the mapping loop adds up per view (train_gaussians.py:215-216).  All arithmetic is in
libsplatloc_b200.so (`sl_mapping_loss`); there is no PyTorch fallback.
"""
from __future__ import annotations

import torch

from . import _capi
from ._C import _on_device, _ptr, _stream

_workspaces: dict = {}


def _workspace(dev: torch.device) -> torch.Tensor:
    """One zero-initialised scratch block per (device, stream): calls on one stream are ordered."""
    key = (dev.index, torch.cuda.current_stream(dev).cuda_stream)
    ws = _workspaces.get(key)
    if ws is None:
        ws = _workspaces[key] = torch.zeros(int(_capi.load().sl_mapping_loss_workspace_bytes()), dtype=torch.uint8, device=dev)
    return ws


def _as_f32(t: torch.Tensor, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (the mapping loss has no CPU path)")
    return t.detach().to(torch.float32).contiguous()


class _MappingLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, depth, exposure_a, exposure_b, gt_image, gt_depth, gt_kp, threshold, initialization):
        lib = _capi.load()
        dev = image.device
        C, H, W = (int(v) for v in image.shape)
        img, dep = _as_f32(image, "image"), _as_f32(depth, "depth")
        gt_i, gt_d = _as_f32(gt_image, "gt_image"), _as_f32(gt_depth, "gt_depth")
        gt_k = _as_f32(gt_kp, "gt_kp") if gt_kp is not None else None
        if dep.numel() != H * W or gt_i.numel() != 3 * H * W or gt_d.numel() != H * W or (gt_k is not None and gt_k.numel() != H * W):
            raise RuntimeError("mapping loss: image (C,H,W), depth (1,H,W), gt_image (3,H,W), gt_depth (H,W), gt_kp (H,W) expected")
        a = _as_f32(exposure_a, "exposure_a") if exposure_a is not None else None
        b = _as_f32(exposure_b, "exposure_b") if exposure_b is not None else None
        with _on_device(dev):
            grads = torch.empty((C + 1, H, W), dtype=torch.float32, device=dev)   # dL/dimage | dL/ddepth
            out = torch.empty(6, dtype=torch.float32, device=dev)
            _capi.check(lib.sl_mapping_loss(W, H, C, _ptr(img), _ptr(dep), _ptr(gt_i), _ptr(gt_d), _ptr(gt_k), _ptr(a), _ptr(b),
                                            float(threshold), int(bool(initialization)), grads.data_ptr(),
                                            grads[C:].data_ptr(), out.data_ptr(), _workspace(dev).data_ptr(), _stream(dev)))
        ctx.save_for_backward(grads, out)
        ctx.C = C
        ctx.terms = out
        return out[0]

    @staticmethod
    def backward(ctx, g):
        grads, out = ctx.saved_tensors
        C = ctx.C
        scaled = grads * g
        ga = gb = None
        if ctx.needs_input_grad[2]:
            ga = (out[4] * g).reshape(1)
        if ctx.needs_input_grad[3]:
            gb = (out[5] * g).reshape(1)
        return scaled[:C], scaled[C:], ga, gb, None, None, None, None, None


def fused_mapping_loss(image, depth, gt_image, gt_depth, gt_kp=None, exposure_a=None, exposure_b=None,
                       rgb_boundary_threshold=0.01, initialization=False):
    """loss = colour L1 + depth L1 (+ key-point BCE when `gt_kp` is given) for one view, as a scalar
    tensor with gradients to `image` (C,H,W; rgb in channels 0..2, the key-point logit in channel C-1),
    `depth` (1,H,W) and the exposure parameters (1-element CUDA tensors, read on the device)."""
    return _MappingLoss.apply(image, depth, exposure_a, exposure_b, gt_image, gt_depth, gt_kp,
                              rgb_boundary_threshold, initialization)


def get_loss_mapping(config, image, depth, viewpoint, opacity=None, initialization=False):
    """utils/utils.py:55-82 with the same arguments.  `image` is render()'s (3,H,W) "render"."""
    gt_depth = viewpoint.depth
    if not torch.is_tensor(gt_depth):
        gt_depth = torch.from_numpy(gt_depth)
    gt_depth = gt_depth.to(device=image.device, dtype=torch.float32)
    return fused_mapping_loss(image, depth, viewpoint.original_image.to(image.device), gt_depth, None,
                              viewpoint.exposure_a, viewpoint.exposure_b,
                              config["Training"]["rgb_boundary_threshold"], initialization)


class _MarkerLoss(torch.autograd.Function):
    """BCE(sigmoid(pred), gt).mean() alone: the key-point term of the same kernel on a 4-channel
    image whose colour part is masked out by a black ground truth."""

    @staticmethod
    def forward(ctx, pred, gt):
        H, W = (int(v) for v in pred.shape[-2:])
        dev = pred.device
        image = torch.zeros((4, H, W), dtype=torch.float32, device=dev)
        image[3] = pred.detach().reshape(H, W)
        zeros = torch.zeros((4, H, W), dtype=torch.float32, device=dev)   # depth, gt_image (3), all masked
        loss = _MappingLoss.forward(ctx, image, zeros[:1], None, None, zeros[1:], zeros[0], gt.reshape(H, W), 1e30, True)
        ctx.shape = pred.shape
        return loss

    @staticmethod
    def backward(ctx, g):
        grads, _ = ctx.saved_tensors
        return (grads[3] * g).reshape(ctx.shape), None


def get_loss_marker(config, pred, gt):
    """train_gaussians.py:38-42 with the same arguments (`pred`: render()'s (H,W) "kp_prob" logits)."""
    return _MarkerLoss.apply(pred, gt)
