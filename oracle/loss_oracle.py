"""CPU restatement of SplatLoc's per-view mapping loss (test infrastructure: only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline leg may import this; the product path never does).

Follows, term by term:
  * `get_loss_mapping` / `get_loss_mapping_rgbd`, /root/reference/utils/utils.py:55-82
      image_ab = exp(exposure_a) * image + exposure_b          (skipped when `initialization`)
      rgb mask   = gt_image.sum(0) > rgb_boundary_threshold ; depth mask = gt_depth > 0.01
      loss = mean |image_ab * m_rgb - gt_image * m_rgb| + mean |depth * m_d - gt_depth * m_d|
  * `get_loss_marker`, /root/reference/train_gaussians.py:38-42
      loss = binary_cross_entropy(sigmoid(kp_prob), gt_kp, reduction="mean")
    (PyTorch clamps each log at -100; its backward divides by max(p (1 - p), 1e-12).)
as summed per view in the mapping loop (train_gaussians.py:215-216).

Pinned: tests/golden/loss_*.npz hold outputs and autograd gradients of those two reference functions,
run on the CPU from /root/reference by tests/golden/make_loss_golden.py.
"""
from __future__ import annotations

import torch


def mapping_loss(image, depth, gt_image, gt_depth, gt_kp, exposure_a, exposure_b, rgb_boundary_threshold,
                 initialization=False, dtype=torch.float64):
    """image (C,H,W) with C >= 3 (channel C-1 is the key-point logit when gt_kp is given), depth (1,H,W),
    gt_image (3,H,W), gt_depth (H,W), gt_kp (H,W) or None.  Returns
    {loss, loss_rgb, loss_depth, loss_kp, dL_dimage (C,H,W), dL_ddepth (1,H,W), dL_dexposure (2,)}."""
    image = image.detach().to(dtype).clone().requires_grad_(True)
    depth = depth.detach().to(dtype).clone().requires_grad_(True)
    a = torch.as_tensor(exposure_a, dtype=dtype).reshape(1).clone().requires_grad_(True)
    b = torch.as_tensor(exposure_b, dtype=dtype).reshape(1).clone().requires_grad_(True)
    gt_image, gt_depth = gt_image.to(dtype), gt_depth.to(dtype)[None]

    rgb = image[:3]
    rgb_ab = rgb if initialization else torch.exp(a) * rgb + b
    m_rgb = (gt_image.sum(dim=0) > rgb_boundary_threshold).view(*depth.shape)
    m_d = (gt_depth > 0.01).view(*depth.shape)
    loss_rgb = torch.abs(rgb_ab * m_rgb - gt_image * m_rgb).mean()
    loss_depth = torch.abs(depth * m_d - gt_depth * m_d).mean()
    loss_kp = torch.zeros((), dtype=dtype)
    if gt_kp is not None:
        # the probability is a float32 quantity in the reference: logits beyond ~ +-17 saturate to exactly
        # 0 / 1 there and hit binary_cross_entropy's -100 clamp, which is part of the specified behaviour
        p = torch.sigmoid(image[-1].reshape(-1).to(torch.float32)).to(dtype)
        loss_kp = torch.nn.functional.binary_cross_entropy(p, gt_kp.reshape(-1).to(dtype), reduction="mean")
    loss = loss_rgb + loss_depth + loss_kp
    loss.backward()
    return {"loss": loss.detach(), "loss_rgb": loss_rgb.detach(), "loss_depth": loss_depth.detach(),
            "loss_kp": loss_kp.detach(), "dL_dimage": image.grad, "dL_ddepth": depth.grad,
            "dL_dexposure": torch.cat([a.grad if a.grad is not None else torch.zeros(1, dtype=dtype),
                                       b.grad if b.grad is not None else torch.zeros(1, dtype=dtype)])}


# ---- SSIM variant of the photometric loss (colour-refinement loop) -------------------------------------
# Follows /root/reference/gaussian_splatting/utils/loss_utils.py:
#   gaussian / create_window  :40-58   1-D window exp(-(x - 5)^2 / (2 * 1.5^2)) normalised, 2-D = outer product
#   ssim / _ssim              :61-102  depth-wise conv2d with zero padding 5 of {x, y, x^2, y^2, xy};
#                                      sigma = conv(.^2) - mu^2;  map = (2 mu1 mu2 + C1)(2 sigma12 + C2) /
#                                      ((mu1^2 + mu2^2 + C1)(sigma1^2 + sigma2^2 + C2)), C1 = 0.01^2, C2 = 0.03^2; mean
#   l1_loss                            mean |a - b|
# and /root/reference/train_gaussians.py:285: (1 - lambda) * L1 + lambda * (1 - ssim).
# Pinned: tests/golden/ssim_*.npz hold outputs and autograd gradients of the reference's own functions,
# run on the CPU from /root/reference by tests/golden/make_ssim_golden.py.
def ssim_window(dtype=torch.float64):
    import math
    g = torch.tensor([math.exp(-((x - 5) ** 2) / float(2 * 1.5 ** 2)) for x in range(11)], dtype=torch.float32)
    g = (g / g.sum()).to(dtype)          # the reference builds the window in float32
    return (g[:, None] @ g[None, :]).to(torch.float32).to(dtype)     # ... and the outer product too


def l1_ssim(image, gt, lambda_dssim, dtype=torch.float64):
    """image, gt (C, H, W).  Returns {ssim, l1, loss, dL_dimage} with loss = (1 - lambda) l1 + lambda (1 - ssim)."""
    import torch.nn.functional as F
    x = image.detach().to(dtype).clone().requires_grad_(True)
    y = gt.detach().to(dtype)
    C = x.shape[-3]
    w = ssim_window(dtype).expand(C, 1, 11, 11).contiguous()
    conv = lambda t: F.conv2d(t[None], w, padding=5, groups=C)[0]      # noqa: E731
    mu1, mu2 = conv(x), conv(y)
    s1 = conv(x * x) - mu1 * mu1
    s2 = conv(y * y) - mu2 * mu2
    s12 = conv(x * y) - mu1 * mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    m = ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1 * mu1 + mu2 * mu2 + C1) * (s1 + s2 + C2))
    ssim_v = m.mean()
    l1 = (x - y).abs().mean()
    loss = (1.0 - lambda_dssim) * l1 + lambda_dssim * (1.0 - ssim_v)
    loss.backward()
    return {"ssim": ssim_v.detach(), "l1": l1.detach(), "loss": loss.detach(), "dL_dimage": x.grad.detach()}
