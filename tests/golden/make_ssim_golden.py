"""Generate tests/golden/ssim_*.npz by running the REFERENCE's own `ssim` and `l1_loss`
(/root/reference/gaussian_splatting/utils/loss_utils.py, pure PyTorch) on the CPU, combined as the
colour-refinement loop does (/root/reference/train_gaussians.py:285); container with /root/reference only:
the fixtures travel, the reference does not.
Usage: python tests/golden/make_ssim_golden.py
"""
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REF)
from gaussian_splatting.utils.loss_utils import l1_loss, ssim  # noqa: E402


def case(name, seed, C, H, W, lam, noise, flat_patch=False, equal_patch=False):
    g = torch.Generator().manual_seed(seed)
    gt = torch.rand((C, H, W), generator=g)
    # smooth structure so that local variances are not pure noise
    yy, xx = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
    for c in range(C):
        gt[c] = (0.5 + 0.35 * torch.sin(0.11 * (c + 1) * xx + 0.07 * yy) + 0.15 * (gt[c] - 0.5)).clamp(0, 1)
    image = (gt + noise * torch.randn((C, H, W), generator=g)).clamp(0, 1)
    if flat_patch:                       # zero local variance in both images: the C2 term carries the map
        gt[:, : H // 3, : W // 3] = 0.25
        image[:, : H // 3, : W // 3] = 0.30
    if equal_patch:                      # identical pixels: |x - y| has a zero (sub)gradient there
        image[:, -H // 4:, -W // 4:] = gt[:, -H // 4:, -W // 4:]
    x = image.clone().requires_grad_(True)
    Ll1 = l1_loss(x, gt)
    s = ssim(x, gt)
    loss = (1.0 - lam) * Ll1 + lam * (1.0 - s)
    loss.backward()
    np.savez_compressed(os.path.join(HERE, f"ssim_{name}.npz"), image=image.numpy(), gt=gt.numpy(), lambda_dssim=np.float32(lam),
                        ssim=s.detach().numpy(), l1=Ll1.detach().numpy(), loss=loss.detach().numpy(), dL_dimage=x.grad.numpy())
    print(name, float(s), float(Ll1), float(loss), float(x.grad.abs().max()))


if __name__ == "__main__":
    torch.manual_seed(0)
    case("rgb_96x64", 1, 3, 64, 96, 0.2, 0.05)
    case("odd_75x53", 2, 3, 53, 75, 0.2, 0.10, flat_patch=True, equal_patch=True)      # not a tile multiple
    case("small_9x7", 3, 3, 7, 9, 0.5, 0.20)                                            # smaller than the window
    case("c4_40x33", 4, 4, 33, 40, 1.0, 0.02)                                           # ssim term only
