"""sl_ssim_forward / sl_ssim_backward (through splatloc_b200.ssim_loss) against the reference's own outputs
(tests/golden/ssim_*.npz), the float64 oracle, and -- at the bench image size -- calibrated against what the
reference's formulation itself achieves in float32 with PyTorch's convolutions on the same GPU.

Tolerances: ssim / l1 / loss absolute 2e-6 (means of O(1) maps); gradients element-wise within 2e-4 of the
gradient's largest magnitude against the float64 oracle (sigma = conv(x^2) - mu^2 cancels in float32 on both
sides), and no more than 3x the float32 PyTorch formulation's own distance from the oracle."""
import os

import numpy as np
import pytest
import torch

from tests.test_ssim_cpu import GOLDEN, grad_close, load_ssim

pytestmark = pytest.mark.gpu


def _run(image, gt, lam):
    from splatloc_b200.ssim_loss import l1_ssim_loss
    x = image.cuda().requires_grad_(True)
    loss = l1_ssim_loss(x, gt.cuda(), lam)
    loss.backward()
    return loss.detach(), x.grad


@pytest.mark.parametrize("path", GOLDEN, ids=lambda p: os.path.basename(p)[5:-4])
def test_kernel_matches_reference_outputs(path):
    from splatloc_b200.ssim_loss import l1_loss, ssim
    g, image, gt, lam = load_ssim(path)
    loss, grad = _run(image, gt, lam)
    assert abs(float(loss) - float(g["loss"])) <= 2e-6
    assert grad_close(grad.cpu().numpy(), g["dL_dimage"], 2e-4)
    # the reference-named entry points
    x = image.cuda().requires_grad_(True)
    s, l1 = ssim(x, gt.cuda()), l1_loss(x, gt.cuda())
    assert abs(float(s.detach()) - float(g["ssim"])) <= 2e-6
    assert abs(float(l1.detach()) - float(g["l1"])) <= 2e-6 * max(1.0, float(g["l1"]))
    ((1.0 - lam) * l1 + lam * (1.0 - s)).backward()
    assert grad_close(x.grad.cpu().numpy(), g["dL_dimage"], 2e-4)


def test_bench_size_against_oracle_and_float32_formulation():
    import torch.nn.functional as F
    from oracle import loss_oracle
    g = torch.Generator().manual_seed(11)
    H, W = 480, 640
    yy, xx = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
    gt = torch.stack([(0.5 + 0.4 * torch.sin(0.05 * (c + 1) * xx + 0.03 * yy)) for c in range(3)])
    gt = (gt + 0.05 * torch.rand((3, H, W), generator=g)).clamp(0, 1)
    image = (gt + 0.08 * torch.randn((3, H, W), generator=g)).clamp(0, 1)
    lam = 0.2
    o = loss_oracle.l1_ssim(image, gt, lam)
    loss, grad = _run(image, gt, lam)
    assert abs(float(loss) - float(o["loss"])) <= 2e-6
    want = o["dL_dimage"].float().numpy()
    ours_err = float(np.abs(grad.cpu().numpy() - want).max())
    # the same formulas as PyTorch float32 ops on the GPU (what the reference executes)
    x = image.cuda().requires_grad_(True)
    y = gt.cuda()
    w = loss_oracle.ssim_window(torch.float32).cuda().expand(3, 1, 11, 11).contiguous()
    conv = lambda t: F.conv2d(t[None], w, padding=5, groups=3)[0]      # noqa: E731
    mu1, mu2 = conv(x), conv(y)
    s1, s2, s12 = conv(x * x) - mu1 * mu1, conv(y * y) - mu2 * mu2, conv(x * y) - mu1 * mu2
    m = ((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1 * mu1 + mu2 * mu2 + 1e-4) * (s1 + s2 + 9e-4))
    ((1.0 - lam) * (x - y).abs().mean() + lam * (1.0 - m.mean())).backward()
    torch_err = float(np.abs(x.grad.cpu().numpy() - want).max())
    assert ours_err <= 2e-4 * float(np.abs(want).max()), (ours_err, float(np.abs(want).max()))
    assert ours_err <= 3.0 * torch_err + 1e-12, (ours_err, torch_err)


def test_determinism_batch_and_errors():
    from splatloc_b200.ssim_loss import l1_ssim_loss, ssim
    g = torch.Generator().manual_seed(5)
    a, b = torch.rand((3, 100, 130), generator=g).cuda(), torch.rand((3, 100, 130), generator=g).cuda()
    r = []
    for _ in range(3):
        x = a.clone().requires_grad_(True)
        loss = l1_ssim_loss(x, b, 0.2)
        loss.backward()
        r.append((loss.detach().clone(), x.grad.clone()))
    assert all(torch.equal(r[0][0], q[0]) and torch.equal(r[0][1], q[1]) for q in r[1:])
    # (B, C, H, W): depth-wise window, every plane on its own; the mean runs over all elements
    x4 = torch.stack([a, b]).requires_grad_(True)
    y4 = torch.stack([b, a])
    s4 = ssim(x4, y4)
    assert abs(float(s4.detach()) - 0.5 * (float(ssim(a, b)) + float(ssim(b, a)))) <= 2e-6
    with pytest.raises(RuntimeError):
        ssim(a.cpu(), b.cpu())
    with pytest.raises(RuntimeError):
        ssim(a, b[:, :50])
    with pytest.raises(NotImplementedError):
        ssim(a, b, window_size=7)
