"""sl_mapping_loss (through splatloc_b200.mapping_loss) against the reference's own outputs
(tests/golden/loss_*.npz), the float64 oracle, and the reference formulas run with PyTorch ops on
the GPU at the bench image size.  Tolerances: loss rtol 2e-6; gradients rtol 1e-5 (float32 kernels
on both sides; sums are accumulated in double here, in float32 by PyTorch)."""
import os

import numpy as np
import pytest
import torch

from tests.test_loss_cpu import GOLDEN, load_loss

pytestmark = pytest.mark.gpu


def _run(args, dev="cuda"):
    from splatloc_b200.mapping_loss import fused_mapping_loss
    image = args["image"].to(dev).requires_grad_(True)
    depth = args["depth"].to(dev).requires_grad_(True)
    a = torch.tensor([args["exposure_a"]], device=dev, requires_grad=True)
    b = torch.tensor([args["exposure_b"]], device=dev, requires_grad=True)
    loss = fused_mapping_loss(image, depth, args["gt_image"].to(dev), args["gt_depth"].to(dev),
                              args["gt_kp"].to(dev) if args["gt_kp"] is not None else None, a, b,
                              args["rgb_boundary_threshold"], args["initialization"])
    loss.backward()
    z = torch.zeros(1, device=dev)
    return loss.detach(), image.grad, depth.grad, torch.cat([a.grad if a.grad is not None else z, b.grad if b.grad is not None else z])


@pytest.mark.parametrize("path", GOLDEN, ids=lambda p: os.path.basename(p)[5:-4])
def test_kernel_matches_reference_outputs(path):
    g, args = load_loss(path)
    loss, gi, gd, ge = _run(args)
    assert abs(float(loss) - float(g["ref_loss"])) <= 2e-6 * abs(float(g["ref_loss"])) + 1e-7
    np.testing.assert_allclose(gi.cpu().numpy(), g["ref_dL_dimage"], rtol=1e-5, atol=1e-9)
    np.testing.assert_allclose(gd.cpu().numpy(), g["ref_dL_ddepth"], rtol=1e-5, atol=1e-9)
    np.testing.assert_allclose(ge.cpu().numpy(), g["ref_dL_dexposure"], rtol=2e-4, atol=1e-7)


def test_reference_named_entry_points_and_determinism():
    """get_loss_mapping / get_loss_marker take the reference's arguments; two runs give identical bits."""
    from splatloc_b200 import mapping_loss as ML
    g, args = load_loss(GOLDEN[0])

    class Viewpoint:
        pass
    vp = Viewpoint()
    vp.original_image = args["gt_image"].cuda()
    vp.depth = args["gt_depth"].numpy()
    vp.exposure_a = torch.tensor([args["exposure_a"]], device="cuda", requires_grad=True)
    vp.exposure_b = torch.tensor([args["exposure_b"]], device="cuda", requires_grad=True)
    config = {"Training": {"rgb_boundary_threshold": args["rgb_boundary_threshold"]}}
    image = args["image"].cuda().requires_grad_(True)
    depth = args["depth"].cuda().requires_grad_(True)
    total = ML.get_loss_mapping(config, image[:3], depth, vp, None, initialization=args["initialization"]) \
        + ML.get_loss_marker(config, image[-1], args["gt_kp"].cuda())
    total.backward()
    assert abs(float(total) - float(g["ref_loss"])) <= 2e-6 * abs(float(g["ref_loss"]))
    np.testing.assert_allclose(image.grad.cpu().numpy(), g["ref_dL_dimage"], rtol=1e-5, atol=1e-9)
    np.testing.assert_allclose(depth.grad.cpu().numpy(), g["ref_dL_ddepth"], rtol=1e-5, atol=1e-9)
    first = _run(args)
    second = _run(args)
    assert all(torch.equal(x, y) for x, y in zip(first, second))


def test_full_size_against_torch_ops_on_gpu():
    """640x480, C = 4: the reference's formulas as PyTorch ops on the same GPU."""
    H, W = 480, 640
    g = torch.Generator().manual_seed(11)
    image = torch.rand((4, H, W), generator=g)
    image[3] = 4.0 * torch.randn((H, W), generator=g)
    depth = 0.5 + 4.0 * torch.rand((1, H, W), generator=g)
    gt_image = (image[:3] + 0.1 * torch.randn((3, H, W), generator=g)).clamp(0, 1)
    gt_image[:, :, :40] = 0
    gt_depth = depth[0] + 0.05 * torch.randn((H, W), generator=g)
    gt_depth[torch.rand((H, W), generator=g) < 0.1] = 0
    gt_kp = (torch.rand((H, W), generator=g) < 0.03).float()
    args = dict(image=image, depth=depth, gt_image=gt_image, gt_depth=gt_depth, gt_kp=gt_kp, exposure_a=0.07,
                exposure_b=-0.01, rgb_boundary_threshold=0.01, initialization=False)
    loss, gi, gd, ge = _run(args)

    img = image.cuda().requires_grad_(True)
    dep = depth.cuda().requires_grad_(True)
    a = torch.tensor([0.07], device="cuda", requires_grad=True)
    b = torch.tensor([-0.01], device="cuda", requires_grad=True)
    gti, gtd = gt_image.cuda(), gt_depth.cuda()[None]
    m_rgb = (gti.sum(dim=0) > 0.01).view(*dep.shape)
    m_d = (gtd > 0.01).view(*dep.shape)
    ref = torch.abs((torch.exp(a) * img[:3] + b) * m_rgb - gti * m_rgb).mean() + torch.abs(dep * m_d - gtd * m_d).mean() \
        + torch.nn.functional.binary_cross_entropy(torch.sigmoid(img[-1].view(-1)), gt_kp.cuda().view(-1), reduction="mean")
    ref.backward()
    assert abs(float(loss) - float(ref)) <= 5e-6 * abs(float(ref))
    np.testing.assert_allclose(gi.cpu().numpy(), img.grad.cpu().numpy(), rtol=1e-5, atol=1e-10)
    np.testing.assert_allclose(gd.cpu().numpy(), dep.grad.cpu().numpy(), rtol=1e-5, atol=1e-10)
    np.testing.assert_allclose(ge.cpu().numpy(), torch.cat([a.grad, b.grad]).cpu().numpy(), rtol=1e-3, atol=1e-7)
