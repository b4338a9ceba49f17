"""Generate tests/golden/loss_*.npz by running the REFERENCE's own loss functions on the CPU
(container with /root/reference only; the fixtures travel, the reference does not).

  get_loss_mapping           imported from /root/reference/utils/utils.py (it only needs torch)
  get_loss_marker            its five lines are compiled from /root/reference/train_gaussians.py by
                             ast (the module itself imports OpenGL / open3d and cannot be imported)

The reference calls `.cuda()` on the ground-truth image; the stand-in viewpoint returns the CPU tensor.
Usage: python tests/golden/make_loss_golden.py
"""
import ast
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REF)
from utils.utils import get_loss_mapping  # noqa: E402

src = open(os.path.join(REF, "train_gaussians.py")).read()
fn = next(n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == "get_loss_marker")
ns = {"torch": torch}
exec(compile(ast.Module(body=[fn], type_ignores=[]), "train_gaussians.py", "exec"), ns)
get_loss_marker = ns["get_loss_marker"]


class _OnHost:
    def __init__(self, t):
        self.t = t

    def cuda(self):
        return self.t


class Viewpoint:
    pass


def case(name, seed, H, W, a, b, initialization, with_kp=True, black_border=True, holes=True):
    g = torch.Generator().manual_seed(seed)
    image4 = torch.rand((4, H, W), generator=g)
    image4[3] = 6.0 * torch.randn((H, W), generator=g)          # key-point logits, some saturated
    image4[3, 0, :4] = torch.tensor([40.0, -40.0, 120.0, -120.0])[: min(4, W)]
    depth = 0.5 + 4.0 * torch.rand((1, H, W), generator=g)
    gt_image = (image4[:3] + 0.2 * torch.randn((3, H, W), generator=g)).clamp(0, 1)
    gt_image[:, 1, 1] = image4[:3, 1, 1]                        # an exact zero residual
    if black_border:
        gt_image[:, :, : max(1, W // 10)] = 0.0                  # below rgb_boundary_threshold
    gt_depth = (depth[0] + 0.1 * torch.randn((H, W), generator=g)).clamp_min(0.02)
    if holes:
        gt_depth[torch.rand((H, W), generator=g) < 0.2] = 0.0    # invalid depth
    gt_kp = (torch.rand((H, W), generator=g) < 0.05).float()
    thr = 0.01

    vp = Viewpoint()
    vp.original_image = _OnHost(gt_image)
    vp.depth = gt_depth.numpy()
    vp.exposure_a = torch.tensor([a], requires_grad=True)
    vp.exposure_b = torch.tensor([b], requires_grad=True)
    config = {"Training": {"rgb_boundary_threshold": thr}}

    img = image4.clone().requires_grad_(True)
    dep = depth.clone().requires_grad_(True)
    loss_map = get_loss_mapping(config, img[:3], dep, vp, None, initialization=initialization)
    loss_kp = get_loss_marker(config, img[-1], gt_kp) if with_kp else torch.zeros(())
    loss = loss_map + loss_kp
    loss.backward()
    z = torch.zeros(1)
    out = dict(in_image=image4.numpy(), in_depth=depth.numpy(), in_gt_image=gt_image.numpy(), in_gt_depth=gt_depth.numpy(),
               in_gt_kp=gt_kp.numpy() if with_kp else np.zeros((0,), np.float32), in_exposure=np.array([a, b], np.float32),
               in_threshold=np.float32(thr), in_initialization=np.int32(initialization),
               ref_loss=loss.detach().numpy(), ref_loss_mapping=loss_map.detach().numpy(), ref_loss_kp=loss_kp.detach().numpy(),
               ref_dL_dimage=img.grad.numpy(), ref_dL_ddepth=dep.grad.numpy(),
               ref_dL_dexposure=torch.cat([vp.exposure_a.grad if vp.exposure_a.grad is not None else z,
                                           vp.exposure_b.grad if vp.exposure_b.grad is not None else z]).numpy())
    np.savez_compressed(os.path.join(HERE, f"loss_{name}.npz"), **out)
    print(name, float(loss), float(loss_map), float(loss_kp))


if __name__ == "__main__":
    case("map_96x64", 1, 64, 96, 0.0, 0.0, False)
    case("exposure_80x60", 2, 60, 80, 0.15, -0.03, False)
    case("init_50x37", 3, 37, 50, 0.3, 0.1, True)                 # initialization: exposure ignored
    case("nokp_64x48", 4, 48, 64, -0.1, 0.02, False, with_kp=False)
