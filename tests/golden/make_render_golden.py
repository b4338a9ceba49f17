"""Generates tests/golden/render_*.npz by running the REFERENCE's own `render()`
(/root/reference/gaussian_splatting/gaussian_renderer/__init__.py:12-141) and `GaussianModel` getters
(scene/gaussian_model.py:75-110) on the CPU, in this container (no GPU needed):

  * modules the reference imports but this image lacks (open3d, plyfile, simple_knn, tinycudann-backed
    decoders) are stubbed -- none of them is on the path exercised here;
  * `diff_gauss` is replaced by a recording rasterizer: its forward stores the tensors render() hands to
    the rasterizer (what the pre-processing produced), its backward returns seeded gradients for them,
    so autograd runs the reference's own graph back to the raw parameters;
  * explicit `device="cuda"` in the reference is redirected to the CPU by a TorchFunctionMode.

    python tests/golden/make_render_golden.py      (needs /root/reference; writes the .npz next to it)
"""
import os
import sys
import types

import numpy as np
import torch
from torch.overrides import TorchFunctionMode

REF = os.environ.get("SPLATLOC_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))


class CudaToCpu(TorchFunctionMode):
    def __torch_function__(self, func, types_, args=(), kwargs=None):
        kwargs = dict(kwargs or {})
        if "device" in kwargs and str(kwargs["device"]).startswith("cuda"):
            kwargs["device"] = "cpu"
        if func is torch.Tensor.cuda:
            return args[0]
        return func(*args, **kwargs)


RECORD = {}


def _stub_modules():
    for name in ("open3d", "plyfile", "simple_knn", "simple_knn._C", "models", "models.decoders"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["plyfile"].PlyData = sys.modules["plyfile"].PlyElement = object
    sys.modules["simple_knn._C"].distCUDA2 = None
    sys.modules["models.decoders"].FeatureDecoder = object
    dg = types.ModuleType("diff_gauss")
    from typing import NamedTuple

    class GaussianRasterizationSettings(NamedTuple):
        image_height: int
        image_width: int
        tanfovx: float
        tanfovy: float
        bg: torch.Tensor
        scale_modifier: float
        viewmatrix: torch.Tensor
        projmatrix: torch.Tensor
        sh_degree: int
        campos: torch.Tensor
        prefiltered: bool
        debug: bool

    class _Record(torch.autograd.Function):
        @staticmethod
        def forward(ctx, means3D, means2D, colors, opacities, scales, rotations, H, W):
            RECORD["in"] = dict(means3D=means3D.detach().clone(), colors_precomp=colors.detach().clone(),
                                opacities=opacities.detach().clone(), scales=scales.detach().clone(),
                                rotations=rotations.detach().clone())
            P, C = colors.shape
            return torch.zeros(C, H, W), torch.zeros(1, H, W), torch.zeros(1, H, W), torch.zeros(P, dtype=torch.int32)

        @staticmethod
        def backward(ctx, *_):
            g = RECORD["grads"]
            return g["means3D"], g["means2D"], g["colors_precomp"], g["opacities"], g["scales"], g["rotations"], None, None

    class GaussianRasterizer(torch.nn.Module):
        def __init__(self, raster_settings):
            super().__init__()
            self.rs = raster_settings

        def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                    cov3D_precomp=None):
            assert shs is None and cov3D_precomp is None
            RECORD["settings"] = self.rs
            return _Record.apply(means3D, means2D, colors_precomp, opacities, scales, rotations, self.rs.image_height,
                                 self.rs.image_width)

    dg.GaussianRasterizationSettings = GaussianRasterizationSettings
    dg.GaussianRasterizer = GaussianRasterizer
    sys.modules["diff_gauss"] = dg


def make_case(name, P, seed, isotropic):
    from gaussian_splatting.gaussian_renderer import render
    from gaussian_splatting.scene.gaussian_model import GaussianModel
    g = torch.Generator().manual_seed(seed)
    pc = object.__new__(GaussianModel)              # __init__ allocates CUDA tensors; the getters need none of it
    pc.active_sh_degree, pc.max_sh_degree = 0, 0
    pc.scaling_activation, pc.opacity_activation = torch.exp, torch.sigmoid
    pc.rotation_activation = torch.nn.functional.normalize
    raw = dict(
        _xyz=4.0 * torch.rand((P, 3), generator=g) - 2.0,
        _features_dc=2.5 * torch.randn((P, 1, 3), generator=g),        # wide enough that clamp_min(. + 0.5, 0) clamps some
        _features_rest=torch.zeros((P, 0, 3)),
        _kp_score=torch.rand((P, 1), generator=g),
        _opacity=3.0 * torch.randn((P, 1), generator=g),
        _scaling=torch.log(torch.tensor(0.02)) + 0.7 * torch.randn((P, 1 if isotropic else 3), generator=g),
        _rotation=torch.randn((P, 4), generator=g) * torch.exp(torch.randn((P, 1), generator=g)),   # far from unit norm
    )
    for k, v in raw.items():
        setattr(pc, k, v.clone().requires_grad_(True))
    cam = types.SimpleNamespace(FoVx=1.4, FoVy=1.1, image_height=48, image_width=64,
                                world_view_transform=torch.eye(4), full_proj_transform=torch.eye(4), camera_center=torch.zeros(3))
    pipe = types.SimpleNamespace(convert_SHs_python=True, compute_cov3D_python=False)
    RECORD["grads"] = dict(means3D=torch.randn((P, 3), generator=g), means2D=torch.randn((P, 3), generator=g),
                           colors_precomp=torch.randn((P, 4), generator=g), opacities=torch.randn((P, 1), generator=g),
                           scales=torch.randn((P, 3), generator=g), rotations=torch.randn((P, 4), generator=g))
    with CudaToCpu():
        out = render(cam, pc, pipe, torch.zeros(3))
        loss = out["render"].sum() + out["kp_prob"].sum() + out["depth"].sum() + out["opacity"].sum()
        loss.backward()
    save = {"raw" + k: v.numpy() for k, v in raw.items() if k != "_features_rest"}
    save.update({"act_" + k: v.numpy() for k, v in RECORD["in"].items()})
    save.update({"gact_" + k: v.numpy() for k, v in RECORD["grads"].items()})
    save.update({"graw" + k: getattr(pc, k).grad.numpy() for k in raw if k != "_features_rest"})
    save["viewspace_grad"] = out["viewspace_points"].grad.numpy()
    save["tanfov"] = np.array([RECORD["settings"].tanfovx, RECORD["settings"].tanfovy], dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, f"render_{name}.npz"), **save)
    print(name, {k: v.shape for k, v in save.items()})


if __name__ == "__main__":
    sys.path.insert(0, REF)
    _stub_modules()
    make_case("aniso_300", 300, 11, False)
    make_case("iso_200", 200, 12, True)
