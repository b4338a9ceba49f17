"""Generates tests/golden/store_*.npz by running the REFERENCE's own GaussianModel
(/root/reference/gaussian_splatting/scene/gaussian_model.py: training_setup :250-300 with
torch.optim.Adam, add_densification_stats :677-679, densify_and_prune :650-675 with
densify_and_clone / densify_and_split / prune_points and their optimizer-state surgery) on the CPU:
three optimiser steps on seeded gradients (with the key-primitive position freeze of
train_gaussians.py:232-234), densification statistics from two seeded views, then one
densify_and_prune.  `torch.normal` is replaced by mean + std * z with a recorded z so that the split
can be replayed; missing modules are stubbed and device="cuda" is redirected as in
make_render_golden.py.

    python tests/golden/make_store_golden.py       (needs /root/reference)
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_render_golden import REF, CudaToCpu, _stub_modules  # noqa: E402

NAMES = ("_xyz", "_features_dc", "_features_rest", "_opacity", "_marker", "_kp_score", "_scaling", "_rotation")
GROUP_OF = dict(_xyz="xyz", _features_dc="f_dc", _features_rest="f_rest", _opacity="opacity", _marker="marker",
                _kp_score="kp_score", _scaling="scaling", _rotation="rotation")


def make_case(name, P, seed, isotropic, max_screen_size, primitive_reg):
    from gaussian_splatting.scene.gaussian_model import GaussianModel
    g = torch.Generator().manual_seed(seed)
    pc = object.__new__(GaussianModel)
    pc.active_sh_degree = pc.max_sh_degree = 0
    pc.scaling_activation, pc.scaling_inverse_activation = torch.exp, torch.log
    pc.opacity_activation, pc.rotation_activation = torch.sigmoid, torch.nn.functional.normalize
    pc.primitive_reg, pc.spatial_lr_scale = primitive_reg, 1.0
    raw = dict(
        _xyz=4.0 * torch.rand((P, 3), generator=g) - 2.0,
        _features_dc=torch.randn((P, 1, 3), generator=g),
        _features_rest=torch.zeros((P, 0, 3)),
        _opacity=2.5 * torch.randn((P, 1), generator=g),
        _marker=(torch.rand((P, 1), generator=g) < 0.15).float() * torch.rand((P, 1), generator=g),
        _kp_score=torch.rand((P, 1), generator=g),
        _scaling=torch.log(torch.tensor(0.006 if not isotropic else 0.012)) + 1.0 * torch.randn((P, 1 if isotropic else 3), generator=g),
        _rotation=torch.randn((P, 4), generator=g),
    )
    for k, v in raw.items():
        setattr(pc, k, torch.nn.Parameter(v.clone()))
    pc.max_radii2D = torch.zeros(P)
    args = types.SimpleNamespace(percent_dense=0.01, position_lr_init=0.00016, position_lr_final=0.0000016,
                                 position_lr_delay_mult=0.01, position_lr_max_steps=30000, feature_lr=0.0025,
                                 opacity_lr=0.05, marker_lr=0.05, kp_score_lr=0.05, scaling_lr=0.001, rotation_lr=0.001)
    save = {"in" + k: v.numpy() for k, v in raw.items()}
    save["lrs"] = np.array([0.00016, 0.0025, 0.0025 / 20.0, 0.05, 0.05, 0.05, 0.001, 0.001], dtype=np.float64)
    noise = []

    def fake_normal(mean, std):
        z = torch.randn(mean.shape, generator=g)
        noise.append(z)
        return mean + std * z

    with CudaToCpu():
        pc.training_setup(args)
        for step in range(3):
            for k in NAMES:
                p = getattr(pc, k)
                p.grad = torch.randn(p.shape, generator=g) * (1e-3 if k == "_xyz" else 1e-2)
                save[f"grad{step}{k}"] = p.grad.numpy().copy()
            key_mask = pc.get_marker.detach().cpu().squeeze() > 0.005          # train_gaussians.py:232-234
            pc.get_xyz.grad[key_mask] = 0
            pc.optimizer.step()
            pc.optimizer.zero_grad(set_to_none=True)
        for k in NAMES:
            save["adam" + k] = getattr(pc, k).detach().numpy().copy()
            st = pc.optimizer.state.get(getattr(pc, k), None)
            if st is not None and "exp_avg" in st:
                save["adam_m" + k] = st["exp_avg"].numpy().copy()
                save["adam_v" + k] = st["exp_avg_sq"].numpy().copy()
        # two views of statistics (train_gaussians.py:239-245)
        V = 2
        radii = (torch.rand((V, P), generator=g) < 0.6).int() * torch.randint(1, 40, (V, P), generator=g).int()
        g2d = torch.randn((V, P, 3), generator=g) * 3e-4
        save["stat_radii"], save["stat_grad2D"] = radii.numpy(), g2d.numpy()
        for v in range(V):
            vis = radii[v] > 0
            pc.max_radii2D[vis] = torch.max(pc.max_radii2D[vis], radii[v][vis].float())
            pc.add_densification_stats(types.SimpleNamespace(grad=g2d[v]), vis)
        save["stat_accum"], save["stat_denom"] = pc.xyz_gradient_accum.numpy().copy(), pc.denom.numpy().copy()
        save["stat_max_radii2D"] = pc.max_radii2D.numpy().copy()
        torch_normal, torch.normal = torch.normal, fake_normal
        try:
            pc.densify_and_prune(0.0002, 0.7, 1.0, max_screen_size)
        finally:
            torch.normal = torch_normal
    save["dense_args"] = np.array([0.0002, 0.7, 1.0, 0.0 if max_screen_size is None else float(max_screen_size), 0.01,
                                   float(primitive_reg)], dtype=np.float64)
    save["noise"] = noise[0].numpy() if noise else np.zeros((0, 3), dtype=np.float32)
    for k in NAMES:
        p = getattr(pc, k)
        save["out" + k] = p.detach().numpy().copy()
        st = pc.optimizer.state.get(p, None)
        if st is not None and "exp_avg" in st:
            save["out_m" + k] = st["exp_avg"].numpy().copy()
            save["out_v" + k] = st["exp_avg_sq"].numpy().copy()
    save["out_accum"], save["out_denom"] = pc.xyz_gradient_accum.numpy().copy(), pc.denom.numpy().copy()
    save["out_max_radii2D"] = pc.max_radii2D.numpy().copy()
    np.savez_compressed(os.path.join(HERE, f"store_{name}.npz"), **save)
    print(name, "P", P, "->", save["out_xyz"].shape[0], "noise rows", save["noise"].shape[0])


if __name__ == "__main__":
    sys.path.insert(0, REF)
    _stub_modules()
    make_case("aniso_400", 400, 21, False, 20, True)
    make_case("iso_300", 300, 22, True, None, False)
