"""GPU: Adam with the step count on the device (sl_adam_step_counted) against the host-counted step, and the
colour-refinement iteration replayed as one CUDA graph (splatloc_b200.graphs.GraphedRefinementIteration)
against the same iteration made of eager calls -- render(fuse=True), l1_ssim_loss, backward, opt.step --
which tests/test_gpu_render.py, test_gpu_ssim.py and test_gpu_store.py pin to the reference."""
import math
import types

import pytest
import torch

pytestmark = pytest.mark.gpu

LRS = dict(xyz=1.6e-4, f_dc=2.5e-3, opacity=0.05, kp_score=0.05, scaling=1e-3, rotation=1e-3)
KEY = dict(xyz="_xyz", f_dc="_features_dc", opacity="_opacity", kp_score="_kp_score", scaling="_scaling", rotation="_rotation")


def _raw_store(P, seed=0):
    from splatloc_b200 import scenes
    sc = {k: v.cuda() for k, v in scenes.make_scene(P, seed=seed, channels=4).items()}
    C0 = 0.28209479177387814
    o = sc["opacities"].clamp(1e-4, 1 - 1e-4)
    return dict(xyz=sc["means3D"].clone(), f_dc=((sc["colors_precomp"][:, :3] - 0.5) / C0).view(P, 1, 3).contiguous(),
                kp_score=sc["colors_precomp"][:, 3:4].contiguous(), opacity=torch.log(o / (1 - o)), scaling=torch.log(sc["scales"]),
                rotation=sc["rotations"] * 1.7)


def test_adam_with_device_side_step_count_equals_host_count():
    from splatloc_b200.gaussian_optimizer import GaussianStoreOptimizer
    g = torch.Generator(device="cuda").manual_seed(1)
    P = 50000
    a = GaussianStoreOptimizer({n: t.clone() for n, t in _raw_store(P).items()}, LRS)
    b = GaussianStoreOptimizer({n: t.clone() for n, t in _raw_store(P).items()}, LRS)
    for _ in range(4):
        grads = {n: torch.randn(t.shape, device="cuda", generator=g) * 1e-3 for n, t in a.params.items()}
        a.step(grads)
        b.step_counted(grads)
    assert int(b.device_step) == 4
    for n in a.params:
        d = (a.params[n] - b.params[n]).abs().max()
        assert float(d) <= 2e-7 * max(1.0, float(a.params[n].abs().max())), (n, float(d))
        assert torch.allclose(a.exp_avg[n], b.exp_avg[n], rtol=0, atol=0) and torch.equal(a.exp_avg_sq[n], b.exp_avg_sq[n])


def test_graphed_refinement_iteration_equals_eager_iterations():
    from splatloc_b200 import scenes
    from splatloc_b200.gaussian_optimizer import GaussianStoreOptimizer
    from splatloc_b200.gaussian_renderer import render
    from splatloc_b200.graphs import GraphedRefinementIteration
    from splatloc_b200.ssim_loss import l1_ssim_loss
    P, lam, n_it = 60000, 0.2, 4
    cams = [scenes.make_camera(seed=i, size=(320, 240)).to("cuda") for i in range(2)]
    H, W = cams[0].H, cams[0].W
    g = torch.Generator().manual_seed(3)
    gts = [torch.rand((3, H, W), generator=g).cuda() for _ in cams]
    bg = torch.zeros(3, device="cuda")
    pipe = types.SimpleNamespace(convert_SHs_python=True, compute_cov3D_python=False)

    # eager: the drop-in pieces
    opt_e = GaussianStoreOptimizer({n: t.clone() for n, t in _raw_store(P).items()}, LRS)
    pc = types.SimpleNamespace(active_sh_degree=0, max_sh_degree=0, _features_rest=torch.zeros((P, 0, 3), device="cuda"))
    for n, k in KEY.items():
        setattr(pc, k, opt_e.params[n].requires_grad_(True))
    losses_e = []
    for i in range(n_it):
        c = cams[i % 2]
        view = types.SimpleNamespace(FoVx=2 * math.atan(c.tanfovx), FoVy=2 * math.atan(c.tanfovy), image_height=H, image_width=W,
                                     world_view_transform=c.viewmatrix, full_proj_transform=c.projmatrix, camera_center=c.campos)
        pkg = render(view, pc, pipe, bg)
        loss = l1_ssim_loss(pkg["render"], gts[i % 2], lam)
        loss.backward()
        losses_e.append(float(loss.detach()))
        opt_e.step({n: getattr(pc, k).grad for n, k in KEY.items() if getattr(pc, k).grad is not None and n != "kp_score"})
        for k in KEY.values():
            getattr(pc, k).grad = None

    # one replay per iteration
    opt_g = GaussianStoreOptimizer({n: t.clone() for n, t in _raw_store(P).items()}, LRS)
    it = GraphedRefinementIteration(opt_g, H, W, cams[0].tanfovx, cams[0].tanfovy, bg, max_pairs=1_500_000, lambda_dssim=lam)
    assert int(opt_g.device_step) == 0
    before = {n: opt_g.params[n].clone() for n in opt_g.params}
    losses_g = []
    for i in range(n_it):
        c = cams[i % 2]
        it.set_view(c.viewmatrix, c.projmatrix, c.campos, gts[i % 2])
        it.replay()
        losses_g.append(float(it.loss))
        assert int(it.overflow) == 0
    assert int(opt_g.device_step) == n_it
    assert torch.equal(before["kp_score"], opt_g.params["kp_score"])              # not in this loss
    assert not torch.equal(before["xyz"], opt_g.params["xyz"])
    for a, b in zip(losses_e, losses_g):
        assert abs(a - b) <= 2e-5 * max(1.0, abs(a)), (losses_e, losses_g)
    # Adam divides by sqrt(v): where a gradient is at rounding level its sign is, too -- allow 0.2 % of the elements
    for n in GraphedRefinementIteration.GROUPS:
        x, y = opt_e.params[n].detach(), opt_g.params[n]
        tol = 1e-4 * (x - before[n]).abs().max() + 1e-7 * x.abs().max()
        bad = ((x - y).abs() > tol).float().mean()
        assert float(bad) <= 2e-3, (n, float(bad))
