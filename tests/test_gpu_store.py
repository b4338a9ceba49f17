"""GPU: the optimiser step and densification kernels (splatloc_b200.gaussian_optimizer, csrc/optimizer.cu;
SURVEY 8(f) row 4) against (1) outputs of the reference's own GaussianModel run on the CPU
(tests/golden/store_*.npz) and (2) the oracle (oracle/store_oracle.py, pinned to the same fixtures)
on a seeded 200 k-Gaussian store.  Row order, counts and which rows keep their Adam moments must be
exact; values within float32 rounding of a handful of operations (rtol 5e-6)."""
import numpy as np
import pytest
import torch

from oracle import store_oracle as S
from tests.test_store_cpu import CASES, KEY, initial_state, load

pytestmark = pytest.mark.gpu


def _close(a, b, rtol=5e-6, atol=None):
    """rtol on the element, plus an absolute floor of a few float32 ulps of the tensor's largest magnitude
    (an update of size lr lands on parameters near zero with the rounding of the O(1) operands)."""
    a, b = a.detach().double().cpu(), torch.as_tensor(b).double()
    if a.shape != b.shape:
        return False
    if atol is None:
        atol = 4e-7 * max(float(b.abs().max()), 1e-30) if b.numel() else 0.0
    return bool(((a - b).abs() <= atol + rtol * b.abs()).all())


@pytest.mark.parametrize("name", CASES)
def test_store_kernels_reproduce_the_reference_fixtures(name):
    from splatloc_b200.gaussian_optimizer import GaussianStoreOptimizer
    g = load(name)
    params, _, _, lrs = initial_state(g, device="cuda")
    max_grad, min_opacity, extent, max_screen, pd, preg = (float(x) for x in g["dense_args"])
    opt = GaussianStoreOptimizer(params, lrs, percent_dense=pd, primitive_reg=True)   # the fixtures freeze key primitives
    for step in range(3):
        opt.step({n: torch.from_numpy(g[f"grad{step}{KEY[n]}"]).cuda() for n in S.GROUPS})
    for n in S.GROUPS:
        assert _close(opt.params[n], g["adam" + KEY[n]]), f"adam {n}"
        if opt.params[n].numel():
            assert _close(opt.exp_avg[n], g["adam_m" + KEY[n]]) and _close(opt.exp_avg_sq[n], g["adam_v" + KEY[n]]), n
    opt.add_densification_stats(torch.from_numpy(g["stat_grad2D"]).cuda(), torch.from_numpy(g["stat_radii"]).cuda())
    assert _close(opt.xyz_gradient_accum, g["stat_accum"]) and _close(opt.denom, g["stat_denom"])
    assert _close(opt.max_radii2D, g["stat_max_radii2D"])
    opt.primitive_reg = bool(preg)
    new_P = opt.densify_and_prune(max_grad, min_opacity, extent, max_screen or None, noise=torch.from_numpy(g["noise"]).cuda())
    assert new_P == g["out_xyz"].shape[0], (new_P, opt.last_densify)
    assert 2 * opt.last_densify["split"] == g["noise"].shape[0]
    for n in S.GROUPS:
        assert _close(opt.params[n], g["out" + KEY[n]]), f"densify {n}"
        if opt.params[n].numel():
            assert _close(opt.exp_avg[n], g["out_m" + KEY[n]]) and _close(opt.exp_avg_sq[n], g["out_v" + KEY[n]]), n
    assert not bool(opt.xyz_gradient_accum.any()) and opt.max_radii2D.shape == (new_P,)


@pytest.mark.parametrize("isotropic,max_screen,preg", [(False, 20, True), (True, None, False)])
def test_store_kernels_match_oracle_on_a_200k_store(isotropic, max_screen, preg):
    from splatloc_b200.gaussian_optimizer import GaussianStoreOptimizer
    P = 200_000
    gen = torch.Generator().manual_seed(5)
    r = lambda *s: torch.randn(s, generator=gen)   # noqa: E731
    params = dict(xyz=4 * torch.rand((P, 3), generator=gen) - 2, f_dc=r(P, 1, 3), f_rest=torch.zeros(P, 0, 3),
                  opacity=2.5 * r(P, 1), marker=(torch.rand((P, 1), generator=gen) < 0.1).float() * torch.rand((P, 1), generator=gen),
                  kp_score=torch.rand((P, 1), generator=gen),
                  scaling=float(np.log(0.012 if isotropic else 0.006)) + r(P, 1 if isotropic else 3), rotation=r(P, 4))
    lrs = dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=1.25e-4, opacity=0.05, marker=0.05, kp_score=0.05, scaling=1e-3, rotation=1e-3)
    cpu = {n: t.clone() for n, t in params.items()}
    m = {n: torch.zeros_like(t) for n, t in cpu.items()}
    v = {n: torch.zeros_like(t) for n, t in cpu.items()}
    opt = GaussianStoreOptimizer({n: t.cuda() for n, t in params.items()}, lrs, primitive_reg=preg)
    for step in range(1, 4):
        grads = {n: 1e-2 * r(*t.shape) for n, t in cpu.items()}
        S.adam_step(cpu, grads, m, v, lrs, step, marker=cpu["marker"] if preg else None)
        opt.step({n: gt.cuda() for n, gt in grads.items()})
    for n in cpu:
        assert _close(opt.params[n], cpu[n]), f"adam {n}"
        # continue from identical states so that later threshold decisions cannot differ by rounding
        opt.params[n].copy_(cpu[n]); opt.exp_avg[n].copy_(m[n]); opt.exp_avg_sq[n].copy_(v[n])
    V = 3
    radii = ((torch.rand((V, P), generator=gen) < 0.5).int() * torch.randint(1, 60, (V, P), generator=gen)).int()
    g2d = 3e-4 * r(V, P, 3)
    accum, denom, maxr = torch.zeros(P, 1), torch.zeros(P, 1), torch.zeros(P)
    S.densification_stats(accum, denom, maxr, g2d, radii)
    opt.add_densification_stats(g2d.cuda(), radii.cuda())
    assert _close(opt.xyz_gradient_accum, accum, rtol=2e-6) and _close(opt.denom, denom) and _close(opt.max_radii2D, maxr)
    opt.xyz_gradient_accum.copy_(accum)
    n_split_guess = 2 * P
    noise = r(n_split_guess, 3)
    # the oracle needs the exact noise block: take the split count from the plan by a dry run of the oracle
    out, m2, v2, counts = S.densify_and_prune(cpu, m, v, accum, denom, noise, 0.0002, 0.7, 1.0, max_screen, 0.01, preg)
    # (the oracle indexed noise[k * n_split + rank]: hand the kernel exactly that block)
    new_P = opt.densify_and_prune(0.0002, 0.7, 1.0, max_screen, noise=noise[:2 * counts[3]].cuda())
    assert (opt.last_densify["kept"], opt.last_densify["clones"], opt.last_densify["split_children_kept"],
            opt.last_densify["split"]) == counts
    assert new_P == out["xyz"].shape[0] and counts[1] > 1000 and counts[2] > 1000
    for n in cpu:
        assert _close(opt.params[n], out[n]), f"densify {n}"
        if out[n].numel():
            assert _close(opt.exp_avg[n], m2[n]) and _close(opt.exp_avg_sq[n], v2[n]), f"moments {n}"


def test_adam_step_skips_groups_without_gradient_and_counts_steps_per_group():
    from splatloc_b200.gaussian_optimizer import GaussianStoreOptimizer
    P = 1000
    params = dict(xyz=torch.randn(P, 3).cuda(), opacity=torch.randn(P, 1).cuda(), scaling=torch.randn(P, 3).cuda(),
                  rotation=torch.randn(P, 4).cuda())
    before = {n: t.clone() for n, t in params.items()}
    opt = GaussianStoreOptimizer(params, dict(xyz=1e-2, opacity=1e-2, scaling=1e-2, rotation=1e-2))
    opt.step({"xyz": torch.ones(P, 3).cuda()})
    opt.step({"xyz": torch.ones(P, 3).cuda(), "opacity": torch.ones(P, 1).cuda()})
    assert opt.steps == dict(xyz=2, opacity=1, scaling=0, rotation=0)
    assert torch.equal(opt.params["scaling"], before["scaling"]) and not torch.equal(opt.params["xyz"], before["xyz"])
    # first Adam step with a constant gradient moves every element by exactly lr (up to rounding)
    assert torch.allclose(opt.params["opacity"], before["opacity"] - 1e-2, rtol=0, atol=1e-6)
    with pytest.raises(RuntimeError):
        GaussianStoreOptimizer({k: v.cpu() for k, v in params.items()}, {})
