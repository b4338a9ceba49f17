"""GPU: render() with its pre-processing inside the kernels (sl_forward_render / sl_backward_render,
SURVEY 8(f) row 1) against the unfused route -- the reference's pre-processing as PyTorch ops
(oracle/render_oracle.py, pinned to the reference's own render() by tests/test_render_cpu.py) in front
of this library's plain rasterizer, which tests/test_gpu_parity.py pins to the reference CUDA build.

The activations are evaluated by different code on the two routes (device expf / division here,
PyTorch's CUDA kernels there), so a scale or a quaternion may differ in the last bit and, rarely, move
a radius across an integer or a splat across a tile edge.  Budgets, stated here: radii may differ for
at most 0.1 % of the Gaussians, images must agree within rtol 1e-4 / atol 1e-5 on all but 0.1 % of
the pixels, gradients of the raw parameters within rtol 1e-4 / atol 1e-5 x 32 x max|g| on all but
0.1 % of the elements."""
import types

import pytest
import torch

from oracle import render_oracle
from tests.test_render_cpu import CASES, load, raw_tensors

pytestmark = pytest.mark.gpu

EMPTY = torch.Tensor([])
C0 = 0.28209479177387814


def _frac_bad(a, b, rtol, atol):
    return float(((a - b).abs() > atol + rtol * b.abs()).float().mean())


def _raw_scene(P, seed, isotropic=False):
    """A room-shaped scene (splatloc_b200.scenes) expressed in the optimiser's raw parameters."""
    from splatloc_b200 import scenes
    sc = scenes.make_scene(P, seed, channels=4)
    g = torch.Generator().manual_seed(seed + 99)
    o = sc["opacities"].clamp(1e-4, 1 - 1e-4)
    rgb = sc["colors_precomp"][:, :3]
    fdc = ((rgb - 0.5) / C0 + 3.0 * (torch.rand((P, 3), generator=g) < 0.05) * torch.randn((P, 3), generator=g))
    scaling = torch.log(sc["scales"])
    if isotropic:
        scaling = scaling.mean(1, keepdim=True)
    rot = sc["rotations"] * torch.exp(0.5 * torch.randn((P, 1), generator=g))
    return [t.contiguous().cuda() for t in (sc["means3D"], fdc.view(P, 1, 3), sc["colors_precomp"][:, 3:4].clone(),
                                            torch.log(o / (1 - o)), scaling, rot)]


def _both_routes(raw, cam, grads, bg):
    from splatloc_b200 import _C
    xyz, fdc, kp, rop, rsc, rrot = raw
    H, W = cam.H, cam.W
    vm, pm, cp = cam.viewmatrix[None].contiguous(), cam.projmatrix[None].contiguous(), cam.campos[None].contiguous()
    # fused
    fw = _C.render_forward(bg, xyz, fdc, kp, rop, rsc, rrot, 1.0, None, vm, pm, cam.tanfovx, cam.tanfovy, H, W, cp)
    bw = _C.render_backward(bg, xyz, fw[4], fdc, kp, rop, rsc, rrot, 1.0, vm, pm, cam.tanfovx, cam.tanfovy,
                            grads[0][None], grads[1][None], grads[2][None], cp, fw[5], fw[0], fw[6], fw[7], fw[3])
    # unfused: the reference's pre-processing as PyTorch ops, then the plain rasterizer, then autograd back
    leaves = [t.clone().requires_grad_(True) for t in raw]
    m3, colors, opac, scales, rots = render_oracle.preprocess(*leaves)
    ufw = _C.rasterize_gaussians(bg, m3.detach(), colors.detach(), opac.detach(), scales.detach(), rots.detach(), 1.0,
                                 EMPTY, EMPTY.bool(), cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, H, W, EMPTY,
                                 0, cam.campos, False, False)
    ubw = _C.rasterize_gaussians_backward(bg, m3.detach(), ufw[4], colors.detach(), scales.detach(), rots.detach(), 1.0,
                                          EMPTY, cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, *grads, EMPTY, 0,
                                          cam.campos, ufw[5], ufw[0], ufw[6], ufw[7], ufw[3], False)
    (g_abs2D, g_m2D, g_col, g_op, g_m3, g_cov, g_sh, g_sc, g_rot) = ubw
    torch.autograd.backward([m3, colors, opac, scales, rots], [g_m3, g_col, g_op, g_sc, g_rot])
    torch.cuda.synchronize()
    return fw, bw, ufw, (g_abs2D, g_m2D), leaves


def _check(fw, bw, ufw, u2d, leaves, P):
    assert float((fw[4][0] != ufw[4]).float().mean()) <= 1e-3, "radii flip budget"
    assert abs(fw[0] - ufw[0]) <= max(8, 2e-3 * ufw[0]), (fw[0], ufw[0])
    for k in (1, 2, 3):
        assert _frac_bad(fw[k][0], ufw[k], 1e-4, 1e-5) <= 1e-3, f"image {k}"
    g_abs2D, g_m2D, g_xyz, g_fdc, g_kp, g_op, g_sc, g_rot = bw
    pairs = [("abs2D", g_abs2D[0], u2d[0]), ("means2D", g_m2D[0], u2d[1]), ("xyz", g_xyz, leaves[0].grad),
             ("features_dc", g_fdc.view(P, 1, 3), leaves[1].grad), ("kp_score", g_kp, leaves[2].grad),
             ("opacity", g_op, leaves[3].grad), ("scaling", g_sc, leaves[4].grad), ("rotation", g_rot, leaves[5].grad)]
    # isotropic Gaussians have an analytically zero rotation gradient: both routes return rounding noise
    # there, so no gradient is compared below 1e-4 of the scaling gradient's magnitude
    floor = 1e-4 * float(leaves[4].grad.abs().max())
    if leaves[4].shape[1] == 1:
        assert float(g_rot.abs().max()) <= floor and float(leaves[5].grad.abs().max()) <= floor, "isotropic: rotation gradient"
        pairs = pairs[:-1]
    for name, a, b in pairs:
        scale = max(float(b.abs().max()), floor, 1e-30)
        assert _frac_bad(a, b, 1e-4, 1e-5 * 32 * scale) <= 1e-3, f"{name}: max err {float((a - b).abs().max()):.3e} of {scale:.3e}"
        assert torch.isfinite(a).all(), name


@pytest.mark.parametrize("name", CASES)
def test_fused_render_on_the_reference_fixture_parameters(name):
    """The raw parameters of the CPU fixtures (clamped colours, quaternions far from unit norm, isotropic scales)."""
    from splatloc_b200 import scenes
    g = load(name)
    raw = raw_tensors(g, device="cuda")
    raw[0] = (raw[0] + torch.tensor([3.0, 1.2, -0.2], device="cuda")).contiguous()     # into the synthetic room
    raw[4] = raw[4] + 1.5                                                               # large enough to cover pixels
    P = raw[0].shape[0]
    cam = scenes.make_camera(3, size=(160, 120)).to("cuda")
    grads = tuple(t.cuda() for t in scenes.make_pixel_grads(1, 4, cam.H, cam.W, normalised=False))
    bg = torch.tensor([0.1, 0.2, 0.3, 0.0], device="cuda")
    fw, bw, ufw, u2d, leaves = _both_routes(raw, cam, grads, bg)
    assert ufw[0] > 100
    _check(fw, bw, ufw, u2d, leaves, P)


@pytest.mark.parametrize("P,isotropic", [(100000, False), (100000, True), (500000, False)])
def test_fused_render_matches_unfused_route(P, isotropic):
    from splatloc_b200 import scenes
    raw = _raw_scene(P, seed=0 if P == 500000 else 4, isotropic=isotropic)
    cam = scenes.make_camera(2).to("cuda")
    grads = tuple(t.cuda() for t in scenes.make_pixel_grads(2, 4, cam.H, cam.W, normalised=False))
    bg = torch.zeros(4, device="cuda")
    fw, bw, ufw, u2d, leaves = _both_routes(raw, cam, grads, bg)
    assert ufw[0] > 100000
    _check(fw, bw, ufw, u2d, leaves, P)


def test_render_function_keeps_the_reference_interface():
    """`splatloc_b200.gaussian_renderer.render`: the reference's signature and result dict; the fused
    route and the PyTorch-op route (forced here with an all-true mask) must agree."""
    from splatloc_b200 import scenes
    from splatloc_b200.gaussian_renderer import render
    import math
    P = 50000
    raw = _raw_scene(P, seed=7)
    cam = scenes.make_camera(1, size=(320, 240)).to("cuda")

    def model():
        pc = types.SimpleNamespace(active_sh_degree=0, max_sh_degree=0)
        for k, t in zip(("_xyz", "_features_dc", "_kp_score", "_opacity", "_scaling", "_rotation"), raw):
            setattr(pc, k, t.clone().requires_grad_(True))
        pc._features_rest = torch.zeros((P, 0, 3), device="cuda")
        return pc

    view = types.SimpleNamespace(FoVx=2 * math.atan(cam.tanfovx), FoVy=2 * math.atan(cam.tanfovy), image_height=cam.H,
                                 image_width=cam.W, world_view_transform=cam.viewmatrix, full_proj_transform=cam.projmatrix,
                                 camera_center=cam.campos)
    pipe = types.SimpleNamespace(convert_SHs_python=True, compute_cov3D_python=False)
    bg = torch.zeros(3, device="cuda")            # SplatLoc passes 3 floats to its 4-channel build
    gt = torch.rand(3, cam.H, cam.W, device="cuda")
    outs = []
    for mask in (None, torch.ones(P, dtype=torch.bool, device="cuda")):
        pc = model()
        pkg = render(view, pc, pipe, bg, mask=mask)
        assert set(pkg) == {"render", "kp_prob", "viewspace_points", "visibility_filter", "radii", "depth", "opacity"}
        assert pkg["render"].shape == (3, cam.H, cam.W) and pkg["kp_prob"].shape == (cam.H, cam.W)
        assert pkg["depth"].shape == (1, cam.H, cam.W) and pkg["radii"].shape == (P,)
        loss = (pkg["render"] - gt).abs().mean() + pkg["kp_prob"].mean() + 0.1 * pkg["depth"].mean() + pkg["opacity"].mean()
        loss.backward()
        outs.append((pkg, pc))
    (fa, pa), (fb, pb) = outs
    assert fa["viewspace_points"].grad is not None and hasattr(fa["viewspace_points"], "absgrad")
    assert _frac_bad(fa["render"], fb["render"], 1e-4, 1e-5) <= 1e-3
    assert float((fa["visibility_filter"] != fb["visibility_filter"]).float().mean()) <= 1e-3
    for k in ("_xyz", "_features_dc", "_kp_score", "_opacity", "_scaling", "_rotation"):
        a, b = getattr(pa, k).grad, getattr(pb, k).grad
        scale = max(float(b.abs().max()), 1e-30)
        assert _frac_bad(a, b, 1e-4, 1e-5 * 32 * scale) <= 1e-3, k
