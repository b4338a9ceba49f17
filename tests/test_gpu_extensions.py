"""GPU tests of the two opt-in extensions the reference does not have:
the camera-pose gradient (theta / rho) and n_touched.  Parity is against the float64 oracle only
("unpinned by the reference"); everything reference-facing must stay unchanged when they are on."""
import pytest
import torch

from tests import helpers as H
from tests.test_golden_cpu import load, run_oracle
from tests.test_gpu_golden import _inputs

pytestmark = pytest.mark.gpu


def _oracle_pose(g):
    from oracle import raster_oracle as O
    t = lambda k: torch.from_numpy(g["in_" + k]) if ("in_" + k) in g else None  # noqa: E731
    fw, _, (W, Hh) = run_oracle(g)
    tfx, tfy = (float(v) for v in g["in_tanfov"])
    kw = dict(shs=t("shs"), sh_degree=int(g["in_sh_degree"]), cov3D_precomp=t("cov3D_precomp"))
    if kw["cov3D_precomp"] is None:
        kw.update(scales=t("scales"), rotations=t("rotations"))
    bw = O.backward(fw, t("means3D"), t("viewmatrix"), t("projmatrix"), t("campos"), W, Hh, tfx, tfy,
                    t("grad0"), t("grad1"), t("grad2"), pose=True, **kw)
    return fw, bw


@pytest.mark.parametrize("name", ["c4_precomp", "sh3_bg", "cov3d_mask"])
def test_pose_gradient_and_n_touched_match_oracle(name):
    from splatloc_b200.diff_gauss import GaussianRasterizationSettings, GaussianRasterizer
    g = load(name)
    d = _inputs(g)
    cam = d["cam"]
    fw, bw = _oracle_pose(g)
    theta = torch.zeros(3, device="cuda", requires_grad=True)
    rho = torch.zeros(3, device="cuda", requires_grad=True)
    means3D = d["means3D"].clone().requires_grad_(True)
    means2D = torch.zeros_like(means3D, requires_grad=True)
    rs = GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, d["bg"], 1.0, cam.viewmatrix,
                                       cam.projmatrix, d["sh_degree"], cam.campos, False, False)
    kw = dict(means3D=means3D, means2D=means2D, opacities=d["opacities"], theta=theta, rho=rho)
    if "shs" in d:
        kw["shs"] = d["shs"]
    else:
        kw["colors_precomp"] = d["colors_precomp"]
    if "cov3D_precomp" in d:
        kw["cov3D_precomp"] = d["cov3D_precomp"]
    else:
        kw.update(scales=d["scales"], rotations=d["rotations"])
    if "tile_mask" in d:
        kw["tile_mask"] = d["tile_mask"]
    out = GaussianRasterizer(rs, return_n_touched=True)(**kw)
    assert len(out) == 5
    color, depth, alpha, radii, n_touched = out
    gc, gd, ga = d["grads"]
    ((color * gc).sum() + (depth * gd).sum() + (alpha * ga).sum()).backward()

    # reference-facing results are exactly what the plain call gives
    from splatloc_b200 import _C
    plain = H.run_backend(_C, d)
    assert torch.equal(color.detach().cpu(), plain["color"]) and torch.equal(radii.cpu(), plain["radii"])
    assert torch.allclose(means3D.grad.cpu(), plain["g_means3D"], rtol=1e-4, atol=1e-5 * float(plain["g_means3D"].abs().max()))

    # pose gradient vs float64 autograd (rho first, then theta, in the oracle's tau)
    g_tau = torch.cat([rho.grad, theta.grad]).double().cpu()
    ref = bw["g_tau"]
    scale = float(ref.abs().max())
    assert torch.allclose(g_tau, ref, rtol=2e-3, atol=2e-4 * scale), (g_tau, ref)

    # n_touched vs oracle: equal except where a float32/float64 threshold decision flips
    nt, nt_ref = n_touched.cpu().long(), fw["n_touched"].long()
    assert int(nt.sum()) > 0
    assert float((nt != nt_ref).float().mean()) <= 5e-3
    assert abs(int(nt.sum()) - int(nt_ref.sum())) <= 2e-3 * int(nt_ref.sum()) + 4


def test_extensions_are_off_by_default():
    from splatloc_b200.diff_gauss import GaussianRasterizationSettings, GaussianRasterizer
    d = H.scene_inputs(5000, 3, size=(96, 64))
    cam = d["cam"]
    rs = GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, d["bg"], 1.0, cam.viewmatrix,
                                       cam.projmatrix, 0, cam.campos, False, False)
    out = GaussianRasterizer(rs)(means3D=d["means3D"], means2D=torch.zeros_like(d["means3D"]), opacities=d["opacities"],
                                 colors_precomp=d["colors_precomp"], scales=d["scales"], rotations=d["rotations"])
    assert len(out) == 4


def test_pose_refinement_converges():
    """Config-3-shaped check at small scale: render a target from the true pose, start from a pose that is
    off by up to 5 cm / 3 degrees, refine through the rasterizer; the pose error must shrink several-fold."""
    from splatloc_b200 import scenes
    from splatloc_b200.pose_refine import PoseRefiner, perturb_pose, pose_error, se3_exp
    torch.manual_seed(0)
    assert torch.allclose(se3_exp(torch.zeros(6)), torch.eye(4, dtype=torch.float64))
    gaussians = scenes.make_scene(60000, seed=4, channels=4)
    cam = scenes.make_camera(seed=4, intrinsics=scenes.SCENES12_INTRINSICS, size=(320, 240))
    k = dict(W=320, H=240, fx=572.0 * 0.5, fy=572.0 * 0.5, cx=(320.0 + 0.5) * 0.5 - 0.5, cy=(240.0 + 0.5) * 0.5 - 0.5)
    w2c_gt = cam.viewmatrix.transpose(0, 1).double()
    ref = PoseRefiner(gaussians, k)
    with torch.no_grad():
        color, depth, alpha, _ = ref.render(w2c_gt.cuda())
    w2c0 = perturb_pose(w2c_gt, seed=1)
    e0 = pose_error(w2c0, w2c_gt)
    w2c, losses = ref.refine(w2c0, color.detach(), depth.detach(), iters=120)
    e1 = pose_error(w2c, w2c_gt)
    assert float(losses[-1]) < 0.5 * float(losses[0]), (float(losses[0]), float(losses[-1]))
    assert e1[0] < 0.4 * e0[0] + 1e-4 and e1[1] < 0.4 * e0[1] + 1e-3, (e0, e1)

    # the same refinement with every iteration replayed from one CUDA graph: same trajectory up to the
    # order of the atomic sums, for this frame and for a second one through the cached graph
    for seed in (1, 2):
        start = perturb_pose(w2c_gt, seed=seed)
        a, la = ref.refine(start, color.detach(), depth.detach(), iters=120)
        b, lb = ref.refine_graphed(start, color.detach(), depth.detach(), iters=120)
        assert abs(float(la[0]) - float(lb[0])) <= 1e-5 * abs(float(la[0]))
        d = pose_error(a, b)
        assert d[0] < 1e-3 and d[1] < 0.2, d      # 120 Adam steps amplify the summation-order noise
        eb = pose_error(b, w2c_gt)
        assert eb[0] < 0.4 * pose_error(start, w2c_gt)[0] + 1e-4, eb
