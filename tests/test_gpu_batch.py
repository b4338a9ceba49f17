"""GPU: several cameras per call (sl_forward_batch / sl_backward_batch) against the single-view entry
points of the same library -- which tests/test_gpu_parity.py pins to the reference CUDA rasterizer.
Per view, images and every integer output must be bit-identical to the single-view call; the
per-Gaussian gradients of a batch are the sum over its views."""
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu

EMPTY = torch.Tensor([])


def _cams(seeds, **kw):
    from splatloc_b200 import scenes
    return [scenes.make_camera(s, **kw).to("cuda") for s in seeds]


def _stack(cams):
    return (torch.stack([c.viewmatrix for c in cams]).contiguous(), torch.stack([c.projmatrix for c in cams]).contiguous(),
            torch.stack([c.campos for c in cams]).contiguous())


def _single(d, cam, grads, want_pose=False):
    from splatloc_b200 import _C
    colors, shs = d.get("colors_precomp", EMPTY), d.get("shs", EMPTY)
    tm = d.get("tile_mask", EMPTY.bool())
    fw = _C.rasterize_gaussians(d["bg"], d["means3D"], colors, d["opacities"], d["scales"], d["rotations"], 1.0, EMPTY, tm,
                                cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, cam.H, cam.W, shs,
                                d["sh_degree"], cam.campos, False, False, True)
    bw = _C.rasterize_gaussians_backward(d["bg"], d["means3D"], fw[4], colors, d["scales"], d["rotations"], 1.0, EMPTY,
                                         cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, *grads, shs,
                                         d["sh_degree"], cam.campos, fw[5], fw[0], fw[6], fw[7], fw[3], False, want_pose)
    return fw, bw


def _batch(d, cams, grads, want_pose=False, max_pairs=None, pose_only=False):
    from splatloc_b200 import _C
    colors, shs = d.get("colors_precomp", EMPTY), d.get("shs", EMPTY)
    tm = d.get("tile_mask", EMPTY.bool())
    vm, pm, cp = _stack(cams)
    c0 = cams[0]
    fw = _C.rasterize_gaussians_batch(d["bg"], d["means3D"], colors, d["opacities"], d["scales"], d["rotations"], 1.0,
                                      EMPTY, tm, vm, pm, c0.tanfovx, c0.tanfovy, c0.H, c0.W, shs, d["sh_degree"], cp,
                                      False, False, True, max_pairs)
    g = tuple(torch.stack([gr[k] for gr in grads]).contiguous() for k in range(3))
    bw = _C.rasterize_gaussians_backward_batch(d["bg"], d["means3D"], fw[4], colors, d["scales"], d["rotations"], 1.0,
                                               EMPTY, vm, pm, c0.tanfovx, c0.tanfovy, *g, shs, d["sh_degree"], cp,
                                               fw[5], fw[0], fw[6], fw[7], fw[3], False, want_pose, pose_only)
    return fw, bw


def _close(a, b, scale_from=None, k=32.0):
    ref = b if scale_from is None else scale_from
    tol = 1e-5 * k * max(float(ref.abs().max()), 1e-30) + 1e-4 * b.abs()
    return bool(((a - b).abs() <= tol).all())


CASES = [
    dict(P=30000, seed=2, size=(320, 240), cams=(0, 1, 2)),
    dict(P=30000, seed=3, size=(300, 200), cams=(3, 4), tile_mask_seed=7),       # not a tile multiple, masked
    dict(P=20000, seed=8, size=(320, 240), cams=(5, 6, 7, 8), sh_degree=3),
    dict(P=60000, seed=7, size=(320, 240), cams=(1, 9), channels=3, bg=[1.0, 1.0, 1.0]),
    dict(P=200000, seed=0, cams=(0, 1, 2, 3, 4, 5, 6, 7)),                       # bench-shaped: 8 views at 640x480
]


@pytest.mark.parametrize("mode", ["tile_buckets", "radix_fallback"])
@pytest.mark.parametrize("case", CASES, ids=lambda c: "-".join(f"{k}{v}" for k, v in c.items() if k != "bg"))
def test_batch_equals_single_view_calls(case, mode):
    from splatloc_b200 import _capi, scenes
    from splatloc_b200.diff_gauss import interpret_binningBuffer, interpret_imgBuffer
    case = dict(case)
    seeds = case.pop("cams")
    d = H.scene_inputs(**case)
    size = case.get("size")
    cams = _cams(seeds, size=size)
    V, P, C = len(cams), d["means3D"].shape[0], d["C"]
    Hh, W = cams[0].H, cams[0].W
    grads = [tuple(t.cuda() for t in scenes.make_pixel_grads(50 + s, C, Hh, W, normalised=False)) for s in seeds]
    _capi.load().sl_set_binning_mode({"tile_buckets": 2, "radix_fallback": 1}[mode])
    try:
        fwb, bwb = _batch(d, cams, grads, want_pose=True)
        singles = [_single(d, cam, g, want_pose=True) for cam, g in zip(cams, grads)]
    finally:
        _capi.load().sl_set_binning_mode(0)
    torch.cuda.synchronize()
    n_tiles = ((Hh + 15) // 16) * ((W + 15) // 16)
    Rb = fwb[0]
    assert Rb == sum(s[0][0] for s in singles)
    bb = interpret_binningBuffer(fwb[6], Rb)
    ib = interpret_imgBuffer(fwb[7], V * Hh * W, V * n_tiles)
    off = 0
    for v, (fw, bw) in enumerate(singles):
        R = fw[0]
        for k in (1, 2, 3):                                   # colour, depth, alpha
            assert torch.equal(fwb[k][v], fw[k]), (v, k)
        assert torch.equal(fwb[4][v], fw[4])                  # radii
        assert torch.equal(fwb[8][v], fw[8])                  # n_touched
        bs, is_ = interpret_binningBuffer(fw[6], R), interpret_imgBuffer(fw[7], Hh * W, n_tiles)
        assert torch.equal(bb["point_list"][off:off + R], bs["point_list"])
        assert torch.equal(bb["point_list_keys"][off:off + R] - ((v * n_tiles) << 32), bs["point_list_keys"])
        rng = ib["ranges"][v * n_tiles:(v + 1) * n_tiles].clone()
        rng[rng[:, 1] > 0] -= off                             # empty tiles stay (0, 0)
        assert torch.equal(rng, is_["ranges"])
        assert torch.equal(ib["n_contrib"][v * Hh * W:(v + 1) * Hh * W], is_["n_contrib"])
        off += R
        # per-view gradients
        assert _close(bwb[0][v], bw[0]) and _close(bwb[1][v], bw[1]), f"view {v}: screen-space gradients"
        assert _close(bwb[9][v], bw[9], k=64.0), f"view {v}: pose gradient {bwb[9][v]} vs {bw[9]}"
        if "shs" in d:
            assert _close(bwb[2][v], bw[2])
    # per-Gaussian gradients: sums over the views
    names = ("abs2D", "means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rot")
    for i in range(2 if "shs" not in d else 3, 9):
        if bwb[i].numel() == 0:
            continue
        total = sum(s[1][i].double() for s in singles)
        biggest = max(s[1][i].abs().max() for s in singles)
        assert _close(bwb[i].double(), total, scale_from=biggest, k=64.0), names[i]


def test_batch_without_readback_and_pose_only():
    from splatloc_b200 import scenes
    d = H.scene_inputs(100000, 4)
    cams = _cams((0, 1, 2, 3))
    grads = [tuple(t.cuda() for t in scenes.make_pixel_grads(s, 4, cams[0].H, cams[0].W, normalised=False)) for s in range(4)]
    fa, ba = _batch(d, cams, grads, want_pose=True)
    R = fa[0]
    fb, bb = _batch(d, cams, grads, want_pose=True, max_pairs=R + 777)
    assert fb[0] == R + 777 and int(fb[-1]) == 0
    for k in (1, 2, 3, 4, 8):
        assert torch.equal(fa[k], fb[k]), k
    for x, y in zip(ba, bb):
        if x is not None and x.numel():
            assert _close(x, y), "gradients of the no-read-back batch"
    # capacity exceeded -> flagged, background only
    fc, _ = _batch(d, cams, grads, max_pairs=R // 2)
    assert int(fc[-1]) == 1 and bool((fc[3] == 0).all())
    # pose only: same dL_dtau, nothing per Gaussian
    fp, bp = _batch(d, cams, grads, pose_only=True)
    assert bp[4] is None and bp[8] is None
    assert _close(bp[9], ba[9], k=64.0)


def test_batched_autograd_front_end_and_single_theta():
    from splatloc_b200 import scenes
    from splatloc_b200.diff_gauss import (BatchedGaussianRasterizer, GaussianRasterizationSettings, GaussianRasterizer)
    d = H.scene_inputs(30000, 12, size=(320, 240))
    cams = _cams((0, 1, 2), size=(320, 240))
    vm, pm, cp = _stack(cams)
    c0 = cams[0]
    V, P = 3, 30000
    leaves = {k: d[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "colors_precomp", "scales", "rotations")}
    m2d = torch.zeros(V, P, 3, device="cuda", requires_grad=True)
    theta = torch.zeros(V, 3, device="cuda", requires_grad=True)
    rs = GaussianRasterizationSettings(c0.H, c0.W, c0.tanfovx, c0.tanfovy, d["bg"], 1.0, vm, pm, 0, cp, False, False)
    color, depth, alpha, radii = BatchedGaussianRasterizer(rs)(
        means3D=leaves["means3D"], means2D=m2d, opacities=leaves["opacities"], colors_precomp=leaves["colors_precomp"],
        scales=leaves["scales"], rotations=leaves["rotations"], theta=theta)          # rho not given
    assert color.shape == (V, 4, c0.H, c0.W) and radii.shape == (V, P)
    g = [tuple(t.cuda() for t in scenes.make_pixel_grads(s, 4, c0.H, c0.W, normalised=False)) for s in range(V)]
    loss = sum((color[v] * g[v][0]).sum() + (depth[v] * g[v][1]).sum() + (alpha[v] * g[v][2]).sum() for v in range(V))
    loss.backward()
    assert theta.grad.shape == (V, 3) and m2d.grad.shape == (V, P, 3) and m2d.absgrad.shape == (V, P, 3)
    # against three single-view autograd calls (theta alone there too: the round-1 advisor's case)
    ref = {k: d[k].clone().requires_grad_(True) for k in leaves}
    th = []
    for v, cam in enumerate(cams):
        t1 = torch.zeros(3, device="cuda", requires_grad=True)
        m1 = torch.zeros(P, 3, device="cuda", requires_grad=True)
        rs1 = GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, d["bg"], 1.0, cam.viewmatrix,
                                            cam.projmatrix, 0, cam.campos, False, False)
        c1, d1, a1, r1 = GaussianRasterizer(rs1)(means3D=ref["means3D"], means2D=m1, opacities=ref["opacities"],
                                                 colors_precomp=ref["colors_precomp"], scales=ref["scales"],
                                                 rotations=ref["rotations"], theta=t1)
        assert torch.equal(c1, color[v]) and torch.equal(r1, radii[v])
        ((c1 * g[v][0]).sum() + (d1 * g[v][1]).sum() + (a1 * g[v][2]).sum()).backward()
        th.append(t1.grad)
        assert _close(m2d.grad[v], m1.grad)
    assert _close(theta.grad, torch.stack(th), k=64.0)
    for k in leaves:
        assert _close(leaves[k].grad, ref[k].grad, k=64.0), k
