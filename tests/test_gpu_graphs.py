"""sl_forward_nosync (no read-back of the pair count) and the CUDA-graph replay built on it must give
exactly what the synchronising path gives; a frame that does not fit the stated capacity must be
flagged and leave background-only images and zero gradients."""
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu

EMPTY = torch.Tensor([])


def _forward(d, max_pairs=None):
    from splatloc_b200 import _C
    cam = d["cam"]
    return _C.rasterize_gaussians(
        d["bg"], d["means3D"], d["colors_precomp"], d["opacities"], d["scales"], d["rotations"], 1.0, EMPTY, EMPTY.bool(),
        cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, cam.H, cam.W, EMPTY, 0, cam.campos, False, False,
        False, max_pairs)


def _backward(d, fw, R):
    from splatloc_b200 import _C
    cam = d["cam"]
    gc, gd, ga = d["grads"]
    return _C.rasterize_gaussians_backward(
        d["bg"], d["means3D"], fw[4], d["colors_precomp"], d["scales"], d["rotations"], 1.0, EMPTY, cam.viewmatrix,
        cam.projmatrix, cam.tanfovx, cam.tanfovy, gc, gd, ga, EMPTY, 0, cam.campos, fw[5], R, fw[6], fw[7], fw[3], False)


@pytest.mark.parametrize("P,size", [(20000, (320, 240)), (200000, (640, 480))])
def test_nosync_forward_backward_equal_the_synchronising_path(P, size):
    d = H.scene_inputs(P, 3, size=size)
    a = _forward(d)
    R = a[0]
    b = _forward(d, max_pairs=R + 1000)
    assert b[0] == R + 1000 and int(b[-1]) == 0
    for x, y in zip(a[1:5], b[1:5]):                      # colour, depth, alpha, radii
        assert torch.equal(x, y)
    from splatloc_b200.diff_gauss import interpret_binningBuffer, interpret_imgBuffer
    n_tiles = ((size[1] + 15) // 16) * ((size[0] + 15) // 16)
    ia, ib = interpret_imgBuffer(a[7], size[0] * size[1], n_tiles), interpret_imgBuffer(b[7], size[0] * size[1], n_tiles)
    assert torch.equal(ia["ranges"], ib["ranges"]) and torch.equal(ia["n_contrib"], ib["n_contrib"])
    ba, bb = interpret_binningBuffer(a[6], R), interpret_binningBuffer(b[6], R + 1000)
    assert torch.equal(ba["point_list"], bb["point_list"][:R])
    assert torch.equal(ba["point_list_keys"], bb["point_list_keys"][:R])
    ga, gb = _backward(d, a, R), _backward(d, b, R + 1000)
    for x, y in zip(ga, gb):
        if x.numel():
            assert float((x - y).abs().max()) <= 2e-5 * 16 * float(x.abs().max()) + 1e-30


def test_overflow_is_flagged_and_harmless():
    d = H.scene_inputs(20000, 4, size=(320, 240))
    R = _forward(d)[0]
    for cap in (R - 1, R // 3, 0):
        b = _forward(d, max_pairs=cap)
        assert int(b[-1]) == 1
        assert bool((b[3] == 0).all()) and bool((b[1] == d["bg"].view(-1, 1, 1)).all())
        g = _backward(d, b, cap)
        assert all(float(x.abs().max()) == 0.0 for x in g if x.numel())
    assert int(_forward(d, max_pairs=R)[-1]) == 0     # exactly enough is enough


def test_graph_replay_matches_eager_steps():
    from splatloc_b200 import scenes
    from splatloc_b200.diff_gauss import GaussianRasterizationSettings, GaussianRasterizer
    from splatloc_b200.graphs import GraphedViewStep
    P, C = 100000, 4
    sc = {k: v.cuda() for k, v in scenes.make_scene(P, 5, channels=C).items()}
    params = {k: sc[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "colors_precomp", "scales", "rotations")}
    cams = [scenes.make_camera(s).to("cuda") for s in (0, 1, 2)]
    Hh, W = cams[0].H, cams[0].W
    bg = torch.zeros(C, device="cuda")
    host = torch.empty(C + 2, Hh, W).pin_memory()
    step = GraphedViewStep(params, Hh, W, cams[0].tanfovx, cams[0].tanfovy, bg, max_pairs=2_000_000, host_images=host)
    for i, cam in enumerate(cams):
        up = torch.cat(scenes.make_pixel_grads(i, C, Hh, W), 0).cuda()
        step.set_camera(cam.viewmatrix, cam.projmatrix, cam.campos)
        step.upstream.copy_(up)
        step.replay()
        torch.cuda.synchronize()
        assert int(step.overflow) == 0
        got = {k: v.grad.clone() for k, v in params.items()}
        got_img = (step.color.clone(), step.depth.clone(), step.alpha.clone(), step.radii.clone())
        assert torch.equal(host.cuda(), torch.cat(got_img[:3], 0))

        ref = {k: sc[k].clone().requires_grad_(True) for k in params}
        rs = GaussianRasterizationSettings(Hh, W, cam.tanfovx, cam.tanfovy, bg, 1.0, cam.viewmatrix, cam.projmatrix, 0,
                                           cam.campos, False, False)
        m2d = torch.zeros_like(ref["means3D"], requires_grad=True)
        color, depth, alpha, radii = GaussianRasterizer(rs)(means3D=ref["means3D"], means2D=m2d, opacities=ref["opacities"],
                                                            colors_precomp=ref["colors_precomp"], scales=ref["scales"],
                                                            rotations=ref["rotations"])
        torch.autograd.backward([color, depth, alpha], [up[:C], up[C:C + 1], up[C + 1:]])
        for x, y in zip(got_img, (color, depth, alpha, radii)):
            assert torch.equal(x, y)
        for k in params:
            scale = float(ref[k].grad.abs().max())
            assert float((got[k] - ref[k].grad).abs().max()) <= 2e-5 * 16 * scale, k
