"""CPU tests of the oracle itself (no GPU): internal consistency of the float64 restatement
(explicit backward == autograd through the forward) and of the kNN restatement."""
import numpy as np
import torch

from oracle import knn_oracle
from oracle import raster_oracle as O
from splatloc_b200 import scenes


def _small(P=1500, seed=3, C=3, size=(96, 64)):
    sc = scenes.make_scene(P, seed, channels=C)
    cam = scenes.make_camera(seed, size=size)
    return sc, cam


def test_explicit_blend_backward_matches_autograd():
    """Without the fork's channel mask (C=3) the hand-derived blend backward must equal reverse-mode
    differentiation of the forward compositing, to float64 round-off."""
    sc, cam = _small()
    W, H = cam.W, cam.H
    bg = torch.tensor([0.2, 0.4, 0.6], dtype=torch.float64)
    fw = O.forward(sc["means3D"], sc["opacities"], cam.viewmatrix, cam.projmatrix, cam.campos, W, H, cam.tanfovx,
                   cam.tanfovy, bg, colors_precomp=sc["colors_precomp"], scales=sc["scales"], rotations=sc["rotations"])
    assert fw["R"] > 200
    gc, gd, ga = (t.double() for t in scenes.make_pixel_grads(3, 3, H, W, normalised=False))
    pre, pairs = fw["pre"], fw["pairs"]
    gb = O.blend_backward(pre, pairs, pre["features"], bg, W, H, fw["alpha"], fw["n_contrib"], gc, gd[0], ga[0])

    leaves = {k: pre[k].clone().requires_grad_(True) for k in ("means2D", "conic_opacity", "depths", "features")}
    pre2 = dict(pre, **leaves)
    color, depth, alpha, _ = O.blend_forward(pre2, pairs, leaves["features"], bg, W, H)
    ((color * gc).sum() + (depth * gd[0]).sum() + (alpha * ga[0]).sum()).backward()

    def close(a, b):
        return torch.allclose(a, b, rtol=1e-9, atol=1e-9 * float(b.abs().max() + 1))
    # pixel-space mean gradient = NDC-space gradient / (0.5 W, 0.5 H)
    scale = torch.tensor([0.5 * W, 0.5 * H], dtype=torch.float64)
    assert close(gb["mean2D"] / scale, leaves["means2D"].grad)
    g_co = leaves["conic_opacity"].grad
    assert close(gb["conic"][:, 0], g_co[:, 0]) and close(gb["conic"][:, 2], g_co[:, 2])
    assert close(2 * gb["conic"][:, 1], g_co[:, 1])       # the stored xy term is half of d/d(conic_xy)
    assert close(gb["opacity"], g_co[:, 3])
    assert close(gb["color"], leaves["features"].grad)
    assert close(gb["depth"], leaves["depths"].grad)
    assert (gb["absmean2D"] >= gb["mean2D"].abs() - 1e-12).all()


def test_forward_invariants():
    sc, cam = _small(P=4000, seed=5, C=4, size=(128, 96))
    bg = torch.zeros(4, dtype=torch.float64)
    fw = O.forward(sc["means3D"], sc["opacities"], cam.viewmatrix, cam.projmatrix, cam.campos, cam.W, cam.H,
                   cam.tanfovx, cam.tanfovy, bg, colors_precomp=sc["colors_precomp"], scales=sc["scales"],
                   rotations=sc["rotations"])
    pairs, pre = fw["pairs"], fw["pre"]
    assert int(pre["tiles_touched"].sum()) == fw["R"] == int(pairs["point_offsets"][-1])
    ks = pairs["keys_sorted"]
    assert bool((ks[1:] >= ks[:-1]).all())                      # sortedness
    assert torch.equal(torch.sort(pairs["vals_unsorted"].long()).values, torch.sort(pairs["point_list"].long()).values)
    r = pairs["ranges"].long()
    assert int((r[:, 1] - r[:, 0]).sum()) == fw["R"]            # ranges partition the list
    assert float(fw["alpha"].min()) >= 0 and float(fw["alpha"].max()) <= 1
    # a tile mask that keeps everything changes nothing
    fw2 = O.forward(sc["means3D"], sc["opacities"], cam.viewmatrix, cam.projmatrix, cam.campos, cam.W, cam.H,
                    cam.tanfovx, cam.tanfovy, bg, colors_precomp=sc["colors_precomp"], scales=sc["scales"],
                    rotations=sc["rotations"], tile_mask=torch.ones((cam.H + 15) // 16, (cam.W + 15) // 16, dtype=torch.bool))
    assert torch.equal(fw2["color"], fw["color"]) and torch.equal(fw2["pairs"]["keys_sorted"], ks)


def test_knn_oracle_kdtree_path_equals_brute_force():
    pts = scenes.make_knn_points(3000, seed=1).numpy()
    a = knn_oracle.dist_cuda2(pts)                 # k-d tree candidates (P > 2048)
    b = np.empty_like(a)
    for i in range(0, 3000, 500):                  # brute force on blocks against the full set
        for j in range(i, i + 500):
            d = knn_oracle.pair_dist(pts[j], pts)
            d[j] = knn_oracle.FLT_MAX
            b[j] = knn_oracle.mean3(np.sort(d)[:3])
    assert np.array_equal(a.view(np.int32), b.view(np.int32))
    with np.errstate(over="ignore"):
        three = knn_oracle.dist_cuda2(pts[:3])         # 2 neighbours: (d1 + d2 + FLT_MAX) / 3 stays finite
        two = knn_oracle.dist_cuda2(pts[:2])           # 1 neighbour: d1 + FLT_MAX + FLT_MAX overflows
    assert np.isfinite(three).all() and (three > 1e37).all() and np.isinf(two).all()


def test_pose_gradient_oracle_matches_finite_differences():
    """Extension with no reference counterpart: dL/d(rho, theta) of the oracle (autograd through the
    per-Gaussian chain with the camera written as Exp(tau) W2C) against central differences of the
    complete float64 forward.  A small step keeps the differences away from the compositing's
    discontinuities (alpha / transmittance thresholds, tile rectangles)."""
    sc = scenes.make_scene(600, 5, channels=3)
    cam = scenes.make_camera(5, size=(64, 48))
    W, H = cam.W, cam.H
    bg = torch.tensor([0.1, 0.2, 0.3], dtype=torch.float64)
    gc, gd, ga = (t.double() for t in scenes.make_pixel_grads(5, 3, H, W, normalised=False))
    v0, p0 = cam.viewmatrix.double().reshape(-1), cam.projmatrix.double().reshape(-1)

    def run(tau):
        v, p, c = O.perturbed_camera(v0, p0, tau)
        fw = O.forward(sc["means3D"], sc["opacities"], v, p, c, W, H, cam.tanfovx, cam.tanfovy, bg,
                       colors_precomp=sc["colors_precomp"], scales=sc["scales"], rotations=sc["rotations"])
        return float((fw["color"] * gc).sum() + (fw["depth"] * gd[0]).sum() + (fw["alpha"] * ga[0]).sum()), fw

    _, fw = run(torch.zeros(6, dtype=torch.float64))
    bw = O.backward(fw, sc["means3D"], cam.viewmatrix, cam.projmatrix, cam.campos, W, H, cam.tanfovx, cam.tanfovy,
                    gc, gd, ga, scales=sc["scales"], rotations=sc["rotations"], pose=True)
    eps = 1e-7
    for k in range(6):
        e = torch.zeros(6, dtype=torch.float64)
        e[k] = eps
        fd = (run(e)[0] - run(-e)[0]) / (2 * eps)
        assert abs(fd - float(bw["g_tau"][k])) <= 1e-4 * abs(fd) + 1e-3, (k, fd, float(bw["g_tau"][k]))
    # n_touched is a subset count of the blended pairs
    assert int(fw["n_touched"].sum()) > 0 and int(fw["n_touched"].max()) <= W * H
