"""GPU parity: libsplatloc_b200 (through the C ABI) against the unmodified reference CUDA
rasterizer compiled into oracle/_ref (same inputs, same GPU).  Integer outputs and the
floats that feed them must be bit-identical; fp32 images and gradients within
rtol 1e-4 / atol 1e-5 (north_star)."""
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu


def _ref():
    from oracle import ref_cuda
    if not ref_cuda.available():
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    return ref_cuda


def _ours():
    from splatloc_b200 import _C
    return _C


CASES = [
    dict(P=3000, seed=1, size=(160, 120)),
    dict(P=20000, seed=2, size=(320, 240)),
    dict(P=20000, seed=3, size=(300, 200)),            # image sizes that are not tile multiples
    dict(P=20000, seed=4, size=(320, 240), use_cov3d=True),
    dict(P=20000, seed=5, size=(320, 240), tile_mask_seed=7),
    dict(P=20000, seed=6, size=(320, 240), bg=[0.3, 0.5, 0.7, 0.2]),
    dict(P=60000, seed=7, channels=3, size=(320, 240), bg=[1.0, 1.0, 1.0]),
    dict(P=20000, seed=8, sh_degree=3, size=(320, 240)),
    dict(P=20000, seed=9, sh_degree=1, size=(320, 240), bg=[0.1, 0.2, 0.3]),
    dict(P=200000, seed=0),                            # config-1-sized, C=4
]


BINNING_MODES = {"radix_fallback": 1, "tile_buckets": 2}


@pytest.fixture(params=list(BINNING_MODES))
def binning_mode(request):
    """Both binning strategies of sl_forward must give the reference's sorted lists."""
    from splatloc_b200 import _capi
    _capi.load().sl_set_binning_mode(BINNING_MODES[request.param])
    yield request.param
    _capi.load().sl_set_binning_mode(0)


@pytest.mark.parametrize("case", CASES, ids=lambda c: "-".join(f"{k}{v}" for k, v in c.items() if k != "bg"))
def test_matches_reference_cuda(case, binning_mode):
    ref, ours = _ref(), _ours()
    d = H.scene_inputs(**case)
    b = H.run_backend(ref, d)
    a = H.run_backend(ours, d)
    assert b["R"] > 0
    bad = H.compare(a, b)
    if binning_mode == "radix_fallback":
        # this path has the reference's pipeline shape: per-Gaussian prefix sum and Gaussian-major emission,
        # so the UNSORTED arrays must equal the reference's element for element too
        for k in ("point_offsets", "keys_unsorted_raw", "vals_unsorted_raw"):
            if not torch.equal(a[k], b[k]):
                bad.append(f"{k} differ")
    assert not bad, "\n".join(bad)


@pytest.mark.parametrize("mode", ["tile_buckets"])
def test_long_tile_lists(mode):
    """Many Gaussians stacked on a few tiles.  Tile buckets: segments beyond 4096 pairs use the
    1024-thread segment sort, beyond 16384 the radix fallback.  All must still match the reference
    bit for bit."""
    from splatloc_b200 import _capi
    ref, ours = _ref(), _ours()
    _capi.load().sl_set_binning_mode(BINNING_MODES[mode])
    try:
        _long_tile_lists(ref, ours)
    finally:
        _capi.load().sl_set_binning_mode(0)


def _long_tile_lists(ref, ours):
    for P, spread in ((60000, 0.15), (120000, 0.05)):
        d = H.scene_inputs(P, 21, size=(128, 96))
        cam = d["cam"]
        fwd, right, down = cam.viewmatrix[:3, 2], cam.viewmatrix[:3, 0], cam.viewmatrix[:3, 1]
        g = torch.Generator().manual_seed(P)
        u = (spread * torch.randn((P, 2), generator=g)).cuda()
        z = (2.0 + 2.0 * torch.rand((P, 1), generator=g)).cuda()
        d["means3D"] = (cam.campos[None, :] + z * fwd[None, :] + u[:, :1] * right[None, :] + u[:, 1:] * down[None, :]).contiguous()
        d["scales"] = (d["scales"] * 0.3).contiguous()
        b = H.run_backend(ref, d)
        a = H.run_backend(ours, d)
        longest = int((b["ranges"][:, 1] - b["ranges"][:, 0]).max())
        assert (4096 < longest <= 16384) if P == 60000 else (longest > 16384), longest
        bad = H.compare(a, b)
        assert not bad, f"longest list {longest}\n" + "\n".join(bad)


@pytest.mark.parametrize("name,case", [
    ("config1_200k_sh3", dict(P=200000, seed=0, sh_degree=3)),
    ("config2_500k_c4", dict(P=500000, seed=0)),            # the bench scene (all 8 bench cameras: test_gpu_bench_parity.py)
    ("config3_1200k_scenes12", dict(P=1200000, seed=1, intrinsics="scenes12")),
    ("config4_3M_1296x968", dict(P=3000000, seed=2, size=(1296, 968), scale=1.5)),
])
def test_baseline_configs_match_reference_cuda(name, case):
    """BASELINE.json's configurations at full size, one view each, against the reference CUDA build."""
    from splatloc_b200 import scenes
    ref, ours = _ref(), _ours()
    case = dict(case)
    if case.get("intrinsics") == "scenes12":
        case["intrinsics"] = scenes.SCENES12_INTRINSICS
    d = H.scene_inputs(**case)
    b = H.run_backend(ref, d)
    a = H.run_backend(ours, d)
    assert b["R"] > 100000
    bad = H.compare(a, b)
    assert not bad, f"{name}: R={b['R']}\n" + "\n".join(bad)


@pytest.mark.parametrize("case", [CASES[2], CASES[4], CASES[8], dict(P=200000, seed=1)],
                         ids=lambda c: "-".join(f"{k}{v}" for k, v in c.items() if k != "bg"))
def test_bulk_copy_record_stream_gives_the_same_frame(case):
    """sl_set_record_stream(1): the forward blend stages whole record chunks with cp.async.bulk instead of
    per-lane loads of the survivors (kept for the A/B in profiles/); every output must be identical."""
    from splatloc_b200 import _capi
    ours = _ours()
    d = H.scene_inputs(**case)
    a = H.run_backend(ours, d, backward=False)
    _capi.load().sl_set_record_stream(1)
    try:
        b = H.run_backend(ours, d, backward=False)
    finally:
        _capi.load().sl_set_record_stream(0)
    for k in ("color", "depth", "alpha", "n_contrib", "radii"):
        assert torch.equal(a[k], b[k]), k


@pytest.mark.parametrize("case", [CASES[2], CASES[4], dict(P=200000, seed=1)],
                         ids=lambda c: "-".join(f"{k}{v}" for k, v in c.items() if k != "bg"))
def test_chained_launches_do_not_change_results(case):
    """sl_set_chained_launches: with programmatic dependent launches every stage kernel may become resident
    before the previous stage has finished and blocks on entry until it has; plain launches must give the
    same frame, the same sorted lists, and gradients that differ by atomic ordering only."""
    from splatloc_b200 import _capi
    ours = _ours()
    d = H.scene_inputs(**case)
    a = H.run_backend(ours, d)
    _capi.load().sl_set_chained_launches(0)
    try:
        b = H.run_backend(ours, d)
    finally:
        _capi.load().sl_set_chained_launches(1)
    for k in ("color", "depth", "alpha", "n_contrib", "radii", "point_list", "point_list_keys", "ranges"):
        if k in a:
            assert torch.equal(a[k], b[k]), k
    assert a["R"] == b["R"]
    assert not H.compare(a, b)


def test_empty_and_invisible():
    ours = _ours()
    # P == 0 -> zero outputs, R == 0
    d = H.scene_inputs(10, 0, size=(64, 48))
    for k in ("means3D", "scales", "rotations", "opacities", "colors_precomp"):
        d[k] = d[k][:0]
    a = H.run_backend(ours, d)
    assert a["R"] == 0 and float(a["color"].abs().max()) == 0.0
    # Gaussians behind the camera are NOT culled by the reference's NDC test (z_ndc ~ 1.002 for any
    # z_cam < 0 with getProjectionMatrix2): whatever it renders, we must render the same.
    ref = _ref()
    d = H.scene_inputs(500, 0, size=(64, 48), bg=[0.25, 0.5, 0.75, 1.0])
    cam = d["cam"]
    fwd, right = cam.viewmatrix[:3, 2], cam.viewmatrix[:3, 0]
    base = d["means3D"].clone()
    d["means3D"] = (cam.campos[None, :] - 5.0 * fwd[None, :] + 0.3 * torch.randn_like(base)).contiguous()
    a, b = H.run_backend(ours, d), H.run_backend(ref, d)
    assert not H.compare(a, b)
    # every Gaussian far outside the frustum sideways -> R == 0, background everywhere, zero gradients
    d["means3D"] = (cam.campos[None, :] + 1.0 * fwd[None, :] + 50.0 * right[None, :] + 0.01 * torch.randn_like(base)).contiguous()
    a, b = H.run_backend(ours, d), H.run_backend(ref, d)
    assert a["R"] == 0 and b["R"] == 0
    assert torch.equal(a["color"], torch.tensor([0.25, 0.5, 0.75, 1.0]).view(4, 1, 1).expand(4, 48, 64))
    assert float(a["g_means3D"].abs().max()) == 0.0
    assert not H.compare(a, b)


def test_mark_visible_matches_reference():
    ref, ours = _ref(), _ours()
    d = H.scene_inputs(50000, 11)
    cam = d["cam"]
    a = ours.mark_visible(d["means3D"], cam.viewmatrix, cam.projmatrix)
    b = ref.mark_visible(d["means3D"], cam.viewmatrix, cam.projmatrix)
    assert a.dtype == torch.bool and torch.equal(a, b)
    assert 0 < int(a.sum()) < a.numel()


@pytest.mark.parametrize("P", [1, 2, 3, 5, 1000, 100000, 1000000, 5000000])
def test_knn_bit_exact(P):
    from splatloc_b200 import scenes
    ref, ours = _ref(), _ours()
    pts = scenes.make_knn_points(P, seed=P % 7).cuda()
    a = ours.distCUDA2(pts)
    b = ref.distCUDA2(pts)
    torch.cuda.synchronize()
    assert torch.equal(a.view(torch.int32), b.view(torch.int32)), \
        f"{int((a.view(torch.int32) != b.view(torch.int32)).sum())}/{P} differ"


@pytest.mark.parametrize("what", ["translated", "outlier", "outliers_1pct"])
def test_knn_bit_exact_far_from_origin_and_with_outliers(what):
    """The grid of the search is laid over a trimmed bounding box and its stopping bound is formed in
    box-relative coordinates: a cloud far from the origin, or one with far outliers (depth
    back-projection produces them), must still give the reference's bits."""
    from splatloc_b200 import scenes
    ref, ours = _ref(), _ours()
    pts = scenes.make_knn_points(200000, seed=3)
    if what == "translated":
        pts = pts + 1000.0
    elif what == "outlier":
        pts[12345] = torch.tensor([1.0e4, -2.0e4, 3.0e3])
    else:
        g = torch.Generator().manual_seed(1)
        idx = torch.randperm(pts.shape[0], generator=g)[:2000]
        pts[idx] = pts[idx] + 200.0 * torch.randn((2000, 3), generator=g)
    pts = pts.contiguous().cuda()
    a, b = ours.distCUDA2(pts), ref.distCUDA2(pts)
    torch.cuda.synchronize()
    assert torch.equal(a.view(torch.int32), b.view(torch.int32)), \
        f"{int((a.view(torch.int32) != b.view(torch.int32)).sum())} differ"


def test_autograd_front_end():
    """The diff_gauss-compatible module: 4-tuple, gradients on the leaf tensors, absgrad side channel."""
    from splatloc_b200.diff_gauss import GaussianRasterizationSettings, GaussianRasterizer
    d = H.scene_inputs(20000, 12, size=(320, 240))
    cam = d["cam"]
    leaves = {k: d[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "colors_precomp", "scales", "rotations")}
    means2D = torch.zeros_like(leaves["means3D"], requires_grad=True)
    rs = GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, d["bg"], 1.0, cam.viewmatrix,
                                       cam.projmatrix, 0, cam.campos, False, False)
    out = GaussianRasterizer(rs)(means3D=leaves["means3D"], means2D=means2D, opacities=leaves["opacities"],
                                 colors_precomp=leaves["colors_precomp"], scales=leaves["scales"],
                                 rotations=leaves["rotations"])
    assert len(out) == 4
    color, depth, alpha, radii = out
    assert color.shape == (4, cam.H, cam.W) and depth.shape == (1, cam.H, cam.W) and radii.dtype == torch.int32
    gc, gd, ga = d["grads"]
    ((color * gc).sum() + (depth * gd).sum() + (alpha * ga).sum()).backward()
    b = H.run_backend(_ours(), d)
    assert torch.allclose(leaves["means3D"].grad.cpu(), b["g_means3D"], rtol=1e-4, atol=1e-5 * float(b["g_means3D"].abs().max()))
    assert torch.allclose(means2D.grad.cpu(), b["g_means2D"], rtol=1e-4, atol=1e-5 * float(b["g_means2D"].abs().max()))
    assert hasattr(means2D, "absgrad") and means2D.absgrad.shape == (20000, 3)
    with pytest.raises(Exception):
        GaussianRasterizer(rs)(means3D=leaves["means3D"], means2D=means2D, opacities=leaves["opacities"])
