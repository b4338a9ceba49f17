"""Size-independent properties of the path at the bench configuration (500 k Gaussians, 640x480,
C = 4), checked without the reference: a sorted and complete binning, run-to-run identical forward
outputs, a backward that is linear in the upstream gradients, a tile mask that changes nothing on
the tiles it keeps, and independence from the order in which the Gaussians are stored."""
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu

P_FULL = 500_000


@pytest.fixture(scope="module")
def full():
    from splatloc_b200 import _C
    d = H.scene_inputs(P_FULL, 0)
    return d, H.run_backend(_C, d)


def test_binning_is_sorted_complete_and_consistent(full):
    d, a = full
    keys, ranges, plist = a["keys_sorted"].long(), a["ranges"].long(), a["point_list"].long()
    R = a["R"]
    assert R == keys.numel() == int(a["tiles_touched"].long().sum())
    tile, depth_bits = keys >> 32, keys & 0xFFFFFFFF
    # total order (tile, depth bits, Gaussian id): strictly increasing composite
    comp_hi = (tile[1:] > tile[:-1]) | ((tile[1:] == tile[:-1]) & (depth_bits[1:] > depth_bits[:-1]))
    tie = (tile[1:] == tile[:-1]) & (depth_bits[1:] == depth_bits[:-1])
    assert bool((comp_hi | (tie & (plist[1:] > plist[:-1]))).all())
    # ranges partition the list by tile
    n_tiles = ranges.shape[0]
    counts = torch.bincount(tile, minlength=n_tiles)
    lens = ranges[:, 1] - ranges[:, 0]
    assert torch.equal(lens, counts)
    nz = lens > 0
    assert torch.equal(tile[ranges[nz, 0]], torch.nonzero(nz).flatten())
    assert torch.equal(ranges[nz, 1], torch.cumsum(counts, 0)[nz])
    # the key's depth word is the Gaussian's view depth, its id a visible Gaussian
    assert bool((a["radii"][plist] > 0).all())
    # n_contrib never exceeds the pixel's tile list
    cam = d["cam"]
    gx = (cam.W + 15) // 16
    ys, xs = torch.meshgrid(torch.arange(cam.H), torch.arange(cam.W), indexing="ij")
    pix_tile = (ys // 16) * gx + xs // 16
    assert bool((a["n_contrib"].long().view(cam.H, cam.W) <= lens[pix_tile]).all())


def test_forward_is_run_to_run_identical(full):
    from splatloc_b200 import _C
    d, a = full
    b = H.run_backend(_C, d, backward=False)
    for k in ("color", "depth", "alpha", "n_contrib", "keys_sorted", "point_list", "ranges", "radii"):
        assert torch.equal(a[k], b[k]), k


def test_backward_is_linear_in_the_upstream_gradients(full):
    from splatloc_b200 import _C
    d, a = full
    d2 = dict(d)
    d2["grads"] = tuple(2.0 * g for g in d["grads"])        # exact in float32
    b = H.run_backend(_C, d2)
    for k in H.GRAD_KEYS:
        scale = float(a[k].abs().max()) if a[k].numel() else 0.0
        # scaling by 2 is exact per term; what remains is the order of the atomic sums
        err = (b[k] - 2.0 * a[k]).abs().max() if a[k].numel() else 0.0
        assert float(err) <= 2e-5 * 16 * max(scale, 1e-30), (k, float(err), scale)


def test_tile_mask_keeps_unmasked_tiles_bit_identical(full):
    from splatloc_b200 import _C
    d, a = full
    cam = d["cam"]
    ty, tx = (cam.H + 15) // 16, (cam.W + 15) // 16
    g = torch.Generator().manual_seed(5)
    mask = torch.rand((ty, tx), generator=g) < 0.5
    dm = dict(d)
    dm["tile_mask"] = mask.cuda()
    b = H.run_backend(_C, dm, backward=False)
    keep = mask.repeat_interleave(16, 0).repeat_interleave(16, 1)[: cam.H, : cam.W]
    for k in ("color", "depth", "alpha"):
        assert torch.equal(a[k][:, keep], b[k][:, keep]), k
    assert torch.equal(a["n_contrib"].view(cam.H, cam.W)[keep], b["n_contrib"].view(cam.H, cam.W)[keep])
    # masked-out tiles are empty: background only
    assert bool((b["alpha"][:, ~keep] == 0).all())
    lens = (b["ranges"][:, 1] - b["ranges"][:, 0]).view(ty, tx)
    assert bool((lens[~mask] == 0).all())


def test_storage_order_of_the_gaussians_does_not_matter(full):
    """A permuted map renders the same image bit for bit (blend order is by depth; only exact depth
    ties fall back to the storage order) and returns permuted gradients."""
    from splatloc_b200 import _C
    d, a = full
    g = torch.Generator().manual_seed(9)
    perm = torch.randperm(P_FULL, generator=g)
    dp = dict(d)
    for k in ("means3D", "opacities", "scales", "rotations", "colors_precomp"):
        dp[k] = d[k][perm.cuda()].contiguous()
    b = H.run_backend(_C, dp)
    # exact (tile, depth bits) ties fall back to the storage order: leave those tiles out
    keys = a["keys_sorted"].long()
    cam = d["cam"]
    ty, tx = (cam.H + 15) // 16, (cam.W + 15) // 16
    tied = torch.zeros(ty * tx, dtype=torch.bool)
    tied[(keys[1:] >> 32)[keys[1:] == keys[:-1]]] = True
    assert int(tied.sum()) < ty * tx // 4
    ok = ~tied.view(ty, tx).repeat_interleave(16, 0).repeat_interleave(16, 1)[: cam.H, : cam.W]
    for k in ("color", "depth", "alpha"):
        assert torch.equal(a[k][:, ok], b[k][:, ok]), k
    assert torch.equal(a["n_contrib"].view(cam.H, cam.W)[ok], b["n_contrib"].view(cam.H, cam.W)[ok])
    assert torch.equal(a["radii"][perm], b["radii"])
    for k in ("g_means3D", "g_opacity", "g_colors", "g_scales"):
        scale = float(a[k].abs().max())
        assert float((a[k][perm] - b[k]).abs().max()) <= 1e-5 * 16 * scale, k
