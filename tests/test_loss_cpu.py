"""oracle/loss_oracle.py against the fixtures the reference's own loss functions produced
(tests/golden/make_loss_golden.py ran utils/utils.py:get_loss_mapping and
train_gaussians.py:get_loss_marker from /root/reference on the CPU)."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import loss_oracle

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "loss_*.npz")))


def load_loss(path):
    g = np.load(path)
    gt_kp = torch.from_numpy(g["in_gt_kp"]) if g["in_gt_kp"].size else None
    args = dict(image=torch.from_numpy(g["in_image"]), depth=torch.from_numpy(g["in_depth"]),
                gt_image=torch.from_numpy(g["in_gt_image"]), gt_depth=torch.from_numpy(g["in_gt_depth"]), gt_kp=gt_kp,
                exposure_a=float(g["in_exposure"][0]), exposure_b=float(g["in_exposure"][1]),
                rgb_boundary_threshold=float(g["in_threshold"]), initialization=bool(g["in_initialization"]))
    return g, args


def test_fixtures_exist():
    assert len(GOLDEN) >= 4


@pytest.mark.parametrize("path", GOLDEN, ids=lambda p: os.path.basename(p)[5:-4])
def test_oracle_matches_reference_loss(path):
    g, args = load_loss(path)
    # float32 like the reference ran: same values up to summation order; float64: within float32 rounding
    for dtype, rtol in ((torch.float32, 2e-6), (torch.float64, 2e-6)):
        o = loss_oracle.mapping_loss(dtype=dtype, **args)
        assert abs(float(o["loss"]) - float(g["ref_loss"])) <= rtol * abs(float(g["ref_loss"])) + 1e-7
        assert abs(float(o["loss_rgb"] + o["loss_depth"]) - float(g["ref_loss_mapping"])) <= rtol * float(g["ref_loss_mapping"]) + 1e-7
        assert abs(float(o["loss_kp"]) - float(g["ref_loss_kp"])) <= rtol * abs(float(g["ref_loss_kp"])) + 1e-7
        np.testing.assert_allclose(o["dL_dimage"].float().numpy(), g["ref_dL_dimage"], rtol=1e-5, atol=1e-9)
        np.testing.assert_allclose(o["dL_ddepth"].float().numpy(), g["ref_dL_ddepth"], rtol=1e-5, atol=1e-9)
        np.testing.assert_allclose(o["dL_dexposure"].float().numpy(), g["ref_dL_dexposure"], rtol=2e-4, atol=1e-7)
