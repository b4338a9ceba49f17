"""oracle/loss_oracle.py:l1_ssim against the fixtures the reference's own `ssim` and `l1_loss` produced
(tests/golden/make_ssim_golden.py ran gaussian_splatting/utils/loss_utils.py from /root/reference on the CPU,
combined as train_gaussians.py:285 does)."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import loss_oracle

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "ssim_*.npz")))


def load_ssim(path):
    g = np.load(path)
    return g, torch.from_numpy(g["image"]), torch.from_numpy(g["gt"]), float(g["lambda_dssim"])


def grad_close(got, want, rel_of_max):
    """Element-wise within `rel_of_max` of the gradient's largest magnitude: SSIM gradients come out of
    sigma = conv(x^2) - mu^2, whose cancellation is what bounds float32 agreement, not the element's size."""
    return float(np.abs(got - want).max()) <= rel_of_max * float(np.abs(want).max())


def test_fixtures_exist():
    assert len(GOLDEN) >= 4


@pytest.mark.parametrize("path", GOLDEN, ids=lambda p: os.path.basename(p)[5:-4])
def test_oracle_matches_reference_ssim(path):
    g, image, gt, lam = load_ssim(path)
    for dtype, tol in ((torch.float32, 2e-6), (torch.float64, 2e-6)):
        o = loss_oracle.l1_ssim(image, gt, lam, dtype=dtype)
        assert abs(float(o["ssim"]) - float(g["ssim"])) <= tol
        assert abs(float(o["l1"]) - float(g["l1"])) <= tol * max(float(g["l1"]), 1.0)
        assert abs(float(o["loss"]) - float(g["loss"])) <= tol
        assert grad_close(o["dL_dimage"].float().numpy(), g["dL_dimage"], 2e-4)


def test_window_is_the_reference_window():
    from splatloc_b200.ssim_loss import gaussian_window
    w = gaussian_window()
    assert w.dtype == torch.float32 and w.shape == (11,) and abs(float(w.sum()) - 1.0) < 1e-6
    w2 = loss_oracle.ssim_window(torch.float32)
    assert torch.equal(w2, (w[:, None] @ w[None, :]))
