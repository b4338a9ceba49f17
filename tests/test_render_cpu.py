"""CPU: the oracle of render()'s pre-processing (oracle/render_oracle.py) against outputs of the
reference's own render() and GaussianModel getters (tests/golden/render_*.npz, generated on the CPU
from /root/reference by tests/golden/make_render_golden.py)."""
import os

import numpy as np
import pytest
import torch

from oracle import render_oracle

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["aniso_300", "iso_200"]


def load(name):
    return dict(np.load(os.path.join(GOLDEN, f"render_{name}.npz")))


def raw_tensors(g, requires_grad=False, device="cpu"):
    names = ("_xyz", "_features_dc", "_kp_score", "_opacity", "_scaling", "_rotation")
    return [torch.from_numpy(g["raw" + k]).to(device).requires_grad_(requires_grad) for k in names]


@pytest.mark.parametrize("name", CASES)
def test_oracle_reproduces_the_reference_preprocessing(name):
    g = load(name)
    raw = raw_tensors(g, requires_grad=True)
    act = render_oracle.preprocess(*raw)
    for t, k in zip(act, ("means3D", "colors_precomp", "opacities", "scales", "rotations")):
        if k in ("means3D", "colors_precomp"):        # copies, exactly rounded multiplies and adds: same bits anywhere
            assert np.array_equal(t.detach().numpy(), g["act_" + k]), k
        else:                                          # exp / sigmoid / normalize: the host's vector maths library may differ by an ulp
            assert np.allclose(t.detach().numpy(), g["act_" + k], rtol=5e-7, atol=0), k
    grads = [torch.from_numpy(g["gact_" + k]) for k in ("means3D", "colors_precomp", "opacities", "scales", "rotations")]
    torch.autograd.backward(list(act), grads)
    for t, k in zip(raw, ("_xyz", "_features_dc", "_kp_score", "_opacity", "_scaling", "_rotation")):
        ref = g["graw" + k]
        assert np.allclose(t.grad.numpy(), ref, rtol=1e-6, atol=1e-7 * max(1.0, np.abs(ref).max())), k
    # the test data must exercise the clamp and the far-from-unit quaternions
    assert (g["act_colors_precomp"][:, :3] == 0).any() and (g["act_colors_precomp"][:, :3] > 0).any()
    assert np.abs(np.linalg.norm(g["raw_rotation"], axis=1) - 1).max() > 0.5
