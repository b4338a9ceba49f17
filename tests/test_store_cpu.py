"""CPU: the oracle of the optimiser step and densification (oracle/store_oracle.py) against outputs of
the reference's own GaussianModel (tests/golden/store_*.npz, generated on the CPU from /root/reference
by tests/golden/make_store_golden.py)."""
import os

import numpy as np
import pytest
import torch

from oracle import store_oracle as S

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["aniso_400", "iso_300"]
KEY = dict(xyz="_xyz", f_dc="_features_dc", f_rest="_features_rest", opacity="_opacity", marker="_marker",
           kp_score="_kp_score", scaling="_scaling", rotation="_rotation")


def load(name):
    return dict(np.load(os.path.join(GOLDEN, f"store_{name}.npz")))


def initial_state(g, device="cpu"):
    params = {n: torch.from_numpy(g["in" + KEY[n]]).to(device).clone() for n in S.GROUPS}
    m = {n: torch.zeros_like(p) for n, p in params.items()}
    v = {n: torch.zeros_like(p) for n, p in params.items()}
    lrs = dict(zip(S.GROUPS, (float(x) for x in g["lrs"])))
    return params, m, v, lrs


def close(a, b, rtol=2e-6, atol=None):
    """rtol on the element plus a floor of a few float32 ulps of the tensor's largest magnitude (an update of
    size lr lands on parameters near zero with the rounding of O(1) operands; hosts differ by an ulp there)."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    if a.shape != b.shape:
        return False
    if atol is None:
        atol = 4e-7 * float(np.abs(b).max()) if b.size else 0.0
    return bool(np.all(np.abs(a - b) <= atol + rtol * np.abs(b)))


@pytest.mark.parametrize("name", CASES)
def test_oracle_reproduces_the_reference_store_updates(name):
    g = load(name)
    params, m, v, lrs = initial_state(g)
    for step in range(3):
        grads = {n: torch.from_numpy(g[f"grad{step}{KEY[n]}"]) for n in S.GROUPS}
        S.adam_step(params, grads, m, v, lrs, step + 1, marker=params["marker"])
    for n in S.GROUPS:
        assert close(params[n].numpy(), g["adam" + KEY[n]]), f"adam {n}"
        if params[n].numel():
            assert close(m[n].numpy(), g["adam_m" + KEY[n]]) and close(v[n].numpy(), g["adam_v" + KEY[n]]), f"moments {n}"
    P = params["xyz"].shape[0]
    accum, denom, maxr = torch.zeros(P, 1), torch.zeros(P, 1), torch.zeros(P)
    S.densification_stats(accum, denom, maxr, torch.from_numpy(g["stat_grad2D"]), torch.from_numpy(g["stat_radii"]))
    assert close(accum.numpy(), g["stat_accum"]) and np.array_equal(denom.numpy(), g["stat_denom"])
    assert np.array_equal(maxr.numpy(), g["stat_max_radii2D"])
    max_grad, min_opacity, extent, max_screen, pd, preg = (float(x) for x in g["dense_args"])
    out, m2, v2, counts = S.densify_and_prune(params, m, v, accum, denom, torch.from_numpy(g["noise"]), max_grad, min_opacity,
                                              extent, max_screen or None, pd, bool(preg))
    assert g["noise"].shape[0] == 2 * counts[3]
    assert counts[0] + counts[1] + 2 * counts[2] == g["out_xyz"].shape[0]
    assert counts[1] > 0 and counts[2] > 0 and counts[0] < P, counts           # every branch exercised
    for n in S.GROUPS:
        assert close(out[n].numpy(), g["out" + KEY[n]]), f"densify {n}"
        if out[n].numel():
            assert close(m2[n].numpy(), g["out_m" + KEY[n]]) and close(v2[n].numpy(), g["out_v" + KEY[n]]), f"densify moments {n}"
    assert not g["out_accum"].any() and not g["out_denom"].any() and not g["out_max_radii2D"].any()
