"""TEST INFRASTRUCTURE -- CPU restatement (float32 torch, no autograd) of what SplatLoc's mapping loop
does to the Gaussian store between rasterizer calls:

    gaussian_splatting/scene/gaussian_model.py:250-300   training_setup: torch.optim.Adam(lr per group, eps=1e-15)
    train_gaussians.py:232-234                           key primitives (marker > 0.005): zero position gradient
    gaussian_model.py:677-679, train_gaussians.py:241-244   densification statistics, max_radii2D
    gaussian_model.py:602-675                            densify_and_clone, densify_and_split (N = 2), final prune,
    gaussian_model.py:493-551                            with the optimizer-state surgery of prune / cat

written as ONE pass over the rows (which rows survive, which are cloned or split, where each lands)
instead of the reference's sequence of cat / mask operations.  Pinned by tests/golden/store_*.npz,
which tests/golden/make_store_golden.py produced by running the reference's own GaussianModel on the
CPU (tests/test_store_cpu.py).  Only tests/, smoke() and bench.py may import this module.
"""
import math

import torch

GROUPS = ("xyz", "f_dc", "f_rest", "opacity", "marker", "kp_score", "scaling", "rotation")


def adam_step(params, grads, exp_avg, exp_avg_sq, lrs, step, beta1=0.9, beta2=0.999, eps=1e-15, marker=None):
    """In place.  `params`, `grads`, moments: dicts name -> tensor (P, ...); lrs: dict name -> float;
    `marker` (P, 1): rows with marker > 0.005 use a zero xyz gradient."""
    bc1, bc2 = 1 - beta1 ** step, 1 - beta2 ** step
    for name, p in params.items():
        g = grads.get(name)
        if g is None or p.numel() == 0:
            continue
        if name == "xyz" and marker is not None:
            g = g.clone()
            g[marker.squeeze() > 0.005] = 0
        exp_avg[name].lerp_(g, 1 - beta1)
        exp_avg_sq[name].mul_(beta2).addcmul_(g, g, value=1 - beta2)
        denom = (exp_avg_sq[name].sqrt() / math.sqrt(bc2)).add_(eps)
        p.addcdiv_(exp_avg[name], denom, value=-(lrs[name] / bc1))


def densification_stats(accum, denom, max_radii2D, grad2D, radii):
    """In place; grad2D (V, P, 3), radii (V, P)."""
    for v in range(radii.shape[0]):
        vis = radii[v] > 0
        max_radii2D[vis] = torch.max(max_radii2D[vis], radii[v][vis].float())
        accum[vis] += torch.norm(grad2D[v][vis, :2], dim=-1, keepdim=True)
        denom[vis] += 1


def densify_and_prune(params, exp_avg, exp_avg_sq, accum, denom, noise, max_grad, min_opacity, extent,
                      max_screen_size, percent_dense=0.01, primitive_reg=False):
    """-> (new params, new exp_avg, new exp_avg_sq, (kept, clones, split-with-children, split)).
    `noise` (2 * n_split, 3): the standard normal draws of the split, child k of the j-th split
    Gaussian at row k * n_split + j."""
    grads = accum / denom
    grads[grads.isnan()] = 0.0
    grad = grads.squeeze(-1)
    s_act = torch.exp(params["scaling"])
    s_max = s_act.max(dim=1).values
    hot = grad >= max_grad
    clone = hot & (s_max <= percent_dense * extent)
    split = hot & (s_max > percent_dense * extent)
    opacity = torch.sigmoid(params["opacity"]).squeeze(-1)
    prune_self, prune_child = opacity < min_opacity, opacity < min_opacity
    if max_screen_size:
        prune_self = prune_self | (s_max > 0.1 * extent)
        prune_child = prune_child | ((s_max / (0.8 * 2)) > 0.1 * extent)
    if primitive_reg:
        plain = params["marker"].squeeze(-1) <= 0.005
        prune_self, prune_child = prune_self & plain, prune_child & plain
    keepA, keepB, keepC = ~split & ~prune_self, clone & ~prune_self, split & ~prune_child
    n_split = int(split.sum())
    rank = torch.cumsum(split.long(), 0) - 1

    # children of the split Gaussians
    q = params["rotation"]
    q = q / torch.sqrt((q * q).sum(1, keepdim=True))
    r, x, y, z = q.unbind(1)
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                     2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                     2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1).view(-1, 3, 3)
    out, m_out, v_out = {}, {}, {}
    for name in params:
        p = params[name]
        rows = [p[keepA], p[keepB]]
        for k in range(2):
            sel = keepC
            if name == "xyz":
                zz = noise[k * n_split + rank[sel]]
                rows.append(torch.bmm(R[sel], (s_act[sel].expand(-1, 3) * zz).unsqueeze(-1)).squeeze(-1) + p[sel])
            elif name == "scaling":
                rows.append(torch.log(s_act[sel] / (0.8 * 2)))
            else:
                rows.append(p[sel])
        out[name] = torch.cat(rows, 0)
        if name in exp_avg:
            n_new = out[name].shape[0] - int(keepA.sum())
            m_out[name] = torch.cat([exp_avg[name][keepA], torch.zeros((n_new,) + tuple(p.shape[1:]))], 0)
            v_out[name] = torch.cat([exp_avg_sq[name][keepA], torch.zeros((n_new,) + tuple(p.shape[1:]))], 0)
    return out, m_out, v_out, (int(keepA.sum()), int(keepB.sum()), int(keepC.sum()), n_split)
