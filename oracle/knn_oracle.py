"""TEST INFRASTRUCTURE -- CPU restatement of simple-knn's distCUDA2 result
(submodules/simple-knn/simple_knn.cu:132-184): for every point the mean of the three smallest
squared distances to the other points, each distance evaluated in float32 as the reference's
`d.x*d.x + d.y*d.y + d.z*d.z` compiles (nvcc contracts it to fma(dz, dz, fma(dx, dx, dy*dy))),
inserted in ascending order (updateKBest, :132-146) and averaged as `(b0 + b1 + b2) / 3.0f` (:183).
The Morton sort and box pruning of the reference (:46-130,148-181) only accelerate the search.

Parity status: PINNED bit-for-bit against the reference CUDA kernel's output on a B200
(tests/golden/knn_*.npz, made by tests/golden/make_golden.py).
"""
from __future__ import annotations

import numpy as np

FLT_MAX = np.float32(3.4028234663852886e38)


def _fma32(a, b, c):
    """float32 fused multiply-add emulated in float64 (a*b is exact in float64)."""
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(np.float32)


def pair_dist(ref, pts):
    """Squared distances from `ref` (3,) float32 to `pts` (n,3) float32, reference rounding."""
    d = pts - ref[None, :]                      # float32 subtraction
    dx, dy, dz = d[:, 0], d[:, 1], d[:, 2]
    yy = (dy * dy).astype(np.float32)
    return _fma32(dz, dz, _fma32(dx, dx, yy))


def mean3(best):
    """(b0 + b1 + b2) / 3.0f in float32, b ascending."""
    b = np.asarray(best, dtype=np.float32)
    return np.float32(np.float32(np.float32(b[0] + b[1]) + b[2]) / np.float32(3.0))   # overflow -> inf, as on the GPU


def dist_cuda2(points: np.ndarray, k_candidates: int = 16) -> np.ndarray:
    """points (P,3) float32 -> (P,) float32.  Candidates come from an exact float64 k-d tree query
    (k_candidates nearest), the float32 distances are then re-evaluated with the reference's rounding
    and the three smallest kept; for P <= 2048 the search is brute force."""
    pts = np.ascontiguousarray(points, dtype=np.float32)
    P = pts.shape[0]
    out = np.empty(P, dtype=np.float32)
    with np.errstate(over="ignore"):
        if P <= 2048:
            for i in range(P):
                d = pair_dist(pts[i], pts)
                d[i] = FLT_MAX
                best = np.sort(d)[:3]
                best = np.concatenate([best, np.full(3 - best.size, FLT_MAX, np.float32)]) if best.size < 3 else best
                if P - 1 < 3:
                    best[P - 1:] = FLT_MAX
                out[i] = mean3(best.astype(np.float32))
            return out
        from scipy.spatial import cKDTree
        tree = cKDTree(pts.astype(np.float64))
        k = min(P, k_candidates + 1)
        _, nn = tree.query(pts.astype(np.float64), k=k)
        for i in range(P):
            cand = nn[i][nn[i] != i]
            d = np.sort(pair_dist(pts[i], pts[cand]))[:3]
            out[i] = mean3(d)
    return out
