"""Diagnostic (GPU box): run-to-run noise of the reference's atomics vs our difference to it."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tests import helpers as H
from oracle import ref_cuda
from splatloc_b200 import _C

cases = [dict(P=20000, seed=5, size=(320, 240), tile_mask_seed=7), dict(P=20000, seed=2, size=(320, 240)),
         dict(P=200000, seed=0)]
for case in cases:
    d = H.scene_inputs(**case)
    b1 = H.run_backend(ref_cuda, d)
    b2 = H.run_backend(ref_cuda, d)
    a1 = H.run_backend(_C, d)
    a2 = H.run_backend(_C, d)
    print(case, "R", b1["R"])
    for k in H.GRAD_KEYS + ("color", "depth", "alpha"):
        y = b1[k].double()
        if y.numel() == 0:
            continue
        def e(u, v):
            return float((u[k].double() - v[k].double()).abs().max())
        def worst(u, v):
            err = (u[k].double() - v[k].double()).abs()
            i = int(err.argmax())
            return f"{float(err.flatten()[i]):.2e}@|y|={float(v[k].double().abs().flatten()[i]):.2e}"
        print(f"  {k:10s} max|ref|={float(y.abs().max()):.3e} ref-ref={e(b1,b2):.2e} ours-ours={e(a1,a2):.2e} ours-ref={worst(a1,b1)}")
