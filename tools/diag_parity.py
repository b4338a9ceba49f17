"""Diagnostic (GPU box): run-to-run noise of the reference's atomics vs our difference to it, on the
bench scene; prints, per camera and gradient tensor, the worst element and what the reference's own
runs say about that element."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tests import helpers as H
from tests.test_gpu_bench_parity import _bench_inputs
from oracle import ref_cuda
from splatloc_b200 import _C, scenes

sc = {k: v.cuda() for k, v in scenes.make_scene(500_000, seed=0, channels=4).items()}
for cam in range(8):
    d = _bench_inputs(sc, cam)
    refs = [H.run_backend(ref_cuda, d) for _ in range(4)]
    ours = [H.run_backend(_C, d) for _ in range(2)]
    print(f"camera {cam} R={refs[0]['R']}")
    for k in H.GRAD_KEYS:
        y = refs[0][k].double()
        if y.numel() == 0:
            continue
        stack = torch.stack([r[k].double() for r in refs])
        spread = (stack.max(0).values - stack.min(0).values)
        err = (ours[0][k].double() - y).abs()
        oo = (ours[0][k].double() - ours[1][k].double()).abs()
        i = int(err.argmax())
        strict = 1e-5 + 1e-4 * y.abs()
        n_out = int((err > strict).sum())
        print(f"  {k:10s} max|ref|={float(y.abs().max()):.3e} ref-spread max={float(spread.max()):.2e} rms={float(spread.pow(2).mean().sqrt()):.2e}"
              f" | ours-ref max={float(err.max()):.2e} rms={float(err.pow(2).mean().sqrt()):.2e} ours-ours max={float(oo.max()):.2e}"
              f" | outside strict: {n_out} | worst elem: ours={float(ours[0][k].flatten()[i]):.6e} refs={[float(r[k].flatten()[i]) for r in refs]}"
              f" local spread={float(spread.flatten()[i]):.2e}")
