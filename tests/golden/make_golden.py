"""Generate tests/golden/*.npz from the UNMODIFIED reference CUDA code (oracle/_ref), on a GPU box:

    gpurun -- python tests/golden/make_golden.py          # writes gpurun_out/golden/*.npz
    cp gpurun_out/golden/*.npz tests/golden/

Each file holds the exact inputs (so nothing depends on RNG reproducibility) and every output,
gradient and integer intermediate of the reference for one small case.  The reference ships no
golden vectors of its own; these are its outputs, run here.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import ref_cuda  # noqa: E402
from splatloc_b200 import scenes  # noqa: E402
from tests import helpers as H  # noqa: E402

CASES = {
    "c4_precomp": dict(P=3000, seed=1, size=(160, 120)),
    "sh3_bg": dict(P=2500, seed=8, sh_degree=3, size=(128, 96), bg=[0.1, 0.2, 0.3]),
    "cov3d_mask": dict(P=3000, seed=4, size=(150, 100), use_cov3d=True, tile_mask_seed=7, bg=[0.3, 0.5, 0.7, 0.2]),
}


def main():
    out_dir = os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out_dir, exist_ok=True)
    for name, case in CASES.items():
        d = H.scene_inputs(**case)
        ref = H.run_backend(ref_cuda, d)
        cam = d["cam"]
        rec = {"in_" + k: d[k].cpu().numpy() for k in ("means3D", "scales", "rotations", "opacities", "bg")}
        for k in ("colors_precomp", "shs", "cov3D_precomp", "tile_mask"):
            if k in d:
                rec["in_" + k] = d[k].cpu().numpy()
        rec.update(in_viewmatrix=cam.viewmatrix.cpu().numpy(), in_projmatrix=cam.projmatrix.cpu().numpy(),
                   in_campos=cam.campos.cpu().numpy(), in_tanfov=np.array([cam.tanfovx, cam.tanfovy]),
                   in_size=np.array([cam.W, cam.H]), in_sh_degree=np.array(d["sh_degree"]),
                   in_visible=ref_cuda.mark_visible(d["means3D"], cam.viewmatrix, cam.projmatrix).cpu().numpy())
        for i, g in enumerate(d["grads"]):
            rec[f"in_grad{i}"] = g.cpu().numpy()
        for k, v in ref.items():
            rec["ref_" + k] = v.numpy() if isinstance(v, torch.Tensor) else np.array(v)
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), **rec)
        print(name, "R", ref["R"], {k: tuple(v.shape) for k, v in ref.items() if isinstance(v, torch.Tensor)})
    for P in (7, 5000):
        pts = scenes.make_knn_points(P, seed=2).cuda()
        out = ref_cuda.distCUDA2(pts)
        np.savez_compressed(os.path.join(out_dir, f"knn_{P}.npz"), in_points=pts.cpu().numpy(), ref_dist=out.cpu().numpy())
        print("knn", P, float(out[:P].median()))


if __name__ == "__main__":
    main()
