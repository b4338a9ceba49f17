"""GPU box: A/B of the two ways the forward blend kernel brings pair records on chip (sl_set_record_stream):
per-lane loads of mask survivors (0) against bulk-copy staging of whole chunks (1); checks that the
results are identical and prints the blend_forward stage time for 1 and 8 views per launch."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from splatloc_b200 import _C, _capi, scenes

dev = torch.device("cuda", 0)
lib = _capi.load()
E = torch.Tensor([])
for P in (500_000, 1_200_000):
    sc = {k: v.to(dev) for k, v in scenes.make_scene(P, seed=0, channels=4).items()}
    bg = torch.zeros(4, device=dev)
    cams = [scenes.make_camera(seed=i).to(dev) for i in range(8)]
    H, W = cams[0].H, cams[0].W
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ref = {}
    for V in (1, 8):
        sets = [cams[s * V:(s + 1) * V] for s in range(8 // V)]
        for mode in (0, 1):
            lib.sl_set_record_stream(mode)
            def run(cs, touch=False):
                vm = torch.stack([c.viewmatrix for c in cs]); pm = torch.stack([c.projmatrix for c in cs]); cp = torch.stack([c.campos for c in cs])
                return _C.rasterize_gaussians_batch(bg, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"], sc["rotations"],
                                                    1.0, E, E.bool(), vm, pm, cs[0].tanfovx, cs[0].tanfovy, H, W, E, 0, cp, False, False, touch)
            for _ in range(2):
                for cs in sets: run(cs)
            outs = [run(cs, True) for cs in sets]
            key = (V,)
            if mode == 0:
                ref[key] = outs
            else:
                for a, b in zip(ref[key], outs):
                    for k in (1, 2, 3, 8):
                        assert torch.equal(a[k], b[k]), "bulk-copy variant differs"
            torch.cuda.synchronize(); lib.sl_profile_enable(1)
            for _ in range(5):
                for cs in sets:
                    flush.zero_(); run(cs)
            prof = _capi.profile_read(); lib.sl_profile_enable(0)
            t, n = prof["blend_forward"]
            print(f"P={P} views/launch {V} record_stream={mode}: blend_forward {t / n / V * 1e3:.1f} us per view (results identical)")
    lib.sl_set_record_stream(0)
