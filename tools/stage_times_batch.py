"""GPU box: per-stage device times (library stage timers) of the config-2 step for 1, 2, 4 and 8 views per launch."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from splatloc_b200 import _C, _capi, scenes

dev = torch.device("cuda", 0)
P, C = int(os.environ.get("STAGE_P", 500_000)), 4
sc = {k: v.to(dev) for k, v in scenes.make_scene(P, seed=0, channels=C).items()}
bg = torch.zeros(C, device=dev)
E = torch.Tensor([])
cams = [scenes.make_camera(seed=i).to(dev) for i in range(8)]
H, W = cams[0].H, cams[0].W
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
lib = _capi.load()
for V in (1, 2, 4, 8):
    sets = [cams[s * V:(s + 1) * V] for s in range(8 // V)]
    def run(cs):
        vm = torch.stack([c.viewmatrix for c in cs]); pm = torch.stack([c.projmatrix for c in cs]); cp = torch.stack([c.campos for c in cs])
        fw = _C.rasterize_gaussians_batch(bg, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"], sc["rotations"], 1.0, E,
                                          E.bool(), vm, pm, cs[0].tanfovx, cs[0].tanfovy, H, W, E, 0, cp, False, False)
        g = [torch.randn_like(fw[k]) / (H * W) for k in (1, 2, 3)]
        _C.rasterize_gaussians_backward_batch(bg, sc["means3D"], fw[4], sc["colors_precomp"], sc["scales"], sc["rotations"], 1.0, E, vm, pm,
                                              cs[0].tanfovx, cs[0].tanfovy, *g, E, 0, cp, fw[5], fw[0], fw[6], fw[7], fw[3], False)
    for _ in range(3):
        for cs in sets: run(cs)
    torch.cuda.synchronize()
    lib.sl_profile_enable(1)
    for _ in range(5):
        for cs in sets:
            flush.zero_(); run(cs)
    prof = _capi.profile_read(); lib.sl_profile_enable(0)
    tot = sum(t / n for t, n in prof.values() if n)
    print(f"views/launch {V}: " + "  ".join(f"{k} {t / n / V * 1e3:.1f}" for k, (t, n) in prof.items() if n) + f"  | sum per view {tot / V * 1e3:.1f} us")
