"""GPU box: per-stage device times (library stage timers) of one 640x480 fwd+bwd view against the number of
Gaussians (the sweep of bench.py), with how many tiles exceed the 256-thread sort's segment length."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from splatloc_b200 import _C, _capi, scenes
from splatloc_b200.diff_gauss import interpret_imgBuffer

dev = torch.device("cuda", 0)
C = 4
E = torch.Tensor([])
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
lib = _capi.load()
lib.sl_set_binning_mode(int(os.environ.get("BINNING_MODE", "0")))
for P in [int(x) for x in os.environ.get("SWEEP_P", "500000,1200000,3000000").split(",")]:
    sc = {k: v.to(dev) for k, v in scenes.make_scene(P, seed=0, channels=C).items()}
    bg = torch.zeros(C, device=dev)
    cams = [scenes.make_camera(seed=i).to(dev) for i in range(4)]
    H, W = cams[0].H, cams[0].W
    n_tiles = ((H + 15) // 16) * ((W + 15) // 16)
    def run(c):
        fw = _C.rasterize_gaussians(bg, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"], sc["rotations"], 1.0, E,
                                    E.bool(), c.viewmatrix, c.projmatrix, c.tanfovx, c.tanfovy, H, W, E, 0, c.campos, False, False)
        g = [torch.randn_like(fw[k]) / (H * W) for k in (1, 2, 3)]
        _C.rasterize_gaussians_backward(bg, sc["means3D"], fw[4], sc["colors_precomp"], sc["scales"], sc["rotations"], 1.0, E,
                                        c.viewmatrix, c.projmatrix, c.tanfovx, c.tanfovy, *g, E, 0, c.campos, fw[5], fw[0], fw[6],
                                        fw[7], fw[3], False)
        return fw
    for _ in range(2):
        for c in cams: fw = run(c)
    torch.cuda.synchronize()
    rng = interpret_imgBuffer(fw[7], H * W, n_tiles)["ranges"].long()
    seg = rng[:, 1] - rng[:, 0]
    lib.sl_profile_enable(1)
    for _ in range(4):
        for c in cams:
            flush.zero_(); run(c)
    prof = _capi.profile_read(); lib.sl_profile_enable(0)
    tot = sum(t / n for t, n in prof.values() if n)
    print(f"P {P}: pairs {int(fw[0])}  longest tile {int(seg.max())}  tiles > 4096: {int((seg > 4096).sum())} of {n_tiles} | "
          + "  ".join(f"{k} {t / n * 1e3:.1f}" for k, (t, n) in prof.items() if n) + f"  | sum {tot * 1e3:.1f} us", flush=True)
