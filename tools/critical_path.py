"""GPU box: how long does the blend of the single longest tile take on its own (the serial critical path of a
one-view launch), against the whole frame?  Renders the bench scene through a tile mask that keeps only
the k longest tiles."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from splatloc_b200 import _C, _capi, scenes
from splatloc_b200.diff_gauss import interpret_imgBuffer

dev = torch.device("cuda", 0)
P, C = 500_000, 4
sc = {k: v.to(dev) for k, v in scenes.make_scene(P, seed=0, channels=C).items()}
bg = torch.zeros(C, device=dev); E = torch.Tensor([])
lib = _capi.load()
for ci in (0, 3):
    cam = scenes.make_camera(seed=ci).to(dev)
    H, W = cam.H, cam.W
    nt = ((H + 15) // 16) * ((W + 15) // 16)
    def run(mask):
        fw = _C.rasterize_gaussians(bg, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"], sc["rotations"], 1.0, E,
                                    mask, cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, H, W, E, 0, cam.campos, False, False)
        g = [torch.randn_like(fw[k]) / (H * W) for k in (1, 2, 3)]
        _C.rasterize_gaussians_backward(bg, sc["means3D"], fw[4], sc["colors_precomp"], sc["scales"], sc["rotations"], 1.0, E,
                                        cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, *g, E, 0, cam.campos, fw[5], fw[0], fw[6], fw[7], fw[3], False)
        return fw
    fw = run(E.bool())
    rg = interpret_imgBuffer(fw[7], H * W, nt)["ranges"]
    lens = (rg[:, 1] - rg[:, 0]).long()
    order = torch.argsort(lens, descending=True)
    print(f"camera {ci}: R={fw[0]} tiles {nt} longest {lens[order[:5]].tolist()} mean {float(lens.float().mean()):.0f}")
    for k in (1, 8, 64, nt):
        mask = torch.zeros(nt, dtype=torch.bool, device=dev); mask[order[:k]] = True
        for _ in range(3): run(mask)
        torch.cuda.synchronize(); lib.sl_profile_enable(1)
        for _ in range(10): run(mask)
        prof = _capi.profile_read(); lib.sl_profile_enable(0)
        print(f"   {k:5d} longest tiles: " + "  ".join(f"{n} {t / c * 1e3:.1f}" for n, (t, c) in prof.items() if c and n in ("sort_pairs", "blend_forward", "blend_backward", "emit_pairs")))
