"""GPU box: per-camera pair statistics and stage times of the bench scene (which views are heavy)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from splatloc_b200 import scenes, _C, _capi
from splatloc_b200.diff_gauss import interpret_imgBuffer
P = int(sys.argv[1]) if len(sys.argv) > 1 else 500000
sc = {k: v.cuda() for k, v in scenes.make_scene(P, 0, channels=4).items()}
E = torch.Tensor([])
lib = _capi.load()
for cs in range(8):
    cam = scenes.make_camera(cs).to("cuda")
    g = [t.cuda() for t in scenes.make_pixel_grads(cs, 4, cam.H, cam.W)]
    def step():
        R, color, depth, alpha, radii, geom, binning, img = _C.rasterize_gaussians(
            torch.zeros(4, device="cuda"), sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"], sc["rotations"],
            1.0, E, E.bool(), cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, cam.H, cam.W, E, 0, cam.campos, False, False)
        _C.rasterize_gaussians_backward(torch.zeros(4, device="cuda"), sc["means3D"], radii, sc["colors_precomp"], sc["scales"],
                                        sc["rotations"], 1.0, E, cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy,
                                        g[0], g[1], g[2], E, 0, cam.campos, geom, R, binning, img, alpha, False)
        return R, radii, img
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    lib.sl_profile_enable(1)
    for _ in range(10):
        R, radii, img = step()
    torch.cuda.synchronize()
    prof = _capi.profile_read()
    lib.sl_profile_enable(0)
    ranges = interpret_imgBuffer(img, cam.H * cam.W, 1200)["ranges"]
    ln = (ranges[:, 1] - ranges[:, 0]).float()
    st = " ".join(f"{k[:9]} {1e3 * t / max(c, 1):4.0f}" for k, (t, c) in prof.items() if c)
    print(f"cam {cs}: R={R} V={int((radii > 0).sum())} max tile {int(ln.max())} tiles>1024 {int((ln > 1024).sum())} >2048 {int((ln > 2048).sum())} >4096 {int((ln > 4096).sum())} | {st}")
