"""GPU box: cProfile of the host side of N fwd+bwd steps through the public autograd API (who spends the
CPU time between kernel launches).  Usage: python tools/host_profile.py [steps]"""
import cProfile, os, pstats, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from splatloc_b200 import scenes
from splatloc_b200.diff_gauss import GaussianRasterizationSettings as Settings, GaussianRasterizer as Rasterizer

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
dev = torch.device("cuda:0")
sc = {k: v.to(dev) for k, v in scenes.make_scene(500000, 0, channels=4).items()}
cam = scenes.make_camera(0).to(dev)
leaves = {k: sc[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "colors_precomp", "scales", "rotations")}
means2D = torch.zeros_like(leaves["means3D"], requires_grad=True)
g = [t.to(dev) for t in scenes.make_pixel_grads(0, 4, cam.H, cam.W)]
bg = torch.zeros(4, device=dev)


def step():
    rs = Settings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, bg, 1.0, cam.viewmatrix, cam.projmatrix, 0, cam.campos, False, False)
    for t in leaves.values():
        t.grad = None
    means2D.grad = None
    color, depth, alpha, radii = Rasterizer(rs)(means3D=leaves["means3D"], means2D=means2D, opacities=leaves["opacities"],
                                                colors_precomp=leaves["colors_precomp"], scales=leaves["scales"],
                                                rotations=leaves["rotations"])
    torch.autograd.backward([color, depth, alpha], g)


for _ in range(20):
    step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(steps):
    step()
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(35)
