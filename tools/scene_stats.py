"""GPU box: per-tile list length / n_contrib statistics of the bench scene (load-balance evidence)
and how many (8x4 patch, Gaussian) evaluations actually contain a contributing pixel (culling headroom)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from splatloc_b200 import scenes, _C
from splatloc_b200.diff_gauss import interpret_imgBuffer, interpret_geomBuffer, interpret_binningBuffer
P = int(sys.argv[1]) if len(sys.argv) > 1 else 500000
sc = {k: v.cuda() for k, v in scenes.make_scene(P, 0, channels=4).items()}
E = torch.Tensor([])
for cs in range(3):
    cam = scenes.make_camera(cs).to("cuda")
    R, color, depth, alpha, radii, geom, binning, img = _C.rasterize_gaussians(
        torch.zeros(4, device="cuda"), sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"], sc["rotations"],
        1.0, E, E.bool(), cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, cam.H, cam.W, E, 0, cam.campos, False, False)
    im = interpret_imgBuffer(img, cam.H * cam.W, 1200)
    ge = interpret_geomBuffer(geom, P)
    bi = interpret_binningBuffer(binning, R)
    ranges = im["ranges"]
    ln = (ranges[:, 1] - ranges[:, 0]).float()
    nc = im["n_contrib"].float().view(cam.H, cam.W)
    t = nc.view(30, 16, 40, 16).permute(0, 2, 1, 3).reshape(1200, 256)
    tmax = t.max(1).values
    q = lambda x, p: float(torch.quantile(x, p))
    print(f"cam {cs}: R={R} vis={int((radii>0).sum())} len mean {float(ln.mean()):.0f} p50 {q(ln,.5):.0f} p90 {q(ln,.9):.0f} p99 {q(ln,.99):.0f} max {float(ln.max()):.0f} | "
          f"tile max n_contrib mean {float(tmax.mean()):.0f} p90 {q(tmax,.9):.0f} max {float(tmax.max()):.0f} | pixel n_contrib mean {float(nc.mean()):.0f} | alpha mean {float(alpha.mean()):.3f}")
    rad = radii[radii > 0].float()
    print(f"    radius mean {float(rad.mean()):.1f} p50 {q(rad,.5):.0f} p90 {q(rad,.9):.0f} p99 {q(rad,.99):.0f} max {float(rad.max()):.0f}; sum len*256 = {float(ln.sum())*256:.3e}, sum n_contrib = {float(nc.sum()):.3e}")
    # culling headroom on a sample of tiles
    g = torch.Generator().manual_seed(0)
    tiles = torch.randperm(1200, generator=g)[:150].tolist()
    tot_wg = any_wg = tot_pg = any_pg = aabb_wg = 0
    lane_hist = torch.zeros(33, dtype=torch.long)
    for tl in tiles:
        lo, hi = int(ranges[tl, 0]), int(ranges[tl, 1])
        if hi <= lo:
            continue
        ids = bi["point_list"][lo:hi].long()
        ty, tx = divmod(tl, 40)
        ys = torch.arange(ty * 16, ty * 16 + 16, device="cuda").float()
        xs = torch.arange(tx * 16, tx * 16 + 16, device="cuda").float()
        py, px = torch.meshgrid(ys, xs, indexing="ij")
        xy = ge["means2D"][ids]; co = ge["conic_opacity"][ids]
        dx = xy[None, None, :, 0] - px[:, :, None]; dy = xy[None, None, :, 1] - py[:, :, None]
        power = -0.5 * (co[:, 0] * dx * dx + co[:, 2] * dy * dy) - co[:, 1] * dx * dy
        a = torch.clamp_max(co[:, 3] * torch.exp(power), 0.99)
        keep = (power <= 0) & (a >= 1 / 255.0)                       # (16,16,L)
        pos = torch.arange(1, hi - lo + 1, device="cuda")
        live = pos[None, None, :] <= t[tl].view(16, 16)[:, :, None]      # entries up to the pixel's last contributor
        k2 = keep & live
        w = k2.view(4, 4, 2, 8, -1).permute(0, 2, 1, 3, 4).reshape(8, 32, -1)   # 8 patches of 8x4
        lw = live.view(4, 4, 2, 8, -1).permute(0, 2, 1, 3, 4).reshape(8, 32, -1)
        tot_wg += int(lw.any(1).sum()); any_wg += int(w.any(1).sum())
        lane_hist += torch.bincount(w.sum(1).reshape(-1).cpu(), minlength=33)[:33]
        # AABB of the alpha >= 1/255 ellipse vs each 8x4 patch
        t2 = 2 * torch.log(255 * co[:, 3])
        det = co[:, 0] * co[:, 2] - co[:, 1] ** 2
        hx = torch.sqrt(torch.clamp_min(t2, 0) * co[:, 2] / det); hy = torch.sqrt(torch.clamp_min(t2, 0) * co[:, 0] / det)
        for wq in range(8):
            x0 = tx * 16 + (wq & 1) * 8; y0 = ty * 16 + (wq >> 1) * 4
            hit = (xy[:, 0] + hx >= x0) & (xy[:, 0] - hx <= x0 + 7) & (xy[:, 1] + hy >= y0) & (xy[:, 1] - hy <= y0 + 3) & (t2 > 0)
            aabb_wg += int((hit & lw[wq].any(0)).sum())
        tot_pg += int(live.sum()); any_pg += int(k2.sum())
    lh = lane_hist[1:].float(); lh = lh / lh.sum()
    print("    blending lanes per any-blend (patch,Gaussian): 1:%.2f 2:%.2f 3-4:%.2f 5-8:%.2f 9-16:%.2f 17-31:%.2f 32:%.2f" % (
        lh[0], lh[1], lh[2:4].sum(), lh[4:8].sum(), lh[8:16].sum(), lh[16:31].sum(), lh[31]))
    print(f"    sample: (patch,Gaussian) evaluated {tot_wg}, with >=1 blending pixel {any_wg} ({100*any_wg/max(tot_wg,1):.1f}%), AABB survivors {aabb_wg} ({100*aabb_wg/max(tot_wg,1):.1f}%); "
          f"(pixel,Gaussian) evaluated {tot_pg}, blended {any_pg} ({100*any_pg/max(tot_pg,1):.1f}%)")
