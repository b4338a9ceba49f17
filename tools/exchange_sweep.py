"""GPU box, under torchrun (N >= 2): cost of the deferred gradient push (GraphedViewStep exchange_mode="next_step")
against the number of CTAs of its push kernel (exchange_push_ctas; -1: no push, protocol only; SWEEP_CTAS=a,b,c picks the points).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29503 tools/exchange_sweep.py
"""
import datetime
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from splatloc_b200 import scenes
from splatloc_b200.exchange import NvlsGradientExchange
from splatloc_b200.graphs import GraphedViewStep


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("NCCL_DEBUG", "WARN")
    dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=120))
    P, C = 500_000, 4
    sc = {k: v.to(dev) for k, v in scenes.make_scene(P, seed=0, channels=C).items()}
    params = {k: sc[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "colors_precomp", "scales", "rotations")}
    bg = torch.zeros(C, device=dev)
    cams = [scenes.make_camera(seed=100 * rank + i).to(dev) for i in range(4)]
    H, W = cams[0].H, cams[0].W
    ups = [torch.cat(scenes.make_pixel_grads(100 * rank + i, C, H, W, normalised=False), 0).to(dev) for i in range(4)]
    xch = NvlsGradientExchange(P, dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    mk = lambda **kw: GraphedViewStep(params, H, W, cams[0].tanfovx, cams[0].tanfovy, bg, max_pairs=2_500_000, **kw)   # noqa: E731

    def timed(fn, n=40):
        for i in range(6): fn(i)
        torch.cuda.synchronize(); dist.barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
        for i in range(n):
            flush.zero_(); ev[i][0].record(); fn(i); ev[i][1].record()
        torch.cuda.synchronize(); dist.barrier()
        t = torch.tensor([sum(a.elapsed_time(b) for a, b in ev) / n], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    def load(st, i):
        st.viewmatrices[0].copy_(cams[i % 4].viewmatrix); st.projmatrices[0].copy_(cams[i % 4].projmatrix)
        st.cam_positions[0].copy_(cams[i % 4].campos); st.upstream.copy_(ups[i % 4])

    plain = mk()
    t0 = timed(lambda i: (load(plain, i), plain.replay()))
    out = [f"no exchange {t0:.4f}"]
    for ctas in [int(x) for x in os.environ.get("SWEEP_CTAS", "-1,8,32,74,148,296,1184").split(",")]:
        pair = [mk(exchange=xch, exchange_buffer=b, exchange_mode="next_step", exchange_push_ctas=ctas) for b in (0, 1)]
        xch.reset()
        t = timed(lambda i: (load(pair[i % 2], i), pair[i % 2].replay()))
        xch.flush_deferred(1); torch.cuda.synchronize()
        assert not xch.timed_out
        xch.reset()
        out.append(f"{ctas} CTAs {t:.4f}")
        del pair
    if rank == 0:
        print(f"{world} ranks, ms per step, rows pushed beside the next forward: " + " | ".join(out), flush=True)
    torch.cuda.synchronize(); dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
