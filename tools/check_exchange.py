"""GPU box, under torchrun (N >= 2): the gradient exchange fused into the backward (splatloc_b200.exchange,
sl_backward_batch_push) against an NCCL all-reduce of the same gradients, then its cost per step.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29503 tools/check_exchange.py
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
    quick = os.environ.get("CHECK_QUICK")      # timing of the deferred form only
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("NCCL_DEBUG", "WARN")
    dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=120))
    P, C = int(os.environ.get("CHECK_P", 500_000)), 4
    sc = {k: v.to(dev) for k, v in scenes.make_scene(P, seed=0, channels=C).items()}
    params = {k: sc[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "colors_precomp", "scales", "rotations")}
    bg = torch.zeros(C, device=dev)
    cams = [scenes.make_camera(seed=100 * rank + i).to(dev) for i in range(4)]
    H, W = cams[0].H, cams[0].W
    ups = [torch.cat(scenes.make_pixel_grads(100 * rank + i, C, H, W, normalised=False), 0).to(dev) for i in range(4)]
    xch = NvlsGradientExchange(P, dev)
    steps = [GraphedViewStep(params, H, W, cams[0].tanfovx, cams[0].tanfovy, bg, max_pairs=2_500_000,
                             exchange_multicast=xch.multicast_ptr(b)) for b in (0, 1)]
    plain = GraphedViewStep(params, H, W, cams[0].tanfovx, cams[0].tanfovy, bg, max_pairs=2_500_000)
    xch.zero(0); xch.zero(1)
    torch.cuda.synchronize(); dist.barrier()
    worst = 0.0
    for i in range(4):
        b = i % 2
        for st in (steps[b], plain):
            st.set_camera(cams[i].viewmatrix, cams[i].projmatrix, cams[i].campos)
            st.upstream.copy_(ups[i])
        xch.zero(1 - b)
        steps[b].replay()
        xch.barrier()
        plain.replay()
        ref = plain.trainable_grads().clone()
        seen = (plain.radii > 0).float()
        dist.all_reduce(ref)
        dist.all_reduce(seen)
        got = NvlsGradientExchange.unpack(xch.rows(b))
        off = 0
        for name, w in (("means3D", 3), ("opacities", 1), ("scales", 3), ("rotations", 4), ("colors_precomp", C)):
            n = (P * w + 3) // 4 * 4
            want = ref[off:off + P * w].view(P, w)
            off += n
            err = float((got[name] - want).abs().max()); scale = float(want.abs().max())
            worst = max(worst, err / max(scale, 1e-30))
            assert err <= 2e-5 * 16 * scale, (name, err, scale)
        assert torch.equal(got["seen"], seen), "visibility counts"
        assert int(steps[b].overflow) == 0
    if rank == 0:
        print(f"{world} ranks: fused exchange == NCCL all-reduce of the same gradients (worst relative difference {worst:.2e}, "
              f"visibility counts exact)", flush=True)

    # split-phase form: forward | wait for the previous step | consume + zero | backward + push | arrive, all inside
    # the replay; the sums of step i are read during step i + 1.  Ranks are skewed on purpose.
    snap = torch.zeros(P, 16, device=dev)
    split = [GraphedViewStep(params, H, W, cams[0].tanfovx, cams[0].tanfovy, bg, max_pairs=2_500_000, exchange=xch,
                             exchange_buffer=b, consume=lambda rows: snap.copy_(rows)) for b in (0, 1)]
    xch.reset()
    n_steps, refs, seens, snaps = 7, [], [], []
    for i in range(n_steps):
        plain.set_camera(cams[i % 4].viewmatrix, cams[i % 4].projmatrix, cams[i % 4].campos)
        plain.upstream.copy_(ups[i % 4])
        plain.replay()
        ref, seen = plain.trainable_grads().clone(), (plain.radii > 0).float()
        dist.all_reduce(ref); dist.all_reduce(seen)
        refs.append(ref); seens.append(seen)
    torch.cuda.synchronize(); dist.barrier()
    for i in range(n_steps):
        st = split[i % 2]
        st.set_camera(cams[i % 4].viewmatrix, cams[i % 4].projmatrix, cams[i % 4].campos)
        st.upstream.copy_(ups[i % 4])
        torch.cuda._sleep(int(3e5) * ((rank + i) % 3))           # up to ~0.3 ms of skew between the ranks
        st.replay()
        snaps.append(snap.clone())                                # sums of step i - 1 (zeros at i = 0)
    xch.wait_all()
    snaps.append(xch.rows((n_steps - 1) % 2).clone())
    torch.cuda.synchronize()
    assert not xch.timed_out, "a split-phase wait timed out"
    assert float(snaps[0].abs().max()) == 0.0
    worst2 = 0.0
    for i in range(n_steps):
        got = NvlsGradientExchange.unpack(snaps[i + 1])
        off = 0
        for name, w in (("means3D", 3), ("opacities", 1), ("scales", 3), ("rotations", 4), ("colors_precomp", C)):
            n = (P * w + 3) // 4 * 4
            want = refs[i][off:off + P * w].view(P, w)
            off += n
            err = float((got[name] - want).abs().max()); scale = float(want.abs().max())
            worst2 = max(worst2, err / max(scale, 1e-30))
            assert err <= 2e-5 * 16 * scale, ("split-phase", i, name, err, scale)
        assert torch.equal(got["seen"], seens[i]), ("split-phase visibility counts", i)
    if rank == 0:
        print(f"{world} ranks: split-phase exchange == NCCL all-reduce over {n_steps} skewed steps (worst relative difference "
              f"{worst2:.2e}, visibility counts exact)", flush=True)
    # deferred form: the rows of step i leave in their own kernel beside the forward of step i + 1
    snap2 = torch.zeros(P, 16, device=dev)
    later = [GraphedViewStep(params, H, W, cams[0].tanfovx, cams[0].tanfovy, bg, max_pairs=2_500_000, exchange=xch,
                             exchange_buffer=b, exchange_mode="next_step", consume=lambda rows: snap2.copy_(rows)) for b in (0, 1)]
    xch.reset()
    snaps = []
    for i in range(n_steps):
        st = later[i % 2]
        st.set_camera(cams[i % 4].viewmatrix, cams[i % 4].projmatrix, cams[i % 4].campos)
        st.upstream.copy_(ups[i % 4])
        torch.cuda._sleep(int(3e5) * ((rank + 2 * i) % 3))
        st.replay()
        snaps.append(snap2.clone())                               # sums of step i - 1 (i = 0: whatever the arrays held)
    xch.flush_deferred((n_steps - 1) % 2)
    snaps.append(xch.rows((n_steps - 1) % 2).clone())
    torch.cuda.synchronize()
    assert not xch.timed_out, "a deferred-push wait timed out"
    worst3 = 0.0
    for i in range(n_steps):
        got = NvlsGradientExchange.unpack(snaps[i + 1])
        off = 0
        for name, w in (("means3D", 3), ("opacities", 1), ("scales", 3), ("rotations", 4), ("colors_precomp", C)):
            n = (P * w + 3) // 4 * 4
            want = refs[i][off:off + P * w].view(P, w)
            off += n
            err = float((got[name] - want).abs().max()); scale = float(want.abs().max())
            worst3 = max(worst3, err / max(scale, 1e-30))
            assert err <= 2e-5 * 16 * scale, ("deferred push", i, name, err, scale)
        assert torch.equal(got["seen"], seens[i]), ("deferred push visibility counts", i)
    if rank == 0:
        print(f"{world} ranks: deferred-push exchange == NCCL all-reduce over {n_steps} skewed steps (worst relative difference "
              f"{worst3:.2e}, visibility counts exact)", flush=True)
    # two views per GPU per step (batch entry points): the rows' visibility slot counts VIEWS
    def load2(st, i):
        for v in range(2):
            c = cams[(2 * i + v) % 4]
            st.set_camera(c.viewmatrix, c.projmatrix, c.campos, view=v)
        u = [ups[(2 * i + v) % 4] for v in range(2)]
        st.upstream_color.copy_(torch.stack([x[:C] for x in u]))
        st.upstream_depth.copy_(torch.stack([x[C:C + 1] for x in u]))
        st.upstream_alpha.copy_(torch.stack([x[C + 1:] for x in u]))
    plain2 = GraphedViewStep(params, H, W, cams[0].tanfovx, cams[0].tanfovy, bg, max_pairs=5_000_000, n_views=2)
    later2 = [GraphedViewStep(params, H, W, cams[0].tanfovx, cams[0].tanfovy, bg, max_pairs=5_000_000, n_views=2, exchange=xch,
                              exchange_buffer=b, exchange_mode="next_step", consume=lambda rows: snap2.copy_(rows)) for b in (0, 1)]
    fused2 = [GraphedViewStep(params, H, W, cams[0].tanfovx, cams[0].tanfovy, bg, max_pairs=5_000_000, n_views=2, exchange=xch,
                              exchange_buffer=b, consume=lambda rows: snap.copy_(rows)) for b in (0, 1)]
    refs2, seens2 = [], []
    for i in range(3):
        load2(plain2, i)
        plain2.replay()
        ref, seen = plain2.trainable_grads().clone(), (plain2.radii > 0).float().sum(0)
        dist.all_reduce(ref); dist.all_reduce(seen)
        refs2.append(ref); seens2.append(seen)
    for name, pair, sn, flush in (("push in the backward", fused2, snap, lambda: xch.wait_all()),
                                  ("push beside the next forward", later2, snap2, lambda: xch.flush_deferred(0))):
        for st in pair:
            if st.exchange_mode == "next_step":
                st.bind_exchange_source()
        xch.reset()
        got2 = []
        for i in range(3):
            load2(pair[i % 2], i)
            pair[i % 2].replay()
            got2.append(sn.clone())
        flush()
        got2.append(xch.rows(0).clone())            # the last step (i = 2) used buffer 0
        torch.cuda.synchronize()
        assert not xch.timed_out
        for i in range(3):
            got = NvlsGradientExchange.unpack(got2[i + 1])
            off = 0
            for nm, w in (("means3D", 3), ("opacities", 1), ("scales", 3), ("rotations", 4), ("colors_precomp", C)):
                n = (P * w + 3) // 4 * 4
                want = refs2[i][off:off + P * w].view(P, w)
                off += n
                err = float((got[nm] - want).abs().max()); scale = float(want.abs().max())
                assert err <= 2e-5 * 16 * scale, ("two views per step", name, i, nm, err, scale)
            assert torch.equal(got["seen"], seens2[i]), ("two views per step: view counts", name, i)
    if rank == 0:
        print(f"{world} ranks: two views per GPU per step == NCCL all-reduce, view counts exact (both push forms)", flush=True)
    del plain2, later2, fused2
    torch.cuda.empty_cache()
    later_plain = [GraphedViewStep(params, H, W, cams[0].tanfovx, cams[0].tanfovy, bg, max_pairs=2_500_000, exchange=xch,
                                   exchange_buffer=b, exchange_mode="next_step") for b in (0, 1)]
    xch.reset()
    split_plain = [GraphedViewStep(params, H, W, cams[0].tanfovx, cams[0].tanfovy, bg, max_pairs=2_500_000, exchange=xch,
                                   exchange_buffer=b) for b in (0, 1)]
    xch.reset()

    # cost: one view per step, L2 flushed; (a) no exchange, (b) fused push + barrier, (c) NCCL all-reduce after the step
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    def timed(fn, n=30):
        for i in range(5): fn(i)
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
    def local(i):
        load(plain, i); plain.replay()
    def fused(i):
        b = i % 2; load(steps[b], i); xch.zero(1 - b); steps[b].replay(); xch.barrier()
    def nccl(i):
        load(plain, i); plain.replay(); dist.all_reduce(plain.trainable_grads())
    def fused_split(i):
        b = i % 2; load(split_plain[b], i); split_plain[b].replay()
    def fused_later(i):
        b = i % 2; load(later_plain[b], i); later_plain[b].replay()
    def push_only(i):
        b = i % 2; load(steps[b], i); steps[b].replay()
    def barrier_only(i):
        xch.barrier()
    def zero_only(i):
        xch.zero(i % 2)
    t0, t1, t2 = timed(local), timed(fused), timed(nccl)
    t6 = timed(fused_split)
    xch.wait_all(); torch.cuda.synchronize(); assert not xch.timed_out
    xch.reset()
    t7 = timed(fused_later)
    xch.flush_deferred(1); torch.cuda.synchronize()   # the last timed step (i = 29) used buffer 1; assert not xch.timed_out
    xch.reset()
    t3, t4, t5 = timed(push_only), timed(barrier_only), timed(zero_only)
    if rank == 0:
        print(f"   step with the pushes but no barrier / zeroing {t3:.4f} | barrier alone {t4:.4f} | zeroing one buffer {t5:.4f}", flush=True)
    if rank == 0:
        print(f"{world} ranks, ms per step: no exchange {t0:.4f} | fused push + barrier {t1:.4f} | fused push, split-phase barrier "
              f"{t6:.4f} | rows pushed beside the next forward {t7:.4f} | NCCL all-reduce after the step {t2:.4f}", flush=True)
    torch.cuda.synchronize(); dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
