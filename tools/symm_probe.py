"""GPU box, under torchrun: does torch's symmetric memory (peer pointers over NVLink, NVSwitch
multicast) come up on this box, and what do its all-reduce kernels reach on the 30 MB exchange of the
bench step?  Prints peer / multicast availability and median times next to NCCL's.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29501 tools/symm_probe.py
"""
import datetime
import os

import torch
import torch.distributed as dist


def timed(fn, reset, dev, n=30):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(n):
        reset()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    t = torch.tensor(sorted(ts)[len(ts) // 2], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("NCCL_DEBUG", "WARN")
    dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=120))
    say = (lambda *a: print(*a, flush=True)) if rank == 0 else (lambda *a: None)
    n = 7_500_000                       # 30 MB of float32: 500 k Gaussians x 15 trainable floats
    buf = torch.ones(n, device=dev)
    ms = timed(lambda: dist.all_reduce(buf), lambda: buf.fill_(1.0), dev)
    say(f"{world} ranks nccl all_reduce 30 MB: {ms * 1e3:.1f} us, bus bw {2 * (world - 1) / world * n * 4 / ms / 1e6:.1f} GB/s")
    try:
        import torch.distributed._symmetric_memory as symm
        t = symm.empty(n, dtype=torch.float32, device=dev)
        hdl = symm.rendezvous(t, dist.group.WORLD.group_name)
        say("symmetric memory: rendezvous ok; peers", len(hdl.buffer_ptrs), "multicast_ptr", hex(hdl.multicast_ptr or 0),
            "signal pads", len(hdl.signal_pad_ptrs))
        for name in ("one_shot_all_reduce", "two_shot_all_reduce_", "multimem_all_reduce_", "multimem_one_shot_all_reduce"):
            op = getattr(torch.ops.symm_mem, name, None)
            if op is None:
                say(f"  {name}: not in this torch")
                continue
            try:
                ms = timed(lambda: op(t, "sum", dist.group.WORLD.group_name), lambda: t.fill_(1.0), dev)
                say(f"  symm_mem.{name} 30 MB: {ms * 1e3:.1f} us")
            except Exception as e:      # noqa: BLE001
                say(f"  {name}: failed: {type(e).__name__}: {str(e)[:200]}")
    except Exception as e:              # noqa: BLE001
        say(f"symmetric memory unavailable: {type(e).__name__}: {str(e)[:300]}")
    torch.cuda.synchronize()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
