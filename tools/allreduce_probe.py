"""GPU box, under torchrun: time torch.distributed all-reduce (NCCL) of float32 buffers of the sizes the
multi-view step exchanges, on its own and right after a busy kernel, and print the bus bandwidth.
Round 1 measured the 30 MB reduce of the bench step at ~0.09 / 0.14 / 0.46 ms on 2 / 4 / 8 B200 -- the
8-rank figure is ~100 GB/s of bus bandwidth, far below NVSwitch; this is the tool to take that apart.

    NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,TUNING python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 \\
        --master-addr 127.0.0.1 --master-port 29500 tools/allreduce_probe.py [MB ...]
"""
import datetime
import os
import sys

import torch
import torch.distributed as dist


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=60))
    sizes_mb = [float(a) for a in sys.argv[1:]] or [1, 4, 8, 16, 30, 64]
    busy = torch.empty(64 << 20, dtype=torch.float32, device=dev)
    for mb in sizes_mb:
        n = int(mb * 1e6 / 4)
        buf = torch.ones(n, dtype=torch.float32, device=dev)
        for label, prelude in (("alone", None), ("after a 256 MB fill", lambda: busy.fill_(1.0))):
            for _ in range(5):
                dist.all_reduce(buf)
            torch.cuda.synchronize()
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            times = []
            for _ in range(30):
                if prelude is not None:
                    prelude()
                e0.record()
                dist.all_reduce(buf)
                e1.record()
                torch.cuda.synchronize()
                times.append(e0.elapsed_time(e1))
                buf.fill_(1.0)
            t = torch.tensor(sorted(times)[len(times) // 2], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if rank == 0:
                ms = float(t)
                busbw = 2 * (world - 1) / world * n * 4 / (ms * 1e-3) / 1e9
                print(f"{world} ranks  {mb:6.1f} MB  {label:22s} median {ms * 1e3:8.1f} us  bus bandwidth {busbw:7.1f} GB/s", flush=True)
    torch.cuda.synchronize()
    dist.barrier()
    sys.stdout.flush()
    os._exit(0)


if __name__ == "__main__":
    main()
