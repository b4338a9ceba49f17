"""CPU, world_size 2, gloo: the host logic of the multi-GPU path (view sharding, packed gradient
all-reduce, frame gather) gives the single-process result."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from splatloc_b200.multiview import PackedGrads, gather_frames, multiview_step, shard_range

P, C, N_VIEWS, N_FRAMES = 257, 4, 7, 9


def fake_render_backward(v):
    """Deterministic stand-in for one view's rasterizer fwd+bwd (the kernels need a GPU)."""
    g = torch.Generator().manual_seed(1234 + v)
    r = lambda *s: torch.randn(*s, generator=g)  # noqa: E731
    radii = (torch.rand(P, generator=g) < 0.6).to(torch.int32) * 5
    return r(P, 3), r(P, 1), r(P, 3), r(P, 4), r(P, C), r(P, 3), radii


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    packed = multiview_step(fake_render_backward, N_VIEWS, P, C, "cpu")
    local = torch.stack([torch.full((6,), float(f)) for f in shard_range(N_FRAMES, rank, world)])
    frames = gather_frames(local, N_FRAMES)
    torch.save((packed.flat, frames), os.path.join(out_dir, f"r{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_range_partitions():
    for n in (0, 1, 7, 32, 256):
        for world in (1, 2, 3, 8):
            got = [i for r in range(world) for i in shard_range(n, r, world)]
            assert got == list(range(n))
            sizes = [len(shard_range(n, r, world)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1


def test_world2_matches_single_process(tmp_path):
    single = multiview_step(fake_render_backward, N_VIEWS, P, C, "cpu", rank=0, world=1)
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        flat, frames = torch.load(os.path.join(tmp_path, f"r{r}.pt"))
        assert torch.allclose(flat, single.flat, rtol=1e-6, atol=1e-6)
        assert torch.equal(frames[:, 0], torch.arange(N_FRAMES, dtype=torch.float32))
    # per-view statistics were accumulated before the reduce
    seen = single.view("seen")
    assert float(seen.max()) <= N_VIEWS and float(seen.sum()) > 0
    assert PackedGrads.width(C) == 17
