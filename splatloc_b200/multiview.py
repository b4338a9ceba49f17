"""Multi-GPU host logic for the rasterizer path (one process per GPU, torch.distributed).

The path shards over *views* (training) and *query frames* (pose refinement); the Gaussian map is
replicated on every GPU.  The reference has no distributed code at all, so there
is no counterpart to cite; the step mirrors what SplatLoc's mapping loop does per view
(`train_gaussians.py:195-245`): render, backward, accumulate densification statistics from
`viewspace_points.grad` / `absgrad` (`gaussian_model.py:677-679`), and -- new here -- sum the
per-Gaussian gradients of all views with ONE all-reduce of a packed buffer.

Works with NCCL (GPU) and gloo (CPU tensors; used by the world_size-2 tests).
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.distributed as dist


def shard_range(n_items: int, rank: int, world: int) -> range:
    """Contiguous, balanced split: the first n_items % world ranks get one extra item."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(n_items, world)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


@dataclass
class PackedGrads:
    """One flat float32 buffer holding, per Gaussian, the gradients a training step consumes plus
    the densification statistics, so that a multi-view step needs a single collective.

    Layout (P rows): means3D 3 | opacity 1 | scales 3 | rotations 4 | colours C | grad2D_norm 1 | seen 1
    `grad2D_norm` is the sum over views of ||dL/dmean2D[:, :2]|| and `seen` the number of views in
    which the Gaussian was visible (radii > 0): both are per-view quantities and are therefore
    accumulated locally BEFORE the reduce."""
    P: int
    C: int
    flat: torch.Tensor

    FIELDS = (("means3D", 3), ("opacity", 1), ("scales", 3), ("rotations", 4))

    @classmethod
    def zeros(cls, P: int, C: int, device) -> "PackedGrads":
        return cls(P, C, torch.zeros(P * cls.width(C), dtype=torch.float32, device=device))

    @staticmethod
    def width(C: int) -> int:
        return 3 + 1 + 3 + 4 + C + 1 + 1

    def _offsets(self):
        off, out = 0, {}
        for name, w in (*self.FIELDS, ("colors", self.C), ("grad2D_norm", 1), ("seen", 1)):
            out[name] = (off, w)
            off += w * self.P
        return out

    def view(self, name: str) -> torch.Tensor:
        off, w = self._offsets()[name]
        return self.flat[off:off + w * self.P].view(self.P, w)

    def accumulate(self, g_means3D, g_opacity, g_scales, g_rot, g_colors, g_means2D, radii) -> None:
        """Add one view's gradients and statistics."""
        self.view("means3D").add_(g_means3D)
        self.view("opacity").add_(g_opacity.reshape(self.P, 1))
        self.view("scales").add_(g_scales)
        self.view("rotations").add_(g_rot)
        self.view("colors").add_(g_colors)
        vis = (radii > 0).reshape(self.P, 1).to(torch.float32)
        self.view("grad2D_norm").add_(torch.norm(g_means2D[:, :2], dim=-1, keepdim=True) * vis)
        self.view("seen").add_(vis)

    def all_reduce(self, group=None) -> None:
        """Sum over ranks (NCCL over NVLink on the GPU box, gloo in the CPU tests)."""
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)


def multiview_step(render_backward, n_views: int, P: int, C: int, device, rank: int | None = None,
                   world: int | None = None, group=None) -> PackedGrads:
    """Run `render_backward(view_index)` for this rank's share of the views and return the packed,
    all-reduced gradients.  `render_backward` returns
    `(g_means3D, g_opacity, g_scales, g_rot, g_colors, g_means2D, radii)` for one view, e.g. from
    `splatloc_b200._C.rasterize_gaussians` + `rasterize_gaussians_backward`."""
    if rank is None:
        rank = dist.get_rank(group) if dist.is_initialized() else 0
    if world is None:
        world = dist.get_world_size(group) if dist.is_initialized() else 1
    packed = PackedGrads.zeros(P, C, device)
    for v in shard_range(n_views, rank, world):
        packed.accumulate(*render_backward(v))
    packed.all_reduce(group)
    return packed


def gather_frames(local: torch.Tensor, n_frames: int, group=None) -> torch.Tensor:
    """Pose refinement: every rank refines `shard_range(n_frames, ...)` frames independently (no
    collective per iteration); at the end the per-frame results (rows of `local`) are gathered in
    frame order on every rank."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    sizes = [len(shard_range(n_frames, r, world)) for r in range(world)]
    pad = max(sizes)
    buf = torch.zeros((pad,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    buf[: local.shape[0]] = local
    out = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(out, buf, group=group)
    return torch.cat([o[:s] for o, s in zip(out, sizes)], dim=0)
