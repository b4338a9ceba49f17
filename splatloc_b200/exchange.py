"""Multi-GPU gradient exchange fused into the backward (SURVEY 8(e) "phase 2"; no reference counterpart:
the reference has no distributed code).

A multi-view training step on N GPUs needs, on every GPU, the sum over all GPUs' views of the
per-Gaussian gradients.  The plain way is an NCCL all-reduce of the packed gradient block after the
backward (30 MB at 500 k Gaussians; 95 us on 2 and 161 us on 8 B200, serialised with the step or
competing with the next step for SMs).  Here the last kernel of the backward pushes each finished row
-- only the rows this GPU's views saw, ~25 % of the block -- straight into EVERY GPU's exchange buffer
through the NVSwitch (`multimem.red` on a multicast address, NVLS): the transfer rides on the kernel
that produces the data, no extra kernel, no SMs on the receiving side.

    xch = NvlsGradientExchange(P, device)                       # after init_process_group("nccl")
    steps = [GraphedViewStep(..., exchange_multicast=xch.multicast_ptr(b)) for b in (0, 1)]
    for i, view in enumerate(views):
        b = i % 2
        xch.zero(1 - b)            # the other buffer: everyone has read it, nobody pushes to it before the barrier
        steps[b].replay()          # forward + backward; rows land in buffer b on all GPUs
        xch.barrier()              # device-side, on the current stream: all GPUs' pushes have landed
        grads = xch.rows(b)        # (P, 16): mean xyz, opacity | scale xyz, seen | rot wxyz | colour c0..c3

Two buffers alternate because a GPU that leaves the barrier early starts pushing the next step while a
slower one still reads this step's sums.  Built on torch.distributed._symmetric_memory (allocation,
rendezvous, multicast mapping, signal-pad barrier); raises if the box has no multicast support.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

ROW = 16      # floats per Gaussian (csrc/common.cuh: EXCHANGE_ROW)


class NvlsGradientExchange:
    def __init__(self, P: int, device, group=None, timeout_ms: int = 20000):
        import torch.distributed._symmetric_memory as symm
        self.P, self.device, self.timeout_ms = int(P), torch.device(device), int(timeout_ms)
        group = group if group is not None else dist.group.WORLD
        self._buf = symm.empty(2 * self.P * ROW, dtype=torch.float32, device=self.device)
        self._hdl = symm.rendezvous(self._buf, group.group_name)
        if not self._hdl.multicast_ptr:
            raise RuntimeError("symmetric memory came up without a multicast mapping (no NVSwitch / NVLS on this box)")
        self._buf.zero_()
        torch.cuda.synchronize(self.device)
        self._hdl.barrier(channel=0, timeout_ms=self.timeout_ms)

    def multicast_ptr(self, b: int) -> int:
        """Multicast device address of buffer `b` (0 or 1): what sl_backward_batch_push takes."""
        return int(self._hdl.multicast_ptr) + b * self.P * ROW * 4

    def rows(self, b: int) -> torch.Tensor:
        """This GPU's copy of buffer `b`, (P, 16); complete after barrier()."""
        return self._buf[b * self.P * ROW:(b + 1) * self.P * ROW].view(self.P, ROW)

    def zero(self, b: int) -> None:
        self.rows(b).zero_()

    def barrier(self) -> None:
        """Cross-GPU barrier on the current stream (no host synchronisation)."""
        self._hdl.barrier(channel=0, timeout_ms=self.timeout_ms)

    @staticmethod
    def unpack(rows: torch.Tensor) -> dict:
        """Views of the exchanged rows under the gradient names."""
        return dict(means3D=rows[:, 0:3], opacities=rows[:, 3:4], scales=rows[:, 4:7], seen=rows[:, 7], rotations=rows[:, 8:12],
                    colors_precomp=rows[:, 12:16])
