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
slower one still reads this step's sums.

Split-phase form (`arrive` / `wait_previous`; `GraphedViewStep(..., exchange=xch, exchange_buffer=b)`
places them): the full barrier at the end of a step makes every GPU wait for the step's slowest view
and squeezes all incoming rows through the link at once.  Instead, a step ends with `arrive()` ("my rows
have been sent", one release-ordered multimem.red on a per-rank counter) and the NEXT step calls
`wait_previous()` after its forward pass, right before its own backward pushes: the sums of step i are
read during step i + 1 -- like an all-reduce overlapped with the next step -- and a GPU only ever waits
when it is more than a forward pass ahead of the slowest one.

    steps = [GraphedViewStep(..., exchange=xch, exchange_buffer=b, consume=apply_sums) for b in (0, 1)]
    for i, view in enumerate(views):
        steps[i % 2].replay()   # forward | wait for step i-1 | consume(rows(1-b)), zero them | backward + push | arrive
    xch.wait_all()              # the last step's sums: xch.rows((len(views) - 1) % 2)

Why two buffers are enough: rank r zeroes buffer 1-b before its arrive of step i, and nobody pushes step
i + 1 (into 1-b) before having waited for everyone's arrive of step i.

Deferred form (`GraphedViewStep(..., exchange=xch, exchange_buffer=b, exchange_mode="next_step")`): the
rows of step i leave in a kernel of their own (`push_source`, sl_exchange_push) on a second stream at the
start of step i + 1, under that step's latency-bound binning kernels, followed by `arrive`; step i + 1
ends with `wait_current()`, after which rows(i % 2) hold the sums of step i.  The backward is the plain
one; nothing of the exchange is on the critical path except the final wait.  `flush_deferred(b)` sends
and waits for the last step's rows.  Built on torch.distributed._symmetric_memory (allocation,
rendezvous, multicast mapping, signal-pad barrier); raises if the box has no multicast support.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import _capi

ROW = 16      # floats per Gaussian (csrc/common.cuh: EXCHANGE_ROW)
_FLAG_WORDS = 1024   # per-rank arrival counters behind the two buffers (one uint32 each; world <= 1024)


class NvlsGradientExchange:
    def __init__(self, P: int, device, group=None, timeout_ms: int = 20000):
        import torch.distributed._symmetric_memory as symm
        self.P, self.device, self.timeout_ms = int(P), torch.device(device), int(timeout_ms)
        group = group if group is not None else dist.group.WORLD
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if self.world > _FLAG_WORDS:
            raise ValueError(f"at most {_FLAG_WORDS} ranks")
        self._buf = symm.empty(2 * self.P * ROW + _FLAG_WORDS, dtype=torch.float32, device=self.device)
        self._hdl = symm.rendezvous(self._buf, group.group_name)
        if not self._hdl.multicast_ptr:
            raise RuntimeError("symmetric memory came up without a multicast mapping (no NVSwitch / NVLS on this box)")
        self._buf.zero_()
        self._flags = self._buf[2 * self.P * ROW:]                      # this GPU's copy of the arrival counters
        self._waits = torch.zeros(1, dtype=torch.int32, device=self.device)   # waits this GPU has done
        self._sources = torch.zeros(2, 8, dtype=torch.int64, device=self.device)   # two sl_exchange_source structs
        self._channels = 0
        self._status = torch.zeros(1, dtype=torch.int32, device=self.device)  # 1: a wait timed out
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

    def reset(self) -> None:
        """Collective: all ranks idle, both buffers zero (after warm-up or capture runs left partial sums)."""
        torch.cuda.synchronize(self.device)
        self._hdl.barrier(channel=0, timeout_ms=self.timeout_ms)
        self.zero(0)
        self.zero(1)
        torch.cuda.synchronize(self.device)
        self._hdl.barrier(channel=0, timeout_ms=self.timeout_ms)

    # ---- split-phase barrier (csrc/exchange.cu) -------------------------------------------------
    def arrive(self) -> None:
        """On the current stream, after a step's push: tell every GPU this rank's rows of the step are out."""
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _capi.check(_capi.load().sl_exchange_arrive(int(self._hdl.multicast_ptr) + 2 * self.P * ROW * 4, self.rank, stream))

    def wait_previous(self) -> None:
        """On the current stream: wait until every rank has arrived at all steps before this one (the n-th
        call waits for n - 1 arrivals per rank; the first returns at once)."""
        self._wait(0, 1)

    def wait_current(self) -> None:
        """On the current stream: wait until every rank has arrived at this step too (the n-th call waits for
        n arrivals per rank): the deferred form, whose arrive comes first in a step."""
        self._wait(1, 1)

    def wait_all(self) -> None:
        """On the current stream: wait until every rank has arrived as often as this GPU has waited."""
        self._wait(0, 0)

    def _wait(self, ahead: int, advance: int) -> None:
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _capi.check(_capi.load().sl_exchange_wait(self._flags.data_ptr(), self.world, self._waits.data_ptr(), ahead, advance,
                                                  self.timeout_ms, self._status.data_ptr(), stream))

    # ---- deferred push (sl_exchange_push) --------------------------------------------------------
    def bind_source(self, b: int, radii: torch.Tensor, dL_dmean3D: torch.Tensor, dL_dopacity: torch.Tensor,
                    dL_dscale: torch.Tensor, dL_drot: torch.Tensor, dL_dcolor: torch.Tensor) -> None:
        """The gradient arrays whose rows `push_source(b)` sends into buffer b: radii (V, P) int32 and the
        outputs of the batch backward (colours precomputed).  Launches captured earlier see the binding."""
        P = self.P
        if radii.dtype != torch.int32 or radii.numel() % P or dL_dcolor.numel() % P or dL_dcolor.numel() // P not in (3, 4):
            raise ValueError("bind_source: radii must be (V, P) int32 and dL_dcolor (P, 3 or 4)")
        for t in (radii, dL_dmean3D, dL_dopacity, dL_dscale, dL_drot, dL_dcolor):
            if not t.is_contiguous() or t.device != self.device:
                raise ValueError("bind_source: contiguous tensors on the exchange's device")
        self._channels = dL_dcolor.numel() // P
        words = [radii.data_ptr(), dL_dmean3D.data_ptr(), dL_dopacity.data_ptr(), dL_dscale.data_ptr(), dL_drot.data_ptr(),
                 dL_dcolor.data_ptr(), (radii.numel() // P) | ((dL_dscale.numel() // P) << 32), 0]
        self._sources[b].copy_(torch.tensor(words, dtype=torch.int64))
        torch.cuda.synchronize(self.device)

    def push_source(self, b: int, max_ctas: int = 0) -> None:
        """On the current stream: add the rows of source b (nothing if unbound) into buffer b on all GPUs."""
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _capi.check(_capi.load().sl_exchange_push(self._sources[b].data_ptr(), self.P, self._channels or 4, self.multicast_ptr(b),
                                                  max_ctas, stream))

    def flush_deferred(self, b: int) -> None:
        """After the last step of a deferred sequence (it used buffer b): send its rows, wait for everyone's."""
        self.push_source(b)
        self.arrive()
        self._wait(1, 1)

    @property
    def timed_out(self) -> bool:
        """True if a wait gave up after `timeout_ms` (synchronises the stream's device)."""
        return bool(int(self._status))

    @staticmethod
    def unpack(rows: torch.Tensor) -> dict:
        """Views of the exchanged rows under the gradient names."""
        return dict(means3D=rows[:, 0:3], opacities=rows[:, 3:4], scales=rows[:, 4:7], seen=rows[:, 7], rotations=rows[:, 8:12],
                    colors_precomp=rows[:, 12:16])
