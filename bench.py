#!/usr/bin/env python
"""bench.py -- headline benchmark of the rasterizer hot path (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Metric: fwd+bwd Mpix/s at 640x480.  One "step" = one forward + backward pass of
one 640x480 view per GPU over the config-2 scene (500 k Gaussians, C=4
precomputed colours [rgb, kp_score], depth and alpha outputs and gradients):
SplatLoc's `train_gaussians.py` mapping step as seen by the rasterizer.  With
N > 1 every rank renders its own views of the replicated map and the packed
per-Gaussian gradient block is all-reduced over NCCL inside the step (weak
scaling).  `value` is device-resident (inputs in HBM when the timer starts);
`e2e` goes through the public API with the per-view inputs (camera, upstream
pixel gradients) copied from pinned host memory and the rendered images copied
back inside the timed region.

The same JSON line carries, as bounded sub-records, the rest of BASELINE.json's
metric: `sweep` (fwd+bwd Mpix/s against the number of Gaussians, N = 1 only),
`pose_refine` (config 3: pose-refine iterations/s, query frames sharded over
the ranks), `multiview` (config 2 with 4 and 8 cameras per launch) and
`train_window` (config 4: 1296x968, 3 M Gaussians, 4 views per GPU between
exchanges) and `mapping_iteration` (render() pre-processing and the optimiser
step fused into kernels).  `--sub none` skips them.

--impl reference runs the UNMODIFIED reference CUDA rasterizer
(oracle/_ref/libref_raster_c4.so, compiled from /root/reference for sm_100a)
on the same scene, cameras and protocol; the reference has no CPU
implementation of this path, so when that library or a GPU is missing the arm
falls back to the float64 CPU port under oracle/ on a bounded sample.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from splatloc_b200 import scenes  # noqa: E402

WORKLOAD = "replica_train_step_500k_640x480_c4"
P_DEFAULT, C_DEFAULT = 500_000, 4
N_CAMERAS = 8          # views cycled per rank, so consecutive steps see different inputs
L2_FLUSH_BYTES = 256 << 20
EMPTY = torch.Tensor([])


def env_int(name, default):
    return int(os.environ.get(name, default))


class ClockSampler:
    """Samples SM clocks / throttle reasons of the job's GPUs with ONE nvidia-smi process (rank 0)
    while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_indices):
        self.gpus = ",".join(str(i) for i in gpu_indices)
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", self.gpus], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def survey_step_bytes(P, V, R, N, Tn, C, M_sh=0):
    """SURVEY.md 8(d), normative: compulsory HBM bytes of one fwd+bwd step, every stage of the
    REFERENCE pipeline shape (K1..K9) reading its inputs and writing its outputs exactly once."""
    fwd = (P * (44 + 12 * M_sh) + 8 * P + 52 * V          # K1
           + 8 * P                                        # K2
           + (16 * V + 8 * P) + 12 * R                    # K3
           + 24 * R                                       # K4
           + 8 * R + 8 * Tn                               # K5
           + (32 + 4 * C) * R + (12 + 4 * C) * N)         # K6
    bwd = ((32 + 4 * C) * R + (16 + 4 * C) * N + (36 + 4 * C) * R                    # K7
           + 4 * P + 164 * V + ((12 * M_sh + 15) * V if M_sh else 0)                 # K8 + K9 inputs
           + (92 + 4 * C + 12 * M_sh) * P)                                           # dense gradient outputs
    return fwd + bwd


def stage_bytes(P, V, R, N, Tn, C, M_sh=0):
    """Compulsory HBM bytes of each stage of THIS library's default pipeline (tile-bucket binning,
    `sl_set_binning_mode` 2): inputs read once, outputs written once.  Keys are the library's stage
    timer names.  The record gather (28 + 4C in, 48 + 1 out per pair; no reference counterpart) rides
    on the segment sort, so the sum of these is a little more than survey_step_bytes()."""
    return {
        "preprocess": P * (44 + 12 * M_sh) + 8 * P + 52 * V + 4 * R,        # + one counter increment per pair
        "tile_scan": 8 * Tn + 8 * Tn,
        "emit_pairs": 16 * V + 8 * P + 12 * R,                              # scatter into tile segments
        "sort_pairs": 12 * R + 12 * R + (28 + 4 * C) * R + (48 + 1) * R,    # segment in, sorted keys/ids out, gather, records out
        "pair_records": 0,                                                  # radix fallback only
        "blend_forward": (48 + 1) * R + (12 + 4 * C) * N,
        "blend_backward": (48 + 1) * R + (16 + 4 * C) * N + (36 + 4 * C) * R,
        "preprocess_backward": 4 * P + 164 * V + (92 + 4 * C + 12 * M_sh) * P,
    }


def measured_traffic(kernel):
    """DRAM bytes per launch of `kernel` from the committed `ncu --set full` capture (profiles/)."""
    for name in ("r2_traffic.json", "r1_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                v = json.load(f).get(kernel)
            if v is not None:
                return v
        except Exception:
            pass
    return None


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


# ------------------------------------------------------------------------------------ back ends
def load_backend(impl):
    if impl == "ours":
        from splatloc_b200 import _C
        from splatloc_b200.diff_gauss import GaussianRasterizationSettings, GaussianRasterizer
        return _C, GaussianRasterizationSettings, GaussianRasterizer
    from oracle import ref_cuda, ref_module
    return ref_cuda, ref_module.GaussianRasterizationSettings, ref_module.GaussianRasterizer


class Job:
    """Rank / device / process-group plumbing shared by the workloads."""

    def __init__(self, impl):
        self.impl = impl
        self.rank, self.world = env_int("RANK", 0), env_int("WORLD_SIZE", 1)
        self.local = env_int("LOCAL_RANK", 0)
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.dist = None
        if impl == "reference":
            self.world = 1      # rank 0 alone times the reference arm; no process group is formed
        if self.world > 1:
            import datetime
            import torch.distributed as dist
            os.environ.setdefault("NCCL_DEBUG", "WARN")
            os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")      # NCCL's version banner and warnings: off stdout (one JSON line)
            dist.init_process_group("nccl", device_id=self.dev, timeout=datetime.timedelta(seconds=300))
            self.dist = dist
        self.flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=self.dev)

    def barrier(self):
        torch.cuda.synchronize()
        if self.dist is not None:
            self.dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, x: float) -> float:
        if self.dist is None:
            return x
        t = torch.tensor([x], device=self.dev, dtype=torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def timed(self, fn, steps, warmup, after=None):
        """W untimed steps, then EXACTLY `steps` timed ones: one CUDA-event pair per step on the current
        stream, the L2 flushed (256 MiB memset, outside the event pair) between steps, barrier +
        synchronize on both sides.  `after()` runs inside the bracket after the last step (e.g. wait
        for an exchange still in flight) and its time is charged to the last step.  Returns (sum of the
        per-step times in ms, max over ranks; this rank's per-step list; wall seconds)."""
        for i in range(warmup):
            fn(i)
            self.flush.zero_()
        if after is not None:
            after()
        self.barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        t0 = time.perf_counter()
        for i in range(steps):
            self.flush.zero_()
            ev[i][0].record()
            fn(warmup + i)
            if after is not None and i == steps - 1:
                after()
            ev[i][1].record()
        self.barrier()
        wall = time.perf_counter() - t0
        per_step = [a.elapsed_time(b) for a, b in ev]
        return self.max_over_ranks(sum(per_step)), per_step, wall

    def finish(self):
        """Explicit teardown, then a normal interpreter exit (the driver's exit hook records which
        shared libraries the process loaded)."""
        torch.cuda.synchronize()
        if self.dist is not None:
            self.dist.barrier()
            torch.cuda.synchronize()
            self.dist.destroy_process_group()
            self.dist = None


def spin_up(step, seconds_min=1.0, seconds_max=10.0):
    """Untimed: a fresh process on a fresh box needs a moment until step times settle (idle GPU
    clocks, first-touch paging).  Runs batches of `step` -- which must contain NO collective, the
    loop's length is decided per rank -- until two consecutive batches agree within 3 %."""
    if os.environ.get("BENCH_NO_SPINUP"):     # (profilers set this: they only want a few launches)
        return
    t_spin, prev = time.perf_counter(), None
    while True:
        t0 = time.perf_counter()
        for i in range(20):
            step(i)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        el = time.perf_counter() - t_spin
        if el > seconds_max or (el > seconds_min and prev is not None and abs(dt - prev) <= 0.03 * prev):
            return
        prev = dt


class Exchange:
    """The multi-view training step's only collective: the sum over the ranks of the gradients an
    optimiser consumes (means3D, opacity, scales, rotations, colours: the leading span of the packed
    gradient block).  `overlap`: the span is copied to one of two staging buffers (a device copy at
    HBM speed) and all-reduced asynchronously, so the exchange of step i runs under the forward and
    blend of step i + 1 -- what a mapping iteration that renders a window of views per optimiser step
    (train_gaussians.py:195-229) allows; wait() closes the last one.  Otherwise the all-reduce is
    issued in place and the next step waits for it."""

    def __init__(self, job, numel, overlap):
        self.dist, self.overlap = job.dist, overlap and job.dist is not None
        self.work = [None, None]
        self.stage = [torch.empty(numel, device=job.dev) for _ in range(2)] if self.overlap else None
        self.i = 0

    def __call__(self, flat):
        if self.dist is None:
            return
        if not self.overlap:
            self.dist.all_reduce(flat)
            return
        b = self.i % 2
        self.i += 1
        if self.work[b] is not None:
            self.work[b].wait()              # staging buffer b is free again (reduced two steps ago)
        self.stage[b].copy_(flat)
        self.work[b] = self.dist.all_reduce(self.stage[b], async_op=True)

    def wait(self):
        for b in range(2):
            if self.work[b] is not None:
                self.work[b].wait()
                self.work[b] = None


def run_train_step(args, impl):
    job = Job(impl)
    if impl == "reference" and job.rank != 0:
        return None
    rank, world, dev = job.rank, job.world, job.dev
    be, Settings, Rasterizer = load_backend(impl)
    P, C = args.points, C_DEFAULT
    W, H = scenes.REPLICA_INTRINSICS["W"], scenes.REPLICA_INTRINSICS["H"]
    N, Tn = W * H, ((W + 15) // 16) * ((H + 15) // 16)
    sc = {k: v.to(dev) for k, v in scenes.make_scene(P, seed=0, channels=C).items()}
    bg = torch.zeros(C, device=dev)
    cams_host = [scenes.make_camera(seed=100 * rank + i) for i in range(N_CAMERAS)]
    cams = [c.to(dev) for c in cams_host]
    grads_host = [tuple(t.pin_memory() for t in scenes.make_pixel_grads(100 * rank + i, C, H, W)) for i in range(N_CAMERAS)]
    grads = [tuple(t.to(dev) for t in g) for g in grads_host]
    use_dist = job.dist is not None

    def step_eager(i, reduce=True):
        """Resident step through the operator layer: forward, backward, (exchange)."""
        cam, (gc, gd, ga) = cams[i % N_CAMERAS], grads[i % N_CAMERAS]
        R, color, depth, alpha, radii, geom, binning, img = be.rasterize_gaussians(
            bg, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"], sc["rotations"], 1.0, EMPTY,
            EMPTY.bool(), cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, H, W, EMPTY, 0, cam.campos,
            False, False)
        out = be.rasterize_gaussians_backward(
            bg, sc["means3D"], radii, sc["colors_precomp"], sc["scales"], sc["rotations"], 1.0, EMPTY,
            cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, gc, gd, ga, EMPTY, 0, cam.campos, geom, R,
            binning, img, alpha, False)
        if use_dist and reduce:
            flat = out[0]._base
            job.dist.all_reduce(flat[:flat.trainable_numel])
        return R, radii

    # ---- e2e, eager: public autograd API, per-view inputs from pinned host memory, images read back
    leaves = {k: sc[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "colors_precomp", "scales", "rotations")}
    means2D = torch.zeros_like(leaves["means3D"], requires_grad=True)
    cam_host = [(c.viewmatrix.pin_memory(), c.projmatrix.pin_memory(), c.campos.pin_memory()) for c in cams_host]
    cam_packed = [torch.cat([t.reshape(-1) for t in c]).pin_memory() for c in cam_host]   # 16 + 16 + 3 floats
    cam_pack = torch.empty(35, device=dev)
    cam_dev = (cam_pack[:16].view(4, 4), cam_pack[16:32].view(4, 4), cam_pack[32:35])
    # upstream gradients travel as one (C + 2, H, W) block: [dL/dcolour | dL/ddepth | dL/dalpha]
    grads_packed = [torch.cat(g, 0).pin_memory() for g in grads_host]
    g_pack = torch.empty(C + 2, H, W, device=dev)
    g_dev = (g_pack[:C], g_pack[C:C + 1], g_pack[C + 1:])
    img_host = (torch.empty(C, H, W).pin_memory(), torch.empty(1, H, W).pin_memory(), torch.empty(1, H, W).pin_memory())
    h2d = sum(t.numel() * 4 for t in cam_dev) + sum(t.numel() * 4 for t in g_dev)
    d2h = sum(t.numel() * 4 for t in img_host)
    copy_stream = torch.cuda.Stream(device=dev)
    ev_grads, ev_fwd, ev_imgs = torch.cuda.Event(), torch.cuda.Event(), torch.cuda.Event()

    def step_e2e_eager(i):
        """Copies ride a second stream: the upstream gradients arrive while the forward runs, the
        images leave while the backward runs; everything still completes inside the timed step."""
        k = i % N_CAMERAS
        main = torch.cuda.current_stream(dev)
        cam_pack.copy_(cam_packed[k], non_blocking=True)
        copy_stream.wait_stream(main)                  # previous step's consumers of g_dev / img_host are done
        with torch.cuda.stream(copy_stream):
            g_pack.copy_(grads_packed[k], non_blocking=True)
            ev_grads.record(copy_stream)
        c = cams_host[k]
        rs = Settings(H, W, c.tanfovx, c.tanfovy, bg, 1.0, cam_dev[0], cam_dev[1], 0, cam_dev[2], False, False)
        for t in leaves.values():
            t.grad = None
        means2D.grad = None
        color, depth, alpha, radii = Rasterizer(rs)(
            means3D=leaves["means3D"], means2D=means2D, opacities=leaves["opacities"],
            colors_precomp=leaves["colors_precomp"], scales=leaves["scales"], rotations=leaves["rotations"])
        ev_fwd.record(main)
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(ev_fwd)
            for dst, src in zip(img_host, (color, depth, alpha)):
                dst.copy_(src.detach(), non_blocking=True)
            ev_imgs.record(copy_stream)
        main.wait_event(ev_grads)
        torch.autograd.backward([color, depth, alpha], list(g_dev))
        if use_dist:
            for t in leaves.values():
                job.dist.all_reduce(t.grad)
        main.wait_event(ev_imgs)                       # the step ends when its images are on the host

    sampler = None
    if rank == 0 and not os.environ.get("BENCH_NO_SAMPLER"):
        sampler = ClockSampler(range(world))
        sampler.start()       # started before the spin-up: nvidia-smi's own start-up (NVML init) perturbs launches
    spin_up(lambda i: step_eager(i, reduce=False))

    # ---- this library: the step captured once in a CUDA graph (splatloc_b200.graphs) -- possible
    # because the no-read-back forward never waits for the device -- and replayed with one launch.
    graphed = None
    if impl == "ours" and not os.environ.get("BENCH_E2E_EAGER"):
        from splatloc_b200.graphs import GraphedViewStep
        max_R = max(step_eager(i, reduce=False)[0] for i in range(N_CAMERAS))
        img_host_packed = torch.empty(C + 2, H, W).pin_memory()
        c0 = cams_host[0]
        graphed = GraphedViewStep(leaves, H, W, c0.tanfovx, c0.tanfovy, bg, max_pairs=int(1.25 * max_R) + 4096,
                                  host_images=img_host_packed)
        exchange = Exchange(job, graphed.grad_block.trainable_numel, overlap=not os.environ.get("BENCH_SYNC_EXCHANGE"))
        exchange_sync = Exchange(job, graphed.grad_block.trainable_numel, overlap=False)
        stage = [torch.empty(C + 2, H, W, device=dev) for _ in range(2)]
        copy_in = torch.cuda.Stream(device=dev)
        ev_in = [torch.cuda.Event(), torch.cuda.Event()]      # stage[b] holds its step's gradients
        ev_free = [torch.cuda.Event(), torch.cuda.Event()]    # stage[b] has been consumed
        pipe = {"sent": None}

        def send_grads(step):
            b = step % 2
            with torch.cuda.stream(copy_in):
                copy_in.wait_event(ev_free[b])
                stage[b].copy_(grads_packed[step % N_CAMERAS], non_blocking=True)
                ev_in[b].record(copy_in)

        def set_camera_from_host(k):
            graphed.viewmatrix.copy_(cam_host[k][0], non_blocking=True)
            graphed.projmatrix.copy_(cam_host[k][1], non_blocking=True)
            graphed.campos.copy_(cam_host[k][2], non_blocking=True)

        def step_e2e_graph(i):
            """Host buffers, software-pipelined: the upstream gradients of step i + 1 are sent on a copy
            stream while step i runs, then moved into the graph's input by a device copy; the images are
            copied to pinned host memory inside the graph, overlapped with the backward."""
            k, b = i % N_CAMERAS, i % 2
            main = torch.cuda.current_stream(dev)
            if pipe["sent"] != i:
                send_grads(i)
            set_camera_from_host(k)
            main.wait_event(ev_in[b])
            graphed.upstream.copy_(stage[b], non_blocking=True)
            ev_free[b].record(main)
            send_grads(i + 1)
            pipe["sent"] = i + 1
            graphed.replay()
            exchange(graphed.trainable_grads())

        def step_e2e_graph_unpipelined(i):
            """Host buffers, nothing prefetched: this step's own camera and upstream gradients cross PCIe
            inside its own event pair, then the replay (which copies the images back)."""
            k = i % N_CAMERAS
            set_camera_from_host(k)
            graphed.upstream.copy_(grads_packed[k], non_blocking=True)
            graphed.replay()
            exchange(graphed.trainable_grads())

        def load_resident(i):
            cam, g3 = cams[i % N_CAMERAS], grads[i % N_CAMERAS]
            graphed.viewmatrix.copy_(cam.viewmatrix)
            graphed.projmatrix.copy_(cam.projmatrix)
            graphed.campos.copy_(cam.campos)
            graphed.upstream[:C].copy_(g3[0])
            graphed.upstream[C:C + 1].copy_(g3[1])
            graphed.upstream[C + 1:].copy_(g3[2])

        def step_resident_graph(i):
            """The resident step submitted as one graph replay: camera and upstream gradients are device
            tensors, copied into the graph's static inputs by device copies."""
            load_resident(i)
            graphed.replay()
            exchange(graphed.trainable_grads())

        def step_resident_graph_sync(i):
            load_resident(i)
            graphed.replay()
            exchange_sync(graphed.trainable_grads())

        # The exchange fused into the backward (splatloc_b200.exchange): the last kernel of the step adds
        # the rows this GPU saw into every GPU's buffer through the switch; a device-side barrier closes
        # the step.  Two graphs, one per exchange buffer; needs NVSwitch multicast, else it stays off.
        fused = None
        if use_dist and not os.environ.get("BENCH_NO_FUSED_EXCHANGE"):
            try:
                from splatloc_b200.exchange import NvlsGradientExchange
                xch = NvlsGradientExchange(P, dev)
                fused = [GraphedViewStep(leaves, H, W, c0.tanfovx, c0.tanfovy, bg, max_pairs=int(1.25 * max_R) + 4096,
                                         exchange_multicast=xch.multicast_ptr(b)) for b in (0, 1)]
                # the same with the barrier split in two inside the replay: arrive at the end of a step, wait
                # after the next step's forward (the sums of step i are read during step i + 1, as with the
                # overlapped all-reduce)
                fused_split = [GraphedViewStep(leaves, H, W, c0.tanfovx, c0.tanfovy, bg, max_pairs=int(1.25 * max_R) + 4096,
                                               exchange=xch, exchange_buffer=b) for b in (0, 1)]
                # and with the rows sent by a kernel of their own (sl_exchange_push) beside the NEXT step's forward:
                # plain backward, nothing of the exchange on the critical path but the wait that ends a step
                fused_later = [GraphedViewStep(leaves, H, W, c0.tanfovx, c0.tanfovy, bg, max_pairs=int(1.25 * max_R) + 4096,
                                               exchange=xch, exchange_buffer=b, exchange_mode="next_step") for b in (0, 1)]
                img_host_later = [torch.empty(C + 2, H, W).pin_memory() for _ in range(2)]
                fused_later_e2e = [GraphedViewStep(leaves, H, W, c0.tanfovx, c0.tanfovy, bg, max_pairs=int(1.25 * max_R) + 4096,
                                                   host_images=img_host_later[b], exchange=xch, exchange_buffer=b,
                                                   exchange_mode="next_step") for b in (0, 1)]
                side = torch.cuda.Stream(device=dev)
            except Exception as e:      # noqa: BLE001
                fused = None
                fused_error = f"{type(e).__name__}: {str(e)[:160]}"
            ok = job.max_over_ranks(0.0 if fused is not None else 1.0) == 0.0      # all ranks or none
            if not ok:
                fused = None

        def load_into(st, i):
            cam, g3 = cams[i % N_CAMERAS], grads[i % N_CAMERAS]
            st.viewmatrix.copy_(cam.viewmatrix)
            st.projmatrix.copy_(cam.projmatrix)
            st.campos.copy_(cam.campos)
            st.upstream[:C].copy_(g3[0])
            st.upstream[C:C + 1].copy_(g3[1])
            st.upstream[C + 1:].copy_(g3[2])

        def step_resident_graph_fused_split(i):
            st = fused_split[i % 2]
            load_into(st, i)
            st.replay()          # forward | wait for step i - 1's rows | zero them | backward + push | arrive

        later = {"last": 0}

        def step_resident_graph_fused_later(i):
            st = fused_later[i % 2]
            load_into(st, i)
            st.replay()          # [zero | push step i - 1's rows | arrive] beside [forward | backward], then wait
            later["last"] = i % 2

        def step_e2e_graph_later(i):
            """step_e2e_graph with the gradient exchange by sl_exchange_push beside the next step's forward."""
            k, b = i % N_CAMERAS, i % 2
            st = fused_later_e2e[b]
            main = torch.cuda.current_stream(dev)
            if pipe["sent"] != i:
                send_grads(i)
            st.viewmatrix.copy_(cam_host[k][0], non_blocking=True)
            st.projmatrix.copy_(cam_host[k][1], non_blocking=True)
            st.campos.copy_(cam_host[k][2], non_blocking=True)
            main.wait_event(ev_in[b])
            st.upstream.copy_(stage[b], non_blocking=True)
            ev_free[b].record(main)
            send_grads(i + 1)
            pipe["sent"] = i + 1
            st.replay()
            later["last"] = b

        def step_resident_graph_fused(i):
            b = i % 2
            st = fused[b]
            load_into(st, i)
            main = torch.cuda.current_stream(dev)
            side.wait_stream(main)                   # the other buffer's sums have been consumed
            with torch.cuda.stream(side):
                xch.zero(1 - b)                      # ... nobody pushes into it before the barrier below
            st.replay()                              # forward + backward; this GPU's rows land in buffer b everywhere
            main.wait_stream(side)
            xch.barrier()                            # all GPUs' rows have landed: xch.rows(b) is the sum

    n_ranks = world
    if sampler is not None:
        sampler.lines.clear()    # keep only samples taken during the timed region
    if impl == "ours":
        from splatloc_b200 import _capi
        lib = _capi.load()
    if graphed is not None:
        ms_total, per_step, wall = job.timed(step_resident_graph, args.steps, args.warmup, after=exchange.wait)
        launches = graphed.kernels_per_replay * args.steps     # counted by the library while the graph was captured
    else:
        n0 = lib.sl_launch_count() if impl == "ours" else 0
        ms_total, per_step, wall = job.timed(step_eager, args.steps, args.warmup)
        launches = ((lib.sl_launch_count() - n0) * args.steps // (args.steps + args.warmup)) if impl == "ours" else 0
    sync_ms = None
    if graphed is not None:
        e2e_ms, _, _ = job.timed(step_e2e_graph, args.steps, args.warmup, after=exchange.wait)
        e2e_unpiped_ms, _, _ = job.timed(step_e2e_graph_unpipelined, args.steps, args.warmup, after=exchange.wait)
        assert int(graphed.overflow) == 0, "bench: the graphed step's pair capacity was exceeded"
        fused_ms = fused_split_ms = fused_later_ms = e2e_later_ms = e2e_nccl_ms = None
        headline_exchange = "nccl_overlapped"
        if use_dist:
            sync_ms, _, _ = job.timed(step_resident_graph_sync, args.steps, args.warmup)
            if fused is not None:
                xch.reset()
                fused_ms, fused_per_step, _ = job.timed(step_resident_graph_fused, args.steps, args.warmup)
                assert int(fused[0].overflow) == 0 and int(fused[1].overflow) == 0
                xch.reset()
                fused_split_ms, fused_split_per_step, _ = job.timed(step_resident_graph_fused_split, args.steps, args.warmup,
                                                                    after=xch.wait_all)
                assert int(fused_split[0].overflow) == 0 and int(fused_split[1].overflow) == 0
                assert not xch.timed_out, "bench: a split-phase exchange wait timed out"
                for st in fused_later:
                    st.bind_exchange_source()
                xch.reset()
                fused_later_ms, fused_later_per_step, _ = job.timed(step_resident_graph_fused_later, args.steps, args.warmup,
                                                                    after=lambda: xch.flush_deferred(later["last"]))
                assert int(fused_later[0].overflow) == 0 and int(fused_later[1].overflow) == 0
                assert not xch.timed_out, "bench: a deferred-push exchange wait timed out"
                for st in fused_later_e2e:
                    st.bind_exchange_source()
                xch.reset()
                pipe["sent"] = None
                e2e_later_ms, _, _ = job.timed(step_e2e_graph_later, args.steps, args.warmup,
                                               after=lambda: xch.flush_deferred(later["last"]))
                assert int(fused_later_e2e[0].overflow) == 0 and int(fused_later_e2e[1].overflow) == 0
                assert not xch.timed_out, "bench: a deferred-push exchange wait timed out (e2e)"
                e2e_nccl_ms = e2e_ms
                e2e_ms = min(e2e_ms, e2e_later_ms)
                nccl_ms = ms_total
                # the headline is the fastest of the exchanges (same on all ranks: max-reduced times)
                if fused_ms < ms_total:
                    ms_total, per_step, headline_exchange = fused_ms, fused_per_step, "fused_push"
                    launches = fused[0].kernels_per_replay * args.steps
                if fused_split_ms < ms_total:
                    ms_total, per_step, headline_exchange = fused_split_ms, fused_split_per_step, "fused_push_split_barrier"
                    launches = fused_split[0].kernels_per_replay * args.steps
                if fused_later_ms < ms_total:
                    ms_total, per_step, headline_exchange = fused_later_ms, fused_later_per_step, "push_beside_next_forward"
                    launches = fused_later[0].kernels_per_replay * args.steps
    else:
        e2e_ms, _, _ = job.timed(step_e2e_eager, args.steps, args.warmup)
    e2e_eager_ms, _, _ = job.timed(step_e2e_eager, args.steps, args.warmup) if graphed is not None else (None, None, None)
    clocks = sampler.stop() if sampler is not None else None      # sampled over the resident and the e2e timed loops

    # per-stage device times: same steps again with the library's event timers on (kept apart from
    # the headline loop because reading the events back blocks the host between stages)
    prof = None
    if impl == "ours":
        lib.sl_profile_enable(1)
        job.timed(lambda i: step_eager(i, reduce=False), max(5, args.steps // 2), 1)
        prof = _capi.profile_read()
        lib.sl_profile_enable(0)

    # visible Gaussians and pairs of each of the cycled views; the byte counts below use their means,
    # like the times they are divided by (steps cycle through the N_CAMERAS views)
    per_cam = []
    for i in range(N_CAMERAS):
        Ri, radii = step_eager(i, reduce=False)
        per_cam.append((int((radii > 0).sum()), int(Ri)))
    V = round(sum(v for v, _ in per_cam) / N_CAMERAS)
    R = round(sum(r for _, r in per_cam) / N_CAMERAS)
    torch.cuda.synchronize()
    ms_step = ms_total / args.steps
    mpix = lambda ms_sum: round(N * n_ranks / (ms_sum / args.steps * 1e-3) / 1e6, 3)   # noqa: E731

    parallelism = f"views sharded over {n_ranks} GPU(s)"
    if use_dist and graphed is not None and headline_exchange == "push_beside_next_forward":
        parallelism += (", gradient exchange by this library's own NVLS kernel: every Gaussian a GPU's view saw adds its row "
                        "(15 trainable floats + a visibility count) into all GPUs' exchange buffers with multimem.red through "
                        "the NVSwitch (sl_exchange_push, splatloc_b200.exchange), launched inside the replayed step on a second "
                        "stream beside the NEXT step's forward; split-phase device-side barrier (arrive after the push, wait at "
                        "the end of that step), so the sums of step i are complete at the end of step i + 1 (as with the "
                        "overlapped all-reduce); the last step's push and wait are inside the timed region; NCCL only for the "
                        "comparison figures in `exchange`")
    elif use_dist and graphed is not None and headline_exchange != "nccl_overlapped":
        parallelism += (", gradient exchange fused into the per-Gaussian backward: every Gaussian a GPU's view saw adds its "
                        "row (15 trainable floats + a visibility count) into all GPUs' exchange buffers with multimem.red "
                        "through the NVSwitch (splatloc_b200.exchange, sl_backward_batch_push); ")
        parallelism += ("a device-side cross-GPU barrier inside the timed step; sums complete at the end of every step"
                        if headline_exchange == "fused_push" else
                        "split-phase device-side barrier inside the replayed step: arrive after the push, wait after the next "
                        "step's forward, so the sums of step i are read during step i + 1 (as with the overlapped all-reduce); "
                        "the last step's wait is inside the timed region")
        parallelism += "; NCCL only for the comparison figures in `exchange`"
    elif use_dist:
        parallelism += (", one NCCL all-reduce per step of the trainable-gradient span of the packed block (means3D, "
                        "opacity, scales, rotations, colours: 15 floats per Gaussian)")
        parallelism += (", issued asynchronously from a double-buffered copy so that it runs under the next step's "
                        "forward and blend; the last one is waited for inside the timed region"
                        if graphed is not None and exchange.overlap else ", issued in place, the next step waits for it")
    out = {
        "metric": "fwd+bwd Mpix/s at 640x480", "value": mpix(ms_total), "unit": "Mpix/s",
        "n_gpus": n_ranks, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_step, 4),
        "ms_per_step_median": round(statistics.median(per_step), 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "impl": impl,
        "config": {"workload": WORKLOAD, "gaussians": P, "channels": C, "image": [W, H], "views_per_gpu_per_step": 1,
                   "visible": V, "pairs": R, "visible_pairs_per_view": per_cam,
                   "submission": "one CUDA-graph replay per step (splatloc_b200.graphs)" if graphed is not None else "eager calls",
                   "l2": f"flushed between timed steps ({L2_FLUSH_BYTES >> 20} MiB memset, untimed)",
                   "parallelism": parallelism},
        "e2e": {"value": mpix(e2e_ms), "unit": "Mpix/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "api": ("splatloc_b200.graphs.GraphedViewStep: forward (no pair-count read-back) + backward captured once "
                        "in a CUDA graph, one replay per step; camera + upstream pixel gradients from pinned host memory "
                        "(step i+1's gradients are sent on a copy stream while step i runs), images copied to pinned "
                        "host memory inside the graph; the Gaussian map is resident state.  unpipelined_value: the same "
                        "without the prefetch (every step's own transfers inside its own event pair); eager_value: the "
                        "drop-in GaussianRasterizer autograd API, no graph") if graphed is not None else
                       ("GaussianRasterizer autograd fwd+bwd; camera + upstream pixel gradients from pinned host memory, "
                        "images read back (copies on a second stream, inside the timed step); the Gaussian map is "
                        "resident state")},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "wall_s": round(wall, 3),
    }
    if graphed is not None:
        out["e2e"]["unpipelined_value"] = mpix(e2e_unpiped_ms)
        out["e2e"]["eager_value"] = mpix(e2e_eager_ms)
        if e2e_later_ms is not None:     # value = the faster of the two exchanges, as for the resident figure
            out["e2e"]["exchange"] = "push_beside_next_forward" if e2e_later_ms <= e2e_nccl_ms else "nccl_overlapped"
        if sync_ms is not None:
            out["exchange"] = {"nccl_overlapped_ms_per_step": round((nccl_ms if fused_ms is not None else ms_total) / args.steps, 4),
                               "nccl_serialised_ms_per_step": round(sync_ms / args.steps, 4),
                               "fused_push_ms_per_step": round(fused_ms / args.steps, 4) if fused_ms is not None else None,
                               "fused_push_split_barrier_ms_per_step": (round(fused_split_ms / args.steps, 4)
                                                                        if fused_split_ms is not None else None),
                               "push_beside_next_forward_ms_per_step": (round(fused_later_ms / args.steps, 4)
                                                                        if fused_later_ms is not None else None),
                               "headline": headline_exchange,
                               "e2e_nccl_overlapped_value": mpix(e2e_nccl_ms) if e2e_nccl_ms is not None else None,
                               "e2e_push_beside_next_forward_value": mpix(e2e_later_ms) if e2e_later_ms is not None else None,
                               "nccl_bytes": int(graphed.grad_block.trainable_numel) * 4,
                               "fused_push_bytes_sent": "64 B per Gaussian this GPU's view saw (multimem.red through the switch)"}
    if impl == "ours":
        peak, how = measured_peaks()
        alg = stage_bytes(P, V, R, N, Tn, C)
        stages = {}
        for name, (tot, calls) in prof.items():
            if calls:
                ms = tot / calls
                stages[name] = {"ms": round(ms, 4), "alg_MB": round(alg.get(name, 0) / 1e6, 2),
                                "GBps": round(alg.get(name, 0) / (ms * 1e-3) / 1e9, 1)}
        top = max(stages, key=lambda k: stages[k]["ms"])
        ach = stages[top]["GBps"]
        step_alg = survey_step_bytes(P, V, R, N, Tn, C)
        out["roofline"] = {"kernel": top, "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s",
                           "frac": round(ach / peak, 4), "traffic": measured_traffic(top), "peak_source": how,
                           "note": "blend kernels are FP32-issue-bound, not HBM-bound; see profiles/ for pipe utilisation",
                           "step_alg_MB": round(step_alg / 1e6, 1),
                           "step_frac": round(step_alg / (ms_step * 1e-3) / 1e9 / peak, 4),
                           "step_alg_source": "SURVEY.md 8(d) B_fwd + B_bwd with the measured V and R"}
        out["stages"] = stages
        out["parity"] = ("integers and images bit-exact vs the reference CUDA build on this scene and these cameras "
                         "(tests/test_gpu_bench_parity.py); extensions (n_touched, pose gradient, pose refinement): "
                         "float64 oracle only, parity unpinned")
    else:
        out["cpu_baseline"] = {"value": out["value"], "unit": "Mpix/s", "cores": 0, "kind": "reference",
                               "sample": "full workload; the reference path is CUDA-only, so this arm is the reference's own "
                                         "CUDA rasterizer (oracle/_ref, sm_100a) on the GPU, not host cores"}
    return job, out, dict(sc=sc, graphed=graphed)


# ------------------------------------------------------------------------------------ sub-records
def _window_cameras(job, n, seed0, dev, **kw):
    cams = [scenes.make_camera(seed=seed0 + i, **kw).to(dev) for i in range(n)]
    return (torch.stack([c.viewmatrix for c in cams]), torch.stack([c.projmatrix for c in cams]),
            torch.stack([c.campos for c in cams]), cams[0])


def _graphed_window(job, sc, V, W, H, cam_sets, steps, warmup, exchange=None, push_exchange=False):
    """fwd+bwd of V views per replay (batch entry points inside one CUDA graph), cycling over `cam_sets`;
    returns (ms per window: max over ranks, pairs of the first window, kernels per replay).
    `exchange`: called on the trainable gradients after every replay (NCCL).  `push_exchange`: instead, two
    alternating steps whose rows leave through sl_exchange_push beside the next window's forward
    (splatloc_b200.exchange, GraphedViewStep exchange_mode="next_step"); the last push and wait are timed."""
    from splatloc_b200 import _C
    from splatloc_b200.graphs import GraphedViewStep
    dev, C = job.dev, int(sc["colors_precomp"].shape[1])
    bg = torch.zeros(C, device=dev)
    c0 = cam_sets[0][3]
    Rs = []
    for vm, pm, cp, _ in cam_sets:
        Rs.append(_C.rasterize_gaussians_batch(bg, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"],
                                               sc["rotations"], 1.0, EMPTY, EMPTY.bool(), vm, pm, c0.tanfovx, c0.tanfovy,
                                               H, W, EMPTY, 0, cp, False, False)[0])
    params = {k: sc[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "colors_precomp", "scales", "rotations")}
    xch = None
    if push_exchange:
        from splatloc_b200.exchange import NvlsGradientExchange
        xch = NvlsGradientExchange(int(sc["means3D"].shape[0]), dev)
        pair = [GraphedViewStep(params, H, W, c0.tanfovx, c0.tanfovy, bg, max_pairs=int(1.25 * max(Rs)) + 4096, n_views=V,
                                exchange=xch, exchange_buffer=b, exchange_mode="next_step") for b in (0, 1)]
        for st in pair:
            st.bind_exchange_source()
        xch.reset()
        step = pair[0]
    else:
        step = GraphedViewStep(params, H, W, c0.tanfovx, c0.tanfovy, bg, max_pairs=int(1.25 * max(Rs)) + 4096, n_views=V)
    ups = [torch.cat([torch.stack([scenes.make_pixel_grads(7 * s + v, C, H, W)[k] for v in range(V)]).reshape(-1)
                      for k in range(3)]).to(dev) for s in range(len(cam_sets))]

    last = {"b": 0}

    def fn(i):
        st = step if xch is None else pair[i % 2]
        vm, pm, cp, _ = cam_sets[i % len(cam_sets)]
        st.viewmatrices.copy_(vm)
        st.projmatrices.copy_(pm)
        st.cam_positions.copy_(cp)
        st.upstream_flat.copy_(ups[i % len(cam_sets)])
        st.replay()
        last["b"] = i % 2
        if exchange is not None:
            exchange(st.trainable_grads())

    total, per, _ = job.timed(fn, steps, warmup, after=(lambda: xch.flush_deferred(last["b"])) if xch is not None else None)
    assert int(step.overflow) == 0, "bench: pair capacity of the multi-view step exceeded"
    if xch is not None:
        assert int(pair[1].overflow) == 0 and not xch.timed_out, "bench: multi-view exchange overflow / timeout"
        del pair
    k = step.kernels_per_replay
    del step, xch
    return total / steps, Rs[0], k


def sub_multiview(job, args, sc):
    """Config 2 with several cameras per launch (sl_forward_batch / sl_backward_batch): what batching buys
    the latency-bound per-Gaussian and binning stages.  No exchange inside (per-GPU figure x ranks)."""
    W, H = scenes.REPLICA_INTRINSICS["W"], scenes.REPLICA_INTRINSICS["H"]
    out = []
    for V in (4, 8):
        sets = [_window_cameras(job, V, 100 * job.rank + V * s, job.dev) for s in range(2)]
        ms, R, k = _graphed_window(job, sc, V, W, H, sets, max(5, args.steps // 2), 3)
        out.append({"views_per_launch": V, "ms_per_launch": round(ms, 4), "ms_per_view": round(ms / V, 4),
                    "Mpix_per_s": round(W * H * V * job.world / (ms * 1e-3) / 1e6, 1), "pairs": R, "kernels_per_launch": k})
    return out


def sub_train_window(job, args):
    """BASELINE.json config 4: 3 M Gaussians, 1296x968, 4 views per GPU accumulated into one packed
    gradient block per optimiser step, then ONE all-reduce of its trainable span (N > 1)."""
    P, W, H, V = 3_000_000, 1296, 968, 4
    sc = {k: v.to(job.dev) for k, v in scenes.make_scene(P, seed=2, channels=4, scale=1.5).items()}
    sets = [_window_cameras(job, V, 1000 * job.rank + V * s, job.dev, size=(W, H), scale=1.5) for s in range(2)]
    from splatloc_b200._C import trainable_grad_numel
    ex = Exchange(job, trainable_grad_numel(P, 4, 0), overlap=False) if job.dist is not None else None
    ms, R, k = _graphed_window(job, sc, V, W, H, sets, max(4, args.steps // 4), 2, exchange=ex)
    rec = {"workload": "multiview_window_3M_1296x968_c4", "gaussians": P, "image": [W, H], "views_per_gpu_per_exchange": V,
           "views_total": V * job.world, "ms_per_window": round(ms, 3), "Mpix_per_s": round(W * H * V * job.world / (ms * 1e-3) / 1e6, 1),
           "pairs_per_window": R, "exchange_bytes": trainable_grad_numel(P, 4, 0) * 4 if ex is not None else 0}
    if job.dist is not None and not os.environ.get("BENCH_NO_FUSED_EXCHANGE"):
        # the same window with this library's own exchange: visible rows pushed through the switch beside the next
        # window's forward (the sums of window i are complete at the end of window i + 1); headline = the faster
        ok = 1.0
        try:
            torch.cuda.empty_cache()
            ms_push, _, _ = _graphed_window(job, sc, V, W, H, sets, max(4, args.steps // 4), 2, push_exchange=True)
        except Exception as e:      # noqa: BLE001
            ok, ms_push = 0.0, None
            rec["push_exchange_error"] = f"{type(e).__name__}: {str(e)[:160]}"
        if job.max_over_ranks(1.0 - ok) == 0.0:
            rec["exchange"] = {"nccl_after_the_window_ms": round(ms, 3), "push_beside_next_forward_ms": round(ms_push, 3),
                               "headline": "push_beside_next_forward" if ms_push < ms else "nccl_after_the_window"}
            if ms_push < ms:
                rec["ms_per_window"] = round(ms_push, 3)
                rec["Mpix_per_s"] = round(W * H * V * job.world / (ms_push * 1e-3) / 1e6, 1)
    del sc
    torch.cuda.empty_cache()
    return rec


def sub_pose_refine(job, args, frames_per_gpu=8, iters=100):
    """BASELINE.json config 3 (bounded): query frames refined independently against a 1.2 M-Gaussian map,
    12-Scenes intrinsics, frames sharded over the ranks, no collective per iteration; all frames of a
    rank are rendered by one launch per iteration (batch entry points, pose-only backward)."""
    from splatloc_b200.multiview import gather_frames, shard_range
    from splatloc_b200.pose_refine import PoseRefiner, perturb_pose, pose_error
    dev, P = job.dev, 1_200_000
    k = scenes.SCENES12_INTRINSICS
    ref = PoseRefiner(scenes.make_scene(P, seed=0, channels=4), k, device=dev)
    n_frames = frames_per_gpu * job.world
    mine = list(shard_range(n_frames, job.rank, job.world))
    gt = [scenes.make_camera(seed=500 + f, intrinsics=k).viewmatrix.transpose(0, 1).double() for f in mine]
    init = torch.stack([perturb_pose(w, seed=f) for w, f in zip(gt, mine)]).to(dev)
    with torch.no_grad():
        tg = [ref.render(w.to(dev)) for w in gt]
    color = torch.stack([t[0] for t in tg]).detach()
    depth = torch.stack([t[1] for t in tg]).detach()
    ref.refine_batch(init, color, depth, iters=iters)            # capture + warm-up
    job.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    res = ref.refine_batch(init, color, depth, iters=iters)
    e1.record()
    torch.cuda.synchronize()
    ms = job.max_over_ranks(e0.elapsed_time(e1))
    rec = {"workload": "scenes12_pose_refine_1200k_640x480_c4", "gaussians": P, "frames": n_frames, "frames_per_launch": frames_per_gpu,
           "iters_per_frame": iters, "parity": "unpinned (no reference implementation; float64 oracle only)"}
    # (every rank must take the same branch: gather_frames below is a collective)
    if job.max_over_ranks(1.0 if res is None else 0.0) > 0:
        rec["error"] = "pair capacity exceeded on some rank"
        return rec
    poses = res[0]
    allposes = gather_frames(poses.reshape(len(mine), 16), n_frames)
    errs0 = [pose_error(a, b) for a, b in zip(init, gt)]
    errs1 = [pose_error(a, b) for a, b in zip(poses, gt)]
    # one frame per launch, for comparison (2 frames, same iteration count)
    ref.refine_graphed(init[0], color[0], depth[0], iters=iters)
    torch.cuda.synchronize()
    e0.record()
    for f in range(2):
        ref.refine_graphed(init[f], color[f], depth[f], iters=iters)
    e1.record()
    torch.cuda.synchronize()
    single_ms = e0.elapsed_time(e1)
    rec.update({"iters_per_s": round(n_frames * iters / (ms * 1e-3), 1), "ms_per_iteration_of_all_frames": round(ms / iters, 4),
                "iters_per_s_one_frame_per_launch": round(job.world * 2 * iters / (single_ms * 1e-3), 1),
                "gathered_poses": list(allposes.shape),
                "mean_trans_err_m": [round(sum(e[0] for e in errs0) / len(errs0), 4), round(sum(e[0] for e in errs1) / len(errs1), 4)],
                "mean_rot_err_deg": [round(sum(e[1] for e in errs0) / len(errs0), 3), round(sum(e[1] for e in errs1) / len(errs1), 3)]})
    del ref
    torch.cuda.empty_cache()
    return rec


def sub_sweep(job, args, impl, points=(100_000, 200_000, 500_000, 1_200_000, 3_000_000)):
    """fwd+bwd Mpix/s at 640x480 against the number of Gaussians (the first half of BASELINE.json's
    metric), N = 1: eager operator calls for both arms, plus one graph replay per step for this library."""
    be, _, _ = load_backend(impl)
    dev, C = job.dev, C_DEFAULT
    W, H = scenes.REPLICA_INTRINSICS["W"], scenes.REPLICA_INTRINSICS["H"]
    bg = torch.zeros(C, device=dev)
    cams = [scenes.make_camera(seed=i).to(dev) for i in range(4)]
    grads = [tuple(t.to(dev) for t in scenes.make_pixel_grads(i, C, H, W)) for i in range(4)]
    steps, rows = max(8, args.steps // 2), []
    for P in points:
        sc = {k: v.to(dev) for k, v in scenes.make_scene(P, seed=0, channels=C).items()}
        stat = {}

        def eager(i):
            cam, g3 = cams[i % 4], grads[i % 4]
            fw = be.rasterize_gaussians(bg, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"], sc["rotations"],
                                        1.0, EMPTY, EMPTY.bool(), cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy,
                                        H, W, EMPTY, 0, cam.campos, False, False)
            be.rasterize_gaussians_backward(bg, sc["means3D"], fw[4], sc["colors_precomp"], sc["scales"], sc["rotations"], 1.0,
                                            EMPTY, cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, *g3, EMPTY, 0,
                                            cam.campos, fw[5], fw[0], fw[6], fw[7], fw[3], False)
            if i % 4 == 0:
                stat["R"], stat["V"] = fw[0], fw[4]
        ms, _, _ = job.timed(eager, steps, 10)     # (the caching allocator settles on this size's chunks first)
        row = {"gaussians": P, "visible": int((stat["V"] > 0).sum()), "pairs": int(stat["R"]),
               "eager_Mpix_per_s": round(W * H / (ms / steps * 1e-3) / 1e6, 1)}
        if impl == "ours":
            sets = [(c.viewmatrix[None], c.projmatrix[None], c.campos[None], c) for c in cams]
            gms, _, _ = _graphed_window(job, sc, 1, W, H, sets, steps, 3)
            row["graph_Mpix_per_s"] = round(W * H / (gms * 1e-3) / 1e6, 1)
        rows.append(row)
        del sc
        torch.cuda.empty_cache()
    return rows


def sub_knn(job, args, impl, points=(200_000, 1_200_000, 5_000_000)):
    """SURVEY 8(d) metric (iii): distCUDA2 (simple_knn: mean squared distance to the 3 nearest neighbours,
    reference: simple_knn.cu:186-221) wall time and Mpoints/s on seeded point clouds, N = 1; device-resident
    points, CUDA events, L2 flushed between calls.  32 algorithmic bytes per point (read xyz, write one float,
    one key/value sort in and out)."""
    be, _, _ = load_backend(impl)
    rows = []
    for P in points:
        pts = scenes.make_knn_points(P, seed=P % 7).to(job.dev)
        out = {}

        def call(i):
            out["d"] = be.distCUDA2(pts)
        ms, _, _ = job.timed(call, 10, 3)
        t = ms / 10
        rows.append({"points": P, "ms": round(t, 4), "Mpoints_per_s": round(P / (t * 1e-3) / 1e6, 1),
                     "GBps_algorithmic": round(32 * P / (t * 1e-3) / 1e9, 1), "mean_dist2": float(out["d"].mean())})
        del pts
        torch.cuda.empty_cache()
    return rows


def sub_mapping_iteration(job, args, sc):
    """SURVEY 8(f) rows 1 and 4 on the config-2 scene, one view, eager calls through the drop-in API: the
    reference's mapping iteration around the rasterizer -- render() with its PyTorch pre-processing,
    backward, torch.optim.Adam over the eight parameter groups -- against render() with the
    pre-processing inside the kernels (sl_forward_render / sl_backward_render) and the fused Adam
    kernel (sl_adam_step).  Both use this library's rasterizer; what differs is the work around it."""
    import math
    import types
    from splatloc_b200.gaussian_optimizer import GaussianStoreOptimizer
    from splatloc_b200.gaussian_renderer import render
    dev, P = job.dev, int(sc["means3D"].shape[0])
    C0 = 0.28209479177387814
    o = sc["opacities"].clamp(1e-4, 1 - 1e-4)
    raw = dict(_xyz=sc["means3D"], _features_dc=((sc["colors_precomp"][:, :3] - 0.5) / C0).view(P, 1, 3).contiguous(),
               _kp_score=sc["colors_precomp"][:, 3:4].contiguous(), _opacity=torch.log(o / (1 - o)),
               _scaling=torch.log(sc["scales"]), _rotation=sc["rotations"] * 1.7)
    lrs = dict(xyz=1.6e-4, f_dc=2.5e-3, opacity=0.05, kp_score=0.05, scaling=1e-3, rotation=1e-3)
    key = dict(xyz="_xyz", f_dc="_features_dc", opacity="_opacity", kp_score="_kp_score", scaling="_scaling", rotation="_rotation")
    cams = [scenes.make_camera(seed=i).to(dev) for i in range(4)]
    views = [types.SimpleNamespace(FoVx=2 * math.atan(c.tanfovx), FoVy=2 * math.atan(c.tanfovy), image_height=c.H, image_width=c.W,
                                   world_view_transform=c.viewmatrix, full_proj_transform=c.projmatrix, camera_center=c.campos)
             for c in cams]
    pipe = types.SimpleNamespace(convert_SHs_python=True, compute_cov3D_python=False)
    bg = torch.zeros(3, device=dev)
    gt = torch.rand(3, cams[0].H, cams[0].W, device=dev)
    steps, out = max(8, args.steps // 2), {}

    # the colour-refinement loop's loss (train_gaussians.py:283-286): PyTorch formulation of loss_utils.py:61-102
    # (five depth-wise 11 x 11 convolutions) against the fused kernel pair (sl_ssim_forward / sl_ssim_backward)
    import torch.nn.functional as F
    from splatloc_b200.ssim_loss import gaussian_window, l1_ssim_loss
    w1 = gaussian_window().to(dev)
    w2 = (w1[:, None] @ w1[None, :]).expand(3, 1, 11, 11).contiguous()

    def torch_l1_ssim(x, y, lam=0.2):
        conv = lambda t: F.conv2d(t[None], w2, padding=5, groups=3)[0]      # noqa: E731
        mu1, mu2 = conv(x), conv(y)
        s1, s2, s12 = conv(x * x) - mu1 * mu1, conv(y * y) - mu2 * mu2, conv(x * y) - mu1 * mu2
        m = ((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1 * mu1 + mu2 * mu2 + 1e-4) * (s1 + s2 + 9e-4))
        return (1.0 - lam) * (x - y).abs().mean() + lam * (1.0 - m.mean())

    losses = {"mapping": {m: (lambda pkg: (pkg["render"] - gt).abs().mean() + pkg["kp_prob"].mean() + 0.1 * pkg["depth"].mean())
                          for m in ("reference_shaped", "fused")},
              "refinement": {"reference_shaped": lambda pkg: torch_l1_ssim(pkg["render"], gt),
                             "fused": lambda pkg: l1_ssim_loss(pkg["render"], gt, 0.2)}}
    for kind, prefix in (("mapping", ""), ("refinement", "refinement_")):
        for mode in ("reference_shaped", "fused"):
            pc = types.SimpleNamespace(active_sh_degree=0, max_sh_degree=0, _features_rest=torch.zeros((P, 0, 3), device=dev))
            for k, t in raw.items():
                setattr(pc, k, t.clone().requires_grad_(True))
            params = {n: getattr(pc, k) for n, k in key.items()}
            if mode == "fused":
                opt = GaussianStoreOptimizer({n: t for n, t in params.items()}, lrs)
                for n, k in key.items():            # the optimiser updates these tensors in place
                    setattr(pc, k, opt.params[n].requires_grad_(True))
                params = {n: getattr(pc, k) for n, k in key.items()}
            else:
                topt = torch.optim.Adam([{"params": [t], "lr": lrs[n], "name": n} for n, t in params.items()], lr=0.0, eps=1e-15)
            loss_fn = losses[kind][mode]

            def it(i):
                pkg = render(views[i % 4], pc, pipe, bg, fuse=mode == "fused")
                loss_fn(pkg).backward()
                if mode == "fused":
                    opt.step({n: t.grad for n, t in params.items() if t.grad is not None})
                    for t in params.values():
                        t.grad = None
                else:
                    topt.step()
                    topt.zero_grad(set_to_none=True)
            ms, _, _ = job.timed(it, steps, 5)
            out[prefix + mode + "_ms_per_iteration"] = round(ms / steps, 4)

    # the refinement iteration replayed as ONE CUDA graph (render without read-back, loss, backward, Adam with the
    # step count on the device): splatloc_b200.graphs.GraphedRefinementIteration
    from splatloc_b200.graphs import GraphedRefinementIteration
    opt = GaussianStoreOptimizer({n: raw[k].clone() for n, k in key.items()}, lrs)
    git = GraphedRefinementIteration(opt, cams[0].H, cams[0].W, cams[0].tanfovx, cams[0].tanfovy, bg, max_pairs=2_500_000,
                                     lambda_dssim=0.2)

    def it_graph(i):
        c = cams[i % 4]
        git.set_view(c.viewmatrix, c.projmatrix, c.campos, gt)
        git.replay()
    ms, _, _ = job.timed(it_graph, steps, 5)
    assert int(git.overflow) == 0
    out["refinement_graphed_ms_per_iteration"] = round(ms / steps, 4)
    out["refinement_graphed_kernels_per_replay"] = git.kernels_per_replay
    del git, opt

    # the loss alone, forward + backward on a (3, H, W) image
    img = torch.rand(3, cams[0].H, cams[0].W, device=dev)
    for name, fn in (("torch_ops", torch_l1_ssim), ("fused", lambda x, y: l1_ssim_loss(x, y, 0.2))):
        def it_loss(i):
            x = img.detach().requires_grad_(True)
            fn(x, gt).backward()
        ms, _, _ = job.timed(it_loss, steps, 5)
        out["l1_ssim_loss_" + name + "_ms"] = round(ms / steps, 4)
    out["note"] = ("render + loss + backward + Adam, eager; reference_shaped = render()'s PyTorch pre-processing "
                   "(gaussian_renderer/__init__.py:73-102), the loss as PyTorch ops and torch.optim.Adam around this library's "
                   "rasterizer; refinement_* = the colour-refinement iteration (train_gaussians.py:270-295) with the "
                   "(1 - 0.2) L1 + 0.2 (1 - SSIM) loss, refinement_graphed = the same iteration replayed as one CUDA graph "
                   "(GraphedRefinementIteration); l1_ssim_loss_*: that loss alone, forward + backward")
    return out


def cpu_baseline(budget_s=20.0):
    """Float64 CPU port (oracle/) on a bounded sample of the same workload: a centred crop of the
    config-2 view, chosen so that the run takes ~10-30 s on the host cores."""
    from oracle import raster_oracle
    torch.set_num_threads(os.cpu_count() or 1)
    return raster_oracle.timed_sample(budget_s=budget_s)


def run_cpu_reference(args):
    cb = cpu_baseline(budget_s=args.cpu_budget)
    cb["kind"] = "port"
    return {"metric": "fwd+bwd Mpix/s at 640x480", "value": cb["value"], "unit": "Mpix/s", "n_gpus": 0,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": None, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "impl": "reference",
            "config": {"workload": WORKLOAD, "note": "reference CUDA library or GPU unavailable: float64 CPU port"},
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def run_mapping_loss(args):
    """Secondary workload (not the default bench line): the per-view mapping loss
    and its gradient seed on one 640x480, C = 4 view.  `--impl ours`: sl_mapping_loss through
    splatloc_b200.mapping_loss (one kernel forward, one scaling multiply backward); `--impl reference`:
    the reference's formulas (utils/utils.py:55-82, train_gaussians.py:38-42) as the PyTorch ops it
    runs, on the same GPU.  Metric: loss fwd+bwd Mpix/s, L2 flushed between steps."""
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    H, W, C = 480, 640, 4
    g = torch.Generator().manual_seed(3)
    image = torch.rand((C, H, W), generator=g)
    image[3] = 4.0 * torch.randn((H, W), generator=g)
    depth = 0.5 + 4.0 * torch.rand((1, H, W), generator=g)
    gt_image = (image[:3] + 0.1 * torch.randn((3, H, W), generator=g)).clamp(0, 1).to(dev)
    gt_depth = (depth[0] + 0.05 * torch.randn((H, W), generator=g)).to(dev)
    gt_kp = (torch.rand((H, W), generator=g) < 0.03).float().to(dev)
    image, depth = image.to(dev).requires_grad_(True), depth.to(dev).requires_grad_(True)
    a = torch.tensor([0.05], device=dev, requires_grad=True)
    b = torch.tensor([-0.01], device=dev, requires_grad=True)
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=dev)

    if args.impl == "ours":
        from splatloc_b200.mapping_loss import fused_mapping_loss

        def loss_fn():
            return fused_mapping_loss(image, depth, gt_image, gt_depth, gt_kp, a, b, 0.01, False)
    else:
        def loss_fn():
            m_rgb = (gt_image.sum(dim=0) > 0.01).view(*depth.shape)
            m_d = (gt_depth[None] > 0.01).view(*depth.shape)
            l1_rgb = torch.abs((torch.exp(a) * image[:3] + b) * m_rgb - gt_image * m_rgb)
            l1_depth = torch.abs(depth * m_d - gt_depth[None] * m_d)
            kp = torch.nn.functional.binary_cross_entropy(torch.sigmoid(image[-1].view(-1)), gt_kp.view(-1), reduction="mean")
            return l1_rgb.mean() + l1_depth.mean() + kp

    def step():
        for t in (image, depth, a, b):
            t.grad = None
        loss_fn().backward()

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    for e0, e1 in ev:
        flush.zero_()
        e0.record()
        step()
        e1.record()
    torch.cuda.synchronize()
    ms = sum(x.elapsed_time(y) for x, y in ev) / args.steps
    N = H * W
    alg = (C + 1 + 3 + 1 + 1 + C + 1) * 4 * N          # reads + gradient writes of the fused forward
    peak, how = measured_peaks()
    out = {"metric": "mapping-loss fwd+bwd Mpix/s at 640x480", "value": round(N / (ms * 1e-3) / 1e6, 2), "unit": "Mpix/s",
           "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 5), "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": args.impl,
           "config": {"workload": "mapping_loss_640x480_c4", "image": [W, H], "channels": C,
                      "l2": f"flushed between timed steps ({L2_FLUSH_BYTES >> 20} MiB memset, untimed)",
                      "note": "SURVEY 8(f) row 2; time includes the autograd call overhead of both arms"}}
    if args.impl == "ours":
        out["roofline"] = {"kernel": "mapping_loss", "bound": "hbm", "achieved": round(alg / (ms * 1e-3) / 1e9, 1), "peak": peak,
                           "unit": "GB/s", "frac": round(alg / (ms * 1e-3) / 1e9 / peak, 4), "traffic": None, "peak_source": how,
                           "note": "whole fwd+bwd step over one kernel's algorithmic bytes: a 3 us kernel inside ~60 us of "
                                   "launch and autograd latency"}
    return out


_RESULT_OUT = None


def claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL's version banner does, whatever
    NCCL_DEBUG_FILE says), so file descriptor 1 is pointed at stderr for the life of the process and the result
    line goes to a private duplicate of the original stdout."""
    global _RESULT_OUT
    if _RESULT_OUT is None:
        sys.stdout.flush()
        _RESULT_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(obj):
    claim_stdout()
    _RESULT_OUT.write(json.dumps(obj) + "\n")
    _RESULT_OUT.flush()


def main():
    world = env_int("WORLD_SIZE", 1)
    if world > 1:     # the ranks share the host cores (scene generation runs on them)
        torch.set_num_threads(max(1, (os.cpu_count() or world) // world))
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--points", type=int, default=P_DEFAULT)
    ap.add_argument("--cpu-budget", type=float, default=20.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", choices=["train_step", "pose_refine", "mapping_loss"], default="train_step")
    ap.add_argument("--frames", type=int, default=8, help="pose_refine: query frames per GPU (one launch renders them all)")
    ap.add_argument("--sub", default="all",
                    help="sub-records of the default line: all, none, or a comma list of sweep,pose_refine,multiview,train_window,mapping_iteration,knn")
    args = ap.parse_args()
    claim_stdout()
    args.warmup = max(args.warmup, 3)
    rank0 = env_int("RANK", 0) == 0
    if args.workload == "mapping_loss":
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
        if rank0:
            emit(run_mapping_loss(args))
        return
    if args.workload == "pose_refine":
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
        job = Job("ours")
        rec = sub_pose_refine(job, args, frames_per_gpu=args.frames, iters=args.steps)
        out = {"metric": "pose-refine iters/s", "value": rec.get("iters_per_s"), "unit": "iters/s", "n_gpus": job.world,
               "steps": args.steps, "warmup": args.warmup, "ms_per_step": rec.get("ms_per_iteration_of_all_frames"),
               "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "impl": "ours", "config": rec}
        if job.rank == 0:
            emit(out)
        job.finish()
        return

    if args.impl == "reference":
        from oracle import ref_cuda
        if not (torch.cuda.is_available() and ref_cuda.available()):
            if rank0:
                emit(run_cpu_reference(args))
            return
        if not rank0:
            return          # rank 0 alone runs the reference arm
    elif not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")

    job, out, state = run_train_step(args, args.impl)

    def note(name):      # progress on stderr: a sub-record that fails later does not hide the earlier ones
        if job.rank == 0:
            print(json.dumps({name: out[name]}), file=sys.stderr, flush=True)
    note("value")
    subs = {"sweep", "pose_refine", "multiview", "train_window", "mapping_iteration", "knn"} if args.sub == "all" else \
        (set() if args.sub == "none" else set(args.sub.split(",")))
    if args.points != P_DEFAULT:
        subs = set()
    state["graphed"] = None         # release the headline step's graph before the sub-records allocate
    if args.impl == "ours":
        if "multiview" in subs:
            out["multiview"] = sub_multiview(job, args, state["sc"])
            note("multiview")
        if "mapping_iteration" in subs:
            out["mapping_iteration"] = sub_mapping_iteration(job, args, state["sc"])
            note("mapping_iteration")
        state.clear()
        torch.cuda.empty_cache()
        if "pose_refine" in subs:
            out["pose_refine"] = sub_pose_refine(job, args)
            note("pose_refine")
        if "train_window" in subs:
            out["train_window"] = sub_train_window(job, args)
            note("train_window")
    state.clear()
    torch.cuda.empty_cache()
    if "sweep" in subs and job.world == 1:
        out["sweep"] = sub_sweep(job, args, args.impl)
    if "knn" in subs and job.world == 1:
        out["knn"] = sub_knn(job, args, args.impl)
    if args.impl == "ours" and job.rank == 0 and job.world == 1 and not args.no_cpu_baseline:   # rank 0 at N = 1 only
        out["cpu_baseline"] = cpu_baseline(budget_s=args.cpu_budget)
    if job.rank == 0:
        emit(out)
    job.finish()


if __name__ == "__main__":
    main()
