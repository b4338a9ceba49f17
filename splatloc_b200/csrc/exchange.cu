// splatloc_b200/csrc/exchange.cu -- split-phase cross-GPU barrier for the gradient exchange that is
// fused into the per-Gaussian backward (preprocess_bwd.cu; no reference counterpart: the
// reference has no distributed code).
//
// A full barrier at the end of every step makes all GPUs wait for the slowest view of that step and
// forces every GPU's incoming rows (N x the visible fraction of the block) through its link in the
// few microseconds between the last kernel and the barrier.  Split in two, "arrive" says "my rows of
// this step have been sent" and "wait" -- one forward pass later, right before the NEXT step's
// backward starts pushing -- says "everyone's rows of the previous step have landed", so a GPU never
// waits unless it is more than a forward pass ahead, and the link traffic spreads over the step.
//
// flags: `world` uint32 counters, one per rank, at the same offset of every GPU's exchange buffer;
// flags[r] = number of steps rank r has arrived at.  An arrive adds 1 to flags[rank] on ALL GPUs with
// one release-ordered multimem.red through the switch (ordered after the rows the preceding kernel
// pushed); a wait spins on the local copy with acquire loads.
#include "kernels.h"

namespace sl {

__global__ void exchange_arrive_kernel(uint32_t* flags_mc, int rank)
{
    if (threadIdx.x == 0) {
        __threadfence_system();
        asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(flags_mc + rank), "r"(1u) : "memory");
    }
}

// One thread per rank.  `waits` (local device memory) counts the waits done so far; the wait needs
// every rank to have arrived `*waits + ahead` times (ahead = 0: at the steps before the one this wait
// is part of; 1: at this step too) and then, if `advance`, counts itself.  Spins for at most
// `timeout_ns`, then sets *status = 1 and gives up.
__global__ void exchange_wait_kernel(const uint32_t* flags, int world, uint32_t* waits, int ahead, int advance,
                                     unsigned long long timeout_ns, int* status)
{
    const uint32_t count = *waits;
    const uint32_t target = count + (uint32_t)ahead;
    if ((int)threadIdx.x < world) {
        unsigned long long t0;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        const uint32_t* f = flags + threadIdx.x;
        for (;;) {
            uint32_t v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
            if ((int32_t)(v - target) >= 0) break;
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t - t0 > timeout_ns) {
                if (status != nullptr) *status = 1;
                break;
            }
            __nanosleep(100);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0 && advance) *waits = count + 1;
}

// The push as a kernel of its own, for a caller that sends step i's rows while step i + 1 renders (on a
// second stream, under the forward's latency-bound binning kernels) instead of from the backward's
// epilogue.  `src` lives in DEVICE memory so that a captured launch can be bound later to gradient
// arrays that did not exist at capture time; src->radii == nullptr: nothing to send.  Same rows and
// the same 64-byte staging as preprocess_bwd.cu: a Gaussian that any of the `n_views` views saw
// (radii [n_views, P]) adds {dL_dmean xyz, dL_dopacity | dL_dscale xyz, views that saw it | dL_drot | dL_dcolour}.
__global__ void __launch_bounds__(256) exchange_push_kernel(const ExchangeSource* src, int P, int C, float* push_mc)
{
    __shared__ __align__(16) float s_rows[256 * EXCHANGE_ROW];
    __shared__ uint32_t s_seen[256];
    const ExchangeSource s = *src;
    if (s.radii == nullptr) return;
    const int lane = threadIdx.x & 31;
    const int warp_row0 = threadIdx.x & ~31;
    float* mine = s_rows + (size_t)threadIdx.x * EXCHANGE_ROW;
    for (size_t base = (size_t)blockIdx.x * 256; base < (size_t)P; base += (size_t)gridDim.x * 256) {
        const size_t idx = base + threadIdx.x;
        int n_seen = 0;                              // views in which this Gaussian was visible
        if (idx < (size_t)P)
            for (int v = 0; v < s.n_views; ++v) n_seen += s.radii[(size_t)v * P + idx] > 0;
        const bool seen = n_seen > 0;
        s_seen[threadIdx.x] = seen ? 1u : 0u;
        if (seen) {
            float sx, sy = 0.f, sz = 0.f, c3 = 0.f;
            if (s.scale_dim == 1) {
                sx = s.dL_dscale[idx];
            } else {
                sx = s.dL_dscale[3 * idx]; sy = s.dL_dscale[3 * idx + 1]; sz = s.dL_dscale[3 * idx + 2];
            }
            if (C == 4) c3 = s.dL_dcolor[idx * C + 3];
            reinterpret_cast<float4*>(mine)[0] = make_float4(s.dL_dmean3D[3 * idx], s.dL_dmean3D[3 * idx + 1],
                                                             s.dL_dmean3D[3 * idx + 2], s.dL_dopacity[idx]);
            reinterpret_cast<float4*>(mine)[1] = make_float4(sx, sy, sz, (float)n_seen);
            reinterpret_cast<float4*>(mine)[2] = reinterpret_cast<const float4*>(s.dL_drot)[idx];
            reinterpret_cast<float4*>(mine)[3] = make_float4(s.dL_dcolor[idx * C], s.dL_dcolor[idx * C + 1],
                                                             s.dL_dcolor[idx * C + 2], c3);
        }
        __syncwarp();
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int r = 8 * k + (lane >> 2), q = lane & 3;       // row within the warp, quarter of the row
            if (s_seen[warp_row0 + r]) {
                const float4 v = reinterpret_cast<const float4*>(s_rows + (size_t)(warp_row0 + r) * EXCHANGE_ROW)[q];
                multimem_red_add_v4(push_mc + (base + warp_row0 + r) * EXCHANGE_ROW + 4 * q, v.x, v.y, v.z, v.w);
            }
        }
        __syncwarp();
    }
}

int launch_exchange_push(const ExchangeSource* src, int P, int C, float* push_mc, int ctas, cudaStream_t stream)
{
    const int need = ceil_div(P, 256);
    exchange_push_kernel<<<ctas < need ? ctas : need, 256, 0, stream>>>(src, P, C, push_mc);
    count_launch();
    return 0;
}

int launch_exchange_arrive(uint32_t* flags_mc, int rank, cudaStream_t stream)
{
    exchange_arrive_kernel<<<1, 32, 0, stream>>>(flags_mc, rank);
    count_launch();
    return 0;
}

int launch_exchange_wait(const uint32_t* flags, int world, uint32_t* waits, int ahead, int advance,
                         unsigned long long timeout_ns, int* status, cudaStream_t stream)
{
    exchange_wait_kernel<<<1, 32 * ceil_div(world, 32), 0, stream>>>(flags, world, waits, ahead, advance, timeout_ns, status);
    count_launch();
    return 0;
}

}  // namespace sl
