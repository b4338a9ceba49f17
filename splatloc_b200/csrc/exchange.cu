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

// One thread per rank.  `waits` (local device memory) counts the waits done so far; wait number w
// (0-based) needs every rank to have arrived w times, i.e. at the steps before the one this wait is
// part of.  advance == 0: wait for everything arrived so far by this GPU's own count (`waits` arrivals
// per rank), without counting.  Spins for at most `timeout_ns`, then sets *status = 1 and gives up.
__global__ void exchange_wait_kernel(const uint32_t* flags, int world, uint32_t* waits, int advance,
                                     unsigned long long timeout_ns, int* status)
{
    const uint32_t target = *waits;
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
    if (threadIdx.x == 0 && advance) *waits = target + 1;
}

int launch_exchange_arrive(uint32_t* flags_mc, int rank, cudaStream_t stream)
{
    exchange_arrive_kernel<<<1, 32, 0, stream>>>(flags_mc, rank);
    count_launch();
    return 0;
}

int launch_exchange_wait(const uint32_t* flags, int world, uint32_t* waits, int advance, unsigned long long timeout_ns,
                         int* status, cudaStream_t stream)
{
    exchange_wait_kernel<<<1, 32 * ceil_div(world, 32), 0, stream>>>(flags, world, waits, advance, timeout_ns, status);
    count_launch();
    return 0;
}

}  // namespace sl
