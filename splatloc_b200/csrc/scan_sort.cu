// splatloc_b200/csrc/scan_sort.cu
//
// K2  inclusive prefix sum of tiles_touched (replaces cub::DeviceScan::InclusiveSum,
//     rasterizer_impl.cu:518): one pass, chained with decoupled look-back.
// K4  stable least-significant-digit radix sort of (u64 key, u32 value) pairs on
//     the low `end_bit` bits (replaces cub::DeviceRadixSort::SortPairs,
//     rasterizer_impl.cu:546-551).  8-bit digits; one histogram kernel for all
//     digit places, then one kernel per digit place that ranks a 4096-key tile
//     in shared memory, resolves its global offsets by decoupled look-back and
//     writes the tile out digit-run by digit-run (onesweep structure).
//     Equal keys keep their input order, so the result is the total order
//     (tile, depth bits, emission order) the reference obtains.
//
// Both kernels are HBM-bound streaming passes: the scan moves 8 B per item, a
// sort pass 24 B per pair.
#include "kernels.h"

namespace sl {

namespace {

constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p)
{
    return *reinterpret_cast<const volatile unsigned long long*>(p);
}
__device__ __forceinline__ void st_volatile_u64(unsigned long long* p, unsigned long long v)
{
    *reinterpret_cast<volatile unsigned long long*>(p) = v;
}
__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* p)
{
    return *reinterpret_cast<const volatile uint32_t*>(p);
}
__device__ __forceinline__ void st_volatile_u32(uint32_t* p, uint32_t v)
{
    *reinterpret_cast<volatile uint32_t*>(p) = v;
}

__device__ __forceinline__ uint32_t warp_inclusive_scan(uint32_t v, int lane)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(FULL, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

// ------------------------------------------------------------------ K2 scan
constexpr unsigned long long SCAN_AGG = 1ull << 32;
constexpr unsigned long long SCAN_INCL = 2ull << 32;

// KEYS_HI: the addends are the high 32-bit words of an array of u64 instead of a plain u32 array.
template <bool KEYS_HI>
__global__ void __launch_bounds__(256)
inclusive_scan_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int n,
                      unsigned long long* desc, uint32_t* ticket)
{
    __shared__ uint32_t s_bid;
    __shared__ uint32_t s_warp[8];
    __shared__ uint32_t s_prefix;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_bid = atomicAdd(ticket, 1u);
    __syncthreads();
    const int bid = s_bid;
    const int base = bid * SCAN_ITEMS_PER_BLOCK + tid * 16;

    uint32_t v[16];
    if (KEYS_HI) {
        const uint2* src = reinterpret_cast<const uint2*>(in);
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] = (base + k < n) ? src[base + k].y : 0u;
    } else if (base + 16 <= n && (reinterpret_cast<uintptr_t>(in) & 15) == 0) {
        const uint4* src = reinterpret_cast<const uint4*>(in + base);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            uint4 q = src[k];
            v[4 * k] = q.x; v[4 * k + 1] = q.y; v[4 * k + 2] = q.z; v[4 * k + 3] = q.w;
        }
    } else {
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] = (base + k < n) ? in[base + k] : 0u;
    }
#pragma unroll
    for (int k = 1; k < 16; ++k) v[k] += v[k - 1];

    uint32_t incl = warp_inclusive_scan(v[15], lane);
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t warp_off = 0, block_total = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
        uint32_t t = s_warp[w];
        if (w < warp) warp_off += t;
        block_total += t;
    }
    const uint32_t thread_excl = warp_off + incl - v[15];

    // Decoupled look-back, 32 predecessors per step.
    if (warp == 0) {
        if (lane == 0 && bid > 0) st_volatile_u64(&desc[bid], SCAN_AGG | block_total);
        uint32_t running = 0;
        int pred = bid - 1 - lane;
        while (true) {
            unsigned long long d = SCAN_INCL;  // virtual predecessor of block 0
            if (pred >= 0) {
                do { d = ld_volatile_u64(&desc[pred]); } while ((d >> 32) == 0);
            }
            unsigned incl_mask = __ballot_sync(FULL, (d >> 32) == 2);
            int first = incl_mask ? (__ffs(incl_mask) - 1) : 31;
            uint32_t c = (lane <= first) ? (uint32_t)d : 0u;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(FULL, c, o);
            running += c;
            if (incl_mask) break;
            pred -= 32;
        }
        if (lane == 0) {
            st_volatile_u64(&desc[bid], SCAN_INCL | (unsigned long long)(running + block_total));
            s_prefix = running;
        }
    }
    __syncthreads();
    const uint32_t off = s_prefix + thread_excl;
    if (base + 16 <= n && (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
        uint4* dst = reinterpret_cast<uint4*>(out + base);
#pragma unroll
        for (int k = 0; k < 4; ++k)
            dst[k] = make_uint4(v[4 * k] + off, v[4 * k + 1] + off, v[4 * k + 2] + off, v[4 * k + 3] + off);
    } else {
#pragma unroll
        for (int k = 0; k < 16; ++k)
            if (base + k < n) out[base + k] = v[k] + off;
    }
}

// ------------------------------------------------------------------ K4 sort
constexpr uint32_t LB_AGG = 1u << 30;
constexpr uint32_t LB_INCL = 2u << 30;
constexpr uint32_t LB_FLAGS = 3u << 30;
constexpr uint32_t LB_VALUE = ~LB_FLAGS;

// Histogram of every digit place in one read of the keys.
__global__ void __launch_bounds__(256)
sort_histogram_kernel(const uint64_t* __restrict__ keys, int n, int passes, int aggregate_from,
                      uint32_t* __restrict__ hist)
{
    __shared__ uint32_t s_hist[SORT_MAX_PASSES * SORT_RADIX];
    for (int i = threadIdx.x; i < passes * SORT_RADIX; i += blockDim.x) s_hist[i] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (long long)(n + 31) / 32 * 32;
         i += (long long)gridDim.x * blockDim.x) {
        const bool valid = i < n;
        const uint64_t k = valid ? keys[i] : 0;
        const unsigned act = __ballot_sync(FULL, valid);
        if (valid) {
            for (int p = 0; p < passes; ++p) {
                const uint32_t d = (uint32_t)(k >> (p * SORT_RADIX_BITS)) & (SORT_RADIX - 1);
                if (p < aggregate_from) {
                    atomicAdd(&s_hist[p * SORT_RADIX + d], 1u);   // low (depth mantissa) digits: spread out
                } else {
                    // Upper digit places (exponent / tile-id bits) are shared by most of a warp:
                    // aggregate equal digits before touching shared memory.
                    const unsigned peers = __match_any_sync(act, d);
                    if (lane == __ffs(peers) - 1) atomicAdd(&s_hist[p * SORT_RADIX + d], (uint32_t)__popc(peers));
                }
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * SORT_RADIX; i += blockDim.x) {
        uint32_t c = s_hist[i];
        if (c) atomicAdd(&hist[i], c);
    }
}

// Turn each 256-bin histogram into exclusive offsets (one block per digit place).
__global__ void __launch_bounds__(SORT_RADIX)
sort_scan_hist_kernel(uint32_t* __restrict__ hist)
{
    __shared__ uint32_t s_warp[SORT_RADIX / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t* h = hist + blockIdx.x * SORT_RADIX;
    const uint32_t c = h[tid];
    const uint32_t incl = warp_inclusive_scan(c, lane);
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t off = 0;
    for (int w = 0; w < warp; ++w) off += s_warp[w];
    h[tid] = off + incl - c;
}

constexpr int SORT_THREADS = 256;
constexpr int SORT_WARPS = SORT_THREADS / 32;

// A CTA ranks SORT_ITEMS keys per thread: 16 (4096 per CTA) for large inputs, 4 (1024 per CTA) for
// small ones, where a pass is bounded by the latency of one CTA rather than by bandwidth.
template <int SORT_ITEMS>
struct SortSmem {
    uint64_t keys[SORT_ITEMS * SORT_THREADS];
    uint32_t vals[SORT_ITEMS * SORT_THREADS];
    uint32_t warp_hist[SORT_WARPS][SORT_RADIX];
    uint32_t bin_excl[SORT_RADIX];
    uint32_t bin_gbase[SORT_RADIX];
    uint32_t warp_tot[SORT_WARPS];
    uint32_t bid;
};

template <int SORT_ITEMS>
__global__ void __launch_bounds__(SORT_THREADS)
sort_pass_kernel(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                 uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int n, int shift,
                 const uint32_t* __restrict__ digit_base, uint32_t* lookback, uint32_t* ticket)
{
    constexpr int SORT_ITEMS_PER_BLOCK = SORT_ITEMS * SORT_THREADS;
    constexpr int SORT_WARP_ITEMS = SORT_ITEMS * 32;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SortSmem<SORT_ITEMS>& s = *reinterpret_cast<SortSmem<SORT_ITEMS>*>(smem_raw);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s.bid = atomicAdd(ticket, 1u);
#pragma unroll
    for (int w = 0; w < SORT_WARPS; ++w) s.warp_hist[w][tid] = 0;
    __syncthreads();
    const int bid = s.bid;
    const int base = bid * SORT_ITEMS_PER_BLOCK;
    const int n_valid = min(SORT_ITEMS_PER_BLOCK, n - base);
    const int wbase = warp * SORT_WARP_ITEMS + lane;

    // Warp-striped load: item i of lane l is element wbase + 32 i (input order =
    // (warp, i, lane), which the ranking below preserves for equal digits).
    uint64_t key[SORT_ITEMS];
    uint32_t rank[SORT_ITEMS];
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const int pos = wbase + 32 * i;
        key[i] = (pos < n_valid) ? keys_in[base + pos] : ~0ull;
    }

    const unsigned lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const bool valid = (wbase + 32 * i) < n_valid;
        const unsigned act = __ballot_sync(FULL, valid);
        if (valid) {
            const uint32_t d = (uint32_t)(key[i] >> shift) & (SORT_RADIX - 1);
            const unsigned peers = __match_any_sync(act, d);
            const int leader = __ffs(peers) - 1;
            uint32_t prev = 0;
            if (lane == leader) {
                prev = s.warp_hist[warp][d];
                s.warp_hist[warp][d] = prev + __popc(peers);
            }
            prev = __shfl_sync(peers, prev, leader);
            rank[i] = prev + __popc(peers & lt_mask);
        }
        __syncwarp();
    }
    __syncthreads();

    // Per-digit: exclusive offsets of the warps inside the tile, tile count.
    uint32_t count = 0;
#pragma unroll
    for (int w = 0; w < SORT_WARPS; ++w) {
        const uint32_t t = s.warp_hist[w][tid];
        s.warp_hist[w][tid] = count;
        count += t;
    }

    // Decoupled look-back for digit `tid`: sum of the counts of all earlier tiles.
    uint32_t prefix = 0;
    uint32_t* mine = lookback + (size_t)bid * SORT_RADIX + tid;
    if (bid == 0) {
        st_volatile_u32(mine, LB_INCL | count);
    } else {
        st_volatile_u32(mine, LB_AGG | count);
        // Walk back over the predecessors LB_WINDOW at a time: the loads of one window are
        // independent, so a window costs one L2 round trip instead of LB_WINDOW of them.
        constexpr int LB_WINDOW = 8;
        int p = bid - 1;
        bool found = false;
        while (!found) {
            uint32_t v[LB_WINDOW];
#pragma unroll
            for (int k = 0; k < LB_WINDOW; ++k)
                v[k] = (p - k >= 0) ? ld_volatile_u32(lookback + (size_t)(p - k) * SORT_RADIX + tid) : LB_INCL;
#pragma unroll
            for (int k = 0; k < LB_WINDOW; ++k) {
                if (!found) {
                    uint32_t x = v[k];
                    while ((x & LB_FLAGS) == 0) x = ld_volatile_u32(lookback + (size_t)(p - k) * SORT_RADIX + tid);
                    prefix += x & LB_VALUE;
                    found = (x & LB_INCL) != 0;
                }
            }
            p -= LB_WINDOW;
        }
        st_volatile_u32(mine, LB_INCL | (prefix + count));
    }

    // Exclusive scan of the tile's digit counts (position of each digit run in the tile).
    const uint32_t incl = warp_inclusive_scan(count, lane);
    if (lane == 31) s.warp_tot[warp] = incl;
    __syncthreads();
    uint32_t woff = 0;
#pragma unroll
    for (int w = 0; w < SORT_WARPS; ++w)
        if (w < warp) woff += s.warp_tot[w];
    const uint32_t excl = woff + incl - count;
    s.bin_excl[tid] = excl;
    s.bin_gbase[tid] = digit_base[tid] + prefix - excl;
    __syncthreads();

    // Reorder the tile in shared memory, then stream it out: consecutive threads
    // write consecutive addresses inside each digit run.
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const int pos = wbase + 32 * i;
        if (pos < n_valid) {
            const uint32_t d = (uint32_t)(key[i] >> shift) & (SORT_RADIX - 1);
            const uint32_t dst = s.bin_excl[d] + s.warp_hist[warp][d] + rank[i];
            s.keys[dst] = key[i];
            s.vals[dst] = vals_in[base + pos];
        }
    }
    __syncthreads();
#pragma unroll 4
    for (int p = tid; p < n_valid; p += SORT_THREADS) {
        const uint64_t k = s.keys[p];
        const uint32_t d = (uint32_t)(k >> shift) & (SORT_RADIX - 1);
        const uint32_t g = s.bin_gbase[d] + p;
        keys_out[g] = k;
        vals_out[g] = s.vals[p];
    }
}

}  // namespace

int launch_inclusive_scan(const uint32_t* in, uint32_t* out, int n, unsigned long long* desc,
                          uint32_t* ticket, cudaStream_t stream)
{
    const int blocks = ceil_div(n, SCAN_ITEMS_PER_BLOCK);
    SL_CUDA(cudaMemsetAsync(desc, 0, (size_t)(blocks + 1) * sizeof(unsigned long long), stream), "scan memset");
    SL_CUDA(cudaMemsetAsync(ticket, 0, sizeof(uint32_t), stream), "scan memset");
    inclusive_scan_kernel<false><<<blocks, 256, 0, stream>>>(in, out, n, desc, ticket); count_launch();
    return 0;
}

size_t SortScratch::bytes(size_t n)
{
    char* p = nullptr;
    SortScratch::from_chunk(p, n);
    return reinterpret_cast<size_t>(p) + 128;
}

template <int ITEMS>
static int sort_passes(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                       SortScratch& sc, int n, int passes, cudaStream_t stream)
{
    SL_CUDA(cudaFuncSetAttribute(sort_pass_kernel<ITEMS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)sizeof(SortSmem<ITEMS>)), "sort smem attribute");
    const int blocks = ceil_div(n, ITEMS * SORT_THREADS);
    const uint64_t* kin = keys_in;
    const uint32_t* vin = vals_in;
    for (int p = 0; p < passes; ++p) {
        // The last digit place must land in the caller's output arrays; alternate with
        // the temporaries backwards from there so the unsorted inputs survive.
        const bool to_final = ((passes - 1 - p) % 2) == 0;
        uint64_t* kout = to_final ? keys_out : sc.keys_tmp;
        uint32_t* vout = to_final ? vals_out : sc.vals_tmp;
        sort_pass_kernel<ITEMS><<<blocks, SORT_THREADS, sizeof(SortSmem<ITEMS>), stream>>>(
            kin, vin, kout, vout, n, p * SORT_RADIX_BITS, sc.histograms + p * SORT_RADIX,
            sc.lookback + (size_t)p * blocks * SORT_RADIX, sc.tickets + p); count_launch();
        kin = kout;
        vin = vout;
    }
    return 0;
}

int sort_prepare(SortScratch& sc, int n, int end_bit, cudaStream_t stream)
{
    const int passes = (end_bit + SORT_RADIX_BITS - 1) / SORT_RADIX_BITS;
    if (passes < 1 || passes > SORT_MAX_PASSES) {
        set_error("sort: %d key bits need 1..%d digit places", end_bit, SORT_MAX_PASSES);
        return -3;
    }
    // histograms | tickets | look-back words of the digit places in use
    const int blocks = ceil_div(n, sort_items_per_block(n));
    const size_t bytes = (size_t)(reinterpret_cast<char*>(sc.lookback) - reinterpret_cast<char*>(sc.histograms)) +
                         (size_t)passes * blocks * SORT_RADIX * sizeof(uint32_t);
    SL_CUDA(cudaMemsetAsync(sc.histograms, 0, bytes, stream), "sort memset");
    return 0;
}

int launch_sort_pairs(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out,
                      uint32_t* vals_out, SortScratch& sc, int n, int end_bit, cudaStream_t stream)
{
    if (n <= 0) return 0;
    const int passes = (end_bit + SORT_RADIX_BITS - 1) / SORT_RADIX_BITS;
    const int rc = sort_prepare(sc, n, end_bit, stream);
    if (rc) return rc;
    const int hist_blocks = min(ceil_div(n, 256 * 8), 148 * 8);
    sort_histogram_kernel<<<hist_blocks, 256, 0, stream>>>(keys_in, n, passes, /*aggregate_from=*/3, sc.histograms); count_launch();
    sort_scan_hist_kernel<<<passes, SORT_RADIX, 0, stream>>>(sc.histograms); count_launch();
    if (sort_items_per_block(n) == SORT_ITEMS_PER_BLOCK)
        return sort_passes<SORT_ITEMS_PER_BLOCK / SORT_THREADS>(keys_in, vals_in, keys_out, vals_out, sc, n, passes, stream);
    return sort_passes<SORT_SMALL_ITEMS_PER_BLOCK / SORT_THREADS>(keys_in, vals_in, keys_out, vals_out, sc, n, passes, stream);
}

}  // namespace sl
