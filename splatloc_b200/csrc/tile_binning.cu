// splatloc_b200/csrc/tile_binning.cu
//
// Tile-bucketed binning: the default replacement for the reference's
//   prefix sum (rasterizer_impl.cu:518) -> duplicateWithKeys (:70-132) -> global 64-bit radix
//   sort (:546-551) -> identifyTileRanges (:137-159)
// chain.  The reference sorts all R pairs globally on (tile << 32 | depth bits), ~6 digit passes
// over 24 B per pair.  But the tile of every pair is known when it is emitted, so:
//   1. preprocess counts the pairs of every tile (preprocess.cu),
//   2. tile_scan_kernel (one CTA) turns the counts into the per-tile ranges, the CTA launch order
//      (longest list first), and the totals the host needs,
//   3. scatter_pairs_kernel writes each pair straight into its tile's segment,
//   4. tile_sort_kernel sorts every segment in shared memory (bitonic network on the 64-bit
//      composite depth bits << 32 | Gaussian id) and, in the same pass, writes the sorted keys and
//      ids and gathers the PairRecords / reach masks the blend kernels stream.
// Pairs move twice instead of seven times and the sort is done by Tn independent CTAs.
// The output is the same total order (tile, depth bits, Gaussian id) the reference's stable sort
// produces -- its emission order within equal keys is ascending Gaussian id -- so point_list,
// sorted keys and ranges are bit-identical.  Segments longer than TILE_SORT_LARGE fall back to the
// global radix sort (capi.cu), which yields the same arrays.
#include "kernels.h"
#include "blend_common.cuh"

namespace sl {

#ifndef SL_SORT_MIN_CTAS
#define SL_SORT_MIN_CTAS 3        // resident 256-thread sort CTAs per SM the register budget is set for
#endif

// ------------------------------------------------------------------ 2. per-tile scan (one CTA)
// `n_tiles` counts the tiles of all views of the call (view-major numbering): one scan gives every
// tile of every view its segment of the single pair list, and one launch order for all of them.
__global__ void __launch_bounds__(1024)
tile_scan_kernel(uint32_t* tile_counts, uint2* __restrict__ ranges, uint32_t* __restrict__ order,
                 uint32_t* __restrict__ large_tiles, BinInfo* __restrict__ info, int n_tiles,
                 uint32_t max_pairs, uint32_t max_tile_pairs, int* status)
{
    chain_wait();        // launched with launch_chained(): nothing before the previous stage is complete
    chain_release();
    constexpr int NT = 1024;
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_hist[NT];
    __shared__ uint32_t s_carry, s_max, s_large, s_huge;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid == 0) { s_carry = 0; s_max = 0; s_large = 0; s_huge = 0; }
    s_hist[tid] = 0;
    __syncthreads();

    // exclusive scan over the tiles, 1024 at a time; counters become the scatter cursors (zeroed)
    uint32_t mx = 0, n_huge = 0;
    for (int base = 0; base < n_tiles; base += NT) {
        const int t = base + tid;
        const uint32_t c = (t < n_tiles) ? tile_counts[(size_t)t * TILE_COUNTER_STRIDE] : 0u;
        uint32_t incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        uint32_t off = s_carry;
        for (int w = 0; w < warp; ++w) off += s_warp[w];
        const uint32_t start = off + incl - c;
        if (t < n_tiles) {
            ranges[t] = c ? make_uint2(start, start + c) : make_uint2(0u, 0u);   // empty tiles stay (0, 0)
            tile_counts[(size_t)t * TILE_COUNTER_STRIDE] = 0u;
        }
        mx = max(mx, c);
        if (c > (uint32_t)TILE_SORT_SMALL) large_tiles[atomicAdd(&s_large, 1u)] = t;   // segment-sort work list
        n_huge += c > (uint32_t)TILE_SORT_LARGE;
        __syncthreads();
        if (tid == NT - 1) s_carry = off + incl;
        __syncthreads();
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        n_huge += __shfl_xor_sync(0xffffffffu, n_huge, o);
    }
    if (lane == 0) {
        atomicMax(&s_max, mx);
        atomicAdd(&s_huge, n_huge);
    }
    __syncthreads();
    // No-readback mode (sl_forward_nosync): the host sized the binning chunk without knowing R.  If
    // the frame does not fit, every later stage of this frame does nothing and the caller is told.
    const bool overflow = s_carry > max_pairs || s_max > max_tile_pairs;
    if (tid == 0) {
        info->num_pairs = s_carry;
        info->n_large = s_large;
        info->n_huge = s_huge;
        info->max_pairs = s_max;
        info->overflow = overflow ? 1u : 0u;
        if (status != nullptr) *status = overflow ? 1 : 0;
    }

    // launch order: counting sort of the tiles by list length, longest first
    const uint32_t width = s_max / NT + 1;
    for (int t = tid; t < n_tiles; t += NT) {
        const uint32_t len = ranges[t].y - ranges[t].x;
        atomicAdd(&s_hist[NT - 1 - len / width], 1u);
    }
    __syncthreads();
    const uint32_t c = s_hist[tid];
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t off = 0;
    for (int w = 0; w < warp; ++w) off += s_warp[w];
    __syncthreads();
    s_hist[tid] = off + incl - c;
    __syncthreads();
    for (int t = tid; t < n_tiles; t += NT) {
        const uint32_t len = ranges[t].y - ranges[t].x;
        order[atomicAdd(&s_hist[NT - 1 - len / width], 1u)] = t;
    }
    if (overflow) {
        __syncthreads();
        for (int t = tid; t < n_tiles; t += NT) ranges[t] = make_uint2(0u, 0u);   // nothing is drawn
    }
}

// ------------------------------------------------------------------ 3. scatter pairs into tile segments
// A Gaussian that covers a handful of tiles is emitted by its own thread; one that covers many
// (a near or very flat splat can cover thousands) is emitted by its whole warp, tile k to lane
// k mod 32, so that no single thread serialises thousands of returning atomics.
__global__ void __launch_bounds__(256)
scatter_pairs_kernel(int P, const float2* __restrict__ means2D, const float* __restrict__ depths,
                     const uint32_t* __restrict__ tiles_touched, const int* __restrict__ radii,
                     const uint8_t* __restrict__ tile_mask, int grid_x, int grid_y, int tiles_per_view,
                     const uint2* __restrict__ ranges, uint32_t* cursors,
                     uint64_t* __restrict__ keys, uint32_t* __restrict__ vals, const BinInfo* __restrict__ info)
{
    chain_wait();        // launched with launch_chained(): nothing before the previous stage is complete
    chain_release();
    constexpr uint32_t SOLO_MAX = 8;
    if (info->overflow) return;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const size_t voff = (size_t)blockIdx.y * P;                    // blockIdx.y: the view
    const uint32_t tile0 = blockIdx.y * (uint32_t)tiles_per_view;  // its first global tile
    uint2 rmin = make_uint2(0, 0), rmax = make_uint2(0, 0);
    uint32_t depth_bits = 0, area = 0;
    if (idx < P) {
        const int r = radii[voff + idx];
        if (r > 0 && tiles_touched[voff + idx] != 0) {
            const float2 xy = means2D[voff + idx];
            tile_rect(xy.x, xy.y, r, grid_x, grid_y, rmin, rmax);
            depth_bits = __float_as_uint(depths[voff + idx]);
            area = (rmax.x - rmin.x) * (rmax.y - rmin.y);
        }
    }
    auto emit = [&](uint32_t t, uint32_t dbits, uint32_t id) {
        if (tile_mask != nullptr && !tile_mask[t]) return;
        const uint32_t gt = tile0 + t;
        const uint32_t slot = ranges[gt].x + atomicAdd(&cursors[(size_t)gt * TILE_COUNTER_STRIDE], 1u);
        keys[slot] = ((uint64_t)gt << 32) | dbits;
        vals[slot] = id;
    };
    if (area > 0 && area <= SOLO_MAX) {
        for (uint32_t y = rmin.y; y < rmax.y; y++)
            for (uint32_t x = rmin.x; x < rmax.x; x++) emit(y * grid_x + x, depth_bits, (uint32_t)idx);
    }
    unsigned big = __ballot_sync(0xffffffffu, area > SOLO_MAX);
    while (big) {
        const int src = __ffs(big) - 1;
        big &= big - 1;
        const uint32_t x0 = __shfl_sync(0xffffffffu, rmin.x, src), y0 = __shfl_sync(0xffffffffu, rmin.y, src);
        const uint32_t w = __shfl_sync(0xffffffffu, rmax.x, src) - x0;
        const uint32_t n = __shfl_sync(0xffffffffu, area, src);
        const uint32_t dbits = __shfl_sync(0xffffffffu, depth_bits, src);
        const uint32_t id = (uint32_t)(idx - lane + src);
        for (uint32_t k = lane; k < n; k += 32) emit((y0 + k / w) * grid_x + x0 + k % w, dbits, id);
    }
}

// ------------------------------------------------------------------ 4. per-tile sort + record gather
// Bitonic network over N = E * THREADS keys held E per thread; key (warp w, slot r, lane l) is
// element w * 32E + 32 r + l.  Compare-exchange distances below 32 are warp shuffles, distances
// 32 .. 16E stay inside a thread's registers, and only the few distances that cross warps go
// through shared memory (6 of the 55 steps at N = 1024) -- the rest of the network needs no barrier.
// take_min ? min(a, b) : max(a, b) for distinct 64-bit keys: one compare, one predicate xor, one
// select (the C++ form compiles to both < and > compares and two predicated selects per half).
__device__ __forceinline__ unsigned long long pick64(unsigned long long a, unsigned long long b, int take_max)
{
    unsigned long long r;
    asm("{\n\t.reg .pred p, q;\n\t"
        "setp.lt.u64 p, %1, %2;\n\t"
        "setp.ne.s32 q, %3, 0;\n\t"
        "xor.pred p, p, q;\n\t"
        "selp.b64 %0, %1, %2, p;\n\t}"
        : "=l"(r) : "l"(a), "l"(b), "r"(take_max));
    return r;
}

// (x, y) -> (min, max), or (max, min) when `descending`: one compare for both selects.
__device__ __forceinline__ void order64(unsigned long long& x, unsigned long long& y, int descending)
{
    unsigned long long lo, hi;
    asm("{\n\t.reg .pred p, q;\n\t"
        "setp.gt.u64 p, %2, %3;\n\t"
        "setp.ne.s32 q, %4, 0;\n\t"
        "xor.pred p, p, q;\n\t"
        "selp.b64 %0, %3, %2, p;\n\t"
        "selp.b64 %1, %2, %3, p;\n\t}"
        : "=&l"(lo), "=l"(hi) : "l"(x), "l"(y), "r"(descending));
    x = lo;
    y = hi;
}

template <int E, int THREADS>
__device__ __forceinline__ void cta_bitonic_sort(unsigned long long (&key)[E], unsigned long long* smem, int tid)
{
    constexpr int N = E * THREADS;
    const int lane = tid & 31;
    const int wbase = (tid >> 5) * 32 * E + lane;
#pragma unroll 1
    for (int k = 2; k <= N; k <<= 1) {
#pragma unroll 1
        for (int j = k >> 1; j > 0; j >>= 1) {
            if (j >= 32 * E) {
                __syncthreads();
#pragma unroll
                for (int r = 0; r < E; ++r) smem[wbase + 32 * r] = key[r];
                __syncthreads();
#pragma unroll
                for (int r = 0; r < E; ++r) {
                    const int g = wbase + 32 * r;
                    const unsigned long long other = smem[g ^ j];
                    key[r] = pick64(key[r], other, ((g & j) == 0) != ((g & k) == 0));
                }
            } else if (j >= 32) {
#pragma unroll
                for (int dr = E / 2; dr >= 1; dr >>= 1) {
                    if (j == 32 * dr) {
#pragma unroll
                        for (int r = 0; r < E; ++r) {
                            if ((r & dr) == 0) {
                                order64(key[r], key[r | dr], ((wbase + 32 * r) & k) != 0);
                            }
                        }
                    }
                }
            } else {
#pragma unroll
                for (int r = 0; r < E; ++r) {
                    const unsigned long long other = __shfl_xor_sync(0xffffffffu, key[r], j);
                    key[r] = pick64(key[r], other, ((lane & j) == 0) != (((wbase + 32 * r) & k) == 0));
                }
            }
        }
    }
}

// Sort one tile's segment (n <= E * THREADS pairs) and write every per-pair output.
template <int C, int E, int THREADS, bool RAW>
__device__ __forceinline__ void sort_and_emit_tile(
    uint32_t tile, uint32_t local_tile, uint2 range, int grid_x, unsigned long long* smem,
    const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
    uint64_t* __restrict__ keys_out, uint32_t* __restrict__ point_list,
    const float2* __restrict__ means2D, const float* __restrict__ depths,
    const float4* __restrict__ conic_opacity, const float* __restrict__ features, const float* __restrict__ kp,
    PairRecord* __restrict__ records, uint8_t* __restrict__ masks)
{
    const int tid = threadIdx.x, lane = tid & 31;
    const uint32_t n = range.y - range.x;
    const uint32_t wbase = (uint32_t)(tid >> 5) * 32 * E + lane;
    unsigned long long key[E];
#pragma unroll
    for (int r = 0; r < E; ++r) {
        const uint32_t g = wbase + 32 * r;
        key[r] = (g < n) ? ((keys_in[range.x + g] << 32) | (unsigned long long)vals_in[range.x + g]) : ~0ull;
    }
    cta_bitonic_sort<E, THREADS>(key, smem, tid);

    const int tile_x0 = (local_tile % grid_x) * SL_TILE_X, tile_y0 = (local_tile / grid_x) * SL_TILE_Y;
#pragma unroll
    for (int r = 0; r < E; ++r) {
        const uint32_t g = wbase + 32 * r;
        if (g < n) {
            const unsigned long long k = key[r];
            const uint32_t id = (uint32_t)k;
            const uint32_t slot = range.x + g;
            point_list[slot] = id;
            keys_out[slot] = ((unsigned long long)tile << 32) | (k >> 32);
            const float2 xy = means2D[id];
            const float4 co = conic_opacity[id];
            const float4 f = gather_features<C>(features, kp, RAW, id);
            PairRecord rec;
            rec.xyd = make_float4(xy.x, xy.y, depths[id], __int_as_float((int)id));
            rec.co = co;
            rec.col = f;
            records[slot] = rec;
            masks[slot] = (uint8_t)patch_mask(xy.x, xy.y, co, tile_x0, tile_y0);
        }
    }
}

// CTAs stride over the work list `tiles` (global, view-major tile numbers); a CTA handles segments
// with min_len < n <= 16 * THREADS and picks the network size from n.  The per-Gaussian arrays are
// those of the tile's view.
template <int C, int THREADS, bool RAW>
__global__ void __launch_bounds__(THREADS, THREADS == 256 ? SL_SORT_MIN_CTAS : 1)
tile_sort_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ tiles, int n_listed, int grid_x,
                 int tiles_per_view, int P, uint32_t min_len,
                 const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                 uint64_t* __restrict__ keys_out, uint32_t* __restrict__ point_list,
                 const float2* __restrict__ means2D, const float* __restrict__ depths,
                 const float4* __restrict__ conic_opacity, const float* __restrict__ features,
                 size_t feature_view_stride, const float* __restrict__ kp,
                 PairRecord* __restrict__ records, uint8_t* __restrict__ masks, const BinInfo* __restrict__ dev_info)
{
    chain_wait();        // launched with launch_chained(): nothing before the previous stage is complete
    chain_release();
    extern __shared__ __align__(16) unsigned long long s_key[];
    if (dev_info->overflow) return;
    // n_listed < 0: the host does not know how long the work list is (no read-back): the device does.
    const uint32_t listed = n_listed < 0 ? dev_info->n_large : (uint32_t)n_listed;
    // The long-segment launch (1024 threads) is a few resident CTAs striding over its list; the
    // 256-thread launch has one CTA per tile (a single trip: the loop would cost it registers).
    for (uint32_t b = blockIdx.x; b < listed; b += (THREADS == 256 ? listed : gridDim.x)) {
        const uint32_t tile = tiles[b];
        const uint2 range = ranges[tile];
        const uint32_t n = range.y - range.x;
        if (n <= min_len || n > 16u * THREADS) continue;
        const uint32_t view = tile / (uint32_t)tiles_per_view, local_tile = tile - view * (uint32_t)tiles_per_view;
        const size_t voff = (size_t)view * P;
        const float* feat = features + view * feature_view_stride;
#define SL_SORT_CASE(EE)                                                                                         \
    sort_and_emit_tile<C, EE, THREADS, RAW>(tile, local_tile, range, grid_x, s_key, keys_in, vals_in, keys_out, \
                                       point_list, means2D + voff, depths + voff, conic_opacity + voff, feat, kp, \
                                       records, masks)
        if (n <= 1u * THREADS) SL_SORT_CASE(1);
        else if (n <= 2u * THREADS) SL_SORT_CASE(2);
        else if (n <= 4u * THREADS) SL_SORT_CASE(4);
        else if (n <= 8u * THREADS) SL_SORT_CASE(8);
        else SL_SORT_CASE(16);
#undef SL_SORT_CASE
    }
}

int launch_tile_scan(uint32_t* tile_counts, uint2* ranges, uint32_t* tile_order, uint32_t* large_tiles, BinInfo* info,
                     int n_tiles, uint32_t max_pairs, uint32_t max_tile_pairs, int* status, cudaStream_t stream)
{
    launch_chained(tile_scan_kernel, dim3(1), dim3(1024), 0, stream, tile_counts, ranges, tile_order, large_tiles, info, n_tiles, max_pairs,
                                             max_tile_pairs, status);
    count_launch();
    return 0;
}

int launch_scatter_pairs(int P, const GeomState& g, const int* radii, const uint8_t* tile_mask,
                         const ViewParams& vp, const uint2* ranges, uint32_t* cursors, uint64_t* keys,
                         uint32_t* vals, const BinInfo* dev_info, cudaStream_t stream)
{
    const dim3 grid(ceil_div(P, 256), vp.views);
    launch_chained(scatter_pairs_kernel, grid, dim3(256), 0, stream, P, g.means2D, g.depths, g.tiles_touched, radii, tile_mask,
                                                   vp.grid_x, vp.grid_y, vp.tiles, ranges, cursors, keys, vals,
                                                   dev_info);
    count_launch();
    return 0;
}

template <int C, bool RAW>
static int tile_sort_launches(int P, uint32_t longest, int n_large_launch, const uint2* ranges,
                              const uint32_t* tile_order, const uint32_t* large_tiles, const ViewParams& vp,
                              const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out,
                              uint32_t* point_list, const GeomState& g, const float* features,
                              size_t feature_view_stride, const float* kp, PairRecord* records, uint8_t* masks,
                              const BinInfo* dev_info, cudaStream_t stream)
{
    const int n_tiles = vp.tiles * vp.views;
    // Long segments (work list written by tile_scan_kernel) go to 1024-thread CTAs with 128 KB of shared
    // memory (one resident per SM) that stride over the list; the 256-thread launch covers all tiles
    // and skips those.  n_large_launch < 0: the host does not know the list's length (no read-back);
    // the CTAs read it from the device.
    if (n_large_launch != 0) {
        constexpr int T = 1024;
        const size_t smem = (size_t)TILE_SORT_LARGE * sizeof(unsigned long long);
        SL_CUDA(cudaFuncSetAttribute(tile_sort_kernel<C, T, RAW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                "tile sort smem attribute");
        const int listed = n_large_launch < 0 ? n_tiles : n_large_launch;
        launch_chained(tile_sort_kernel<C, T, RAW>, dim3(listed < 148 ? listed : 148), dim3(T), smem, stream,
            ranges, large_tiles, n_large_launch < 0 ? -1 : n_large_launch, vp.grid_x, vp.tiles, P,
            (uint32_t)TILE_SORT_SMALL, keys_in, vals_in, keys_out, point_list, g.means2D, g.depths, g.conic_opacity,
            features, feature_view_stride, kp, records, masks, dev_info);
        count_launch();
    }
    constexpr int T = 256;
    // shared memory is only touched by the cross-warp steps: size it for the longest short segment
    if (longest > (uint32_t)TILE_SORT_SMALL) longest = (uint32_t)TILE_SORT_SMALL;
    uint32_t pad = T;
    while (pad < longest) pad <<= 1;
    launch_chained(tile_sort_kernel<C, T, RAW>, dim3(n_tiles), dim3(T), (size_t)pad * sizeof(unsigned long long), stream,
        ranges, tile_order, n_tiles, vp.grid_x, vp.tiles, P, 0u, keys_in, vals_in, keys_out, point_list, g.means2D,
        g.depths, g.conic_opacity, features, feature_view_stride, kp, records, masks, dev_info);
    count_launch();
    return 0;
}

int launch_tile_sort(int P, int C, uint32_t longest, int n_large_launch, const uint2* ranges,
                     const uint32_t* tile_order, const uint32_t* large_tiles, const ViewParams& vp,
                     const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out,
                     uint32_t* point_list, const GeomState& g, const float* features, size_t feature_view_stride,
                     const RawParams* rawp, PairRecord* records, uint8_t* masks, const BinInfo* dev_info,
                     cudaStream_t stream)
{
    const float* kp = rawp != nullptr ? rawp->kp_score : nullptr;
#define SL_SORT_LAUNCH(CC, RR)                                                                                      \
    return tile_sort_launches<CC, RR>(P, longest, n_large_launch, ranges, tile_order, large_tiles, vp, keys_in, vals_in, \
                                      keys_out, point_list, g, features, feature_view_stride, kp, records, masks,    \
                                      dev_info, stream)
    if (rawp != nullptr) {
        if (C == 4) SL_SORT_LAUNCH(4, true);
        SL_SORT_LAUNCH(3, true);
    }
    if (C == 4) SL_SORT_LAUNCH(4, false);
    SL_SORT_LAUNCH(3, false);
#undef SL_SORT_LAUNCH
}

}  // namespace sl
