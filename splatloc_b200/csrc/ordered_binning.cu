// splatloc_b200/csrc/ordered_binning.cu
//
// Depth-ordered emit: the default replacement for the reference's
//   prefix sum (rasterizer_impl.cu:518) -> duplicateWithKeys (:70-132) -> global 64-bit radix
//   sort (:546-551) -> identifyTileRanges (:137-159)
// chain.  The reference sorts all R (tile, Gaussian) pairs on (tile << 32 | depth bits); a Gaussian
// that covers 6 tiles is therefore sorted 6 times.  But inside every tile the order is simply the
// order of the Gaussians by (depth bits, id), so it is enough to sort the V visible Gaussians ONCE
// and to write each pair straight to its final position:
//
//   position of pair (g, t) = ranges[t].x + #{ g' before g in depth order : g' touches t }.
//
//   1. preprocess counts the pairs of every tile and the visible Gaussians of every CTA
//      (preprocess.cu); tile_scan_kernel turns both into offsets (tile_binning.cu),
//   2. compact_visible_kernel: stable compaction of (depth bits, id) of the visible Gaussians,
//   3. stable LSD radix sort of those V entries on the 32 depth bits (scan_sort.cu); equal depths
//      keep ascending ids, which is the reference's tie order (its emission order),
//   4. tile_hist_kernel: the sorted Gaussians are cut into blocks of `block` (one CTA each, one warp
//      per 32 Gaussians).  A warp marks in a shared (warp, tile) bit mask which of its 32 Gaussians
//      touch which tile; the popcounts give block_hist[block][tile],
//   5. column_scan_kernel: exclusive scan of block_hist down every tile column, started at
//      ranges[t].x -> block_base[block][tile],
//   6. ordered_emit_kernel rebuilds the masks; the position of pair (lane k of warp w, tile t) is
//      block_base[block][t] + (pairs of warps < w in t) + popc(mask[w][t] & bits below k).  The pair's
//      sorted key, id, PairRecord and reach mask are written there -- once, in final order.
//
// The result equals the stable global sort on (tile | depth): point_list, sorted keys and ranges
// are bit-identical to the reference's.  The "unsorted" arrays of the binning chunk, whose order the
// reference leaves to its emission loop, receive the same pairs in list order.
#include "kernels.h"
#include "blend_common.cuh"

namespace sl {

namespace {

constexpr unsigned FULL = 0xffffffffu;

// ------------------------------------------------------------------ 2. stable compaction
__global__ void __launch_bounds__(PREPROCESS_THREADS)
compact_visible_kernel(int P, const uint32_t* __restrict__ tiles_touched, const float* __restrict__ depths,
                       const uint32_t* __restrict__ block_off, uint64_t* __restrict__ keys,
                       uint32_t* __restrict__ ids)
{
    __shared__ uint32_t s_warp[PREPROCESS_THREADS / 32];
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool vis = idx < P && tiles_touched[idx] != 0;
    const unsigned b = __ballot_sync(FULL, vis);
    if (lane == 0) s_warp[warp] = __popc(b);
    __syncthreads();
    if (!vis) return;
    uint32_t dst = block_off[blockIdx.x] + __popc(b & ((1u << lane) - 1u));
    for (int w = 0; w < warp; ++w) dst += s_warp[w];
    keys[dst] = (uint64_t)__float_as_uint(depths[idx]);
    ids[dst] = (uint32_t)idx;
}

// Lane k of the warp holds the tile rectangle of the warp's k-th Gaussian (empty for none) and a
// payload.  Calls fn(k, tile, payload of k) once for every (Gaussian, tile) pair of the warp, with
// the pairs dealt to the lanes 32 at a time.
template <typename F>
__device__ __forceinline__ void for_each_pair(uint2 rmin, uint2 rmax, uint32_t payload, int grid_x, int lane, F fn)
{
    const uint32_t w = rmax.x - rmin.x;
    const uint32_t area = w * (rmax.y - rmin.y);
    uint32_t incl = area;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(FULL, incl, o);
        if (lane >= o) incl += v;
    }
    const uint32_t excl = incl - area;
    const uint32_t total = __shfl_sync(FULL, incl, 31);
    for (uint32_t p0 = 0; p0 < total; p0 += 32) {
        const uint32_t p = p0 + lane;
        int k = 0;                        // owner of pair p: the number of lanes whose inclusive sum is <= p
#pragma unroll
        for (int step = 16; step > 0; step >>= 1)
            if (__shfl_sync(FULL, incl, k + step - 1) <= p) k += step;
        const uint32_t q = p - __shfl_sync(FULL, excl, k);
        const uint32_t wk = __shfl_sync(FULL, w, k);
        const uint32_t x0 = __shfl_sync(FULL, rmin.x, k), y0 = __shfl_sync(FULL, rmin.y, k);
        const uint32_t pl = __shfl_sync(FULL, payload, k);
        if (p < total) {
            const uint32_t qy = q / wk;
            fn(k, (y0 + qy) * (uint32_t)grid_x + x0 + (q - qy * wk), pl);
        }
    }
}

// Rectangle of the i-th sorted Gaussian (empty past the end) and its id.
__device__ __forceinline__ void load_rect(int i, int V, const uint32_t* __restrict__ vis_ids,
                                          const float2* __restrict__ means2D, const int* __restrict__ radii,
                                          int grid_x, int grid_y, uint2& rmin, uint2& rmax, uint32_t& id)
{
    rmin = make_uint2(0, 0);
    rmax = make_uint2(0, 0);
    id = 0;
    if (i < V) {
        id = vis_ids[i];
        const float2 xy = means2D[id];
        tile_rect(xy.x, xy.y, radii[id], grid_x, grid_y, rmin, rmax);
    }
}

// mask[w][t] bit k: the k-th Gaussian of warp w touches tile t.  Ends with a CTA barrier.
__device__ __forceinline__ void build_masks(uint32_t* s_mask, int n_tiles, uint2 rmin, uint2 rmax,
                                            const uint8_t* __restrict__ tile_mask, int grid_x)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int words = (blockDim.x >> 5) * n_tiles;
    for (int i = tid; i < words; i += blockDim.x) s_mask[i] = 0u;
    __syncthreads();
    uint32_t* mine = s_mask + (size_t)warp * n_tiles;
    for_each_pair(rmin, rmax, 0u, grid_x, lane, [&](int k, uint32_t t, uint32_t) {
        if (tile_mask == nullptr || tile_mask[t]) atomicOr(&mine[t], 1u << k);
    });
    __syncthreads();
}

// ------------------------------------------------------------------ 4. pairs per (block, tile)
__global__ void __launch_bounds__(256)
tile_hist_kernel(int V, const uint32_t* __restrict__ vis_ids, const float2* __restrict__ means2D,
                 const int* __restrict__ radii, const uint8_t* __restrict__ tile_mask, int grid_x, int grid_y,
                 uint16_t* __restrict__ block_hist)
{
    extern __shared__ __align__(16) uint32_t s_mask[];
    const int n_tiles = grid_x * grid_y, nwarps = blockDim.x >> 5;
    uint2 rmin, rmax;
    uint32_t id;
    load_rect(blockIdx.x * blockDim.x + threadIdx.x, V, vis_ids, means2D, radii, grid_x, grid_y, rmin, rmax, id);
    build_masks(s_mask, n_tiles, rmin, rmax, tile_mask, grid_x);
    uint16_t* row = block_hist + (size_t)blockIdx.x * n_tiles;
    for (int t = threadIdx.x; t < n_tiles; t += blockDim.x) {
        uint32_t c = 0;
        for (int w = 0; w < nwarps; ++w) c += __popc(s_mask[(size_t)w * n_tiles + t]);
        row[t] = (uint16_t)c;
    }
}

// ------------------------------------------------------------------ 5. scan down the tile columns
// CTA = 32 tile columns (lanes) x 16 row segments (warps): every warp sums its segment, the segment
// sums are scanned through shared memory, then every warp writes its rows' exclusive prefixes.
constexpr int COLSCAN_WARPS = 16;

__global__ void __launch_bounds__(COLSCAN_WARPS * 32)
column_scan_kernel(const uint16_t* __restrict__ block_hist, uint32_t* __restrict__ block_base,
                   const uint2* __restrict__ ranges, int rows, int n_tiles)
{
    __shared__ uint32_t s_seg[COLSCAN_WARPS][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int t = blockIdx.x * 32 + lane;
    const int seg = (rows + COLSCAN_WARPS - 1) / COLSCAN_WARPS;
    const int r0 = min(rows, warp * seg), r1 = min(rows, r0 + seg);
    uint32_t sum = 0;
    if (t < n_tiles) {
#pragma unroll 8
        for (int r = r0; r < r1; ++r) sum += block_hist[(size_t)r * n_tiles + t];
    }
    s_seg[warp][lane] = sum;
    __syncthreads();
    if (t >= n_tiles) return;
    uint32_t run = ranges[t].x;
    for (int w = 0; w < warp; ++w) run += s_seg[w][lane];
#pragma unroll 8
    for (int r = r0; r < r1; ++r) {
        const uint32_t c = block_hist[(size_t)r * n_tiles + t];
        block_base[(size_t)r * n_tiles + t] = run;
        run += c;
    }
}

// ------------------------------------------------------------------ 6. write every pair in place
template <int C>
__global__ void __launch_bounds__(256)
ordered_emit_kernel(int V, const uint32_t* __restrict__ vis_ids, const float2* __restrict__ means2D,
                    const float* __restrict__ depths, const float4* __restrict__ conic_opacity,
                    const int* __restrict__ radii, const float* __restrict__ features,
                    const uint8_t* __restrict__ tile_mask, int grid_x, int grid_y,
                    const uint32_t* __restrict__ block_base, uint32_t* __restrict__ point_list,
                    uint64_t* __restrict__ keys_out, uint32_t* __restrict__ point_list_unsorted,
                    uint64_t* __restrict__ keys_unsorted, PairRecord* __restrict__ records,
                    uint8_t* __restrict__ masks_out)
{
    extern __shared__ __align__(16) uint32_t s_mask[];
    const int n_tiles = grid_x * grid_y, nwarps = blockDim.x >> 5;
    uint16_t* s_before = reinterpret_cast<uint16_t*>(s_mask + (size_t)nwarps * n_tiles);   // pairs of earlier warps
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint2 rmin, rmax;
    uint32_t id;
    load_rect(blockIdx.x * blockDim.x + threadIdx.x, V, vis_ids, means2D, radii, grid_x, grid_y, rmin, rmax, id);
    build_masks(s_mask, n_tiles, rmin, rmax, tile_mask, grid_x);
    for (int t = threadIdx.x; t < n_tiles; t += blockDim.x) {
        uint32_t run = 0;
        for (int w = 0; w < nwarps; ++w) {
            s_before[(size_t)w * n_tiles + t] = (uint16_t)run;
            run += __popc(s_mask[(size_t)w * n_tiles + t]);
        }
    }
    __syncthreads();

    const uint32_t* base_row = block_base + (size_t)blockIdx.x * n_tiles;
    const uint32_t* my_mask = s_mask + (size_t)warp * n_tiles;
    const uint16_t* my_before = s_before + (size_t)warp * n_tiles;
    for_each_pair(rmin, rmax, id, grid_x, lane, [&](int k, uint32_t t, uint32_t gid) {
        if (tile_mask != nullptr && !tile_mask[t]) return;
        const uint32_t slot = base_row[t] + my_before[t] + __popc(my_mask[t] & ((1u << k) - 1u));
        const float2 xy = means2D[gid];
        const float4 co = conic_opacity[gid];
        const float depth = depths[gid];
        float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
        if (C == 4) {
            f = reinterpret_cast<const float4*>(features)[gid];
        } else {
            f.x = features[gid * C + 0];
            if (C > 1) f.y = features[gid * C + 1];
            if (C > 2) f.z = features[gid * C + 2];
        }
        const uint64_t key = ((uint64_t)t << 32) | (uint64_t)__float_as_uint(depth);
        point_list[slot] = gid;
        keys_out[slot] = key;
        point_list_unsorted[slot] = gid;
        keys_unsorted[slot] = key;
        PairRecord rec;
        rec.xyd = make_float4(xy.x, xy.y, depth, __int_as_float((int)gid));
        rec.co = co;
        rec.col = f;
        records[slot] = rec;
        const int tile_x0 = (int)(t % (uint32_t)grid_x) * SL_TILE_X, tile_y0 = (int)(t / (uint32_t)grid_x) * SL_TILE_Y;
        masks_out[slot] = (uint8_t)patch_mask(xy.x, xy.y, co, tile_x0, tile_y0);
    });
}

}  // namespace

int ordered_block_size(int n_tiles)
{
    // shared memory per CTA: (4-byte mask + 2-byte count) per (warp, tile)
    for (int block = 256; block >= 32; block >>= 1)
        if ((size_t)(block / 32) * (size_t)n_tiles * 6u <= (size_t)ORDERED_MASK_BYTES) return block;
    return 0;
}

int launch_depth_sort(int P, int V, const GeomState& g, BinningState& bin, cudaStream_t stream)
{
    if (V <= 0) return 0;
    compact_visible_kernel<<<(int)g.pre_blocks, PREPROCESS_THREADS, 0, stream>>>(
        P, g.tiles_touched, g.depths, g.block_vis, bin.vis_keys_in, bin.vis_ids_in);
    count_launch();
    return launch_sort_pairs(bin.vis_keys_in, bin.vis_ids_in, bin.vis_keys, bin.vis_ids, bin.vis_sort, V, 32, stream);
}

int launch_ordered_emit(int V, int C, int block, const GeomState& g, const int* radii, const uint8_t* tile_mask,
                        int grid_x, int grid_y, const uint2* ranges, const float* features, BinningState& bin,
                        cudaStream_t stream)
{
    if (V <= 0) return 0;
    const int n_tiles = grid_x * grid_y;
    const int rows = (int)bin.rows;
    if (C != 3 && C != 4) {
        set_error("ordered emit: unsupported channel count %d (3 or 4)", C);
        return -2;
    }
    const size_t smem_hist = (size_t)(block / 32) * n_tiles * 4u;
    const size_t smem_emit = (size_t)(block / 32) * n_tiles * 6u;
    SL_CUDA(cudaFuncSetAttribute(tile_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_hist),
            "ordered emit smem attribute");
    tile_hist_kernel<<<rows, block, smem_hist, stream>>>(V, bin.vis_ids, g.means2D, radii, tile_mask, grid_x, grid_y,
                                                         bin.block_hist);
    count_launch();
    column_scan_kernel<<<ceil_div(n_tiles, 32), COLSCAN_WARPS * 32, 0, stream>>>(bin.block_hist, bin.block_base, ranges,
                                                                                 rows, n_tiles);
    count_launch();
#define SL_EMIT(CC)                                                                                                \
    do {                                                                                                           \
        SL_CUDA(cudaFuncSetAttribute(ordered_emit_kernel<CC>, cudaFuncAttributeMaxDynamicSharedMemorySize,         \
                                     (int)smem_emit), "ordered emit smem attribute");                            \
        ordered_emit_kernel<CC><<<rows, block, smem_emit, stream>>>(                                               \
            V, bin.vis_ids, g.means2D, g.depths, g.conic_opacity, radii, features, tile_mask, grid_x, grid_y,      \
            bin.block_base, bin.point_list, bin.point_list_keys, bin.point_list_unsorted,                          \
            bin.point_list_keys_unsorted, bin.records, bin.masks);                                                 \
    } while (0)
    if (C == 4) SL_EMIT(4); else SL_EMIT(3);
#undef SL_EMIT
    count_launch();
    return 0;
}

}  // namespace sl
