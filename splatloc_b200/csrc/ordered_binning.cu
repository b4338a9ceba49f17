// splatloc_b200/csrc/ordered_binning.cu
//
// Depth-ordered emit: a replacement for the reference's
//   prefix sum (rasterizer_impl.cu:518) -> duplicateWithKeys (:70-132) -> global 64-bit radix
//   sort (:546-551) -> identifyTileRanges (:137-159)
// chain.  The reference sorts all R (tile, Gaussian) pairs on (tile << 32 | depth bits); a Gaussian
// that covers 6 tiles is therefore sorted 6 times, on a 43-bit key.  But inside every tile the order
// is simply the order of the Gaussians by (depth bits, id), so it is enough to sort the V visible
// Gaussians ONCE on their 32 depth bits and then to run ONE stable counting-sort pass of the pairs,
// taken in that order, on their tile:
//
//   position of pair (g, t) = ranges[t].x + #{ pairs (g', t) with g' before g in depth order }.
//
//   1. preprocess counts the pairs of every tile and the visible Gaussians (those with at least one pair) of every CTA
//      (preprocess.cu); tile_scan_kernel turns both into offsets (tile_binning.cu),
//   2. compact_visible_kernel: stable compaction of (tile count << 32 | depth bits, id) of the
//      visible Gaussians, accumulating the sort's digit histograms on the way,
//   3. stable LSD radix sort of those V entries on the 32 depth bits (scan_sort.cu); equal depths
//      keep ascending ids, which is the reference's tie order (its emission order).  The tile
//      counts ride along in the keys' high words,
//   4. prefix sum of those tile counts: the pairs, Gaussian-major in depth order, get indices 0..A-1,
//   5. pair_hist_kernel: CTA c takes pairs [2048 c, 2048 (c + 1)) -- equal work per CTA however the
//      splat sizes are distributed -- and counts them per tile: block_hist[c][tile],
//   6. column_scan_kernel: exclusive scan of block_hist down every tile column, started at
//      ranges[t].x -> block_base[c][tile],
//   7. pair_emit_kernel: same pair ranges; a warp ranks its 256 pairs per tile in index order
//      (match.any + per-warp counters, the ranking step of a radix sort with the tile as digit),
//      the warps' counts are scanned per tile, and every pair's sorted key, id, PairRecord and reach
//      mask are written to block_base[c][t] + rank -- once, in final order.
//
// The result equals the stable global sort on (tile | depth): point_list, sorted keys and ranges
// are bit-identical to the reference's.  The "unsorted" arrays of the binning chunk, whose order the
// reference leaves to its emission loop, receive the same pairs in list order.
#include "kernels.h"
#include "blend_common.cuh"

namespace sl {

namespace {

constexpr unsigned FULL = 0xffffffffu;

constexpr int OE_THREADS = 256;
constexpr int OE_WARPS = OE_THREADS / 32;
constexpr int OE_ITEMS = 8;                              // pairs per thread
constexpr int OE_CTA_PAIRS = OE_THREADS * OE_ITEMS;      // 2048 pairs per CTA
constexpr int OE_WARP_PAIRS = 32 * OE_ITEMS;

// ------------------------------------------------------------------ 2. stable compaction (+ digit histograms)
// CTA-strided over the preprocess CTAs' 256-Gaussian blocks, whose output offsets tile_scan_kernel
// prepared.  The four 8-bit digit histograms of the depth keys are accumulated in shared memory
// and flushed once per CTA, which saves the radix sort its own histogram pass.
__global__ void __launch_bounds__(PREPROCESS_THREADS)
compact_visible_kernel(int P, int n_blocks, const uint32_t* __restrict__ tiles_touched,
                       const float* __restrict__ depths, const float2* __restrict__ means2D,
                       const int* __restrict__ radii, int grid_x, int grid_y, const uint32_t* __restrict__ block_off,
                       uint64_t* __restrict__ keys, uint32_t* __restrict__ ids, uint32_t* __restrict__ hist)
{
    __shared__ uint32_t s_hist[4 * SORT_RADIX];
    __shared__ uint32_t s_warp[2][PREPROCESS_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < 4 * SORT_RADIX; i += PREPROCESS_THREADS) s_hist[i] = 0u;
    __syncthreads();
    int flip = 0;
    for (int blk = blockIdx.x; blk < n_blocks; blk += gridDim.x, flip ^= 1) {
        const int idx = blk * PREPROCESS_THREADS + tid;
        const bool vis = idx < P && tiles_touched[idx] != 0;
        const unsigned b = __ballot_sync(FULL, vis);
        if (lane == 0) s_warp[flip][warp] = __popc(b);
        __syncthreads();
        if (vis) {
            uint32_t dst = block_off[blk] + __popc(b & ((1u << lane) - 1u));
            for (int w = 0; w < warp; ++w) dst += s_warp[flip][w];
            const uint32_t d = __float_as_uint(depths[idx]);
            // tiles in the rectangle (tiles_touched counts only those a tile mask keeps)
            const float2 xy = means2D[idx];
            uint2 rmin, rmax;
            tile_rect(xy.x, xy.y, radii[idx], grid_x, grid_y, rmin, rmax);
            const uint32_t area = (rmax.x - rmin.x) * (rmax.y - rmin.y);
            keys[dst] = ((uint64_t)area << 32) | d;
            ids[dst] = (uint32_t)idx;
            atomicAdd(&s_hist[d & 255u], 1u);
            atomicAdd(&s_hist[256 + ((d >> 8) & 255u)], 1u);
            // the upper digits (exponent, top of the mantissa) are shared by most of a warp:
            // aggregate equal digits before touching shared memory
#pragma unroll
            for (int p = 2; p < 4; ++p) {
                const uint32_t dg = (d >> (8 * p)) & 255u;
                const unsigned peers = __match_any_sync(b, dg);
                if (lane == __ffs(peers) - 1) atomicAdd(&s_hist[p * 256 + dg], (uint32_t)__popc(peers));
            }
        }
    }
    __syncthreads();
    for (int i = tid; i < 4 * SORT_RADIX; i += PREPROCESS_THREADS) {
        const uint32_t c = s_hist[i];
        if (c) atomicAdd(&hist[i], c);
    }
}

// ------------------------------------------------------------------ shared by 5. and 7.
// The Gaussians whose pairs fall into the CTA's pair range, staged in shared memory.
struct RangeSmem {
    uint32_t incl[OE_CTA_PAIRS];    // inclusive pair offsets of Gaussians g_first, g_first + 1, ...
    uint32_t xy[OE_CTA_PAIRS];      // tile rectangle origin: x | y << 16
    uint32_t id[OE_CTA_PAIRS];      // Gaussian id
    uint16_t w[OE_CTA_PAIRS];       // rectangle width in tiles
    uint32_t g_first, first_excl;
};

// First Gaussian whose inclusive pair offset exceeds `start` (a 32-ary search by one warp; `start`
// is below the total, so it exists).
__device__ __forceinline__ uint32_t first_gaussian_after(const uint32_t* __restrict__ pair_incl, uint32_t V,
                                                         uint32_t start, int lane)
{
    uint32_t lo = 0, hi = V;    // the answer is in [lo, hi) and pair_incl[hi - 1] > start
    while (hi - lo > 32u) {
        const uint32_t step = (hi - lo + 31u) / 32u;
        const uint32_t probe = min(hi - 1u, lo + (uint32_t)(lane + 1) * step - 1u);
        const uint32_t n = __popc(__ballot_sync(FULL, pair_incl[probe] <= start));
        const uint32_t lo_new = min(hi, lo + n * step);
        hi = min(hi, lo + (n + 1u) * step);
        lo = lo_new;
    }
    const bool le = (lo + lane < hi) && pair_incl[lo + lane] <= start;
    return lo + __popc(__ballot_sync(FULL, le));
}

__device__ __forceinline__ void stage_range(RangeSmem& rs, const uint32_t* __restrict__ pair_incl,
                                            const uint32_t* __restrict__ vis_ids,
                                            const float2* __restrict__ means2D, const int* __restrict__ radii,
                                            uint32_t V, uint32_t start, uint32_t end, int grid_x, int grid_y)
{
    const int tid = threadIdx.x;
    if (tid < 32) {
        const uint32_t g = first_gaussian_after(pair_incl, V, start, tid);
        if (tid == 0) {
            rs.g_first = g;
            rs.first_excl = g ? pair_incl[g - 1] : 0u;
        }
    }
    __syncthreads();
    const uint32_t g_first = rs.g_first;
    for (int j = tid; j < OE_CTA_PAIRS; j += OE_THREADS)
        rs.incl[j] = (g_first + j < V) ? pair_incl[g_first + j] : 0xffffffffu;
    __syncthreads();
    for (int j = tid; j < OE_CTA_PAIRS; j += OE_THREADS) {
        const uint32_t excl = j ? rs.incl[j - 1] : rs.first_excl;
        if (g_first + j < V && excl < end) {      // Gaussian j has pairs in [start, end)
            const uint32_t id = vis_ids[g_first + j];
            const float2 xy = means2D[id];
            uint2 rmin, rmax;
            tile_rect(xy.x, xy.y, radii[id], grid_x, grid_y, rmin, rmax);
            rs.xy[j] = rmin.x | (rmin.y << 16);
            rs.w[j] = (uint16_t)(rmax.x - rmin.x);
            rs.id[j] = id;
        }
    }
    __syncthreads();
}

// Pair `p` of the CTA's range: its Gaussian (index into the staged arrays) and tile.
__device__ __forceinline__ void locate_pair(const RangeSmem& rs, uint32_t p, int grid_x, int& j, uint32_t& tile)
{
    int lo = 0;                          // number of staged Gaussians whose inclusive offset is <= p
#pragma unroll
    for (int step = OE_CTA_PAIRS / 2; step > 0; step >>= 1)
        if (rs.incl[lo + step - 1] <= p) lo += step;
    j = lo;
    const uint32_t q = p - (lo ? rs.incl[lo - 1] : rs.first_excl);
    const uint32_t w = rs.w[lo], xy = rs.xy[lo];
    const uint32_t qy = q / w;
    tile = ((xy >> 16) + qy) * (uint32_t)grid_x + (xy & 0xffffu) + (q - qy * w);
}

// ------------------------------------------------------------------ 5. pairs per (CTA range, tile)
__global__ void __launch_bounds__(OE_THREADS)
pair_hist_kernel(uint32_t V, uint32_t A, const uint32_t* __restrict__ pair_incl,
                 const uint32_t* __restrict__ vis_ids, const float2* __restrict__ means2D,
                 const int* __restrict__ radii, const uint8_t* __restrict__ tile_mask, int grid_x, int grid_y,
                 uint16_t* __restrict__ block_hist)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    RangeSmem& rs = *reinterpret_cast<RangeSmem*>(smem_raw);
    uint32_t* s_cnt = reinterpret_cast<uint32_t*>(smem_raw + sizeof(RangeSmem));
    const int n_tiles = grid_x * grid_y, tid = threadIdx.x;
    const uint32_t start = blockIdx.x * (uint32_t)OE_CTA_PAIRS, end = min(A, start + (uint32_t)OE_CTA_PAIRS);
    for (int t = tid; t < n_tiles; t += OE_THREADS) s_cnt[t] = 0u;
    stage_range(rs, pair_incl, vis_ids, means2D, radii, V, start, end, grid_x, grid_y);
#pragma unroll
    for (int i = 0; i < OE_ITEMS; ++i) {
        const uint32_t p = start + (uint32_t)(i * OE_THREADS + tid);
        if (p < end) {
            int j;
            uint32_t t;
            locate_pair(rs, p, grid_x, j, t);
            if (tile_mask == nullptr || tile_mask[t]) atomicAdd(&s_cnt[t], 1u);
        }
    }
    __syncthreads();
    uint16_t* row = block_hist + (size_t)blockIdx.x * n_tiles;
    for (int t = tid; t < n_tiles; t += OE_THREADS) row[t] = (uint16_t)s_cnt[t];
}

// ------------------------------------------------------------------ 6. scan down the tile columns
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

// ------------------------------------------------------------------ 7. write every pair in place
template <int C>
__global__ void __launch_bounds__(OE_THREADS)
pair_emit_kernel(uint32_t V, uint32_t A, const uint32_t* __restrict__ pair_incl,
                 const uint32_t* __restrict__ vis_ids, const float2* __restrict__ means2D,
                 const float* __restrict__ depths, const float4* __restrict__ conic_opacity,
                 const int* __restrict__ radii, const float* __restrict__ features,
                 const uint8_t* __restrict__ tile_mask, int grid_x, int grid_y,
                 const uint32_t* __restrict__ block_base, uint32_t* __restrict__ point_list,
                 uint64_t* __restrict__ keys_out, uint32_t* __restrict__ point_list_unsorted,
                 uint64_t* __restrict__ keys_unsorted, PairRecord* __restrict__ records,
                 uint8_t* __restrict__ masks_out)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    RangeSmem& rs = *reinterpret_cast<RangeSmem*>(smem_raw);
    uint16_t* s_cnt = reinterpret_cast<uint16_t*>(smem_raw + sizeof(RangeSmem));   // [warp][tile]
    const int n_tiles = grid_x * grid_y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t start = blockIdx.x * (uint32_t)OE_CTA_PAIRS, end = min(A, start + (uint32_t)OE_CTA_PAIRS);
    {
        uint32_t* z = reinterpret_cast<uint32_t*>(s_cnt);
        for (int i = tid; i < (OE_WARPS * n_tiles + 1) / 2; i += OE_THREADS) z[i] = 0u;
    }
    stage_range(rs, pair_incl, vis_ids, means2D, radii, V, start, end, grid_x, grid_y);

    // rank of each of the warp's pairs among the warp's pairs of the same tile, in index order
    uint16_t* my_cnt = s_cnt + (size_t)warp * n_tiles;
    const unsigned lt = (1u << lane) - 1u;
    uint32_t tile[OE_ITEMS], gid[OE_ITEMS], rank[OE_ITEMS];
#pragma unroll
    for (int i = 0; i < OE_ITEMS; ++i) {
        const uint32_t p = start + (uint32_t)(warp * OE_WARP_PAIRS + i * 32 + lane);
        tile[i] = 0xffffffffu;
        gid[i] = 0;
        rank[i] = 0;
        if (p < end) {
            int j;
            uint32_t t;
            locate_pair(rs, p, grid_x, j, t);
            if (tile_mask == nullptr || tile_mask[t]) {
                tile[i] = t;
                gid[i] = rs.id[j];
            }
        }
    }
#pragma unroll
    for (int i = 0; i < OE_ITEMS; ++i) {
        const bool valid = tile[i] != 0xffffffffu;
        const unsigned act = __ballot_sync(FULL, valid);
        if (valid) {
            const unsigned peers = __match_any_sync(act, tile[i]);
            const int leader = __ffs(peers) - 1;
            uint32_t prev = 0;
            if (lane == leader) {
                prev = my_cnt[tile[i]];
                my_cnt[tile[i]] = (uint16_t)(prev + __popc(peers));
            }
            prev = __shfl_sync(peers, prev, leader);
            rank[i] = prev + __popc(peers & lt);
        }
        __syncwarp();
    }
    __syncthreads();
    // pairs of the CTA's earlier warps in each tile
    for (int t = tid; t < n_tiles; t += OE_THREADS) {
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < OE_WARPS; ++w) {
            const uint32_t c = s_cnt[(size_t)w * n_tiles + t];
            s_cnt[(size_t)w * n_tiles + t] = (uint16_t)run;
            run += c;
        }
    }
    __syncthreads();

    const uint32_t* base_row = block_base + (size_t)blockIdx.x * n_tiles;
#pragma unroll
    for (int i = 0; i < OE_ITEMS; ++i) {
        const uint32_t t = tile[i];
        if (t == 0xffffffffu) continue;
        const uint32_t slot = base_row[t] + my_cnt[t] + rank[i];
        const uint32_t g = gid[i];
        const float2 xy = means2D[g];
        const float4 co = conic_opacity[g];
        const float depth = depths[g];
        float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
        if (C == 4) {
            f = reinterpret_cast<const float4*>(features)[g];
        } else {
            f.x = features[g * C + 0];
            if (C > 1) f.y = features[g * C + 1];
            if (C > 2) f.z = features[g * C + 2];
        }
        const uint64_t key = ((uint64_t)t << 32) | (uint64_t)__float_as_uint(depth);
        point_list[slot] = g;
        keys_out[slot] = key;
        point_list_unsorted[slot] = g;
        keys_unsorted[slot] = key;
        PairRecord rec;
        rec.xyd = make_float4(xy.x, xy.y, depth, __int_as_float((int)g));
        rec.co = co;
        rec.col = f;
        records[slot] = rec;
        const int tile_x0 = (int)(t % (uint32_t)grid_x) * SL_TILE_X, tile_y0 = (int)(t / (uint32_t)grid_x) * SL_TILE_Y;
        masks_out[slot] = (uint8_t)patch_mask(xy.x, xy.y, co, tile_x0, tile_y0);
    }
}

}  // namespace

size_t ordered_smem_bytes(int n_tiles)
{
    return sizeof(RangeSmem) + (size_t)n_tiles * 4u * (OE_WARPS / 2);   // emit: 2 bytes per (warp, tile); hist: 4 per tile
}

bool ordered_supported(int n_tiles) { return ordered_smem_bytes(n_tiles) <= 200u * 1024u && n_tiles < 65536; }

int ordered_rows(size_t rect_tiles) { return (int)((rect_tiles + OE_CTA_PAIRS - 1) / OE_CTA_PAIRS); }

int launch_depth_sort(int P, int V, const GeomState& g, const int* radii, int grid_x, int grid_y, BinningState& bin,
                      cudaStream_t stream)
{
    if (V <= 0) return 0;
    int rc = sort_prepare(bin.vis_sort, V, 32, stream);
    if (rc) return rc;
    const int grid = min((int)g.pre_blocks, 148 * 4);
    compact_visible_kernel<<<grid, PREPROCESS_THREADS, 0, stream>>>(P, (int)g.pre_blocks, g.tiles_touched, g.depths,
                                                                    g.means2D, radii, grid_x, grid_y, g.block_vis,
                                                                    bin.vis_keys_in, bin.vis_ids_in,
                                                                    bin.vis_sort.histograms);
    count_launch();
    rc = launch_sort_pairs(bin.vis_keys_in, bin.vis_ids_in, bin.vis_keys, bin.vis_ids, bin.vis_sort, V, 32, stream,
                           /*histograms_ready=*/true);
    if (rc) return rc;
    return launch_inclusive_scan_keys_hi(bin.vis_keys, bin.pair_incl, V, g.scan_desc, g.counters, stream);
}

int launch_ordered_emit(int V, size_t rect_tiles, int C, const GeomState& g, const int* radii,
                        const uint8_t* tile_mask, int grid_x, int grid_y, const uint2* ranges, const float* features,
                        BinningState& bin, cudaStream_t stream)
{
    if (V <= 0 || rect_tiles == 0) return 0;
    const int n_tiles = grid_x * grid_y;
    const int rows = ordered_rows(rect_tiles);
    if (C != 3 && C != 4) {
        set_error("ordered emit: unsupported channel count %d (3 or 4)", C);
        return -2;
    }
    const size_t smem_hist = sizeof(RangeSmem) + (size_t)n_tiles * 4u;
    const size_t smem_emit = sizeof(RangeSmem) + ((size_t)OE_WARPS * n_tiles * 2u + 4u);
    SL_CUDA(cudaFuncSetAttribute(pair_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_hist),
            "ordered emit smem attribute");
    pair_hist_kernel<<<rows, OE_THREADS, smem_hist, stream>>>((uint32_t)V, (uint32_t)rect_tiles, bin.pair_incl,
                                                              bin.vis_ids, g.means2D, radii, tile_mask, grid_x, grid_y,
                                                              bin.block_hist);
    count_launch();
    column_scan_kernel<<<ceil_div(n_tiles, 32), COLSCAN_WARPS * 32, 0, stream>>>(bin.block_hist, bin.block_base, ranges,
                                                                                 rows, n_tiles);
    count_launch();
#define SL_EMIT(CC)                                                                                                \
    do {                                                                                                           \
        SL_CUDA(cudaFuncSetAttribute(pair_emit_kernel<CC>, cudaFuncAttributeMaxDynamicSharedMemorySize,            \
                                     (int)smem_emit), "ordered emit smem attribute");                            \
        pair_emit_kernel<CC><<<rows, OE_THREADS, smem_emit, stream>>>(                                             \
            (uint32_t)V, (uint32_t)rect_tiles, bin.pair_incl, bin.vis_ids, g.means2D, g.depths, g.conic_opacity,    \
            radii, features, tile_mask, grid_x, grid_y, bin.block_base, bin.point_list, bin.point_list_keys,        \
            bin.point_list_unsorted, bin.point_list_keys_unsorted, bin.records, bin.masks);                        \
    } while (0)
    if (C == 4) SL_EMIT(4); else SL_EMIT(3);
#undef SL_EMIT
    count_launch();
    return 0;
}

}  // namespace sl
