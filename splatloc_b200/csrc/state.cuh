// splatloc_b200/csrc/state.cuh
//
// Carve-up of the three caller-owned scratch chunks (geometry / binning /
// image).  The leading fields keep the order, element types and 128-byte
// alignment of the reference's GeometryState / BinningState / ImageState
// (cuda_rasterizer/rasterizer_impl.h:33-66, rasterizer_impl.cu:393-431) so
// that the reference's Python buffer interpreters
// (diff_gauss/__init__.py:399-502) read this library's chunks unchanged; the
// space the reference gives to CUB temp storage holds this library's own
// scratch (scan descriptors, radix-sort histograms and ping-pong buffers).
#pragma once
#include "common.cuh"

namespace sl {

constexpr int SCAN_ITEMS_PER_BLOCK = 4096;   // 256 threads x 16 items
constexpr int SORT_ITEMS_PER_BLOCK = 4096;   // 256 threads x 16 keys
constexpr int SORT_SMALL_ITEMS_PER_BLOCK = 1024;   // 256 threads x 4 keys: inputs below SORT_SMALL_LIMIT
constexpr int SORT_SMALL_LIMIT = 1 << 20;
__host__ __device__ inline int sort_items_per_block(size_t n)
{
    return n < (size_t)SORT_SMALL_LIMIT ? SORT_SMALL_ITEMS_PER_BLOCK : SORT_ITEMS_PER_BLOCK;
}
constexpr int SORT_RADIX_BITS = 8;
constexpr int SORT_RADIX = 1 << SORT_RADIX_BITS;
constexpr int SORT_MAX_PASSES = 8;
constexpr int TILE_COUNTER_STRIDE = 32;      // u32 words between per-tile counters: one 128-byte line each
constexpr int TILE_SORT_SMALL = 4096;        // pairs a 256-thread CTA sorts (16 keys per thread, 32 KB of smem)
constexpr int TILE_SORT_LARGE = 16384;       // pairs a 1024-thread CTA sorts in 128 KB of shared memory

constexpr int PREPROCESS_THREADS = 256;      // CTA size of preprocess

// What the host reads back after the per-tile scan (one 32-byte copy).
struct BinInfo {
    uint32_t num_pairs;      // R
    uint32_t n_large;        // tiles with more than TILE_SORT_SMALL pairs
    uint32_t n_huge;         // tiles with more than TILE_SORT_LARGE pairs (-> radix-sort fallback)
    uint32_t max_pairs;      // longest tile list
    uint32_t overflow;       // no-readback mode: R or a tile's list exceeds what the caller provided for
    uint32_t pad[3];
};

template <typename T>
__host__ __device__ inline void carve(char*& chunk, T*& ptr, size_t count, size_t alignment = 128)
{
    size_t off = (reinterpret_cast<uintptr_t>(chunk) + alignment - 1) & ~(alignment - 1);
    ptr = reinterpret_cast<T*>(off);
    chunk = reinterpret_cast<char*>(ptr + count);
}

struct GeomState {
    // A call renders `views` cameras at once (sl_forward_batch; 1 for the reference-shaped entry
    // points).  Per-view arrays are laid out [view][Gaussian]; with views == 1 the chunk is the
    // reference's, field for field.
    // --- reference-compatible prefix ---
    float* depths;            // [views * P]   view-space z (sort-key low word)
    uint8_t* clamped;         // [views * 3P]  SH colour clamp flags
    int* internal_radii;      // [views * P]   used when the caller passes radii == nullptr
    float2* means2D;          // [views * P]   pixel-space centre
    float* cov3D;             // [6P]          upper triangle of the world covariance (the same for every view)
    float4* conic_opacity;    // [views * P]   (conic xx, xy, yy, opacity)
    float* rgb;               // [views * 3P]  SH -> RGB result (C == 3 only)
    uint32_t* tiles_touched;  // [views * P]
    uint32_t* point_offsets;  // [views * P]   inclusive prefix sum of tiles_touched (radix-sort path only)
    // --- own scratch ---
    unsigned long long* scan_desc;  // [scan_blocks] decoupled look-back descriptors
    uint32_t* counters;             // [8] dynamic block tickets etc.
    size_t scan_blocks;

    static GeomState from_chunk(char*& chunk, size_t P, size_t views = 1)
    {
        GeomState g;
        const size_t VP = views * P;
        carve(chunk, g.depths, VP);
        carve(chunk, g.clamped, VP * 3);
        carve(chunk, g.internal_radii, VP);
        carve(chunk, g.means2D, VP);
        carve(chunk, g.cov3D, P * 6);
        carve(chunk, g.conic_opacity, VP);
        carve(chunk, g.rgb, VP * 3);
        carve(chunk, g.tiles_touched, VP);
        carve(chunk, g.point_offsets, VP);
        g.scan_blocks = (VP + SCAN_ITEMS_PER_BLOCK - 1) / SCAN_ITEMS_PER_BLOCK;
        carve(chunk, g.scan_desc, g.scan_blocks + 1);
        carve(chunk, g.counters, 8);
        return g;
    }
};

// `N` and `Tn` are the pixels and tiles of ALL views of the call (views * W * H, views * tiles per view);
// tiles are numbered view-major: global tile = view * tiles_per_view + tile.
struct ImageState {
    uint32_t* n_contrib;  // [N] last contributing list position per pixel
    uint2* ranges;        // [Tn] [first, last+1) in the sorted pair list
    // --- own scratch ---
    uint32_t* tile_order; // [Tn] tile ids, longest list first (launch order of the blend CTAs)
    uint32_t* large_tiles;// [Tn] ids of the tiles whose list needs the large segment sort
    uint32_t* tile_counts;// [Tn * TILE_COUNTER_STRIDE] per-tile pair counters, then scatter cursors
    BinInfo* bin_info;    // [1]

    static ImageState from_chunk(char*& chunk, size_t N, size_t Tn)
    {
        ImageState s;
        carve(chunk, s.n_contrib, N);
        carve(chunk, s.ranges, Tn);
        carve(chunk, s.tile_order, Tn);
        carve(chunk, s.large_tiles, Tn);
        carve(chunk, s.tile_counts, Tn * TILE_COUNTER_STRIDE);
        carve(chunk, s.bin_info, 1);
        return s;
    }
};

// One (tile, Gaussian) pair as the blend kernels consume it: 48 bytes, three 16-byte loads.
struct __align__(16) PairRecord {
    float4 xyd;   // pixel-space x, y, view depth, Gaussian id (bit pattern)
    float4 co;    // conic xx, xy, yy, opacity
    float4 col;   // up to 4 feature channels
};

// Scratch of one radix sort of n (u64 key, u32 value) pairs.
struct SortScratch {
    uint64_t* keys_tmp;      // [n] ping-pong
    uint32_t* vals_tmp;      // [n]
    uint32_t* histograms;    // [SORT_MAX_PASSES * 256] global digit histograms
    uint32_t* tickets;       // [SORT_MAX_PASSES] dynamic block ids
    uint32_t* lookback;      // [SORT_MAX_PASSES * sort_blocks * 256]; zeroed per sort, with the two above
    size_t sort_blocks;

    static SortScratch from_chunk(char*& chunk, size_t n)
    {
        SortScratch b;
        carve(chunk, b.keys_tmp, n);
        carve(chunk, b.vals_tmp, n);
        const size_t per_block = (size_t)sort_items_per_block(n);
        b.sort_blocks = (n + per_block - 1) / per_block;
        carve(chunk, b.histograms, (size_t)SORT_MAX_PASSES * SORT_RADIX);
        carve(chunk, b.tickets, SORT_MAX_PASSES, 4);
        carve(chunk, b.lookback, (size_t)SORT_MAX_PASSES * b.sort_blocks * SORT_RADIX, 4);
        return b;
    }
    static size_t bytes(size_t n);
};

struct BinningState {
    // --- reference-compatible prefix ---
    uint32_t* point_list;             // [R] sorted Gaussian ids
    uint32_t* point_list_unsorted;    // [R]
    uint64_t* point_list_keys;        // [R] sorted (tile << 32 | depth bits)
    uint64_t* point_list_keys_unsorted;  // [R]
    // --- own scratch ---
    PairRecord* records;     // [R] per sorted pair: everything the blend kernels read, gathered once
    uint8_t* masks;          // [R] bit w: the pair can reach the w-th 8x4 patch of its tile
    SortScratch sort;

    static BinningState from_chunk(char*& chunk, size_t R)
    {
        BinningState b;
        carve(chunk, b.point_list, R);
        carve(chunk, b.point_list_unsorted, R);
        carve(chunk, b.point_list_keys, R);
        carve(chunk, b.point_list_keys_unsorted, R);
        carve(chunk, b.records, R);
        carve(chunk, b.masks, R + 64);   // slack: the blend kernels prefetch one 32-entry chunk ahead
        b.sort = SortScratch::from_chunk(chunk, R);
        return b;
    }
};

template <typename S, typename... A>
inline size_t required_bytes(A... a)
{
    char* p = nullptr;
    S::from_chunk(p, a...);
    return reinterpret_cast<size_t>(p) + 128;
}

}  // namespace sl
