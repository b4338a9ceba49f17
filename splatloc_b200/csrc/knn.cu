// splatloc_b200/csrc/knn.cu
//
// K11-K15  distCUDA2: for every point, the mean of the squared distances to its
// three nearest other points (reference: SimpleKNN::knn,
// submodules/simple-knn/simple_knn.cu:186-221, per-pair distance and best-3
// insertion :132-146, output :183).
//
// The reference Morton-sorts the points, forms boxes of 1024 consecutive points
// and lets every point test *all* boxes (O(P^2/1024) box tests, two host syncs,
// six thrust allocations).  Its pruning is conservative, so what it returns is
// the exact set of the three smallest FP32 distances; any exact search gives
// the same bits as long as the per-pair distance is evaluated with the same
// expression and the three values are summed in ascending order.
//
// B200 design: a uniform grid sized to ~P/4..4P cells, filled by a counting
// sort (cell histogram -> prefix sum -> scatter; no comparison sort, no host
// sync, no allocation), then one query thread per point in cell order that
// scans the 3x3x3 block around its cell as 9 contiguous runs of the sorted
// array and widens the block shell by shell until the third-best distance is
// provably inside the searched block.
#include <cfloat>
#include <mutex>
#include "kernels.h"

namespace sl {

namespace {

struct KnnGrid {
    float min_x, min_y, min_z;
    float scale_x, scale_y, scale_z;  // cells per unit length
    float cell_x, cell_y, cell_z;     // unit length per cell
    float ext_x, ext_y, ext_z;
};

constexpr int KNN_HIST_BINS = 1024;   // per-axis bins of the coordinate histogram that trims outliers

struct KnnScratch {
    uint32_t* bbox;        // [6] order-preserving uint encodings of min xyz, max xyz of all points,
                           // then [6] the same for the trimmed box the grid is laid over
    uint32_t* axis_hist;   // [3 * KNN_HIST_BINS] per-axis histogram of the coordinates over the raw box
    uint32_t* cell_of;     // [P]
    uint32_t* counts;      // [n_cells] histogram, consumed by the scatter
    uint32_t* cell_end;    // [n_cells] inclusive prefix sum of counts
    float4* sorted;        // [P] x, y, z, original index (bit pattern)
    unsigned long long* scan_desc;
    uint32_t* ticket;

    static KnnScratch from_chunk(char*& chunk, size_t P, size_t n_cells)
    {
        KnnScratch s;
        carve(chunk, s.bbox, 16);
        carve(chunk, s.axis_hist, 3 * KNN_HIST_BINS);
        carve(chunk, s.cell_of, P);
        carve(chunk, s.counts, n_cells);
        carve(chunk, s.cell_end, n_cells);
        carve(chunk, s.sorted, P);
        carve(chunk, s.scan_desc, (n_cells + SCAN_ITEMS_PER_BLOCK - 1) / SCAN_ITEMS_PER_BLOCK + 1);
        carve(chunk, s.ticket, 8);
        return s;
    }
};

int grid_level(int P)
{
    // G = 2^L cells per axis, with 8^L <= 4 P (at least 1, at most 8 -> 256^3 cells).
    int L = 1;
    while (L < 8 && (1ull << (3 * (L + 1))) <= 4ull * (unsigned long long)P) ++L;
    return L;
}

__device__ __forceinline__ uint32_t f2ord(float f)
{
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float ord2f(uint32_t u)
{
    u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
}

__global__ void knn_bbox_init_kernel(uint32_t* bbox)
{
    if (threadIdx.x < 3) bbox[threadIdx.x] = 0xffffffffu;
    else if (threadIdx.x < 6) bbox[threadIdx.x] = 0u;
}

__global__ void __launch_bounds__(256)
knn_bbox_kernel(int P, const float* __restrict__ pts, uint32_t* bbox)
{
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float v = pts[3 * i + a];
            mn[a] = fminf(mn[a], v);
            mx[a] = fmaxf(mx[a], v);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[a] = fminf(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
            mx[a] = fmaxf(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            atomicMin(&bbox[a], f2ord(mn[a]));
            atomicMax(&bbox[3 + a], f2ord(mx[a]));
        }
    }
}

// The grid is laid over a TRIMMED bounding box: a single far outlier (common in depth
// back-projection) would otherwise stretch the raw box until nearly all points share a handful of
// cells and every query degenerates to brute force.  Per axis, the P / 1000 smallest and largest
// coordinates are cut off with a 1024-bin histogram; points outside the trimmed box are clamped
// into the border cells (cell_coord), which the search handles like any other cell content: the
// stopping bound below only ever uses the INNER faces of the searched block.
__global__ void __launch_bounds__(256)
knn_axis_hist_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ bbox,
                     uint32_t* __restrict__ axis_hist)
{
    __shared__ uint32_t s_hist[3 * KNN_HIST_BINS];
    for (int k = threadIdx.x; k < 3 * KNN_HIST_BINS; k += blockDim.x) s_hist[k] = 0u;
    __syncthreads();
    float mn[3], sc[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        mn[a] = ord2f(bbox[a]);
        const float ext = ord2f(bbox[3 + a]) - mn[a];
        sc[a] = ext > 0.f ? (float)KNN_HIST_BINS / ext : 0.f;
    }
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float f = (pts[3 * i + a] - mn[a]) * sc[a];
            const int b = (f >= 0.f) ? min(KNN_HIST_BINS - 1, (int)fminf(f, (float)(KNN_HIST_BINS - 1))) : 0;
            atomicAdd(&s_hist[a * KNN_HIST_BINS + b], 1u);
        }
    }
    __syncthreads();
    for (int k = threadIdx.x; k < 3 * KNN_HIST_BINS; k += blockDim.x)
        if (s_hist[k]) atomicAdd(&axis_hist[k], s_hist[k]);
}

// One warp per axis: walk the histogram from both ends until P / 1000 points are passed.
__global__ void knn_trim_box_kernel(int P, uint32_t* bbox, const uint32_t* __restrict__ axis_hist)
{
    const int a = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (a >= 3) return;
    const uint32_t trim = (uint32_t)(P / 1000);
    const uint32_t* h = axis_hist + a * KNN_HIST_BINS;
    int lo = 0, hi = KNN_HIST_BINS - 1;
    if (lane == 0) {
        uint32_t acc = 0;
        while (lo < KNN_HIST_BINS - 1 && acc + h[lo] <= trim) acc += h[lo++];
        acc = 0;
        while (hi > lo && acc + h[hi] <= trim) acc += h[hi--];
        const float mn = ord2f(bbox[a]), mx = ord2f(bbox[3 + a]);
        const float w = (mx - mn) / (float)KNN_HIST_BINS;
        // bin edges, never outside the raw box
        const float tmn = lo == 0 ? mn : fminf(mx, mn + (float)lo * w);
        const float tmx = hi == KNN_HIST_BINS - 1 ? mx : fmaxf(tmn, fminf(mx, mn + (float)(hi + 1) * w));
        bbox[6 + a] = f2ord(tmn);
        bbox[9 + a] = f2ord(tmx);
    }
}

__device__ __forceinline__ KnnGrid make_grid(const uint32_t* bbox, int G)
{
    bbox += 6;   // the trimmed box
    KnnGrid g;
    g.min_x = ord2f(bbox[0]); g.min_y = ord2f(bbox[1]); g.min_z = ord2f(bbox[2]);
    g.ext_x = ord2f(bbox[3]) - g.min_x;
    g.ext_y = ord2f(bbox[4]) - g.min_y;
    g.ext_z = ord2f(bbox[5]) - g.min_z;
    g.scale_x = g.ext_x > 0.f ? (float)G / g.ext_x : 0.f;
    g.scale_y = g.ext_y > 0.f ? (float)G / g.ext_y : 0.f;
    g.scale_z = g.ext_z > 0.f ? (float)G / g.ext_z : 0.f;
    g.cell_x = g.ext_x / (float)G;
    g.cell_y = g.ext_y / (float)G;
    g.cell_z = g.ext_z / (float)G;
    return g;
}

__device__ __forceinline__ int cell_coord(float v, float mn, float scale, int G)
{
    const float f = (v - mn) * scale;
    int c = (f >= 0.f) ? (int)fminf(f, (float)(G - 1)) : 0;  // NaN -> 0
    return min(G - 1, max(0, c));
}

__global__ void __launch_bounds__(256)
knn_count_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ bbox, int L,
                 uint32_t* __restrict__ cell_of, uint32_t* counts)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const int G = 1 << L;
    const KnnGrid g = make_grid(bbox, G);
    const int ix = cell_coord(pts[3 * i], g.min_x, g.scale_x, G);
    const int iy = cell_coord(pts[3 * i + 1], g.min_y, g.scale_y, G);
    const int iz = cell_coord(pts[3 * i + 2], g.min_z, g.scale_z, G);
    const uint32_t c = ((uint32_t)iz << (2 * L)) | ((uint32_t)iy << L) | (uint32_t)ix;
    cell_of[i] = c;
    atomicAdd(&counts[c], 1u);
}

__global__ void __launch_bounds__(256)
knn_scatter_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ cell_of,
                   uint32_t* counts, const uint32_t* __restrict__ cell_end, float4* __restrict__ sorted)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const uint32_t c = cell_of[i];
    const uint32_t slot = cell_end[c] - atomicSub(&counts[c], 1u);
    sorted[slot] = make_float4(pts[3 * i], pts[3 * i + 1], pts[3 * i + 2], __uint_as_float((uint32_t)i));
}

__device__ __forceinline__ void best3_insert(const float4& ref, const float4& p, float* best)
{
    const float dx = p.x - ref.x, dy = p.y - ref.y, dz = p.z - ref.z;
    float dist = dx * dx + dy * dy + dz * dz;
#pragma unroll
    for (int j = 0; j < 3; j++) {
        if (best[j] > dist) {
            const float t = best[j];
            best[j] = dist;
            dist = t;
        }
    }
}

__device__ __forceinline__ void scan_run(const float4* __restrict__ sorted, const uint32_t* __restrict__ cell_end,
                                         uint32_t c0, uint32_t c1, int self, const float4& ref, float* best)
{
    const uint32_t begin = c0 == 0 ? 0u : cell_end[c0 - 1];
    const uint32_t end = cell_end[c1];
    for (uint32_t k = begin; k < end; ++k) {
        if ((int)k == self) continue;
        best3_insert(ref, sorted[k], best);
    }
}

__global__ void __launch_bounds__(128)
knn_query_kernel(int P, const float4* __restrict__ sorted, const uint32_t* __restrict__ cell_end,
                 const uint32_t* __restrict__ bbox, int L, float* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const int G = 1 << L;
    const KnnGrid g = make_grid(bbox, G);
    const float4 ref = sorted[i];
    const int ix = cell_coord(ref.x, g.min_x, g.scale_x, G);
    const int iy = cell_coord(ref.y, g.min_y, g.scale_y, G);
    const int iz = cell_coord(ref.z, g.min_z, g.scale_z, G);

    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    for (int r = 1;; ++r) {
        const int x0 = max(ix - r, 0), x1 = min(ix + r, G - 1);
        for (int dz = -r; dz <= r; ++dz) {
            const int z = iz + dz;
            if (z < 0 || z >= G) continue;
            for (int dy = -r; dy <= r; ++dy) {
                const int y = iy + dy;
                if (y < 0 || y >= G) continue;
                const uint32_t row = ((uint32_t)z << (2 * L)) | ((uint32_t)y << L);
                if (r == 1 || max(abs(dy), abs(dz)) == r) {
                    scan_run(sorted, cell_end, row + x0, row + x1, i, ref, best);
                } else {
                    // interior row of the shell: only its two end cells are new
                    if (ix - r >= 0) scan_run(sorted, cell_end, row + ix - r, row + ix - r, i, ref, best);
                    if (ix + r < G) scan_run(sorted, cell_end, row + ix + r, row + ix + r, i, ref, best);
                }
            }
        }
        // Every point outside the searched block is at least `reach` away (with a
        // safety margin for the rounding of the cell assignment).
        // Box-relative coordinates: `ref - min` is the very expression the cell assignment rounds, so
        // the bound is off by a few ulp of the EXTENT, whatever the cloud's distance from the origin
        // (min + k * cell would round at ulp(|coordinate|), which exceeds the pad for a far-away cloud).
        float reach = FLT_MAX;
        const float rx = ref.x - g.min_x, ry = ref.y - g.min_y, rz = ref.z - g.min_z;
        if (g.ext_x > 0.f && ix - r > 0) reach = fminf(reach, rx - (float)(ix - r) * g.cell_x - 4e-6f * g.ext_x);
        if (g.ext_x > 0.f && ix + r < G - 1) reach = fminf(reach, (float)(ix + r + 1) * g.cell_x - rx - 4e-6f * g.ext_x);
        if (g.ext_y > 0.f && iy - r > 0) reach = fminf(reach, ry - (float)(iy - r) * g.cell_y - 4e-6f * g.ext_y);
        if (g.ext_y > 0.f && iy + r < G - 1) reach = fminf(reach, (float)(iy + r + 1) * g.cell_y - ry - 4e-6f * g.ext_y);
        if (g.ext_z > 0.f && iz - r > 0) reach = fminf(reach, rz - (float)(iz - r) * g.cell_z - 4e-6f * g.ext_z);
        if (g.ext_z > 0.f && iz + r < G - 1) reach = fminf(reach, (float)(iz + r + 1) * g.cell_z - rz - 4e-6f * g.ext_z);
        if (reach == FLT_MAX) break;  // the block already covers the whole grid
        if (reach > 0.f && best[2] < reach * reach * 0.9999f) break;
    }
    out[__float_as_uint(ref.w)] = (best[0] + best[1] + best[2]) / 3.0f;
}

}  // namespace

size_t knn_scratch_bytes(int P)
{
    if (P <= 0) return 256;
    const int L = grid_level(P);
    char* p = nullptr;
    KnnScratch::from_chunk(p, (size_t)P, (size_t)1 << (3 * L));
    return reinterpret_cast<size_t>(p) + 256;
}

int launch_knn(int P, const float* points, float* mean_dists, char* scratch, cudaStream_t stream)
{
    if (P <= 0) return 0;
    const int L = grid_level(P);
    const size_t n_cells = (size_t)1 << (3 * L);
    char* chunk = scratch;
    KnnScratch s = KnnScratch::from_chunk(chunk, (size_t)P, n_cells);

    knn_bbox_init_kernel<<<1, 32, 0, stream>>>(s.bbox); count_launch();
    knn_bbox_kernel<<<min(ceil_div(P, 256), 148 * 8), 256, 0, stream>>>(P, points, s.bbox); count_launch();
    SL_CUDA(cudaMemsetAsync(s.axis_hist, 0, 3 * KNN_HIST_BINS * sizeof(uint32_t), stream), "knn memset");
    knn_axis_hist_kernel<<<min(ceil_div(P, 256 * 8), 148 * 4), 256, 0, stream>>>(P, points, s.bbox, s.axis_hist); count_launch();
    knn_trim_box_kernel<<<1, 96, 0, stream>>>(P, s.bbox, s.axis_hist); count_launch();
    SL_CUDA(cudaMemsetAsync(s.counts, 0, n_cells * sizeof(uint32_t), stream), "knn memset");
    knn_count_kernel<<<ceil_div(P, 256), 256, 0, stream>>>(P, points, s.bbox, L, s.cell_of, s.counts); count_launch();
    int rc = launch_inclusive_scan(s.counts, s.cell_end, (int)n_cells, s.scan_desc, s.ticket, stream);
    if (rc != 0) return rc;
    knn_scatter_kernel<<<ceil_div(P, 256), 256, 0, stream>>>(P, points, s.cell_of, s.counts, s.cell_end, s.sorted); count_launch();
    knn_query_kernel<<<ceil_div(P, 128), 128, 0, stream>>>(P, s.sorted, s.cell_end, s.bbox, L, mean_dists); count_launch();
    return 0;
}

}  // namespace sl

namespace sl {

// ---------------------------------------------------------------------------------------------
// sl_knn_static: distCUDA2 for callers that must not allocate.
//
// Who needs it.  sl_knn asks its caller for a scratch buffer, and a caller normally gets one
// from cudaMalloc or from torch's allocator.  Some callers cannot: a keyframe-insertion hook
// running inside a real-time tracking loop, where an allocation may synchronise the whole device
// and blow the frame budget; a plugin loaded into a process whose device memory is entirely owned
// by another allocator, where a fresh cudaMalloc can simply fail; code running while device-wide
// synchronisation is forbidden.  For them this entry point carries its own scratch: a fixed array
// in device memory that belongs to the library image.
//
// What it costs.  The array exists on every GPU that loads the library, from load to process
// exit, whether or not this function is ever called: 3.5 MiB (kKnnStaticScratchBytes), sized
// for at most 65536 points (kKnnStaticMaxPoints; SL_KNN_STATIC_MAX_POINTS in the header) -- ample for SplatLoc, which runs distCUDA2
// on the few thousand points a keyframe adds (gaussian_model.py:178-207).  Larger inputs are
// refused with SL_ERR_CAPACITY; use sl_knn for those.
//
// Rules.  * The first call on each device looks the array's device address up once
//           (cudaGetSymbolAddress, with that device current) and caches it; make that first call
//           before any stream capture begins.
//         * All calls on a device share the one array.  Each call re-initialises the parts that
//           must start from zero (cell counters, scan descriptors, ticket, bounding box), so
//           back-to-back calls on one stream are fine, but two calls whose GPU work can overlap
//           (different streams, different host threads) race on it and give garbage.
//         * No cudaMalloc / cudaFree / allocating library call is made, on any call.
constexpr int kKnnStaticMaxPoints = 65536;
constexpr size_t kKnnStaticScratchBytes = 3670016;   // 3.5 MiB >= knn_scratch_bytes(65536) = 3421216

static __device__ __align__(256) unsigned char g_knn_static_scratch[kKnnStaticScratchBytes];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_knn_static_mutex;
char* g_knn_static_ptr[kMaxDevices] = {};   // device address of g_knn_static_scratch, per device
}  // namespace

}  // namespace sl

extern "C" int sl_knn_static(int P, const float* points, float* mean_dists, void* stream_)
{
    using namespace sl;
    set_error("%s", "");
    if (P < 0) { set_error("sl_knn_static: P < 0"); return -2; }
    if (P == 0) return 0;
    if (!points || !mean_dists) { set_error("sl_knn_static: a required pointer is NULL"); return -2; }
    if (P > kKnnStaticMaxPoints || knn_scratch_bytes(P) > kKnnStaticScratchBytes) {
        set_error("sl_knn_static: %d points exceed the built-in scratch (at most %d points, %zu bytes); "
                  "use sl_knn with caller-provided scratch", P, kKnnStaticMaxPoints, kKnnStaticScratchBytes);
        return -3;
    }
    int dev = -1;
    SL_CUDA(cudaGetDevice(&dev), "sl_knn_static: cudaGetDevice");
    if (dev < 0 || dev >= kMaxDevices) { set_error("sl_knn_static: device index %d out of range", dev); return -2; }
    char* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_knn_static_mutex);
        if (g_knn_static_ptr[dev] == nullptr) {
            // In host code the array's name is only a handle for the runtime, not a device address.
            void* p = nullptr;
            SL_CUDA(cudaGetSymbolAddress(&p, g_knn_static_scratch), "sl_knn_static: cudaGetSymbolAddress");
            g_knn_static_ptr[dev] = static_cast<char*>(p);
        }
        scratch = g_knn_static_ptr[dev];
    }
    cudaStream_t stream = (cudaStream_t)stream_;
    int rc = launch_knn(P, points, mean_dists, scratch, stream);
    if (rc) return rc;
    return check_cuda(cudaGetLastError(), "sl_knn_static");
}

