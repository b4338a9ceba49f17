// splatloc_b200/csrc/capi.cu -- the extern "C" surface declared in
// include/splatloc_b200.h: argument checks, scratch carve-up, stage order.
//
// Stage order of the forward follows CudaRasterizer::Rasterizer::forward
// (cuda_rasterizer/rasterizer_impl.cu:435-582): preprocess -> prefix sum ->
// read R -> emit pairs -> sort -> tile ranges -> blend; the backward follows
// Rasterizer::backward (:586-692): blend backward -> per-Gaussian backward.
// Everything is enqueued on the caller's stream.
#include <atomic>
#include <mutex>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include "kernels.h"
#include "../../include/splatloc_b200.h"

namespace sl {

static thread_local char g_error[512] = "";

void set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

const char* last_error() { return g_error; }

int check_cuda(cudaError_t e, const char* what)
{
    if (e == cudaSuccess) return 0;
    set_error("CUDA error in %s: %s", what, cudaGetErrorString(e));
    return SL_ERR_CUDA;
}

int check_stage(bool debug, cudaStream_t stream, const char* what)
{
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess && debug) e = cudaStreamSynchronize(stream);
    return check_cuda(e, what);
}

// ---------------------------------------------------------------- stage profiler
// Optional per-stage timing with CUDA events on the caller's stream
// (sl_profile_enable / sl_profile_read).  Off by default; when off the hot path
// records nothing.
enum Stage {
    ST_PREPROCESS = 0, ST_SCAN, ST_EMIT, ST_SORT, ST_RANGES, ST_BLEND_FWD, ST_BLEND_BWD,
    ST_PREPROCESS_BWD, ST_COUNT
};
static const char* const kStageNames[ST_COUNT] = {
    "preprocess", "tile_scan", "emit_pairs", "sort_pairs", "pair_records", "blend_forward",
    "blend_backward", "preprocess_backward"};

struct Profiler {
    std::mutex mu;            // guards everything below; taken only while profiling is enabled
    std::atomic<bool> enabled{false};
    int device = -1;          // the device that was current at sl_profile_enable(1): only its calls are timed
    bool created = false;
    cudaEvent_t ev[ST_COUNT][2];
    bool recorded[ST_COUNT] = {};
    double total_ms[ST_COUNT] = {};
    long calls[ST_COUNT] = {};
};
static Profiler g_prof;
static std::atomic<long long> g_launches{0};
static std::atomic<int> g_binning_mode{0};  // sl_set_binning_mode(): 0 default, 1 radix sort, 2 tile buckets
constexpr int kDefaultBinningMode = 2;
static std::atomic<int> g_record_stream{0};  // sl_set_record_stream(): 0 per-lane loads of survivors, 1 bulk-copy staging

static std::atomic<int> g_chained{1};        // sl_set_chained_launches(): programmatic dependent launches between the stages

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
bool chained_launches_enabled() { return g_chained.load(std::memory_order_relaxed) != 0; }

struct StageTimer {
    int st;
    cudaStream_t stream;
    bool on;
    StageTimer(int st_, cudaStream_t s) : st(st_), stream(s), on(g_prof.enabled.load(std::memory_order_relaxed))
    {
        if (!on) return;
        int dev = -1;
        cudaGetDevice(&dev);
        std::lock_guard<std::mutex> lock(g_prof.mu);
        on = dev == g_prof.device;
        if (!on) return;
        if (!g_prof.created) {
            for (int i = 0; i < ST_COUNT; ++i)
                for (int j = 0; j < 2; ++j) cudaEventCreate(&g_prof.ev[i][j]);
            g_prof.created = true;
        }
        if (g_prof.recorded[st]) {  // fold the previous measurement of this stage
            float ms = 0.f;
            if (cudaEventSynchronize(g_prof.ev[st][1]) == cudaSuccess &&
                cudaEventElapsedTime(&ms, g_prof.ev[st][0], g_prof.ev[st][1]) == cudaSuccess) {
                g_prof.total_ms[st] += ms;
                g_prof.calls[st] += 1;
            }
            g_prof.recorded[st] = false;
        }
        cudaEventRecord(g_prof.ev[st][0], stream);
    }
    ~StageTimer()
    {
        if (!on) return;
        std::lock_guard<std::mutex> lock(g_prof.mu);
        cudaEventRecord(g_prof.ev[st][1], stream);
        g_prof.recorded[st] = true;
    }
};

// Number of bits needed for tile ids: same value as the reference's getHigherMsb
// (rasterizer_impl.cu:35-50) -- the position of the highest set bit, plus one.
static int tile_id_bits(uint32_t n)
{
    int bits = 0;
    while (n >> bits) ++bits;
    return bits;
}

static ViewParams make_view(int width, int height, float tan_fovx, float tan_fovy, float scale_modifier, int views)
{
    ViewParams vp;
    vp.W = width;
    vp.H = height;
    vp.grid_x = (width + SL_TILE_X - 1) / SL_TILE_X;
    vp.grid_y = (height + SL_TILE_Y - 1) / SL_TILE_Y;
    vp.tan_fovx = tan_fovx;
    vp.tan_fovy = tan_fovy;
    vp.focal_y = height / (2.0f * tan_fovy);
    vp.focal_x = width / (2.0f * tan_fovx);
    vp.scale_modifier = scale_modifier;
    vp.views = views;
    vp.tiles = vp.grid_x * vp.grid_y;
    return vp;
}

}  // namespace sl

using namespace sl;

extern "C" {

const char* sl_version(void) { return "splatloc_b200 0.1 sm_100a"; }

const char* sl_last_error(void) { return last_error(); }

size_t sl_mapping_loss_workspace_bytes(void) { return mapping_loss_workspace_bytes(); }

int sl_mapping_loss(int width, int height, int channels, const float* image, const float* depth,
                    const float* gt_image, const float* gt_depth, const float* gt_kp, const float* exposure_a,
                    const float* exposure_b, float rgb_boundary_threshold, int initialization, float* dL_dimage,
                    float* dL_ddepth, float* out, void* workspace, void* stream_)
{
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (width <= 0 || height <= 0) { set_error("sl_mapping_loss: empty image %d x %d", width, height); return SL_ERR_ARGUMENT; }
    if (!image || !depth || !gt_image || !gt_depth || !dL_dimage || !dL_ddepth || !out || !workspace) {
        set_error("sl_mapping_loss: null pointer argument");
        return SL_ERR_ARGUMENT;
    }
    int rc = launch_mapping_loss(width, height, channels, image, depth, gt_image, gt_depth, gt_kp, exposure_a, exposure_b,
                                 rgb_boundary_threshold, initialization != 0, dL_dimage, dL_ddepth, out,
                                 static_cast<char*>(workspace), stream);
    if (rc) return rc;
    return check_stage(false, stream, "mapping loss");
}

size_t sl_ssim_workspace_bytes(int width, int height, int channels)
{
    return (width > 0 && height > 0 && channels > 0) ? ssim_workspace_bytes(width, height, channels) : 0;
}

static int check_ssim_shape(const char* who, int width, int height, int channels)
{
    g_error[0] = 0;
    if (width <= 0 || height <= 0 || channels <= 0 || channels > 65535) {
        set_error("%s: bad image shape (%d, %d, %d)", who, channels, height, width);
        return SL_ERR_ARGUMENT;
    }
    return 0;
}

int sl_ssim_forward(int width, int height, int channels, const float* img1, const float* img2, const float* window11,
                    float* maps, float* out, void* workspace, void* stream_)
{
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (int rc = check_ssim_shape("sl_ssim_forward", width, height, channels)) return rc;
    if (!img1 || !img2 || !window11 || !maps || !out || !workspace) {
        set_error("sl_ssim_forward: null pointer argument");
        return SL_ERR_ARGUMENT;
    }
    launch_ssim_forward(width, height, channels, img1, img2, window11, maps, out, static_cast<char*>(workspace), stream);
    return check_stage(false, stream, "ssim forward");
}

int sl_ssim_backward(int width, int height, int channels, const float* img1, const float* img2, const float* window11,
                     const float* maps, const float* upstream, float scale_ssim, float scale_l1, float* dL_dimg1,
                     void* stream_)
{
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (int rc = check_ssim_shape("sl_ssim_backward", width, height, channels)) return rc;
    if (!img1 || !img2 || !window11 || !maps || !dL_dimg1) {
        set_error("sl_ssim_backward: null pointer argument");
        return SL_ERR_ARGUMENT;
    }
    launch_ssim_backward(width, height, channels, img1, img2, window11, maps, upstream, scale_ssim, scale_l1, dL_dimg1, stream);
    return check_stage(false, stream, "ssim backward");
}

long long sl_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

void sl_profile_enable(int on)
{
    std::lock_guard<std::mutex> lock(g_prof.mu);
    int dev = -1;
    cudaGetDevice(&dev);
    if (on != 0 && g_prof.created && dev != g_prof.device) {   // events belong to the device they were made on
        for (int i = 0; i < ST_COUNT; ++i)
            for (int j = 0; j < 2; ++j) cudaEventDestroy(g_prof.ev[i][j]);
        g_prof.created = false;
    }
    if (on != 0) g_prof.device = dev;
    g_prof.enabled.store(on != 0, std::memory_order_relaxed);
    for (int i = 0; i < ST_COUNT; ++i) {
        g_prof.total_ms[i] = 0.0;
        g_prof.calls[i] = 0;
        g_prof.recorded[i] = false;
    }
}

void sl_set_binning_mode(int mode)
{
    g_binning_mode.store(mode >= 1 && mode <= 2 ? mode : 0, std::memory_order_relaxed);
}

void sl_set_chained_launches(int on) { g_chained.store(on ? 1 : 0, std::memory_order_relaxed); }

void sl_set_record_stream(int mode) { g_record_stream.store(mode == 1 ? 1 : 0, std::memory_order_relaxed); }

int sl_profile_stage_count(void) { return ST_COUNT; }

const char* sl_profile_stage_name(int stage) { return (stage >= 0 && stage < ST_COUNT) ? kStageNames[stage] : ""; }

int sl_profile_read(double* total_ms, long long* calls, int n)
{
    std::lock_guard<std::mutex> lock(g_prof.mu);
    for (int i = 0; i < ST_COUNT && i < n; ++i) {
        if (g_prof.recorded[i]) {
            float ms = 0.f;
            if (cudaEventSynchronize(g_prof.ev[i][1]) == cudaSuccess &&
                cudaEventElapsedTime(&ms, g_prof.ev[i][0], g_prof.ev[i][1]) == cudaSuccess) {
                g_prof.total_ms[i] += ms;
                g_prof.calls[i] += 1;
            }
            g_prof.recorded[i] = false;
        }
        total_ms[i] = g_prof.total_ms[i];
        calls[i] = g_prof.calls[i];
    }
    return ST_COUNT;
}

size_t sl_geometry_bytes_batch(int P, int n_views)
{
    return required_bytes<GeomState>((size_t)(P > 0 ? P : 0), (size_t)(n_views > 0 ? n_views : 1));
}
size_t sl_geometry_bytes(int P) { return sl_geometry_bytes_batch(P, 1); }
size_t sl_binning_bytes(int R) { return required_bytes<BinningState>((size_t)(R > 0 ? R : 0)); }
size_t sl_image_bytes_batch(int width, int height, int n_views)
{
    const size_t gx = (width + SL_TILE_X - 1) / SL_TILE_X, gy = (height + SL_TILE_Y - 1) / SL_TILE_Y;
    const size_t v = (size_t)(n_views > 0 ? n_views : 1);
    return required_bytes<ImageState>(v * width * height, v * gx * gy);
}
size_t sl_image_bytes(int width, int height) { return sl_image_bytes_batch(width, height, 1); }

int sl_forward(sl_alloc_fn geometry_alloc, void* geometry_ctx,
               sl_alloc_fn binning_alloc, void* binning_ctx,
               sl_alloc_fn image_alloc, void* image_ctx,
               int P, int D, int M, int C,
               const float* background, int width, int height,
               const float* means3D, const float* shs, const float* colors_precomp,
               const float* opacities, const float* scales, float scale_modifier,
               const float* rotations, const float* cov3D_precomp, const uint8_t* tile_mask,
               const float* viewmatrix, const float* projmatrix, const float* cam_pos,
               float tan_fovx, float tan_fovy, int prefiltered,
               float* out_color, float* out_depth, float* out_alpha, int* radii,
               int debug, void* stream_)
{
    return sl_forward_ext(geometry_alloc, geometry_ctx, binning_alloc, binning_ctx, image_alloc, image_ctx, P, D, M, C,
                          background, width, height, means3D, shs, colors_precomp, opacities, scales, scale_modifier,
                          rotations, cov3D_precomp, tile_mask, viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy,
                          prefiltered, out_color, out_depth, out_alpha, radii, nullptr, debug, stream_);
}

// `n_views` cameras of one image size per call; outputs and per-view state are laid out [view][...].
// `max_pairs` < 0: the reference's protocol -- read R back, size the binning chunk exactly.
// `max_pairs` >= 0 (no read-back): the caller's capacity; no host synchronisation at all.
static int forward_impl(sl_alloc_fn geometry_alloc, void* geometry_ctx,
                        sl_alloc_fn binning_alloc, void* binning_ctx,
                        sl_alloc_fn image_alloc, void* image_ctx,
                        int n_views, int P, int D, int M, int C,
                        const float* background, int width, int height,
                        const float* means3D, const float* shs, const float* colors_precomp,
                        const float* opacities, const float* scales, float scale_modifier,
                        const float* rotations, const float* cov3D_precomp, const uint8_t* tile_mask,
                        const float* viewmatrix, const float* projmatrix, const float* cam_pos,
                        float tan_fovx, float tan_fovy, int prefiltered,
                        float* out_color, float* out_depth, float* out_alpha, int* radii, int* n_touched,
                        int debug, void* stream_, int max_pairs, int* status, const RawParams* raw = nullptr)
{
    g_error[0] = 0;
    const bool nosync = max_pairs >= 0;
    cudaStream_t stream = (cudaStream_t)stream_;
    const bool dbg = debug != 0;
    if (P < 0 || width <= 0 || height <= 0 || (C != 3 && C != 4) || n_views < 1) {
        set_error("sl_forward: bad sizes (P=%d, W=%d, H=%d, C=%d, views=%d; C must be 3 or 4)", P, width, height, C,
                  n_views);
        return SL_ERR_ARGUMENT;
    }
    if (!geometry_alloc || !binning_alloc || !image_alloc || !out_color || !out_depth || !out_alpha ||
        !background || !viewmatrix || !projmatrix) {
        set_error("sl_forward: a required pointer is NULL");
        return SL_ERR_ARGUMENT;
    }
    const size_t N = (size_t)width * height * n_views;     // pixels of all views
    if (P == 0) {
        // Nothing to draw: all-zero outputs, as the reference binding returns
        // (rasterize_points.cu:69-72,84).
        SL_CUDA(cudaMemsetAsync(out_color, 0, N * C * sizeof(float), stream), "forward memset");
        SL_CUDA(cudaMemsetAsync(out_depth, 0, N * sizeof(float), stream), "forward memset");
        SL_CUDA(cudaMemsetAsync(out_alpha, 0, N * sizeof(float), stream), "forward memset");
        if (status != nullptr) SL_CUDA(cudaMemsetAsync(status, 0, sizeof(int), stream), "forward memset");
        return 0;
    }
    if ((long long)P * n_views >= (1ll << 31)) {
        set_error("sl_forward: %d Gaussians x %d views exceed the 2^31 index range", P, n_views);
        return SL_ERR_CAPACITY;
    }
    if (n_touched != nullptr)
        SL_CUDA(cudaMemsetAsync(n_touched, 0, (size_t)P * n_views * sizeof(int), stream), "n_touched memset");
    if (!means3D || !opacities || (!cov3D_precomp && (!scales || !rotations))) {
        set_error("sl_forward: means3D, opacities and either cov3D_precomp or scales+rotations are required");
        return SL_ERR_ARGUMENT;
    }
    if (C != 3 && colors_precomp == nullptr) {
        set_error("For non-RGB, provide precomputed Gaussian colors!");
        return SL_ERR_ARGUMENT;
    }
    if (colors_precomp == nullptr && (shs == nullptr || cam_pos == nullptr || M <= 0)) {
        set_error("sl_forward: without colors_precomp, shs (M > 0) and cam_pos are required");
        return SL_ERR_ARGUMENT;
    }
    if (colors_precomp == nullptr && (D + 1) * (D + 1) > M) {
        set_error("sl_forward: SH degree %d needs %d coefficients, got M=%d", D, (D + 1) * (D + 1), M);
        return SL_ERR_ARGUMENT;
    }

    const ViewParams vp = make_view(width, height, tan_fovx, tan_fovy, scale_modifier, n_views);
    const int n_tiles = vp.tiles * n_views;                 // tiles of all views, numbered view-major

    char* chunk = geometry_alloc(geometry_ctx, sl_geometry_bytes_batch(P, n_views));
    if (!chunk) { set_error("sl_forward: geometry allocation failed"); return SL_ERR_ALLOC; }
    GeomState geom = GeomState::from_chunk(chunk, P, n_views);
    if (radii == nullptr) radii = geom.internal_radii;

    char* img_chunk = image_alloc(image_ctx, sl_image_bytes_batch(width, height, n_views));
    if (!img_chunk) { set_error("sl_forward: image allocation failed"); return SL_ERR_ALLOC; }
    ImageState img = ImageState::from_chunk(img_chunk, N, n_tiles);

    // Per-tile pair counters (one 128-byte line each), filled by preprocess.
    SL_CUDA(cudaMemsetAsync(img.tile_counts, 0, (size_t)n_tiles * TILE_COUNTER_STRIDE * sizeof(uint32_t), stream),
            "tile counter memset");
    int mode = g_binning_mode.load(std::memory_order_relaxed);
    if (mode == 0) mode = kDefaultBinningMode;
    if (nosync) mode = 2;   // the only binning path whose launches do not depend on R
    int rc;
    {
        StageTimer t(ST_PREPROCESS, stream);
        rc = launch_preprocess(P, D, M, C, means3D, scales, rotations, opacities, shs, cov3D_precomp, tile_mask,
                               colors_precomp, viewmatrix, projmatrix, cam_pos, vp, prefiltered != 0, radii,
                               geom, img.tile_counts, raw, stream);
    }
    if (rc) return rc;
    SL_STAGE(dbg, stream, "preprocess");

    {
        StageTimer t(ST_SCAN, stream);
        rc = launch_tile_scan(img.tile_counts, img.ranges, img.tile_order, img.large_tiles, img.bin_info, n_tiles,
                              nosync ? (uint32_t)max_pairs : 0xffffffffu,
                              nosync ? (uint32_t)TILE_SORT_LARGE : 0xffffffffu, status, stream);
    }
    if (rc) return rc;
    SL_STAGE(dbg, stream, "tile scan");

    // The binning chunk is sized by R, which only the device knows: one blocking read.
    BinInfo info;
    if (nosync) {
        // No read-back: the chunk holds `max_pairs` pairs; tile_scan_kernel has flagged the frame on the
        // device if it needs more (or has a tile list beyond what one CTA sorts), in which case the
        // remaining stages do nothing.  Launch shapes
        // below assume the worst admissible case.
        memset(&info, 0, sizeof(info));
        info.num_pairs = (uint32_t)max_pairs;
        info.max_pairs = (uint32_t)TILE_SORT_SMALL;
    } else {
        SL_CUDA(cudaMemcpyAsync(&info, img.bin_info, sizeof(BinInfo), cudaMemcpyDeviceToHost, stream), "read R");
        SL_CUDA(cudaStreamSynchronize(stream), "read R");
    }
    if (info.num_pairs >= (1u << 30)) {
        set_error("sl_forward: %u (tile, Gaussian) pairs exceed the 2^30 limit", info.num_pairs);
        return SL_ERR_CAPACITY;
    }
    const int R = (int)info.num_pairs;

    // Binning, two interchangeable paths that produce the same sorted arrays:
    //   2 tile buckets (tile_binning.cu): scatter pairs into their tile's segment, sort each segment
    //     in registers / shared memory by one CTA;
    //   1 the reference's own pipeline shape -- prefix sum, Gaussian-major emit, global radix sort --
    //     taken when a tile's list exceeds what one CTA sorts.
    const bool radix_path = info.n_huge > 0 || mode == 1;

    char* bin_chunk = binning_alloc(binning_ctx, sl_binning_bytes(R));
    if (!bin_chunk) { set_error("sl_forward: binning allocation failed"); return SL_ERR_ALLOC; }
    BinningState bin = BinningState::from_chunk(bin_chunk, R);
    // colours the blend reads: the caller's (shared by all views) or this call's SH -> RGB results (per view)
    const float* features = colors_precomp != nullptr ? colors_precomp : geom.rgb;
    const size_t feature_view_stride = colors_precomp != nullptr ? 0 : (size_t)3 * P;

    if (R > 0 && !radix_path) {
        {
            StageTimer t(ST_EMIT, stream);
            rc = launch_scatter_pairs(P, geom, radii, tile_mask, vp, img.ranges, img.tile_counts,
                                      bin.point_list_keys_unsorted, bin.point_list_unsorted, img.bin_info, stream);
        }
        if (rc) return rc;
        SL_STAGE(dbg, stream, "scatter pairs");
        {
            StageTimer t(ST_SORT, stream);
            rc = launch_tile_sort(P, C, info.max_pairs, nosync ? -1 : (int)info.n_large, img.ranges,
                                  img.tile_order, img.large_tiles, vp, bin.point_list_keys_unsorted,
                                  bin.point_list_unsorted, bin.point_list_keys, bin.point_list, geom, features,
                                  feature_view_stride, raw, bin.records, bin.masks, img.bin_info, stream);
        }
        if (rc) return rc;
        SL_STAGE(dbg, stream, "tile sort");
    } else if (R > 0) {
        {
            StageTimer t(ST_EMIT, stream);
            rc = launch_inclusive_scan(geom.tiles_touched, geom.point_offsets, P * n_views, geom.scan_desc,
                                       geom.counters, stream);
            if (rc == 0)
                rc = launch_emit_pairs(P, geom, radii, tile_mask, vp, bin.point_list_keys_unsorted,
                                       bin.point_list_unsorted, stream);
        }
        if (rc) return rc;
        SL_STAGE(dbg, stream, "emit pairs");
        {
            StageTimer t(ST_SORT, stream);
            rc = launch_sort_pairs(bin.point_list_keys_unsorted, bin.point_list_unsorted, bin.point_list_keys,
                                   bin.point_list, bin.sort, R, 32 + tile_id_bits(n_tiles), stream);
        }
        if (rc) return rc;
        SL_STAGE(dbg, stream, "sort");
        {
            StageTimer t(ST_RANGES, stream);
            rc = launch_build_records(R, P, C, bin.point_list_keys, bin.point_list, geom, features,
                                      feature_view_stride, raw, vp, bin.records, bin.masks, img.ranges, stream);
        }
        if (rc) return rc;
        SL_STAGE(dbg, stream, "pair records");
    }

    {
        StageTimer t(ST_BLEND_FWD, stream);
        rc = launch_blend_forward(P, C, vp, img.ranges, img.tile_order, bin.records, bin.masks, background, out_color,
                                  out_depth, out_alpha, img.n_contrib, n_touched,
                                  g_record_stream.load(std::memory_order_relaxed), stream);
    }
    if (rc) return rc;
    SL_STAGE(dbg, stream, "blend forward");
    return R;
}

int sl_forward_ext(sl_alloc_fn geometry_alloc, void* geometry_ctx,
                   sl_alloc_fn binning_alloc, void* binning_ctx,
                   sl_alloc_fn image_alloc, void* image_ctx,
                   int P, int D, int M, int C,
                   const float* background, int width, int height,
                   const float* means3D, const float* shs, const float* colors_precomp,
                   const float* opacities, const float* scales, float scale_modifier,
                   const float* rotations, const float* cov3D_precomp, const uint8_t* tile_mask,
                   const float* viewmatrix, const float* projmatrix, const float* cam_pos,
                   float tan_fovx, float tan_fovy, int prefiltered,
                   float* out_color, float* out_depth, float* out_alpha, int* radii, int* n_touched,
                   int debug, void* stream_)
{
    return forward_impl(geometry_alloc, geometry_ctx, binning_alloc, binning_ctx, image_alloc, image_ctx, 1, P, D, M, C,
                        background, width, height, means3D, shs, colors_precomp, opacities, scales, scale_modifier,
                        rotations, cov3D_precomp, tile_mask, viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy,
                        prefiltered, out_color, out_depth, out_alpha, radii, n_touched, debug, stream_, -1, nullptr);
}

int sl_forward_nosync(sl_alloc_fn geometry_alloc, void* geometry_ctx,
                      sl_alloc_fn binning_alloc, void* binning_ctx,
                      sl_alloc_fn image_alloc, void* image_ctx,
                      int P, int D, int M, int C,
                      const float* background, int width, int height,
                      const float* means3D, const float* shs, const float* colors_precomp,
                      const float* opacities, const float* scales, float scale_modifier,
                      const float* rotations, const float* cov3D_precomp, const uint8_t* tile_mask,
                      const float* viewmatrix, const float* projmatrix, const float* cam_pos,
                      float tan_fovx, float tan_fovy, int prefiltered,
                      float* out_color, float* out_depth, float* out_alpha, int* radii, int* n_touched,
                      int max_pairs, int* status, void* stream_)
{
    if (max_pairs < 0 || status == nullptr) {
        set_error("sl_forward_nosync: max_pairs >= 0 and a device status word are required");
        return SL_ERR_ARGUMENT;
    }
    return forward_impl(geometry_alloc, geometry_ctx, binning_alloc, binning_ctx, image_alloc, image_ctx, 1, P, D, M, C,
                        background, width, height, means3D, shs, colors_precomp, opacities, scales, scale_modifier,
                        rotations, cov3D_precomp, tile_mask, viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy,
                        prefiltered, out_color, out_depth, out_alpha, radii, n_touched, /*debug=*/0, stream_, max_pairs,
                        status);
}

int sl_forward_batch(sl_alloc_fn geometry_alloc, void* geometry_ctx,
                     sl_alloc_fn binning_alloc, void* binning_ctx,
                     sl_alloc_fn image_alloc, void* image_ctx,
                     int n_views, int P, int D, int M, int C,
                     const float* background, int width, int height,
                     const float* means3D, const float* shs, const float* colors_precomp,
                     const float* opacities, const float* scales, float scale_modifier,
                     const float* rotations, const float* cov3D_precomp, const uint8_t* tile_mask,
                     const float* viewmatrices, const float* projmatrices, const float* cam_positions,
                     float tan_fovx, float tan_fovy, int prefiltered,
                     float* out_color, float* out_depth, float* out_alpha, int* radii, int* n_touched,
                     int max_pairs, int* status, int debug, void* stream_)
{
    if (max_pairs >= 0 && status == nullptr) {
        set_error("sl_forward_batch: max_pairs >= 0 (no read-back) needs a device status word");
        return SL_ERR_ARGUMENT;
    }
    return forward_impl(geometry_alloc, geometry_ctx, binning_alloc, binning_ctx, image_alloc, image_ctx, n_views, P, D,
                        M, C, background, width, height, means3D, shs, colors_precomp, opacities, scales,
                        scale_modifier, rotations, cov3D_precomp, tile_mask, viewmatrices, projmatrices, cam_positions,
                        tan_fovx, tan_fovy, prefiltered, out_color, out_depth, out_alpha, radii, n_touched,
                        max_pairs >= 0 ? 0 : debug, stream_, max_pairs < 0 ? -1 : max_pairs,
                        max_pairs >= 0 ? status : nullptr);
}

int sl_backward(int P, int D, int M, int R, int C,
                const float* background, int width, int height,
                const float* means3D, const float* shs, const float* colors_precomp,
                const float* scales, float scale_modifier, const float* rotations,
                const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                const float* campos, float tan_fovx, float tan_fovy, const int* radii,
                char* geometry_chunk, char* binning_chunk, char* image_chunk,
                const float* accum_alphas, const float* dL_dpix, const float* dL_dpix_depth,
                const float* dL_dpix_alpha,
                float* dL_dmean2D, float* dL_dabsmean2D, float* dL_dconic, float* dL_dopacity,
                float* dL_dcolor, float* dL_ddepth, float* dL_dmean3D, float* dL_dcov3D,
                float* dL_dsh, float* dL_dscale, float* dL_drot,
                int debug, void* stream_)
{
    return sl_backward_ext(P, D, M, R, C, background, width, height, means3D, shs, colors_precomp, scales,
                           scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy,
                           radii, geometry_chunk, binning_chunk, image_chunk, accum_alphas, dL_dpix, dL_dpix_depth,
                           dL_dpix_alpha, dL_dmean2D, dL_dabsmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_ddepth,
                           dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, nullptr, debug, stream_);
}

int sl_backward_ext(int P, int D, int M, int R, int C,
                    const float* background, int width, int height,
                    const float* means3D, const float* shs, const float* colors_precomp,
                    const float* scales, float scale_modifier, const float* rotations,
                    const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                    const float* campos, float tan_fovx, float tan_fovy, const int* radii,
                    char* geometry_chunk, char* binning_chunk, char* image_chunk,
                    const float* accum_alphas, const float* dL_dpix, const float* dL_dpix_depth,
                    const float* dL_dpix_alpha,
                    float* dL_dmean2D, float* dL_dabsmean2D, float* dL_dconic, float* dL_dopacity,
                    float* dL_dcolor, float* dL_ddepth, float* dL_dmean3D, float* dL_dcov3D,
                    float* dL_dsh, float* dL_dscale, float* dL_drot, float* dL_dtau,
                    int debug, void* stream_)
{
    if (P > 0 && (!dL_dmean3D || !dL_dcov3D)) {
        g_error[0] = 0;
        set_error("sl_backward: a required pointer is NULL");
        return SL_ERR_ARGUMENT;
    }
    return sl_backward_batch(1, P, D, M, R, C, background, width, height, means3D, shs, colors_precomp, scales,
                             scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx,
                             tan_fovy, radii, geometry_chunk, binning_chunk, image_chunk, accum_alphas, dL_dpix,
                             dL_dpix_depth, dL_dpix_alpha, dL_dmean2D, dL_dabsmean2D, dL_dconic, dL_dopacity, dL_dcolor,
                             dL_ddepth, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, dL_dtau, debug, stream_);
}

static int backward_impl(int n_views, int P, int D, int M, int R, int C,
                      const float* background, int width, int height,
                      const float* means3D, const float* shs, const float* colors_precomp,
                      const float* scales, float scale_modifier, const float* rotations,
                      const float* cov3D_precomp, const float* viewmatrices, const float* projmatrices,
                      const float* cam_positions, float tan_fovx, float tan_fovy, const int* radii,
                      char* geometry_chunk, char* binning_chunk, char* image_chunk,
                      const float* accum_alphas, const float* dL_dpix, const float* dL_dpix_depth,
                      const float* dL_dpix_alpha,
                      float* dL_dmean2D, float* dL_dabsmean2D, float* dL_dconic, float* dL_dopacity,
                      float* dL_dcolor, float* dL_ddepth, float* dL_dmean3D, float* dL_dcov3D,
                      float* dL_dsh, float* dL_dscale, float* dL_drot, float* dL_dtau,
                      int debug, void* stream_, const RawParams* raw, const float* raw_opacities,
                      float* dL_dfdc, float* dL_dkp, float* push_mc = nullptr)
{
    g_error[0] = 0;
    cudaStream_t stream = (cudaStream_t)stream_;
    const bool dbg = debug != 0;
    if (P < 0 || R < 0 || width <= 0 || height <= 0 || (C != 3 && C != 4) || n_views < 1) {
        set_error("sl_backward: bad sizes (P=%d, R=%d, W=%d, H=%d, C=%d, views=%d)", P, R, width, height, C, n_views);
        return SL_ERR_ARGUMENT;
    }
    if (P == 0) return 0;
    if (!geometry_chunk || !binning_chunk || !image_chunk || !accum_alphas || !dL_dpix || !dL_dpix_depth ||
        !dL_dpix_alpha || !dL_dmean2D || !dL_dabsmean2D || !dL_dconic || !dL_dopacity || !dL_dcolor ||
        !dL_ddepth || !means3D || !background || !viewmatrices || !projmatrices) {
        set_error("sl_backward: a required pointer is NULL");
        return SL_ERR_ARGUMENT;
    }
    // dL_dmean3D == NULL: the caller only wants the camera-pose gradients (pose refinement against a
    // fixed map); then none of the per-Gaussian outputs of the last stage is written.
    const bool pose_only = dL_dmean3D == nullptr;
    if (pose_only && dL_dtau == nullptr) {
        set_error("sl_backward: neither dL_dmean3D nor dL_dtau is wanted");
        return SL_ERR_ARGUMENT;
    }
    if (!pose_only && !dL_dcov3D) {
        set_error("sl_backward: dL_dcov3D is required with dL_dmean3D");
        return SL_ERR_ARGUMENT;
    }
    if (!pose_only && scales != nullptr && (!rotations || !dL_dscale || !dL_drot)) {
        set_error("sl_backward: scales given without rotations / dL_dscale / dL_drot");
        return SL_ERR_ARGUMENT;
    }
    if (shs != nullptr && ((!pose_only && !dL_dsh) || !cam_positions || C != 3)) {
        set_error("sl_backward: SH gradients need dL_dsh, campos and C == 3");
        return SL_ERR_ARGUMENT;
    }
    const ViewParams vp = make_view(width, height, tan_fovx, tan_fovy, scale_modifier, n_views);
    const int n_tiles = vp.tiles * n_views;

    GeomState geom = GeomState::from_chunk(geometry_chunk, P, n_views);
    BinningState bin = BinningState::from_chunk(binning_chunk, R);
    ImageState img = ImageState::from_chunk(image_chunk, (size_t)width * height * n_views, n_tiles);
    if (radii == nullptr) radii = geom.internal_radii;
    // With precomputed colours dL/dcolour is the same quantity in every view and is summed in place;
    // SH colours depend on the view direction, so each view keeps its own (P, 3) block for the SH backward.
    const size_t color_view_stride = shs != nullptr ? (size_t)3 * P : 0;

    int rc;
    {
        StageTimer t(ST_BLEND_BWD, stream);
        rc = launch_blend_backward(P, C, vp, img.ranges, img.tile_order, bin.records, bin.masks, background,
                                   accum_alphas, img.n_contrib, dL_dpix, dL_dpix_depth, dL_dpix_alpha, dL_dmean2D,
                                   dL_dabsmean2D, dL_dconic, dL_dopacity, dL_dcolor, color_view_stride, dL_ddepth,
                                   stream);
    }
    if (rc) return rc;
    SL_STAGE(dbg, stream, "blend backward");

    const float* cov3D_ptr = (cov3D_precomp != nullptr) ? cov3D_precomp : geom.cov3D;
    {
        StageTimer t(ST_PREPROCESS_BWD, stream);
        rc = launch_preprocess_backward(P, D, M, C, means3D, radii, shs, geom.clamped, pose_only ? nullptr : scales,
                                        rotations, cov3D_ptr, viewmatrices, projmatrices, cam_positions, vp, dL_dmean2D,
                                        dL_dconic, dL_dmean3D, dL_dcolor, color_view_stride, dL_ddepth, dL_dcov3D,
                                        pose_only ? nullptr : dL_dsh, dL_dscale, dL_drot, dL_dtau,
                                        pose_only ? nullptr : raw, raw_opacities, colors_precomp, dL_dopacity, dL_dfdc,
                                        dL_dkp, pose_only ? nullptr : push_mc, stream);
    }
    if (rc) return rc;
    SL_STAGE(dbg, stream, "preprocess backward");
    return 0;
}

int sl_backward_batch(int n_views, int P, int D, int M, int R, int C,
                      const float* background, int width, int height,
                      const float* means3D, const float* shs, const float* colors_precomp,
                      const float* scales, float scale_modifier, const float* rotations,
                      const float* cov3D_precomp, const float* viewmatrices, const float* projmatrices,
                      const float* cam_positions, float tan_fovx, float tan_fovy, const int* radii,
                      char* geometry_chunk, char* binning_chunk, char* image_chunk,
                      const float* accum_alphas, const float* dL_dpix, const float* dL_dpix_depth,
                      const float* dL_dpix_alpha,
                      float* dL_dmean2D, float* dL_dabsmean2D, float* dL_dconic, float* dL_dopacity,
                      float* dL_dcolor, float* dL_ddepth, float* dL_dmean3D, float* dL_dcov3D,
                      float* dL_dsh, float* dL_dscale, float* dL_drot, float* dL_dtau,
                      int debug, void* stream_)
{
    return backward_impl(n_views, P, D, M, R, C, background, width, height, means3D, shs, colors_precomp, scales,
                         scale_modifier, rotations, cov3D_precomp, viewmatrices, projmatrices, cam_positions, tan_fovx,
                         tan_fovy, radii, geometry_chunk, binning_chunk, image_chunk, accum_alphas, dL_dpix,
                         dL_dpix_depth, dL_dpix_alpha, dL_dmean2D, dL_dabsmean2D, dL_dconic, dL_dopacity, dL_dcolor,
                         dL_ddepth, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, dL_dtau, debug, stream_, nullptr,
                         nullptr, nullptr, nullptr);
}

int sl_backward_batch_push(int n_views, int P, int D, int M, int R, int C,
                           const float* background, int width, int height,
                           const float* means3D, const float* shs, const float* colors_precomp,
                           const float* scales, float scale_modifier, const float* rotations,
                           const float* cov3D_precomp, const float* viewmatrices, const float* projmatrices,
                           const float* cam_positions, float tan_fovx, float tan_fovy, const int* radii,
                           char* geometry_chunk, char* binning_chunk, char* image_chunk,
                           const float* accum_alphas, const float* dL_dpix, const float* dL_dpix_depth,
                           const float* dL_dpix_alpha,
                           float* dL_dmean2D, float* dL_dabsmean2D, float* dL_dconic, float* dL_dopacity,
                           float* dL_dcolor, float* dL_ddepth, float* dL_dmean3D, float* dL_dcov3D,
                           float* dL_dsh, float* dL_dscale, float* dL_drot, float* dL_dtau,
                           float* exchange_multicast, int debug, void* stream_)
{
    if (exchange_multicast != nullptr && (shs != nullptr || scales == nullptr || colors_precomp == nullptr)) {
        g_error[0] = 0;
        set_error("sl_backward_batch_push: the exchange rows hold colours_precomp / scale / rotation gradients "
                  "(no SH, no precomputed covariance)");
        return SL_ERR_ARGUMENT;
    }
    return backward_impl(n_views, P, D, M, R, C, background, width, height, means3D, shs, colors_precomp, scales,
                         scale_modifier, rotations, cov3D_precomp, viewmatrices, projmatrices, cam_positions, tan_fovx,
                         tan_fovy, radii, geometry_chunk, binning_chunk, image_chunk, accum_alphas, dL_dpix,
                         dL_dpix_depth, dL_dpix_alpha, dL_dmean2D, dL_dabsmean2D, dL_dconic, dL_dopacity, dL_dcolor,
                         dL_ddepth, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, dL_dtau, debug, stream_, nullptr,
                         nullptr, nullptr, nullptr, exchange_multicast);
}

int sl_exchange_arrive(unsigned int* flags_multicast, int rank, void* stream_)
{
    g_error[0] = 0;
    if (flags_multicast == nullptr || rank < 0) {
        set_error("sl_exchange_arrive: flags_multicast is required and rank must be >= 0");
        return SL_ERR_ARGUMENT;
    }
    sl::launch_exchange_arrive(flags_multicast, rank, (cudaStream_t)stream_);
    return check_stage(false, (cudaStream_t)stream_, "sl_exchange_arrive");
}

int sl_exchange_wait(const unsigned int* flags, int world, unsigned int* waits, int ahead, int advance, int timeout_ms,
                     int* status, void* stream_)
{
    g_error[0] = 0;
    if (flags == nullptr || waits == nullptr || world < 1 || world > 1024 || timeout_ms < 0 || ahead < 0 || ahead > 1) {
        set_error("sl_exchange_wait: flags and waits are required, 1 <= world <= 1024, ahead 0 or 1, timeout_ms >= 0");
        return SL_ERR_ARGUMENT;
    }
    sl::launch_exchange_wait(flags, world, waits, ahead, advance, (unsigned long long)timeout_ms * 1000000ull, status,
                             (cudaStream_t)stream_);
    return check_stage(false, (cudaStream_t)stream_, "sl_exchange_wait");
}

static_assert(sizeof(sl_exchange_source) == sizeof(sl::ExchangeSource), "sl_exchange_source mirrors sl::ExchangeSource");

int sl_exchange_push(const sl_exchange_source* source, int P, int C, float* exchange_multicast, int max_ctas, void* stream_)
{
    g_error[0] = 0;
    if (source == nullptr || exchange_multicast == nullptr || P < 0 || (C != 3 && C != 4)) {
        set_error("sl_exchange_push: source and exchange_multicast are required, C must be 3 or 4 (got %d)", C);
        return SL_ERR_ARGUMENT;
    }
    if (P == 0) return 0;
    sl::launch_exchange_push(reinterpret_cast<const sl::ExchangeSource*>(source), P, C, exchange_multicast,
                             max_ctas > 0 ? max_ctas : 48, (cudaStream_t)stream_);
    return check_stage(false, (cudaStream_t)stream_, "sl_exchange_push");
}

// ---- render() with its pre-processing inside --------------------------------
static int check_render_args(const char* who, int scale_dim, const float* features_dc, const float* raw_opacity,
                             const float* raw_scaling, const float* raw_rotation)
{
    if (scale_dim != 1 && scale_dim != 3) {
        set_error("%s: scale_dim must be 3, or 1 for isotropic Gaussians (got %d)", who, scale_dim);
        return SL_ERR_ARGUMENT;
    }
    if (!features_dc || !raw_opacity || !raw_scaling || !raw_rotation) {
        set_error("%s: features_dc, raw_opacity, raw_scaling and raw_rotation are required", who);
        return SL_ERR_ARGUMENT;
    }
    return 0;
}

int sl_forward_render(sl_alloc_fn geometry_alloc, void* geometry_ctx,
                      sl_alloc_fn binning_alloc, void* binning_ctx,
                      sl_alloc_fn image_alloc, void* image_ctx,
                      int n_views, int P, const float* background, int width, int height,
                      const float* means3D, const float* features_dc, const float* kp_score,
                      const float* raw_opacity, const float* raw_scaling, int scale_dim, float scale_modifier,
                      const float* raw_rotation, const uint8_t* tile_mask,
                      const float* viewmatrices, const float* projmatrices, const float* cam_positions,
                      float tan_fovx, float tan_fovy,
                      float* out_color, float* out_depth, float* out_alpha, int* radii,
                      int max_pairs, int* status, int debug, void* stream_)
{
    g_error[0] = 0;
    if (P > 0) {
        const int rc = check_render_args("sl_forward_render", scale_dim, features_dc, raw_opacity, raw_scaling, raw_rotation);
        if (rc) return rc;
    }
    if (max_pairs >= 0 && status == nullptr) {
        set_error("sl_forward_render: max_pairs >= 0 (no read-back) needs a device status word");
        return SL_ERR_ARGUMENT;
    }
    const RawParams raw = {kp_score, scale_dim};
    return forward_impl(geometry_alloc, geometry_ctx, binning_alloc, binning_ctx, image_alloc, image_ctx, n_views, P,
                        /*D=*/0, /*M=*/0, kp_score != nullptr ? 4 : 3, background, width, height, means3D, nullptr,
                        features_dc, raw_opacity, raw_scaling, scale_modifier, raw_rotation, nullptr, tile_mask,
                        viewmatrices, projmatrices, cam_positions, tan_fovx, tan_fovy, /*prefiltered=*/0, out_color,
                        out_depth, out_alpha, radii, nullptr, max_pairs >= 0 ? 0 : debug, stream_,
                        max_pairs < 0 ? -1 : max_pairs, max_pairs >= 0 ? status : nullptr, &raw);
}

int sl_backward_render(int n_views, int P, int R, const float* background, int width, int height,
                       const float* means3D, const float* features_dc, const float* kp_score,
                       const float* raw_opacity, const float* raw_scaling, int scale_dim, float scale_modifier,
                       const float* raw_rotation,
                       const float* viewmatrices, const float* projmatrices, const float* cam_positions,
                       float tan_fovx, float tan_fovy, const int* radii,
                       char* geometry_chunk, char* binning_chunk, char* image_chunk,
                       const float* accum_alphas, const float* dL_dpix, const float* dL_dpix_depth,
                       const float* dL_dpix_alpha,
                       float* dL_dmean2D, float* dL_dabsmean2D, float* dL_dconic, float* dL_dcolor,
                       float* dL_ddepth, float* dL_dcov3D,
                       float* dL_dmean3D, float* dL_dfeatures_dc, float* dL_dkp_score, float* dL_draw_opacity,
                       float* dL_draw_scaling, float* dL_draw_rotation, int debug, void* stream_)
{
    g_error[0] = 0;
    if (P > 0) {
        const int rc = check_render_args("sl_backward_render", scale_dim, features_dc, raw_opacity, raw_scaling, raw_rotation);
        if (rc) return rc;
        if (!dL_dmean3D || !dL_dfeatures_dc || !dL_draw_opacity || !dL_draw_scaling || !dL_draw_rotation ||
            (kp_score != nullptr && !dL_dkp_score)) {
            set_error("sl_backward_render: an output pointer is NULL");
            return SL_ERR_ARGUMENT;
        }
    }
    const RawParams raw = {kp_score, scale_dim};
    return backward_impl(n_views, P, 0, 0, R, kp_score != nullptr ? 4 : 3, background, width, height, means3D, nullptr,
                         features_dc, raw_scaling, scale_modifier, raw_rotation, nullptr, viewmatrices, projmatrices,
                         cam_positions, tan_fovx, tan_fovy, radii, geometry_chunk, binning_chunk, image_chunk,
                         accum_alphas, dL_dpix, dL_dpix_depth, dL_dpix_alpha, dL_dmean2D, dL_dabsmean2D, dL_dconic,
                         dL_draw_opacity, dL_dcolor, dL_ddepth, dL_dmean3D, dL_dcov3D, nullptr, dL_draw_scaling,
                         dL_draw_rotation, nullptr, debug, stream_, &raw, raw_opacity, dL_dfeatures_dc, dL_dkp_score);
}

// ---- optimiser step and densification ---------------------------------------
static int adam_impl(const char* who, int P, int n_groups, const sl_adam_group* groups, int step, int* step_counter,
                     double beta1, double beta2, double eps, const float* marker, float marker_threshold, void* stream_)
{
    g_error[0] = 0;
    if (P < 0 || n_groups < 0 || n_groups > STORE_MAX_GROUPS || (step_counter == nullptr && step < 1) ||
        (n_groups > 0 && groups == nullptr)) {
        set_error("%s: bad arguments (P=%d, groups=%d of at most %d, step=%d)", who, P, n_groups, STORE_MAX_GROUPS, step);
        return SL_ERR_ARGUMENT;
    }
    if (P == 0 || n_groups == 0) return 0;
    AdamArgs a;
    a.n_groups = 0;
    // torch.optim.Adam forms these scalars in double and hands them to float32 tensor ops
    // (hence beta1, beta2, eps arrive as doubles: 1 - (double)0.999f is off by 1.3e-5 of itself)
    const double bc1 = step_counter ? 1.0 : 1.0 - pow(beta1, (double)step);
    const double bc2 = step_counter ? 1.0 : 1.0 - pow(beta2, (double)step);
    for (int k = 0; k < n_groups; ++k) {
        const sl_adam_group& g = groups[k];
        if (g.width <= 0) continue;
        if (!g.param || !g.grad || !g.exp_avg || !g.exp_avg_sq) {
            set_error("%s: group %d has a NULL pointer", who, k);
            return SL_ERR_ARGUMENT;
        }
        AdamGroup& o = a.group[a.n_groups++];
        o.param = g.param; o.grad = g.grad; o.exp_avg = g.exp_avg; o.exp_avg_sq = g.exp_avg_sq;
        o.width = g.width;
        o.step_size = (float)((double)g.lr / bc1);      // (the plain learning rate with a device-side count)
        o.freeze_by_marker = g.freeze_by_marker;
    }
    a.one_minus_beta1 = (float)(1.0 - beta1);
    a.beta2 = (float)beta2;
    a.one_minus_beta2 = (float)(1.0 - beta2);
    a.bias_correction2_sqrt = (float)sqrt(bc2);
    a.eps = (float)eps;
    a.step_counter = step_counter;
    a.beta1_d = beta1;
    a.beta2_d = beta2;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int rc = launch_adam(a, P, marker, marker_threshold, step_counter, stream);
    if (rc) return rc;
    return check_stage(false, stream, "adam step");
}

int sl_adam_step(int P, int n_groups, const sl_adam_group* groups, int step, double beta1, double beta2, double eps,
                 const float* marker, float marker_threshold, void* stream_)
{
    return adam_impl("sl_adam_step", P, n_groups, groups, step, nullptr, beta1, beta2, eps, marker, marker_threshold, stream_);
}

int sl_adam_step_counted(int P, int n_groups, const sl_adam_group* groups, int* step_counter, double beta1, double beta2,
                         double eps, const float* marker, float marker_threshold, void* stream_)
{
    if (step_counter == nullptr) {
        g_error[0] = 0;
        set_error("sl_adam_step_counted: step_counter is required");
        return SL_ERR_ARGUMENT;
    }
    return adam_impl("sl_adam_step_counted", P, n_groups, groups, 0, step_counter, beta1, beta2, eps, marker,
                     marker_threshold, stream_);
}

int sl_densify_stats(int P, int n_views, const float* means2D_grad, const int* radii, float* xyz_gradient_accum,
                     float* denom, float* max_radii2D, void* stream_)
{
    g_error[0] = 0;
    if (P < 0 || n_views < 1) { set_error("sl_densify_stats: bad sizes"); return SL_ERR_ARGUMENT; }
    if (P == 0) return 0;
    if (!means2D_grad || !radii || !xyz_gradient_accum || !denom || !max_radii2D) {
        set_error("sl_densify_stats: a required pointer is NULL");
        return SL_ERR_ARGUMENT;
    }
    cudaStream_t stream = (cudaStream_t)stream_;
    const int rc = launch_densify_stats(P, n_views, means2D_grad, radii, xyz_gradient_accum, denom, max_radii2D, stream);
    if (rc) return rc;
    return check_stage(false, stream, "densify stats");
}

size_t sl_densify_scratch_bytes(int P) { return densify_scratch_bytes(P); }

int sl_densify_plan(int P, const float* xyz_gradient_accum, const float* denom, const float* raw_scaling, int scale_dim,
                    const float* raw_opacity, const float* marker, float grad_threshold, float percent_dense,
                    float extent, float min_opacity, int use_size_test, int primitive_reg, char* scratch,
                    size_t scratch_bytes, uint32_t* totals, void* stream_)
{
    g_error[0] = 0;
    if (P < 0 || (scale_dim != 1 && scale_dim != 3)) { set_error("sl_densify_plan: bad sizes"); return SL_ERR_ARGUMENT; }
    if (!(grad_threshold > 0.f)) {
        set_error("sl_densify_plan: grad_threshold must be positive (fresh clones carry a zero gradient and must not split)");
        return SL_ERR_ARGUMENT;
    }
    if (!totals) { set_error("sl_densify_plan: totals is NULL"); return SL_ERR_ARGUMENT; }
    cudaStream_t stream = (cudaStream_t)stream_;
    if (P == 0) {
        SL_CUDA(cudaMemsetAsync(totals, 0, 4 * sizeof(uint32_t), stream), "densify plan memset");
        return 0;
    }
    if (!xyz_gradient_accum || !denom || !raw_scaling || !raw_opacity || !scratch) {
        set_error("sl_densify_plan: a required pointer is NULL");
        return SL_ERR_ARGUMENT;
    }
    if (scratch_bytes < densify_scratch_bytes(P)) {
        set_error("sl_densify_plan: scratch of %zu bytes is smaller than the %zu required", scratch_bytes, densify_scratch_bytes(P));
        return SL_ERR_CAPACITY;
    }
    DensifyParams dp;
    dp.grad_threshold = grad_threshold;
    dp.dense_extent = percent_dense * extent;
    dp.extent = extent;
    dp.min_opacity = min_opacity;
    dp.use_size_test = use_size_test;
    dp.primitive_reg = primitive_reg;
    dp.scale_dim = scale_dim;
    const int rc = launch_densify_plan(dp, P, xyz_gradient_accum, denom, raw_scaling, raw_opacity, marker, scratch, totals, stream);
    if (rc) return rc;
    return check_stage(false, stream, "densify plan");
}

int sl_densify_apply(int P, int n_groups, const sl_densify_group* groups, int xyz_group, int scaling_group,
                     int rotation_group, const char* scratch, const uint32_t* totals, const float* noise, void* stream_)
{
    g_error[0] = 0;
    if (P < 0 || n_groups < 1 || n_groups > STORE_MAX_GROUPS || !groups) {
        set_error("sl_densify_apply: bad arguments");
        return SL_ERR_ARGUMENT;
    }
    if (P == 0) return 0;
    auto ok = [&](int g, int w) { return g >= 0 && g < n_groups && (w == 0 || groups[g].width == w); };
    if (!ok(xyz_group, 3) || !ok(rotation_group, 4) || !ok(scaling_group, 0) ||
        (groups[scaling_group].width != 1 && groups[scaling_group].width != 3) || !scratch || !totals || !noise) {
        set_error("sl_densify_apply: xyz (3), scaling (1 or 3) and rotation (4) groups, scratch, totals and noise are required");
        return SL_ERR_ARGUMENT;
    }
    DensifyArrays arr;
    arr.n_groups = n_groups;
    arr.xyz_group = xyz_group; arr.scaling_group = scaling_group; arr.rotation_group = rotation_group;
    for (int k = 0; k < n_groups; ++k) {
        const sl_densify_group& g = groups[k];
        if (g.width > 0 && (!g.param_in || !g.param_out || ((g.exp_avg_in != nullptr) != (g.exp_avg_out != nullptr)) ||
                            ((g.exp_avg_in != nullptr) != (g.exp_avg_sq_in != nullptr)) ||
                            ((g.exp_avg_out != nullptr) != (g.exp_avg_sq_out != nullptr)))) {
            set_error("sl_densify_apply: group %d has inconsistent pointers", k);
            return SL_ERR_ARGUMENT;
        }
        DensifyGroup& o = arr.group[k];
        o.param_in = g.param_in; o.param_out = g.param_out; o.exp_avg_in = g.exp_avg_in; o.exp_avg_out = g.exp_avg_out;
        o.exp_avg_sq_in = g.exp_avg_sq_in; o.exp_avg_sq_out = g.exp_avg_sq_out; o.width = g.width;
    }
    cudaStream_t stream = (cudaStream_t)stream_;
    const int rc = launch_densify_apply(arr, P, groups[scaling_group].width, scratch, totals, noise, stream);
    if (rc) return rc;
    return check_stage(false, stream, "densify apply");
}

int sl_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                    uint8_t* present, void* stream_)
{
    g_error[0] = 0;
    if (P < 0) { set_error("sl_mark_visible: P < 0"); return SL_ERR_ARGUMENT; }
    if (P == 0) return 0;
    if (!means3D || !viewmatrix || !projmatrix || !present) {
        set_error("sl_mark_visible: a required pointer is NULL");
        return SL_ERR_ARGUMENT;
    }
    cudaStream_t stream = (cudaStream_t)stream_;
    int rc = launch_mark_visible(P, means3D, viewmatrix, projmatrix, present, stream);
    if (rc) return rc;
    SL_STAGE(false, stream, "mark visible");
    return 0;
}

size_t sl_knn_scratch_bytes(int P) { return knn_scratch_bytes(P); }

int sl_knn(int P, const float* points, float* mean_dists, char* scratch, size_t scratch_bytes, void* stream_)
{
    g_error[0] = 0;
    if (P < 0) { set_error("sl_knn: P < 0"); return SL_ERR_ARGUMENT; }
    if (P == 0) return 0;
    if (!points || !mean_dists || !scratch) { set_error("sl_knn: a required pointer is NULL"); return SL_ERR_ARGUMENT; }
    if (scratch_bytes < knn_scratch_bytes(P)) {
        set_error("sl_knn: scratch of %zu bytes is smaller than the %zu required", scratch_bytes, knn_scratch_bytes(P));
        return SL_ERR_CAPACITY;
    }
    cudaStream_t stream = (cudaStream_t)stream_;
    int rc = launch_knn(P, points, mean_dists, scratch, stream);
    if (rc) return rc;
    SL_STAGE(false, stream, "knn");
    return 0;
}

}  // extern "C"
