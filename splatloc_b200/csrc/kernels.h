// splatloc_b200/csrc/kernels.h -- host-side launchers of every stage.
// Each returns 0 or a negative C-ABI error code (see include/splatloc_b200.h).
//
// Every stage handles `vp.views` cameras of the same image size in one launch (sl_forward_batch /
// sl_backward_batch; the reference-shaped entry points pass 1).  Cameras are `views` consecutive
// 16-float matrices / 3-float centres; per-view arrays are laid out [view][...]; tiles are numbered
// view-major, global tile = view * vp.tiles + tile, and a pair's sort key carries the global number.
#pragma once
#include "common.cuh"
#include "state.cuh"

namespace sl {

struct ViewParams {
    int W, H;
    int grid_x, grid_y;
    float tan_fovx, tan_fovy;
    float focal_x, focal_y;
    float scale_modifier;
    int views;   // cameras rendered by this call
    int tiles;   // tiles per view (grid_x * grid_y)
};

// K1: per-Gaussian forward preprocess (reference: FORWARD::preprocess, forward.cu:995-1050); also
// counts the pairs of every tile for the tile-bucketed binning.
int launch_preprocess(int P, int D, int M, int C, const float* means3D, const float* scales,
                      const float* rotations, const float* opacities, const float* shs,
                      const float* cov3D_precomp, const uint8_t* tile_mask, const float* colors_precomp,
                      const float* viewmatrices, const float* projmatrices, const float* cam_positions,
                      const ViewParams& vp, bool prefiltered, int* radii, GeomState& g,
                      uint32_t* tile_counts, const RawParams* raw, cudaStream_t stream);

// K10: frustum visibility (reference: checkFrustum, rasterizer_impl.cu:54-66).
int launch_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                        uint8_t* present, cudaStream_t stream);

// K2: inclusive prefix sum, single pass with decoupled look-back.
int launch_inclusive_scan(const uint32_t* in, uint32_t* out, int n, unsigned long long* desc,
                          uint32_t* ticket, cudaStream_t stream);

// K3: emit (tile | depth) keys and Gaussian ids (reference: duplicateWithKeys, rasterizer_impl.cu:70-132).
int launch_emit_pairs(int P, const GeomState& g, const int* radii, const uint8_t* tile_mask,
                      const ViewParams& vp, uint64_t* keys, uint32_t* vals, cudaStream_t stream);

// ---- tile-bucketed binning (the default path; no counterpart in the reference, which sorts all
// pairs globally): per-tile counts come out of preprocess, one CTA turns them into ranges, launch
// order and the totals the host reads back; pairs are scattered straight into their tile's segment
// and each segment is sorted on (depth bits, Gaussian id) by one CTA, which also writes the
// PairRecords and masks.  The result equals the stable global sort on (tile | depth).
int launch_tile_scan(uint32_t* tile_counts, uint2* ranges, uint32_t* tile_order, uint32_t* large_tiles, BinInfo* info,
                     int n_tiles, uint32_t max_pairs, uint32_t max_tile_pairs, int* status, cudaStream_t stream);
int launch_scatter_pairs(int P, const GeomState& g, const int* radii, const uint8_t* tile_mask,
                         const ViewParams& vp, const uint2* ranges, uint32_t* cursors, uint64_t* keys,
                         uint32_t* vals, const BinInfo* dev_info, cudaStream_t stream);
// `n_large_launch`: length of the long-segment work list when the host knows it, < 0 in no-read-back
// mode (the CTAs then read it from the device), 0: no long segments.
int launch_tile_sort(int P, int C, uint32_t longest, int n_large_launch, const uint2* ranges,
                     const uint32_t* tile_order, const uint32_t* large_tiles, const ViewParams& vp,
                     const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out,
                     uint32_t* point_list, const GeomState& g, const float* features, size_t feature_view_stride,
                     const RawParams* raw, PairRecord* records, uint8_t* masks, const BinInfo* dev_info,
                     cudaStream_t stream);

// K4: stable LSD radix sort of n (u64 key, u32 value) pairs on key bits [0, end_bit);
// result in (keys_out, vals_out); the inputs are preserved.
int sort_prepare(SortScratch& sc, int n, int end_bit, cudaStream_t stream);
int launch_sort_pairs(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out,
                      uint32_t* vals_out, SortScratch& sc, int n, int end_bit, cudaStream_t stream);

// Gather, once per frame, everything the blend kernels read for each sorted pair (PairRecord) plus
// the per-patch reach mask, and write the per-tile ranges (reference: identifyTileRanges).
int launch_build_records(int R, int P, int C, const uint64_t* sorted_keys, const uint32_t* point_list,
                         const GeomState& g, const float* features, size_t feature_view_stride, const RawParams* raw,
                         const ViewParams& vp, PairRecord* records, uint8_t* masks, uint2* ranges,
                         cudaStream_t stream);

// K6: forward alpha blending per 16x16 tile (reference: renderCUDA, forward.cu:844-963).
// Outputs are [views][...]: out_color [views, C, H, W], out_depth / out_alpha / n_contrib [views, H, W],
// n_touched [views, P].
int launch_blend_forward(int P, int C, const ViewParams& vp, const uint2* ranges, const uint32_t* tile_order,
                         const PairRecord* records, const uint8_t* masks, const float* bg, float* out_color,
                         float* out_depth, float* out_alpha, uint32_t* n_contrib, int* n_touched,
                         int record_stream, cudaStream_t stream);

// K7: backward blending (reference: BACKWARD::render, backward.cu:1018-1209).
// Per view: dL_dmean2D / dL_dabsmean2D [views, P, 3], dL_dconic [views, P, 4], dL_ddepth [views, P],
// dL_dcolor + view * color_view_stride; summed over the views: dL_dopacity [P].
int launch_blend_backward(int P, int C, const ViewParams& vp, const uint2* ranges, const uint32_t* tile_order,
                          const PairRecord* records, const uint8_t* masks, const float* bg,
                          const float* out_alpha, const uint32_t* n_contrib, const float* dL_dpix,
                          const float* dL_dpix_depth, const float* dL_dpix_alpha, float* dL_dmean2D,
                          float* dL_dabsmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                          size_t color_view_stride, float* dL_ddepth, cudaStream_t stream);

// `raw` != nullptr (raw-parameter mode, common.cuh): `scales` / `rotations` / `raw_opacities` are the raw
// parameters, `dL_dcolor` arrives as the gradient of the activated colour (P, C) and every gradient is
// taken through its activation: dL_dscale, dL_drot, dL_dopacity (in place) become gradients of the raw
// parameters, dL_dfdc (P, 3) receives the gradient of f_dc and dL_dkp (P) the 4th colour channel's.
// `push_mc` != nullptr: multicast address of the [P, EXCHANGE_ROW] exchange rows (common.cuh); the summed
// gradients of every Gaussian this call saw are added into all GPUs' buffers from the kernel's epilogue.
// K8+K9 fused: per-Gaussian backward (reference: BACKWARD::preprocess, backward.cu:1810-1876).  One
// thread carries a Gaussian through all views and stores each gradient once (the sum over the views).
// dL_dmean3D == nullptr: only the camera-pose gradients dL_dtau [views, 6] are wanted.
int launch_preprocess_backward(int P, int D, int M, int C, const float* means3D, const int* radii,
                               const float* shs, const uint8_t* clamped, const float* scales,
                               const float* rotations, const float* cov3D, const float* viewmatrices,
                               const float* projmatrices, const float* cam_positions, const ViewParams& vp,
                               const float* dL_dmean2D, const float* dL_dconic, float* dL_dmean3D,
                               const float* dL_dcolor, size_t color_view_stride, const float* dL_ddepth,
                               float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot, float* dL_dtau,
                               const RawParams* raw, const float* raw_opacities, const float* features_dc,
                               float* dL_dopacity, float* dL_dfdc, float* dL_dkp, float* push_mc, cudaStream_t stream);

// K11-K15: mean squared distance to the 3 nearest neighbours (reference: SimpleKNN::knn,
// simple_knn.cu:186-221).  `scratch` must hold knn_scratch_bytes(P) bytes.
size_t knn_scratch_bytes(int P);
int launch_knn(int P, const float* points, float* mean_dists, char* scratch, cudaStream_t stream);

// per-view mapping loss + gradient seed (reference: utils/utils.py:55-82,
// train_gaussians.py:38-42).  `workspace` holds mapping_loss_workspace_bytes() bytes, zero before first use.
size_t mapping_loss_workspace_bytes();
int launch_mapping_loss(int W, int H, int C, const float* image, const float* depth, const float* gt_image,
                        const float* gt_depth, const float* gt_kp, const float* exposure_a, const float* exposure_b,
                        float rgb_threshold, bool initialization, float* dL_dimage, float* dL_ddepth, float* out,
                        char* workspace, cudaStream_t stream);

// SURVEY 8(f) row 2, SSIM variant: (1 - lambda) L1 + lambda (1 - SSIM) of the colour-refinement loop
// (reference: gaussian_splatting/utils/loss_utils.py:61-102, train_gaussians.py:283-286); ssim_loss.cu.
// `window11`: the 11 taps of the 1-D Gaussian window, in HOST memory.  maps: [3, C, H, W] partial derivatives
// of the SSIM map kept for the backward; out: {mean ssim, mean |img1 - img2|} on the device.
size_t ssim_workspace_bytes(int W, int H, int C);
int launch_ssim_forward(int W, int H, int C, const float* img1, const float* img2, const float* window11, float* maps,
                        float* out, char* workspace, cudaStream_t stream);
int launch_ssim_backward(int W, int H, int C, const float* img1, const float* img2, const float* window11,
                         const float* maps, const float* upstream, float scale_ssim, float scale_l1, float* dL_dimg1,
                         cudaStream_t stream);

// ---- optimiser step and densification over the Gaussian store (optimizer.cu;
// reference: gaussian_splatting/scene/gaussian_model.py:250-300, 493-679).
constexpr int STORE_MAX_GROUPS = 8;      // xyz, f_dc, f_rest, opacity, marker, kp_score, scaling, rotation

struct AdamGroup {
    float* param;
    const float* grad;
    float* exp_avg;
    float* exp_avg_sq;
    int width;               // floats per Gaussian
    float step_size;         // lr / (1 - beta1^step)
    int freeze_by_marker;    // zero the gradient of rows whose marker exceeds the threshold (xyz group)
};
struct AdamArgs {
    AdamGroup group[STORE_MAX_GROUPS];
    int n_groups;
    float one_minus_beta1, beta2, one_minus_beta2, bias_correction2_sqrt, eps;
    // step count kept on the device (a captured launch cannot carry a new count per replay): when
    // step_counter != nullptr the kernel forms the bias corrections itself from *step_counter, in double as
    // torch.optim.Adam does, and group[k].step_size holds the plain learning rate
    const int* step_counter;
    double beta1_d, beta2_d;
};
// `step_counter` != nullptr: incremented by one (a one-thread kernel) before the update reads it.
int launch_adam(const AdamArgs& args, int P, const float* marker, float marker_threshold, int* step_counter,
                cudaStream_t stream);

int launch_densify_stats(int P, int n_views, const float* means2D_grad, const int* radii, float* xyz_gradient_accum,
                         float* denom, float* max_radii2D, cudaStream_t stream);

struct DensifyParams {
    float grad_threshold;    // max_grad
    float dense_extent;      // percent_dense * scene extent
    float extent;
    float min_opacity;
    int use_size_test;       // max_screen_size given
    int primitive_reg;       // rows with marker > 0.005 are never pruned
    int scale_dim;
};
struct DensifyGroup {
    const float* param_in;
    float* param_out;
    const float* exp_avg_in;     // nullptr: this group carries no optimiser state
    float* exp_avg_out;
    const float* exp_avg_sq_in;
    float* exp_avg_sq_out;
    int width;
};
struct DensifyArrays {
    DensifyGroup group[STORE_MAX_GROUPS];
    int n_groups;
    int xyz_group, scaling_group, rotation_group;
};
size_t densify_scratch_bytes(int P);
int launch_densify_plan(const DensifyParams& dp, int P, const float* xyz_gradient_accum, const float* denom,
                        const float* raw_scaling, const float* raw_opacity, const float* marker, char* scratch,
                        uint32_t* totals, cudaStream_t stream);
int launch_densify_apply(const DensifyArrays& arr, int P, int scale_dim, const char* scratch, const uint32_t* totals,
                         const float* noise, cudaStream_t stream);

// ---- split-phase cross-GPU barrier of the fused gradient exchange (exchange.cu; no reference counterpart).
int launch_exchange_arrive(uint32_t* flags_multicast, int rank, cudaStream_t stream);
int launch_exchange_wait(const uint32_t* flags, int world, uint32_t* waits, int ahead, int advance,
                         unsigned long long timeout_ns, int* status, cudaStream_t stream);
struct ExchangeSource {      // mirrors sl_exchange_source (include/splatloc_b200.h); lives in device memory
    const int* radii;        // [n_views, P]; nullptr: nothing to send
    const float* dL_dmean3D;
    const float* dL_dopacity;
    const float* dL_dscale;
    const float* dL_drot;
    const float* dL_dcolor;
    int n_views;
    int scale_dim;           // 3, or 1: one scale gradient per Gaussian
};
int launch_exchange_push(const ExchangeSource* src, int P, int C, float* push_mc, int ctas, cudaStream_t stream);

}  // namespace sl
