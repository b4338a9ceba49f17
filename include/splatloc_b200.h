/* include/splatloc_b200.h
 *
 * C ABI of libsplatloc_b200.so: the B200 (sm_100a) implementation of SplatLoc's
 * differentiable Gaussian-splat rasterizer (diff_gauss) and simple-knn's
 * distCUDA2.  Plain pointers and sizes only; every pointer is a DEVICE pointer
 * unless stated otherwise; the caller owns all memory.  A null pointer means
 * "optional input absent", exactly as the reference treats the data_ptr() of an
 * empty tensor (rasterize_points.cu:100-107).
 *
 * Paths below are relative to the reference tree; DGR = submodules/
 * diff-gaussian-rasterization, KNN = submodules/simple-knn.
 *
 * Return convention: >= 0 success (sl_forward returns the number of
 * (tile, Gaussian) pairs R), < 0 failure with the text in sl_last_error():
 *   -1 CUDA runtime error      -2 invalid argument
 *   -3 capacity exceeded       -4 allocation callback returned NULL
 */
#ifndef SPLATLOC_B200_H_INCLUDED
#define SPLATLOC_B200_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SL_ERR_CUDA     (-1)
#define SL_ERR_ARGUMENT (-2)
#define SL_ERR_CAPACITY (-3)
#define SL_ERR_ALLOC    (-4)

/* Growable scratch chunk: must return a device pointer to at least `bytes`
 * bytes that stays valid until the matching sl_backward has run.  Replaces the
 * std::function<char*(size_t)> of DGR/cuda_rasterizer/rasterizer.h:135-137 (in
 * the reference it resizes a torch byte tensor, DGR/rasterize_points.cu:27-33). */
typedef char* (*sl_alloc_fn)(void* ctx, size_t bytes);

/* Library / build identification, e.g. "splatloc_b200 0.1 sm_100a". */
const char* sl_version(void);

/* Text of the last error raised on the calling thread ("" if none). */
const char* sl_last_error(void);

/* Bytes of the three scratch chunks for P Gaussians, R pairs, a W x H image.
 * Replaces CudaRasterizer::required<GeometryState|BinningState|ImageState>
 * (DGR/cuda_rasterizer/rasterizer_impl.h:66-81).  The leading fields of each
 * chunk keep the reference layout (rasterizer_impl.cu:393-431), so
 * diff_gauss.interpret_{geom,binning,img}Buffer read them unchanged.
 *
 * Which fields of the chunks hold defined data after sl_forward (for debug tooling that reads them
 * through interpret_*Buffer):
 *   always:  geometry  depths, means2D, conic_opacity, tiles_touched (and cov3D / rgb / clamped when
 *            they are computed), for Gaussians with radii > 0;
 *            binning   point_list, point_list_keys (sorted);   image  ranges, n_contrib;
 *   binning mode 1 only (sl_set_binning_mode): geometry point_offsets, and point_list_unsorted /
 *            point_list_keys_unsorted in the reference's Gaussian-major emission order.  In the default
 *            mode point_offsets is never written and the two "unsorted" arrays are grouped by tile;
 *   no-read-back calls (sl_forward_nosync, sl_forward_batch with max_pairs >= 0): only the first R
 *            entries of the binning arrays are filled; the rest up to max_pairs is undefined.
 * The *_batch forms size the geometry and image chunks of a call that renders n_views cameras. */
size_t sl_geometry_bytes(int P);
size_t sl_binning_bytes(int R);
size_t sl_image_bytes(int width, int height);
size_t sl_geometry_bytes_batch(int P, int n_views);
size_t sl_image_bytes_batch(int width, int height, int n_views);

/* Forward rasterization.  Replaces CudaRasterizer::Rasterizer::forward
 * (DGR/cuda_rasterizer/rasterizer.h:134-159, rasterizer_impl.cu:435-582), which
 * RasterizeGaussiansCUDA binds (DGR/rasterize_points.cu:35-121).
 *
 *   C           number of feature channels, 3 or 4 (a compile-time constant,
 *               NUM_CHANNELS, in the reference: cuda_rasterizer/config.h:15)
 *   D, M        active SH degree, SH coefficients per Gaussian (0 without shs)
 *   background  [C]
 *   means3D [P,3]  shs [P,M,3]|NULL  colors_precomp [P,C]|NULL  opacities [P]
 *   scales [P,3]|NULL  rotations [P,4] (w,x,y,z)|NULL  cov3D_precomp [P,6]|NULL
 *   tile_mask   [ceil(H/16) * ceil(W/16)] bytes (0/1) | NULL
 *   viewmatrix, projmatrix  [16] floats, element 4*c+r = mathematical M[r][c]
 *   cam_pos [3]
 *   out_color [C,H,W]  out_depth [H,W]  out_alpha [H,W]  radii [P] (may be NULL)
 *   stream      cudaStream_t the work is enqueued on (the reference uses the
 *               legacy default stream)
 * Exactly one of shs / colors_precomp and one of (scales, rotations) /
 * cov3D_precomp must be given; C != 3 requires colors_precomp (same rule and
 * message as rasterizer_impl.cu:482-485).  Outputs are written for every pixel
 * and every Gaussian; they need not be zeroed by the caller.  The call blocks
 * the host once, to read R back after the prefix sum (as the reference does,
 * rasterizer_impl.cu:522).  Returns R. */
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
               int debug, void* stream);

/* Backward rasterization.  Replaces CudaRasterizer::Rasterizer::backward
 * (DGR/cuda_rasterizer/rasterizer.h:161-195, rasterizer_impl.cu:586-692), bound
 * by RasterizeGaussiansBackwardCUDA (DGR/rasterize_points.cu:123-213).
 * The three scratch chunks are the ones sl_forward filled (same P, R, W, H).
 * All dL_d* outputs must be ZERO on entry (the reference binding allocates them
 * with torch::zeros, rasterize_points.cu:160-171):
 *   dL_dmean2D [P,3]  dL_dabsmean2D [P,3]  dL_dconic [P,4]  dL_dopacity [P]
 *   dL_dcolor [P,C]   dL_ddepth [P]        dL_dmean3D [P,3] dL_dcov3D [P,6]
 *   dL_dsh [P,M,3]    dL_dscale [P,3]      dL_drot [P,4]
 * Inputs: dL_dpix [C,H,W], dL_dpix_depth [H,W], dL_dpix_alpha [H,W],
 * accum_alphas = sl_forward's out_alpha, radii = sl_forward's radii. */
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
                int debug, void* stream);

/* ---- extensions named by the task's north_star that the reference does NOT have (its rasterizer
 * returns no camera gradient and no n_touched, DGR/diff_gauss/__init__.py:122,162,170-181); both are
 * MonoGS-style additions, opt-in, and leave every reference-facing result unchanged.
 *
 * sl_forward_ext = sl_forward + n_touched [P] int32 (may be NULL): number of pixels in which the
 * Gaussian is blended while the transmittance behind it is still above 0.5.
 *
 * sl_backward_ext = sl_backward + dL_dtau [6] (may be NULL; must be ZERO on entry): gradient of the loss
 * w.r.t. a left perturbation of the world-to-camera transform, W2C' = Exp(rho, theta) * W2C, evaluated
 * at rho = theta = 0; dL_dtau = (dL/drho xyz, dL/dtheta xyz).  It accounts for the projected centre,
 * the depth, the EWA covariance (through the camera-frame mean and the rotation) and, with SH colours,
 * the view direction. */
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
                   int debug, void* stream);
/* sl_forward_nosync = sl_forward_ext without the reference's one blocking read of the pair count
 * (rasterizer_impl.cu:522; no reference counterpart): the caller states how many (tile, Gaussian)
 * pairs the binning chunk may hold, the chunk is requested with sl_binning_bytes(max_pairs) before
 * anything runs, and nothing in the call waits for the device -- so a forward + backward step can be
 * captured in a CUDA graph and replayed with no host work between its kernels.
 *   returns max_pairs: pass THAT as R to sl_backward[_ext] (it locates the chunk's arrays).
 *   status [1] (device int, required): 0 after the call's work has run, or 1 if the frame had more
 *     than max_pairs pairs or a tile with more than 16384 of them.  Then colour / depth / alpha are
 *     background-only, n_contrib is 0 and the backward yields zero gradients: check the word when
 *     convenient and redo the frame with sl_forward_ext or a larger max_pairs.
 * Always uses the tile-bucket binning (sl_set_binning_mode does not apply).  debug is not available. */
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
                      int max_pairs, int* status, void* stream);
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
                    int debug, void* stream);

/* ---- several cameras per call (no reference counterpart: the reference renders one camera per call
 * and its callers loop, train_gaussians.py:195-229; the vestigial per-camera pose deltas are
 * utils/camera_utils.py:141-143,179-180).  n_views cameras of ONE image size and field of view are
 * rendered against the same Gaussians by the same seven launches a single view takes: tiles are
 * numbered view-major (global tile = view * tiles_per_view + tile; the sorted keys carry the global
 * number), one pair list holds all views.  Every per-view result is bit-identical to what the
 * single-view call returns for that camera.
 *
 * sl_forward_batch: viewmatrices / projmatrices [n_views, 16], cam_positions [n_views, 3];
 *   out_color [n_views, C, H, W], out_depth / out_alpha [n_views, H, W], radii [n_views, P] (may be NULL),
 *   n_touched [n_views, P] | NULL; tile_mask (one, applied to every view) | NULL.
 *   max_pairs < 0: read the total pair count back once (like sl_forward; status may be NULL, debug applies);
 *   max_pairs >= 0: no read-back, capacity for all views together, status as in sl_forward_nosync.
 *   Chunk sizes: sl_geometry_bytes_batch, sl_image_bytes_batch, sl_binning_bytes(total pairs).
 *   Returns the total pair count (or max_pairs): pass it as R to sl_backward_batch.
 *
 * sl_backward_batch: dL_dpix [n_views, C, H, W], dL_dpix_depth / dL_dpix_alpha / accum_alphas [n_views, H, W].
 *   Outputs, all ZERO on entry.  Per view (screen-space statistics and chain intermediates):
 *     dL_dmean2D, dL_dabsmean2D [n_views, P, 3]   dL_dconic [n_views, P, 4]   dL_ddepth [n_views, P]
 *     dL_dtau [n_views, 6] | NULL
 *     dL_dcolor: [P, C] summed over the views with colors_precomp; [n_views, P, 3] with shs (a view's
 *                colour gradient feeds that view's SH backward)
 *   Summed over the views (what an optimiser consumes; each stored once, no atomics across views):
 *     dL_dopacity [P]  dL_dmean3D [P, 3]  dL_dcov3D [P, 6]  dL_dsh [P, M, 3]  dL_dscale [P, 3]  dL_drot [P, 4]
 *   dL_dmean3D == NULL: only dL_dtau is wanted (pose refinement against a fixed map); then dL_dcov3D,
 *     dL_dsh, dL_dscale and dL_drot are not written and may be NULL. */
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
                     int max_pairs, int* status, int debug, void* stream);
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
                      int debug, void* stream);

/* sl_backward_batch_push = sl_backward_batch + the multi-GPU gradient exchange of a multi-view training
 * step fused into its last kernel (no reference counterpart: the reference has no distributed code;
 * SURVEY.md 8(e) "phase 2").  `exchange_multicast` is a MULTICAST device address (NVSwitch / NVLS; e.g.
 * torch.distributed._symmetric_memory's `multicast_ptr`) of an exchange buffer of [P, 16] floats that
 * exists on every GPU of the job: every Gaussian this call saw adds its row
 *     {dL_dmean3D xyz, dL_dopacity | dL_dscale xyz, n | dL_drot wxyz | dL_dcolor c0..c3}   (n: this call's views that saw it)
 * into ALL GPUs' buffers with multimem.red from the epilogue of the per-Gaussian backward, so after a
 * cross-GPU barrier every buffer holds the sum over all GPUs' views (float 7 of a row: in how many
 * views the Gaussian was visible).  The caller zeroes the buffers and places the barriers
 * (splatloc_b200/exchange.py).  NULL: plain sl_backward_batch.  Needs colors_precomp and
 * scales / rotations (no SH, no precomputed covariance). */
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
                           float* exchange_multicast, int debug, void* stream);

/* Split-phase cross-GPU barrier for that exchange (no reference counterpart).  `flags` are `world`
 * uint32 counters at the same offset of every GPU's exchange allocation, zero at start; flags[r] counts
 * the steps rank r has arrived at.  sl_exchange_arrive (enqueue after sl_backward_batch_push on the same
 * stream) adds 1 to flags[rank] on ALL GPUs through `flags_multicast`, ordered after the rows already
 * pushed.  sl_exchange_wait (enqueue before the rows of the PREVIOUS step are read, and before this
 * step's sl_backward_batch_push) makes the stream wait until every rank has arrived as many times as
 * this GPU has waited (`waits`: one uint32 in this GPU's memory, zero at start, incremented by the wait
 * when `advance` != 0; advance == 0 waits for all arrivals up to now without counting: the flush after
 * the last step).  A wait gives up after `timeout_ms` and sets *status = 1 (status may be NULL). */
int sl_exchange_arrive(unsigned int* flags_multicast, int rank, void* stream);
/* `ahead`: 0 as described; 1: every rank must ALSO have arrived at the step this wait is part of (for
 * the deferred push below, whose arrive comes first in a step and its wait last). */
int sl_exchange_wait(const unsigned int* flags, int world, unsigned int* waits, int ahead, int advance,
                     int timeout_ms, int* status, void* stream);

/* The push as a launch of its own (no reference counterpart): sends the rows sl_backward_batch_push would
 * have sent, from finished gradient arrays, so that a caller can push step i's rows on a second stream
 * while step i + 1 renders.  `source` is a sl_exchange_source in DEVICE memory (read when the kernel
 * runs, so a captured launch can be bound afterwards); source->radii == NULL: nothing is sent.
 * Gradients as sl_backward_batch returns them with colors_precomp and scales / rotations.
 * `max_ctas`: CTAs of the push kernel (256 threads each, striding over the Gaussians); 0: 48, which on a
 * B200 sends 500 k rows' visible fraction in well under a 640x480 step while leaving the concurrent
 * forward its SMs (measured: 8 CTAs take longer than the step, 148 and more slow the forward down). */
typedef struct {
    const int* radii;            /* [n_views, P]: a Gaussian is sent if any view has radii > 0 */
    const float* dL_dmean3D;     /* [P, 3] */
    const float* dL_dopacity;    /* [P]    */
    const float* dL_dscale;      /* [P, 3] ([P] if scale_dim == 1) */
    const float* dL_drot;        /* [P, 4], 16-byte aligned */
    const float* dL_dcolor;      /* [P, C] */
    int n_views;
    int scale_dim;
} sl_exchange_source;
int sl_exchange_push(const sl_exchange_source* source, int P, int C, float* exchange_multicast, int max_ctas, void* stream);

/* ---- render() with its pre-processing inside ----------------------------
 * SplatLoc's render() (gaussian_splatting/gaussian_renderer/__init__.py:12-141) turns the optimiser's
 * raw parameters into rasterizer inputs with PyTorch ops on every call -- opacity = sigmoid(_opacity),
 * scales = exp(_scaling) (repeated 3x when isotropic), rotations = normalize(_rotation)
 * (scene/gaussian_model.py:75-110), colours = clamp_min(eval_sh(0, _features_dc) + 0.5, 0)
 * (utils/sh_utils.py:55-73) concatenated with _kp_score (:85-102) -- materialising four (P, .) tensors
 * and their autograd nodes.  These entry points take the raw parameters and apply the activations inside
 * preprocess / the record gather, and the backward returns gradients of the RAW parameters.  For the
 * configuration SplatLoc trains with: SH degree 0 (configs/.../base_config.yaml:88), convert_SHs_python.
 *   features_dc [P, 3]  kp_score [P] | NULL (then 3 channels)  raw_opacity [P]
 *   raw_scaling [P, scale_dim], scale_dim 3 or 1  raw_rotation [P, 4]
 * Camera, image, tile mask, max_pairs / status / debug and the n_views layout: as sl_forward_batch.
 * C = 4 with kp_score, else 3; out_color [n_views, C, H, W].  render()'s "render" is channels 0..2 and
 * "kp_prob" channel C - 1 of it.
 * sl_backward_render: intermediates (ZERO on entry, contents afterwards are scratch):
 *     dL_dmean2D, dL_dabsmean2D [n_views, P, 3] (render()'s viewspace_points.grad / .absgrad),
 *     dL_dconic [n_views, P, 4], dL_dcolor [P, C], dL_ddepth [n_views, P], dL_dcov3D [P, 6];
 *   results (ZERO on entry, summed over the views): dL_dmean3D [P, 3], dL_dfeatures_dc [P, 3],
 *     dL_dkp_score [P] (NULL iff kp_score is), dL_draw_opacity [P], dL_draw_scaling [P, scale_dim],
 *     dL_draw_rotation [P, 4]. */
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
                      int max_pairs, int* status, int debug, void* stream);
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
                       float* dL_draw_scaling, float* dL_draw_rotation, int debug, void* stream);

/* ---- optimiser step and densification over the Gaussian store -----------
 * Replaces, on device arrays the caller owns, what gaussian_splatting/scene/gaussian_model.py does
 * with PyTorch indexing, torch.cat and optimizer-state surgery between two rasterizer calls.
 *
 * sl_adam_step: torch.optim.Adam(l, lr=0.0, eps=1e-15).step() (training_setup, :250-300) for up to 8
 *   parameter groups in one launch; `step` is the 1-based step count (bias correction); beta1, beta2 and
 *   eps are doubles because PyTorch forms 1 - beta and the bias corrections in double.  A group with
 *   freeze_by_marker != 0 uses a zero gradient for rows whose marker[row] > marker_threshold
 *   (train_gaussians.py:232-234 zeroes the position gradient of key primitives).
 * sl_densify_stats: add_densification_stats (:677-679) and the max_radii2D update
 *   (train_gaussians.py:241-244) for the n_views views of a window: means2D_grad [n_views, P, 3],
 *   radii [n_views, P]; xyz_gradient_accum, denom, max_radii2D [P] are updated in place.
 * sl_densify_plan + sl_densify_apply: densify_and_prune(max_grad, min_opacity, extent, max_screen_size)
 *   (:650-675), i.e. densify_and_clone, densify_and_split (N = 2) and the final prune, as one
 *   compaction.  _plan writes the selection flags and their prefix sums into `scratch`
 *   (sl_densify_scratch_bytes(P) bytes) and four device counters totals[4] = {kept originals, clones,
 *   split Gaussians whose children survive, split Gaussians}; the caller reads them back (the one
 *   host synchronisation: the new size P' = totals[0] + totals[1] + 2 totals[2] sizes the output
 *   arrays), provides `noise` [2 * totals[3], 3] standard-normal floats (the reference draws
 *   torch.normal there; rows k * totals[3] + j belong to child k of the j-th split Gaussian), and
 *   _apply writes every group's output rows: kept originals in ascending order, then clones, then first
 *   children, then second children -- the reference's order.  Kept rows keep their Adam moments, new
 *   rows get zeros; a group with exp_avg_in == NULL carries no moments.  The statistics of the new
 *   store are all zero (densification_postfix :583-585): the caller zero-fills them.
 *   use_size_test: max_screen_size was given (the reference then also prunes Gaussians larger than
 *   0.1 * extent; its screen-size test reads statistics it has just reset and never fires).
 *   primitive_reg: rows with marker > 0.005 are never pruned (:671-673).  grad_threshold must be > 0. */
typedef struct {
    float* param; const float* grad; float* exp_avg; float* exp_avg_sq;
    int width; float lr; int freeze_by_marker;
} sl_adam_group;
typedef struct {
    const float* param_in; float* param_out;
    const float* exp_avg_in; float* exp_avg_out; const float* exp_avg_sq_in; float* exp_avg_sq_out;
    int width;
} sl_densify_group;
int sl_adam_step(int P, int n_groups, const sl_adam_group* groups, int step, double beta1, double beta2, double eps,
                 const float* marker, float marker_threshold, void* stream);
/* The same with the step count in DEVICE memory (one int, incremented by the call before the update reads
 * it; zero before the first step): a launch captured in a CUDA graph cannot carry a new count per replay.
 * The bias corrections are then formed on the device, in double. */
int sl_adam_step_counted(int P, int n_groups, const sl_adam_group* groups, int* step_counter, double beta1, double beta2,
                         double eps, const float* marker, float marker_threshold, void* stream);
int sl_densify_stats(int P, int n_views, const float* means2D_grad, const int* radii, float* xyz_gradient_accum,
                     float* denom, float* max_radii2D, void* stream);
size_t sl_densify_scratch_bytes(int P);
int sl_densify_plan(int P, const float* xyz_gradient_accum, const float* denom, const float* raw_scaling, int scale_dim,
                    const float* raw_opacity, const float* marker, float grad_threshold, float percent_dense,
                    float extent, float min_opacity, int use_size_test, int primitive_reg, char* scratch,
                    size_t scratch_bytes, uint32_t* totals, void* stream);
int sl_densify_apply(int P, int n_groups, const sl_densify_group* groups, int xyz_group, int scaling_group,
                     int rotation_group, const char* scratch, const uint32_t* totals, const float* noise, void* stream);

/* Frustum visibility, present[i] in {0,1}.  Replaces
 * CudaRasterizer::Rasterizer::markVisible (DGR/cuda_rasterizer/rasterizer.h:71-76,
 * rasterizer_impl.cu:379-391), bound by markVisible (DGR/rasterize_points.cu:215-234). */
int sl_mark_visible(int P, const float* means3D, const float* viewmatrix,
                    const float* projmatrix, uint8_t* present, void* stream);

/* distCUDA2: mean_dists[i] = mean of the squared distances from point i to its 3
 * nearest other points.  Replaces SimpleKNN::knn (KNN/simple_knn.h:16-20,
 * simple_knn.cu:186-221), bound by distCUDA2 (KNN/spatial.cu:15-25).  The
 * reference allocates its temporaries itself (cudaMalloc / thrust); here the
 * caller passes `scratch` of at least sl_knn_scratch_bytes(P) bytes. */
size_t sl_knn_scratch_bytes(int P);
int sl_knn(int P, const float* points, float* mean_dists,
           char* scratch, size_t scratch_bytes, void* stream);

/* distCUDA2 without any allocation (no counterpart in the reference, whose SimpleKNN::knn
 * allocates with cudaMalloc and thrust on every call).  Same result as sl_knn, but the scratch is
 * a fixed 3.5 MiB array inside the library's device image, so a caller that must not allocate --
 * a real-time keyframe hook that cannot afford an allocation's device-wide synchronisation, a
 * plugin in a process whose device memory is owned by another allocator -- passes only the input
 * and the output.  Limits and rules:
 *   - at most SL_KNN_STATIC_MAX_POINTS = 65536 points; more returns SL_ERR_CAPACITY (use sl_knn);
 *   - the array occupies 3.5 MiB on every GPU that loads the library until the process exits;
 *   - the first call on each device resolves the array's device address once and caches it:
 *     make that first call before any stream capture starts;
 *   - every call on a device uses the same array (each call re-zeroes what it needs): calls whose
 *     GPU work may overlap -- other streams, other threads -- race; serialise them. */
#define SL_KNN_STATIC_MAX_POINTS 65536
int sl_knn_static(int P, const float* points, float* mean_dists, void* stream);

/* ---- per-view mapping loss and its gradient seed ---------------------
 * Replaces, for one rendered view, the reference's
 *   get_loss_mapping(config, image, depth, viewpoint, opacity, initialization)   utils/utils.py:55-82
 *   get_loss_marker(config, pred, gt)                                            train_gaussians.py:38-42
 * as the mapping loop sums them (train_gaussians.py:215-216):
 *   loss = mean |(exp(a) * rgb + b) * m_rgb - gt_rgb * m_rgb| + mean |depth * m_d - gt_depth * m_d|
 *        + mean BCE(sigmoid(kp_logit), gt_kp),
 *   m_rgb = gt_rgb.sum(0) > rgb_boundary_threshold,  m_d = gt_depth > 0.01,
 * and writes d loss / d image, d loss / d depth in the same pass over the pixels.
 *   image     (channels, H, W): the rasterizer's colour output; channels 0..2 are rgb and, when
 *             gt_kp != NULL, channel channels-1 is the key-point logit (render()'s "kp_prob")
 *   depth     (1, H, W); gt_image (3, H, W); gt_depth (H, W); gt_kp (H, W) or NULL (no key-point term)
 *   exposure_a, exposure_b: device pointers to the camera's exposure parameters (one float each; NULL = 0)
 *   initialization != 0: the exposure affine is skipped (utils.py:56-57)
 *   out[6] (device): loss, colour term, depth term, key-point term, d loss / d exposure_a, d loss / d exposure_b
 *   workspace: sl_mapping_loss_workspace_bytes() bytes of device memory, zero-filled before its first
 *             use (the kernel leaves it ready for the next call); one per concurrently running call.
 * Sums are accumulated in double and reduced in a fixed order: results are run-to-run identical. */
size_t sl_mapping_loss_workspace_bytes(void);
int sl_mapping_loss(int width, int height, int channels, const float* image, const float* depth,
                    const float* gt_image, const float* gt_depth, const float* gt_kp, const float* exposure_a,
                    const float* exposure_b, float rgb_boundary_threshold, int initialization, float* dL_dimage,
                    float* dL_ddepth, float* out, void* workspace, void* stream);

/* ---- SSIM + L1 photometric loss (SSIM variant) ---------------------------
 * The loss of SplatLoc's colour-refinement loop (train_gaussians.py:283-286),
 *   loss = (1 - lambda_dssim) * l1_loss(image, gt) + lambda_dssim * (1 - ssim(image, gt)),
 * with ssim as gaussian_splatting/utils/loss_utils.py:61-102 computes it: depth-wise 11 x 11 Gaussian
 * window (sigma 1.5, the outer product of `window11`, the 11 normalised 1-D taps, HOST memory), zero
 * padding, C1 = 0.01^2, C2 = 0.03^2, mean over all channels * H * W elements.
 * sl_ssim_forward: img1, img2 (channels, H, W) on the device -> out[2] (device) = {mean ssim,
 *   mean |img1 - img2|}, and `maps` (3 * channels * H * W floats, device): the partial derivatives of the
 *   SSIM map the backward filters.  workspace: sl_ssim_workspace_bytes() bytes, its first 16 bytes zero
 *   before the first use (the kernel leaves them zero); one per concurrently running call.
 * sl_ssim_backward: dL_dimg1 = upstream * (scale_ssim * d(sum of the ssim map)/d img1
 *                                          + scale_l1 * sign(img1 - img2));
 *   `upstream`: device pointer to one float (NULL = 1).  For the loss above with means:
 *   scale_ssim = -lambda / (channels * H * W), scale_l1 = (1 - lambda) / (channels * H * W).
 * Sums are accumulated in double in a fixed order: results are run-to-run identical. */
size_t sl_ssim_workspace_bytes(int width, int height, int channels);
int sl_ssim_forward(int width, int height, int channels, const float* img1, const float* img2,
                    const float* window11, float* maps, float* out, void* workspace, void* stream);
int sl_ssim_backward(int width, int height, int channels, const float* img1, const float* img2,
                     const float* window11, const float* maps, const float* upstream, float scale_ssim,
                     float scale_l1, float* dL_dimg1, void* stream);

/* ---- instrumentation (no counterpart in the reference, which has no tracing,
 * SURVEY.md section 5) ------------------------------------------------------ */

/* Number of this library's kernels launched by the process so far. */
long long sl_launch_count(void);

/* Binning strategy of sl_forward (no reference counterpart; both strategies produce the reference's
 * sorted lists bit for bit, the switch exists so tests can cover each of them).  The setting is
 * PROCESS-WIDE: it changes the frames of every thread that calls sl_forward afterwards.
 *   0 = the library's default (currently 2);
 *   1 = prefix sum, Gaussian-major emit and one global radix sort (the reference's pipeline shape);
 *   2 = pairs scattered into per-tile segments, each segment sorted by one CTA with a bitonic network
 *       held in registers (warp shuffles; shared memory only for the cross-warp steps); falls back to
 *       1 when a tile's list is longer than the 16384 pairs one CTA sorts. */
void sl_set_binning_mode(int mode);

/* How the forward blend kernel brings a tile's pair records on chip (no reference counterpart; the
 * reference gathers per-Gaussian data into shared memory per 256-entry batch, forward.cu:899-908).
 * PROCESS-WIDE, like the binning mode; both give identical results:
 *   0 = default: every lane loads the record of one list entry that survives the per-patch reach mask
 *       (three 16-byte loads), survivors are parked in a warp-private shared buffer;
 *   1 = the copy engine stages whole 32-record chunks (cp.async.bulk, completion on an mbarrier, two
 *       stages per warp).  Kept for the A/B in profiles/README.md; measured slower on the B200. */
void sl_set_record_stream(int mode);

/* Programmatic dependent launches between the stages of a call (no reference counterpart).  PROCESS-WIDE.
 * 1 (default): every stage kernel is launched so that its CTAs may become resident while the previous
 * stage's last CTAs still run; they block on entry until that stage has completed, so results are
 * unchanged -- what is saved is the drain / launch / ramp-up between the dependent kernels of a step.
 * 0: plain stream-ordered launches. */
void sl_set_chained_launches(int on);

/* Per-stage device timing with CUDA events recorded on the caller's stream
 * around each stage of sl_forward / sl_backward.  sl_profile_enable(1) resets
 * and starts accumulation, sl_profile_enable(0) stops it.  sl_profile_read
 * synchronises on the last recorded events and fills, for stage i <
 * sl_profile_stage_count(), the accumulated milliseconds and number of timed
 * calls.  Host-blocking; not for use inside latency-critical loops.  The profiler is one
 * process-wide object guarded by a mutex: enable it from one thread, for calls on ONE device
 * (the device that is current at sl_profile_enable(1)); calls on other devices are not timed. */
void sl_profile_enable(int on);
int sl_profile_stage_count(void);
const char* sl_profile_stage_name(int stage);
int sl_profile_read(double* total_ms, long long* calls, int n);

#ifdef __cplusplus
}
#endif
#endif /* SPLATLOC_B200_H_INCLUDED */
