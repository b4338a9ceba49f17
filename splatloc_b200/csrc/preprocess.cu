// splatloc_b200/csrc/preprocess.cu
//
// K1  forward per-Gaussian preprocess: frustum cull, EWA projection of the 3D
//     covariance, conic, screen radius, tile rectangle, optional SH -> RGB.
//     Follows the arithmetic specified by the reference's preprocessCUDA
//     (cuda_rasterizer/forward.cu:694-839) with computeCov3D (:394-428),
//     computeCov2D (:347-389) and computeColorFromSH (:20-71).
// K10 frustum visibility (rasterizer_impl.cu:54-66).
//
// One thread per (Gaussian, view), 256 threads per block; the kernel is HBM-bound
// (44 B in, 8 B + 52 B per visible Gaussian out) and carries no shared state.
#include "kernels.h"
#include "sh.cuh"

namespace sl {

// World covariance from scale and (un-normalised, as in the reference) quaternion.
__device__ __forceinline__ void cov3d_from_scale_rot(const f3 s, float mod, const float4 q, float* cov3D)
{
    mat3 S = mat3_cols(1.0f, 0.0f, 0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.0f, 1.0f);
    S.c[0][0] = mod * s.x;
    S.c[1][1] = mod * s.y;
    S.c[2][2] = mod * s.z;

    float r = q.x, x = q.y, y = q.z, z = q.w;
    mat3 R = mat3_cols(
        1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
        2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
        2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));

    mat3 Mx = mul(S, R);
    mat3 Sigma = mul(transpose(Mx), Mx);

    cov3D[0] = Sigma.c[0][0];
    cov3D[1] = Sigma.c[0][1];
    cov3D[2] = Sigma.c[0][2];
    cov3D[3] = Sigma.c[1][1];
    cov3D[4] = Sigma.c[1][2];
    cov3D[5] = Sigma.c[2][2];
}

// Screen-space covariance (EWA, Zwicker et al. 2002, eq. 29/31) + 0.3 px low-pass.
__device__ __forceinline__ f3 cov2d_ewa(const f3& mean, float focal_x, float focal_y, float tan_fovx,
                                        float tan_fovy, const float* cov3D, const float* __restrict__ view)
{
    f3 t = xform_4x3(mean, view);

    const float limx = 1.3f * tan_fovx;
    const float limy = 1.3f * tan_fovy;
    const float txtz = t.x / t.z;
    const float tytz = t.y / t.z;
    t.x = min(limx, max(-limx, txtz)) * t.z;
    t.y = min(limy, max(-limy, tytz)) * t.z;

    mat3 J = mat3_cols(
        focal_x / t.z, 0.0f, -(focal_x * t.x) / (t.z * t.z),
        0.0f, focal_y / t.z, -(focal_y * t.y) / (t.z * t.z),
        0, 0, 0);

    mat3 Wm = mat3_cols(
        view[0], view[4], view[8],
        view[1], view[5], view[9],
        view[2], view[6], view[10]);

    mat3 T = mul(Wm, J);

    mat3 Vrk = mat3_cols(
        cov3D[0], cov3D[1], cov3D[2],
        cov3D[1], cov3D[3], cov3D[4],
        cov3D[2], cov3D[4], cov3D[5]);

    mat3 cov = mul(mul(transpose(T), transpose(Vrk)), T);

    cov.c[0][0] += 0.3f;
    cov.c[1][1] += 0.3f;
    return {cov.c[0][0], cov.c[0][1], cov.c[1][1]};
}

// One Gaussian.  Returns the number of tiles it touches (0 if culled, degenerate or off-screen)
// and its tile rectangle.
template <bool RAW>
__device__ __forceinline__ uint32_t
preprocess_one(int idx, int D, int M, int C,
                  const float* __restrict__ means3D,
                  const float* __restrict__ scales,
                  const float* __restrict__ rotations,
                  const float* __restrict__ opacities,
                  const float* __restrict__ shs,
                  const float* __restrict__ cov3D_precomp,
                  const uint8_t* __restrict__ tile_mask,
                  const float* __restrict__ colors_precomp,
                  const float* __restrict__ viewmatrix,
                  const float* __restrict__ projmatrix,
                  const float* __restrict__ cam_pos,
                  ViewParams vp, bool prefiltered, int scale_dim,
                  int* __restrict__ radii,
                  float2* __restrict__ means2D,
                  float* __restrict__ depths,
                  float* __restrict__ cov3Ds,
                  float* __restrict__ rgb,
                  uint8_t* __restrict__ clamped,
                  float4* __restrict__ conic_opacity,
                  uint2& rect_min, uint2& rect_max)
{
    // A Gaussian that leaves early keeps radius 0 / 0 tiles and is ignored downstream.
    radii[idx] = 0;

    f3 p_orig = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
    f3 p_view = xform_4x3(p_orig, viewmatrix);
    if (!prefiltered && !in_frustum_ndc(p_orig, projmatrix)) return 0;
    // RAW: `opacities`, `scales`, `rotations` hold the optimiser's raw parameters (common.cuh)
    const float opacity = RAW ? act_sigmoid(opacities[idx]) : opacities[idx];
    if (opacity < 0.0001f) return 0;

    float4 p_hom = xform_4x4(p_orig, projmatrix);
    float p_w = 1.0f / (p_hom.w + 0.0000001f);
    f3 p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};

    const float* cov3D;
    if (cov3D_precomp != nullptr) {
        cov3D = cov3D_precomp + idx * 6;
    } else {
        f3 s;
        float4 q = reinterpret_cast<const float4*>(rotations)[idx];
        if (RAW) {
            if (scale_dim == 1) {
                s.x = s.y = s.z = expf(scales[idx]);
            } else {
                s = {expf(scales[3 * idx]), expf(scales[3 * idx + 1]), expf(scales[3 * idx + 2])};
            }
            q = act_normalize(q);
        } else {
            s = {scales[3 * idx], scales[3 * idx + 1], scales[3 * idx + 2]};
        }
        cov3d_from_scale_rot(s, vp.scale_modifier, q, cov3Ds + idx * 6);
        cov3D = cov3Ds + idx * 6;
    }

    f3 cov = cov2d_ewa(p_orig, vp.focal_x, vp.focal_y, vp.tan_fovx, vp.tan_fovy, cov3D, viewmatrix);

    float det = (cov.x * cov.z - cov.y * cov.y);
    if (det <= 0.0f || cov.x <= 0.0f || cov.z <= 0.0f) {
        radii[idx] = -1;  // degenerate footprint: pruned with zero gradients
        return 0;
    }
    float det_inv = 1.f / det;
    f3 conic = {cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv};

    float mid = 0.5f * (cov.x + cov.z);
    float lambda1 = mid + sqrt(max(0.1f, mid * mid - det));
    float lambda2 = mid - sqrt(max(0.1f, mid * mid - det));
    if (lambda1 <= 0.01 || lambda2 <= 0.01 || lambda1 < lambda2 || (lambda1 / lambda2) > 10000.0) {
        radii[idx] = -1;
        return 0;
    }
    float my_radius = ceil(3.f * sqrt(max(lambda1, lambda2)));
    // A NaN covariance (a degenerate camera, e.g. an all-zero matrix, or non-finite inputs) passes every
    // comparison above and arrives here with a radius that converts to 0: without this line the
    // Gaussian would count tiles it never emits pairs for (the reference has that inconsistency and
    // then sorts uninitialised keys).  Finite inputs always give a radius >= 1, so nothing else changes.
    if (!((int)my_radius > 0)) return 0;

    float2 point_image = {ndc_to_pix(p_proj.x, vp.W), ndc_to_pix(p_proj.y, vp.H)};
    uint2 rmin, rmax;
    tile_rect(point_image.x, point_image.y, (int)my_radius, vp.grid_x, vp.grid_y, rmin, rmax);

    uint32_t touched = (rmax.y - rmin.y) * (rmax.x - rmin.x);
    if (touched == 0) return 0;  // off-screen: tiles_touched stays 0, radius stays 0

    if (tile_mask != nullptr) {
        touched = 0;
        for (uint32_t y = rmin.y; y < rmax.y; y++)
            for (uint32_t x = rmin.x; x < rmax.x; x++)
                if (tile_mask[y * vp.grid_x + x]) touched += 1;
        if (touched == 0) return 0;
    }
    rect_min = rmin;
    rect_max = rmax;

    conic_opacity[idx] = {conic.x, conic.y, conic.z, opacity};

    if (colors_precomp == nullptr) {
        f3 campos = {cam_pos[0], cam_pos[1], cam_pos[2]};
        f3 result = sh_to_rgb(idx, D, M, p_orig, campos, shs, clamped);
        rgb[idx * C + 0] = result.x;
        rgb[idx * C + 1] = result.y;
        rgb[idx * C + 2] = result.z;
    }

    depths[idx] = p_view.z;
    radii[idx] = my_radius;
    means2D[idx] = point_image;
    return touched;
}


// K1.  One thread per (Gaussian, view): blockIdx.y is the view.  Afterwards the pairs of every tile
// are counted for the tile-bucketed binning (one counter per 128-byte line, tiles numbered
// view-major).  A Gaussian covering up to 8 tiles adds its own counts; one covering more (near or
// very flat splats cover thousands) is counted by its whole warp.
template <bool RAW>
__global__ void __launch_bounds__(PREPROCESS_THREADS)
preprocess_kernel(int P, int D, int M, int C,
                  const float* __restrict__ means3D,
                  const float* __restrict__ scales,
                  const float* __restrict__ rotations,
                  const float* __restrict__ opacities,
                  const float* __restrict__ shs,
                  const float* __restrict__ cov3D_precomp,
                  const uint8_t* __restrict__ tile_mask,
                  const float* __restrict__ colors_precomp,
                  const float* __restrict__ viewmatrices,
                  const float* __restrict__ projmatrices,
                  const float* __restrict__ cam_positions,
                  ViewParams vp, bool prefiltered, int scale_dim,
                  int* __restrict__ radii,
                  float2* __restrict__ means2D,
                  float* __restrict__ depths,
                  float* __restrict__ cov3Ds,
                  float* __restrict__ rgb,
                  uint8_t* __restrict__ clamped,
                  float4* __restrict__ conic_opacity,
                  uint32_t* __restrict__ tiles_touched,
                  uint32_t* __restrict__ tile_counts)
{
    chain_wait();        // launched with launch_chained(): nothing before the previous stage is complete
    chain_release();
    constexpr uint32_t SOLO_MAX = 8;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    // this view's camera and its slices of the per-view arrays (cov3Ds is view-independent and shared)
    const size_t view = blockIdx.y, voff = view * (size_t)P;
    const float* viewmatrix = viewmatrices + 16 * view;
    const float* projmatrix = projmatrices + 16 * view;
    const float* cam_pos = cam_positions != nullptr ? cam_positions + 3 * view : nullptr;
    tile_counts += view * (size_t)vp.tiles * TILE_COUNTER_STRIDE;
    uint2 rmin = make_uint2(0, 0), rmax = make_uint2(0, 0);
    uint32_t touched = 0;
    if (idx < P) {
        touched = preprocess_one<RAW>(idx, D, M, C, means3D, scales, rotations, opacities, shs, cov3D_precomp, tile_mask,
                                 colors_precomp, viewmatrix, projmatrix, cam_pos, vp, prefiltered, scale_dim, radii + voff,
                                 means2D + voff, depths + voff, cov3Ds, rgb + 3 * voff, clamped + 3 * voff,
                                 conic_opacity + voff, rmin, rmax);
        tiles_touched[voff + idx] = touched;
    }

    auto count = [&](uint32_t t) {
        if (tile_mask == nullptr || tile_mask[t]) atomicAdd(&tile_counts[(size_t)t * TILE_COUNTER_STRIDE], 1u);
    };
    const uint32_t area = touched ? (rmax.x - rmin.x) * (rmax.y - rmin.y) : 0u;
    if (area > 0 && area <= SOLO_MAX) {
        for (uint32_t y = rmin.y; y < rmax.y; y++)
            for (uint32_t x = rmin.x; x < rmax.x; x++) count(y * vp.grid_x + x);
    }
    unsigned big = __ballot_sync(0xffffffffu, area > SOLO_MAX);
    while (big) {
        const int src = __ffs(big) - 1;
        big &= big - 1;
        const uint32_t x0 = __shfl_sync(0xffffffffu, rmin.x, src), y0 = __shfl_sync(0xffffffffu, rmin.y, src);
        const uint32_t w = __shfl_sync(0xffffffffu, rmax.x, src) - x0;
        const uint32_t n = __shfl_sync(0xffffffffu, area, src);
        for (uint32_t k = lane; k < n; k += 32) count((y0 + k / w) * vp.grid_x + x0 + k % w);
    }
}

__global__ void __launch_bounds__(256)
mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ viewmatrix,
                    const float* __restrict__ projmatrix, uint8_t* __restrict__ present)
{
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    f3 p = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
    present[idx] = in_frustum_ndc(p, projmatrix) ? 1 : 0;
}

int launch_preprocess(int P, int D, int M, int C, const float* means3D, const float* scales,
                      const float* rotations, const float* opacities, const float* shs,
                      const float* cov3D_precomp, const uint8_t* tile_mask, const float* colors_precomp,
                      const float* viewmatrices, const float* projmatrices, const float* cam_positions,
                      const ViewParams& vp, bool prefiltered, int* radii, GeomState& g,
                      uint32_t* tile_counts, const RawParams* raw, cudaStream_t stream)
{
    const dim3 grid(ceil_div(P, PREPROCESS_THREADS), vp.views);
#define SL_LAUNCH_PRE(RR, SD)                                                                                     \
    launch_chained(preprocess_kernel<RR>, grid, dim3(PREPROCESS_THREADS), 0, stream,                              \
        P, D, M, C, means3D, scales, rotations, opacities, shs, cov3D_precomp, tile_mask,                         \
        colors_precomp, viewmatrices, projmatrices, cam_positions, vp, prefiltered, SD, radii, g.means2D, g.depths, \
        g.cov3D, g.rgb, g.clamped, g.conic_opacity, g.tiles_touched, tile_counts)
    if (raw != nullptr) SL_LAUNCH_PRE(true, raw->scale_dim); else SL_LAUNCH_PRE(false, 3);
#undef SL_LAUNCH_PRE
    count_launch();
    return 0;
}

int launch_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                        uint8_t* present, cudaStream_t stream)
{
    mark_visible_kernel<<<ceil_div(P, 256), 256, 0, stream>>>(P, means3D, viewmatrix, projmatrix, present); count_launch();
    return 0;
}

}  // namespace sl
