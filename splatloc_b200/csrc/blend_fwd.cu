// splatloc_b200/csrc/blend_fwd.cu
//
// K6  forward alpha blending, one CTA per 16x16 tile, one thread per pixel
//     (reference: renderCUDA, cuda_rasterizer/forward.cu:844-963).
//
// Per (pixel, Gaussian) arithmetic is pinned with explicit round-to-nearest
// intrinsics to the operation sequence the reference compiles to, because the
// alpha / transmittance thresholds decide the integer output n_contrib:
//     power = fma(fma(dx, dx*A, (dy*C)*dy), -0.5, -((dx*B)*dy))
//     alpha = min(0.99, o * __expf(power));   T' = T * (1 - alpha)
//     acc  += T * (alpha * f)                 (fma)
// The Gaussian that drives T below 1e-4 is still blended, then the pixel stops
// (forward.cu:932-949).
//
// Layout: a warp owns an 8x4 pixel patch (not two 16-pixel rows) so that whole
// warps finish -- and skip the inner loop -- earlier; every batch of 256
// Gaussians stages position, depth, conic/opacity *and* features in shared
// memory, so the inner loop touches no global memory; each warp only visits the
// batch entries whose alpha-test bounding box reaches its patch
// (blend_common.cuh).  CTAs take tiles in order of decreasing list length.
#include "kernels.h"
#include "blend_common.cuh"

namespace sl {

template <int C>
__global__ void __launch_bounds__(BLEND_THREADS)
blend_forward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order,
                     const uint32_t* __restrict__ point_list,
                     int W, int H, int grid_x,
                     const float2* __restrict__ means2D, const float* __restrict__ features,
                     const float* __restrict__ depths, const float4* __restrict__ conic_opacity,
                     const float* __restrict__ bg, float* __restrict__ out_color,
                     float* __restrict__ out_depth, float* __restrict__ out_alpha,
                     uint32_t* __restrict__ n_contrib)
{
    __shared__ Staged s_rec[BLEND_BATCH];
    __shared__ uint8_t s_mask[BLEND_BATCH];               // bit w: may reach warp w's patch
    __shared__ uint8_t s_list[BLEND_WARPS][BLEND_BATCH];  // per-warp surviving batch indices

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = tile_order[blockIdx.x / BLEND_SUBS];
    const int patch = (blockIdx.x % BLEND_SUBS) * BLEND_WARPS + warp;   // this warp's patch of the tile
    const PatchGeom pg = patch_geom(tile, grid_x, patch, lane);
    const int tile_x0 = (tile % grid_x) * SL_TILE_X, tile_y0 = (tile / grid_x) * SL_TILE_Y;
    const int px = pg.px, py = pg.py;
    const bool inside = px < W && py < H;
    const uint32_t pix_id = (uint32_t)W * py + px;
    const float pixf_x = (float)px, pixf_y = (float)py;

    const uint32_t list_a = smem_addr(s_list[warp]), rec_a = smem_addr(s_rec);

    const uint2 range = ranges[tile];
    const int total = range.y - range.x;
    const int rounds = (total + BLEND_BATCH - 1) / BLEND_BATCH;

    bool done = !inside;
    float T = 1.0f;
    uint32_t last_contributor = 0;
    float acc[C];
#pragma unroll
    for (int ch = 0; ch < C; ++ch) acc[ch] = 0.f;
    float D = 0.f;

    for (int i = 0; i < rounds; ++i) {
        if (__syncthreads_count(done) == BLEND_THREADS) break;

        const int n = min(BLEND_BATCH, total - i * BLEND_BATCH);
        for (int t = tid; t < n; t += BLEND_THREADS) {
            const int id = point_list[range.x + i * BLEND_BATCH + t];
            const float2 xy = means2D[id];
            const float4 co = conic_opacity[id];
            float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
            if (C == 4) {
                f = reinterpret_cast<const float4*>(features)[id];
            } else {
                f.x = features[id * C + 0];
                if (C > 1) f.y = features[id * C + 1];
                if (C > 2) f.z = features[id * C + 2];
            }
            s_rec[t].xyd = make_float4(xy.x, xy.y, depths[id], __int_as_float(id));
            s_rec[t].co = co;
            s_rec[t].col = f;
            s_mask[t] = (uint8_t)patch_mask(xy.x, xy.y, co, tile_x0, tile_y0);
        }
        __syncthreads();

        // A warp whose 32 pixels are all finished only helps with staging.
        if (__all_sync(0xffffffffu, done)) continue;

        uint8_t* list = s_list[warp];
        const int cnt = compact_batch(list, n, lane, [&](int j) { return (s_mask[j] >> patch) & 1; });
        const uint32_t base = (uint32_t)i * BLEND_BATCH + 1u;   // list position (1-based) of batch entry 0

        for (int k0 = 0; k0 < cnt; k0 += 4) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int k = k0 + u;
                if (k < cnt && !done) {
                    const uint32_t j = lds_u8(list_a + k);
                    const uint32_t ra = rec_a + j * (uint32_t)sizeof(Staged);
                    const float4 xyd = lds128(ra);
                    const float4 co = lds128(ra + 16);
                    const float dx = __fsub_rn(xyd.x, pixf_x);
                    const float dy = __fsub_rn(xyd.y, pixf_y);
                    const float t1 = __fmul_rn(__fmul_rn(dy, co.z), dy);
                    const float t3 = __fmul_rn(__fmul_rn(dx, co.y), dy);
                    const float s = __fmaf_rn(dx, __fmul_rn(dx, co.x), t1);
                    const float power = __fmaf_rn(s, -0.5f, -t3);
                    if (!(power > 0.0f)) {
                        const float alpha = fminf(0.99f, __fmul_rn(co.w, exp_like_reference(power)));
                        if (!(alpha < 1.0f / 255.0f)) {
                            const float test_T = __fmul_rn(T, __fsub_rn(1.0f, alpha));
                            const float4 f = lds128(ra + 32);
                            acc[0] = __fmaf_rn(T, __fmul_rn(alpha, f.x), acc[0]);
                            if (C > 1) acc[1] = __fmaf_rn(T, __fmul_rn(alpha, f.y), acc[1]);
                            if (C > 2) acc[2] = __fmaf_rn(T, __fmul_rn(alpha, f.z), acc[2]);
                            if (C > 3) acc[3] = __fmaf_rn(T, __fmul_rn(alpha, f.w), acc[3]);
                            D = __fmaf_rn(T, __fmul_rn(alpha, xyd.z), D);
                            T = test_T;
                            last_contributor = base + j;
                            if (test_T < 0.0001f) done = true;
                        }
                    }
                }
            }
            if (__all_sync(0xffffffffu, done)) break;
        }
    }

    if (inside) {
        out_alpha[pix_id] = __fsub_rn(1.0f, T);
        n_contrib[pix_id] = last_contributor;
#pragma unroll
        for (int ch = 0; ch < C; ++ch)
            out_color[(size_t)ch * H * W + pix_id] = __fmaf_rn(bg[ch], T, acc[ch]);
        out_depth[pix_id] = D;
    }
}

int launch_blend_forward(int C, const ViewParams& vp, const uint2* ranges, const uint32_t* tile_order,
                         const uint32_t* point_list,
                         const float2* means2D, const float* features, const float* depths,
                         const float4* conic_opacity, const float* bg, float* out_color,
                         float* out_depth, float* out_alpha, uint32_t* n_contrib, cudaStream_t stream)
{
    const int tiles = vp.grid_x * vp.grid_y;
    if (C == 4)
        blend_forward_kernel<4><<<tiles * BLEND_SUBS, BLEND_THREADS, 0, stream>>>(ranges, tile_order, point_list, vp.W, vp.H, vp.grid_x,
            means2D, features, depths, conic_opacity, bg, out_color, out_depth, out_alpha, n_contrib);
    else if (C == 3)
        blend_forward_kernel<3><<<tiles * BLEND_SUBS, BLEND_THREADS, 0, stream>>>(ranges, tile_order, point_list, vp.W, vp.H, vp.grid_x,
            means2D, features, depths, conic_opacity, bg, out_color, out_depth, out_alpha, n_contrib);
    else {
        set_error("blend_forward: unsupported channel count %d (3 or 4)", C);
        return -2;
    }
    count_launch();
    return 0;
}

}  // namespace sl
