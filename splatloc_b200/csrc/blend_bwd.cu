// splatloc_b200/csrc/blend_bwd.cu
//
// K7  backward alpha blending (reference: renderCUDA in
//     cuda_rasterizer/backward.cu:1018-1209).  Same per-pixel recurrences and
//     per-term arithmetic as the reference, walked back to front over the first
//     n_contrib entries of the tile's list; what changes is how the 9 + C
//     per-Gaussian partial gradients leave the CTA.
//
//     The reference issues 9 + C global float atomics per blended
//     (pixel, Gaussian) pair.  Here a warp owns an 8x4 pixel patch and streams
//     the tile's PairRecords back to front on its own (blend_common.cuh; same
//     chunked survivor pipeline as the forward, starting at the patch's largest
//     n_contrib); for every
//     Gaussian that at least one of its 32 pixels blends, the partials are
//     reduced across the warp with a 16-shuffle transpose-reduction (16 values,
//     halving the live set at each of the 5 butterfly levels) and the 9 + C sums
//     leave the warp as one `red.global.add.f32` per value, issued by 9 + C
//     different lanes.  Warps none of whose pixels blend the Gaussian skip it on
//     a ballot.  Global atomic traffic drops from (9+C) per blended pixel pair
//     to at most 8 x (9+C) per (tile, Gaussian) pair.
//
//     Channel rule kept from the fork: channels >= 3 do not feed dL/dalpha
//     (backward.cu:1153-1154) but do receive colour gradients.
#include "kernels.h"
#include "blend_common.cuh"

namespace sl {

namespace {

constexpr unsigned FULL = 0xffffffffu;

// Sum 16 per-lane values across the warp.  On return lane L holds in the result the
// warp-wide total of value (L >> 1).
__device__ __forceinline__ float warp_transpose_reduce16(float (&v)[16], int lane)
{
    {
        const bool up = lane & 16;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float send = up ? v[k] : v[k + 8];
            const float keep = up ? v[k + 8] : v[k];
            v[k] = keep + __shfl_xor_sync(FULL, send, 16);
        }
    }
    {
        const bool up = lane & 8;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float send = up ? v[k] : v[k + 4];
            const float keep = up ? v[k + 4] : v[k];
            v[k] = keep + __shfl_xor_sync(FULL, send, 8);
        }
    }
    {
        const bool up = lane & 4;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const float send = up ? v[k] : v[k + 2];
            const float keep = up ? v[k + 2] : v[k];
            v[k] = keep + __shfl_xor_sync(FULL, send, 4);
        }
    }
    {
        const bool up = lane & 2;
        const float send = up ? v[0] : v[1];
        const float keep = up ? v[1] : v[0];
        v[0] = keep + __shfl_xor_sync(FULL, send, 2);
    }
    return v[0] + __shfl_xor_sync(FULL, v[0], 1);
}

}  // namespace

template <int C>
__global__ void __launch_bounds__(BLEND_THREADS)
blend_backward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order,
                      const PairRecord* __restrict__ records, const uint8_t* __restrict__ masks,
                      int W, int H, int grid_x, const float* __restrict__ bg,
                      const float* __restrict__ out_alpha, const uint32_t* __restrict__ n_contrib,
                      const float* __restrict__ dL_dpix, const float* __restrict__ dL_dpix_depth,
                      const float* __restrict__ dL_dpix_alpha,
                      float* __restrict__ dL_dmean2D,     // (P,3)
                      float* __restrict__ dL_dabsmean2D,  // (P,3)
                      float* __restrict__ dL_dconic,      // (P,4)
                      float* __restrict__ dL_dopacity,    // (P,1)
                      float* __restrict__ dL_dcolor,      // (P,C)
                      float* __restrict__ dL_ddepth)      // (P,1)
{
    __shared__ PairRecord s_rec[BLEND_WARPS][BLEND_CHUNK];
    __shared__ uint8_t s_lane[BLEND_WARPS][BLEND_CHUNK];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = tile_order[blockIdx.x / BLEND_SUBS];
    const int patch = (blockIdx.x % BLEND_SUBS) * BLEND_WARPS + warp;
    const PatchGeom pg = patch_geom(tile, grid_x, patch, lane);
    const int px = pg.px, py = pg.py;
    const bool inside = px < W && py < H;
    const uint32_t pix_id = (uint32_t)W * py + px;
    const float pixf_x = (float)px, pixf_y = (float)py;
    const uint32_t rec_a = smem_addr(&s_rec[warp][0]), lane_a = smem_addr(&s_lane[warp][0]);
    const unsigned lt = (1u << lane) - 1u;

    const uint2 range = ranges[tile];
    const int lo = range.x;

    const float T_final = inside ? (1 - out_alpha[pix_id]) : 0;
    float T = T_final;
    const int last_contributor = inside ? n_contrib[pix_id] : 0;

    float accum_rec[C];
    float last_color[C];
    float dL_dpixel[C];
#pragma unroll
    for (int ch = 0; ch < C; ++ch) {
        accum_rec[ch] = 0.f;
        last_color[ch] = 0.f;
        dL_dpixel[ch] = inside ? dL_dpix[(size_t)ch * H * W + pix_id] : 0.f;
    }
    float accum_red = 0.f, accum_rea = 0.f;
    const float dL_dpixel_depth = inside ? dL_dpix_depth[pix_id] : 0.f;
    const float dL_dpixel_alpha = inside ? dL_dpix_alpha[pix_id] : 0.f;
    float last_alpha = 0.f, last_depth = 0.f;

    float bg_dot_dpixel = 0.f;
#pragma unroll
    for (int ch = 0; ch < C; ++ch) bg_dot_dpixel += bg[ch] * dL_dpixel[ch];
    const float neg_Tfinal_bg = -T_final * bg_dot_dpixel;

    const float ddelx_dx = 0.5 * W;
    const float ddely_dy = 0.5 * H;

    // Which output this lane flushes after a warp reduction: value index lane >> 1.
    //   0..C-1 colour | C depth | C+1,C+2 mean2D x,y | C+3,C+4 |mean2D| x,y |
    //   C+5..C+7 conic xx,xy,yy | C+8 opacity
    float* flush_base = nullptr;
    uint32_t flush_stride = 0;
    {
        const int v = lane >> 1;
        if ((lane & 1) == 0) {
            if (v < C) { flush_base = dL_dcolor + v; flush_stride = C; }
            else if (v == C) { flush_base = dL_ddepth; flush_stride = 1; }
            else if (v == C + 1) { flush_base = dL_dmean2D; flush_stride = 3; }
            else if (v == C + 2) { flush_base = dL_dmean2D + 1; flush_stride = 3; }
            else if (v == C + 3) { flush_base = dL_dabsmean2D; flush_stride = 3; }
            else if (v == C + 4) { flush_base = dL_dabsmean2D + 1; flush_stride = 3; }
            else if (v == C + 5) { flush_base = dL_dconic; flush_stride = 4; }
            else if (v == C + 6) { flush_base = dL_dconic + 1; flush_stride = 4; }
            else if (v == C + 7) { flush_base = dL_dconic + 3; flush_stride = 4; }
            else if (v == C + 8) { flush_base = dL_dopacity; flush_stride = 1; }
        }
    }

    // Entries at list positions >= the patch's largest n_contrib are never blended by it:
    // the walk starts there (positions are 0-based here, like the reference's `contributor`).
    int warp_last = last_contributor;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) warp_last = max(warp_last, __shfl_xor_sync(FULL, warp_last, o));
    if (warp_last == 0) return;

    // Back-propagate through the `cnt` records parked in the warp's buffer; `top_pos` is the list
    // position of the chunk's entry 0 (entries run towards the front of the list).
    auto backprop_parked = [&](int cnt, int top_pos) {
        for (int k = 0; k < cnt; ++k) {
            const uint32_t ra = rec_a + (uint32_t)k * (uint32_t)sizeof(PairRecord);
            const int pos = top_pos - (int)lds_u8(lane_a + k);
            const float4 xyd = lds128(ra);
            const float4 co = lds128(ra + 16);
            const float dx = __fsub_rn(xyd.x, pixf_x);
            const float dy = __fsub_rn(xyd.y, pixf_y);
            const float t1 = __fmul_rn(__fmul_rn(dy, co.z), dy);
            const float t3 = __fmul_rn(__fmul_rn(dx, co.y), dy);
            const float s = __fmaf_rn(dx, __fmul_rn(dx, co.x), t1);
            const float power = __fmaf_rn(s, -0.5f, -t3);
            const float G = exp_like_reference(power);
            const float alpha = fminf(0.99f, __fmul_rn(co.w, G));
            const bool blend = (pos < last_contributor) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
            if (!__any_sync(FULL, blend)) continue;

            float v[16];
#pragma unroll
            for (int q = 0; q < 16; ++q) v[q] = 0.f;

            if (blend) {
                // one reciprocal serves T / (1 - alpha) and T_final / (1 - alpha); 1 - alpha is in
                // [0.01, 1], so MUFU.RCP plus one Newton step is within an ulp of the reference's division
                const float om = 1.f - alpha;
                float inv;
                asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(om));
                inv = __fmaf_rn(__fmaf_rn(-om, inv, 1.f), inv, inv);
                T = T * inv;
                const float w = alpha * T;
                float dL_dalpha = 0.0f;
                const float4 f = lds128(ra + 32);
                const float c4[4] = {f.x, f.y, f.z, f.w};
#pragma unroll
                for (int ch = 0; ch < C; ++ch) {
                    const float c = c4[ch];
                    accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                    last_color[ch] = c;
                    const float dL_dchannel = dL_dpixel[ch];
                    if (ch <= 2) dL_dalpha += (c - accum_rec[ch]) * dL_dchannel;   // channels >= 3 are masked out
                    v[ch] = w * dL_dchannel;
                }
                const float dep = xyd.z;
                accum_red = last_alpha * last_depth + (1.f - last_alpha) * accum_red;
                last_depth = dep;
                dL_dalpha += (dep - accum_red) * dL_dpixel_depth;
                v[C] = w * dL_dpixel_depth;

                accum_rea = last_alpha + (1.f - last_alpha) * accum_rea;
                dL_dalpha += (1 - accum_rea) * dL_dpixel_alpha;

                dL_dalpha *= T;
                last_alpha = alpha;
                dL_dalpha += neg_Tfinal_bg * inv;

                const float dL_dG = co.w * dL_dalpha;
                const float gdx = G * dx;
                const float gdy = G * dy;
                const float dG_ddelx = -gdx * co.x - gdy * co.y;
                const float dG_ddely = -gdy * co.z - gdx * co.y;
                const float gx = dL_dG * dG_ddelx * ddelx_dx;
                const float gy = dL_dG * dG_ddely * ddely_dy;
                v[C + 1] = gx;
                v[C + 2] = gy;
                v[C + 3] = fabsf(gx);
                v[C + 4] = fabsf(gy);
                v[C + 5] = -0.5f * gdx * dx * dL_dG;
                v[C + 6] = -0.5f * gdx * dy * dL_dG;
                v[C + 7] = -0.5f * gdy * dy * dL_dG;
                v[C + 8] = G * dL_dalpha;
            }

            const float total_v = warp_transpose_reduce16(v, lane);
            if (flush_base != nullptr) {
                const int gid = __float_as_int(xyd.w);
                atomicAdd(flush_base + (uint32_t)gid * (uint32_t)flush_stride, total_v);
            }
        }
    };

    int parked = 0, parked_top = 0;
    const int start = lo + warp_last - 1;                 // global index of the first entry to visit
    uint32_t m_next = (start - lane >= lo) ? masks[start - lane] : 0u;

    for (int t = start; t >= lo; t -= BLEND_CHUNK) {
        const uint32_t m = m_next;
        const bool bit = (t - lane >= lo) && ((m >> patch) & 1u);
        float4 r0 = make_float4(0.f, 0.f, 0.f, 0.f), r1 = r0, r2 = r0;
        if (bit) {
            const float4* src = reinterpret_cast<const float4*>(records + (t - lane));
            r0 = __ldg(src);
            r1 = __ldg(src + 1);
            r2 = __ldg(src + 2);
        }
        m_next = (t - BLEND_CHUNK - lane >= lo) ? masks[t - BLEND_CHUNK - lane] : 0u;

        backprop_parked(parked, parked_top);

        __syncwarp();
        const unsigned b = __ballot_sync(0xffffffffu, bit);
        if (bit) {
            const int slot = __popc(b & lt);
            PairRecord* dst = &s_rec[warp][slot];
            dst->xyd = r0;
            dst->co = r1;
            dst->col = r2;
            s_lane[warp][slot] = (uint8_t)lane;
        }
        __syncwarp();
        parked = __popc(b);
        parked_top = t - lo;
    }
    backprop_parked(parked, parked_top);
}

int launch_blend_backward(int C, const ViewParams& vp, const uint2* ranges, const uint32_t* tile_order,
                          const PairRecord* records, const uint8_t* masks, const float* bg,
                          const float* out_alpha, const uint32_t* n_contrib, const float* dL_dpix,
                          const float* dL_dpix_depth, const float* dL_dpix_alpha, float* dL_dmean2D,
                          float* dL_dabsmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                          float* dL_ddepth, cudaStream_t stream)
{
    const int tiles = vp.grid_x * vp.grid_y;
    if (C == 4)
        blend_backward_kernel<4><<<tiles * BLEND_SUBS, BLEND_THREADS, 0, stream>>>(
            ranges, tile_order, records, masks, vp.W, vp.H, vp.grid_x, bg, out_alpha, n_contrib, dL_dpix,
            dL_dpix_depth, dL_dpix_alpha, dL_dmean2D, dL_dabsmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_ddepth);
    else if (C == 3)
        blend_backward_kernel<3><<<tiles * BLEND_SUBS, BLEND_THREADS, 0, stream>>>(
            ranges, tile_order, records, masks, vp.W, vp.H, vp.grid_x, bg, out_alpha, n_contrib, dL_dpix,
            dL_dpix_depth, dL_dpix_alpha, dL_dmean2D, dL_dabsmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_ddepth);
    else {
        set_error("blend_backward: unsupported channel count %d (3 or 4)", C);
        return -2;
    }
    count_launch();
    return 0;
}

}  // namespace sl
