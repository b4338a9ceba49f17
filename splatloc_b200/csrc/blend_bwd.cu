// splatloc_b200/csrc/blend_bwd.cu
//
// K7  backward alpha blending (reference: renderCUDA in
//     cuda_rasterizer/backward.cu:1018-1209).  Same per-pixel recurrences as the
//     reference, walked back to front over the first n_contrib entries of the
//     tile's list; what changes is how the 9 + C per-Gaussian partial gradients
//     are summed over pixels and leave the CTA.
//
//     The reference issues 9 + C global float atomics per blended
//     (pixel, Gaussian) pair.  Here a warp owns an 8x4 pixel patch and streams
//     the tile's PairRecords back to front on its own (blend_common.cuh; same
//     chunked survivor pipeline as the forward, starting at the patch's largest
//     n_contrib).  The work on a Gaussian is split in two phases with the lanes
//     changing roles:
//
//     pixel phase (lane = pixel): evaluate alpha exactly as the forward did,
//       advance the pixel's recurrences (T, accum_rec/red/rea) and reduce
//       everything the pixel contributes to the Gaussian's gradients to two
//       scalars,  w = alpha * T  and  a = G * dL/dalpha;  Gaussians no pixel of
//       the patch blends are dropped on a ballot.  (w, a) of all 32 pixels go
//       to a row of a warp-private shared tile;
//     Gaussian phase (lane = Gaussian x half patch), once BWD_ROWS rows are
//       filled: every gradient is linear in (w, a) with coefficients that depend
//       only on the pixel offset (dx, dy) and on per-Gaussian constants,
//           dL/dcolour = sum w dL/dpix      dL/ddepth  = sum w dL/dpix_depth
//           dL/dmean2D = -o (W/2, H/2) . sum a Q d,     |.| likewise
//           dL/dconic  = -o/2 sum a (dx^2, dx dy, dy^2) dL/dopacity = sum a
//       so a lane walks 16 pixels of its row with plain FMAs into registers --
//       no cross-lane reduction per Gaussian -- the two halves are combined with
//       one shuffle per value and 9 + C `red.global.add.f32` leave per Gaussian.
//
//     Global atomic traffic drops from (9+C) per blended pixel pair to at most
//     8 x (9+C) per (tile, Gaussian) pair, and the pixel sums cost ~25 instead of
//     ~60 (shuffle butterfly) warp instructions per Gaussian.
//
//     Channel rule kept from the fork: channels >= 3 do not feed dL/dalpha
//     (backward.cu:1153-1154) but do receive colour gradients.
#include "kernels.h"
#include "blend_common.cuh"

namespace sl {

namespace {

constexpr unsigned FULL = 0xffffffffu;
constexpr int BWD_ROWS = 16;      // Gaussians per Gaussian-phase batch (x 2 half patches = 32 lanes)
constexpr int BWD_STRIDE = 33;    // row pitch of the (w, a) tiles in floats: conflict-free both ways
#ifndef SL_BWD_MIN_CTAS
#define SL_BWD_MIN_CTAS 4         // register budget 128; measured 206 / 211 / 214 us at 4 / 5 / 6 CTAs per SM
#endif

__device__ __forceinline__ float lds32(uint32_t a)
{
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
    return v;
}

}  // namespace

template <int C>
__global__ void __launch_bounds__(BLEND_THREADS, SL_BWD_MIN_CTAS)
blend_backward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order,
                      const PairRecord* __restrict__ records, const uint8_t* __restrict__ masks,
                      int W, int H, int grid_x, int tiles_per_view, int P, const float* __restrict__ bg,
                      const float* __restrict__ out_alpha, const uint32_t* __restrict__ n_contrib,
                      const float* __restrict__ dL_dpix, const float* __restrict__ dL_dpix_depth,
                      const float* __restrict__ dL_dpix_alpha,
                      float* __restrict__ dL_dmean2D,     // (views,P,3)
                      float* __restrict__ dL_dabsmean2D,  // (views,P,3)
                      float* __restrict__ dL_dconic,      // (views,P,4)
                      float* __restrict__ dL_dopacity,    // (P,1): summed over the views
                      float* __restrict__ dL_dcolor,      // (P,C) + view * color_view_stride
                      size_t color_view_stride,
                      float* __restrict__ dL_ddepth)      // (views,P,1)
{
    chain_wait();        // launched with launch_chained(): nothing before the previous stage is complete
    chain_release();
    __shared__ PairRecord s_rec[BLEND_WARPS][BLEND_CHUNK];
    __shared__ uint8_t s_lane[BLEND_WARPS][BLEND_CHUNK];
    __shared__ float s_w[BLEND_WARPS][BWD_ROWS * BWD_STRIDE];          // alpha * T per (row, pixel)
    __shared__ float s_a[BLEND_WARPS][BWD_ROWS * BWD_STRIDE];          // G * dL/dalpha per (row, pixel)
    __shared__ __align__(16) float4 s_grec[BLEND_WARPS][BWD_ROWS][2];  // {x, y, Qxx, Qxy}, {Qyy, o, id, -}
    __shared__ __align__(16) float4 s_dpix[BLEND_WARPS][32];           // dL/dpix, channels 0..3, per pixel
    __shared__ float s_ddep[BLEND_WARPS][32];                          // dL/dpix_depth per pixel

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = tile_order[blockIdx.x / BLEND_SUBS];          // global (view-major) tile number
    const int view = tile / tiles_per_view;
    const int patch = (blockIdx.x % BLEND_SUBS) * BLEND_WARPS + warp;
    const PatchGeom pg = patch_geom(tile - view * tiles_per_view, grid_x, patch, lane);
    {   // this view's images and per-view gradient accumulators
        const size_t N = (size_t)W * H, vP = (size_t)view * P;
        out_alpha += (size_t)view * N;
        n_contrib += (size_t)view * N;
        dL_dpix += (size_t)view * C * N;
        dL_dpix_depth += (size_t)view * N;
        dL_dpix_alpha += (size_t)view * N;
        dL_dmean2D += 3 * vP;
        dL_dabsmean2D += 3 * vP;
        dL_dconic += 4 * vP;
        dL_ddepth += vP;
        dL_dcolor += (size_t)view * color_view_stride;
    }
    const int px = pg.px, py = pg.py;
    const bool inside = px < W && py < H;
    const uint32_t pix_id = (uint32_t)W * py + px;
    const float pixf_x = (float)px, pixf_y = (float)py;
    const uint32_t rec_a = smem_addr(&s_rec[warp][0]), lane_a = smem_addr(&s_lane[warp][0]);
    const uint32_t w_a = smem_addr(&s_w[warp][0]), a_a = smem_addr(&s_a[warp][0]);
    const uint32_t grec_a = smem_addr(&s_grec[warp][0][0]);
    const uint32_t dpix_a = smem_addr(&s_dpix[warp][0]), ddep_a = smem_addr(&s_ddep[warp][0]);
    const unsigned lt = (1u << lane) - 1u;

    const uint2 range = ranges[tile];
    const int lo = range.x;

    const float T_final = inside ? (1 - out_alpha[pix_id]) : 0;
    float T = T_final;
    const int last_contributor = inside ? n_contrib[pix_id] : 0;

    float dL_dpixel[C];
#pragma unroll
    for (int ch = 0; ch < C; ++ch) dL_dpixel[ch] = inside ? dL_dpix[(size_t)ch * H * W + pix_id] : 0.f;
    const float dL_dpixel_depth = inside ? dL_dpix_depth[pix_id] : 0.f;
    const float dL_dpixel_alpha = inside ? dL_dpix_alpha[pix_id] : 0.f;

    // The reference carries one recurrence per output (accum_rec[ch], accum_red, accum_rea:
    // backward.cu:1139-1170), each the blend of everything behind the current Gaussian normalised
    // by the transmittance after it, and forms dL/dalpha = T sum_ch (c_ch - accum_ch) dL/dpix_ch
    // - T_final / (1 - alpha) bg . dL/dpix.  With  d_i = dL/dpix . (colour_i, depth_i, 1)  and
    // T_i accum_i = (sum_{j behind i} w_j c_j) / (1 - alpha_i)  that is the same as
    //     dL/dalpha_i = T_i d_i - behind_i / (1 - alpha_i),
    //     behind_i = T_final bg . dL/dpix + sum_{j behind i} w_j d_j,   w_j = alpha_j T_j,
    // one scalar recurrence with the same cancellation structure (both terms scale with T_i).
    float bg_dot_dpixel = 0.f;
#pragma unroll
    for (int ch = 0; ch < C; ++ch) bg_dot_dpixel += bg[ch] * dL_dpixel[ch];
    float behind = T_final * bg_dot_dpixel;

    const float ddelx_dx = 0.5 * W;
    const float ddely_dy = 0.5 * H;

    // Entries at list positions >= the patch's largest n_contrib are never blended by it:
    // the walk starts there (positions are 0-based here, like the reference's `contributor`).
    int warp_last = last_contributor;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) warp_last = max(warp_last, __shfl_xor_sync(FULL, warp_last, o));
    if (warp_last == 0) return;

    // per-pixel image gradients, read by every lane in the Gaussian phase
    s_dpix[warp][lane] = make_float4(dL_dpixel[0], dL_dpixel[1], dL_dpixel[2], C > 3 ? dL_dpixel[C - 1] : 0.f);
    s_ddep[warp][lane] = dL_dpixel_depth;

    // Gaussian phase over the n (<= BWD_ROWS) filled rows.  Lane = (row k, half patch h): 16 pixels each.
    auto gaussian_phase = [&](int n) {
        __syncwarp();
        const int k = lane & (BWD_ROWS - 1), half = lane >> 4;
        const float4 ga = lds128(grec_a + (uint32_t)k * 32u), gb = lds128(grec_a + (uint32_t)k * 32u + 16u);
        const float gx = ga.x, gy = ga.y, qxx = ga.z, qxy = ga.w, qyy = gb.x;
        const uint32_t row_w = w_a + (uint32_t)(k * BWD_STRIDE + half * 16) * 4u;
        const uint32_t row_a = a_a + (uint32_t)(k * BWD_STRIDE + half * 16) * 4u;
        const uint32_t dp = dpix_a + (uint32_t)half * 256u, dd = ddep_a + (uint32_t)half * 64u;
        const float x0f = pg.wx0, y0f = pg.wy0 + (float)(2 * half);
        float sc[C];
#pragma unroll
        for (int ch = 0; ch < C; ++ch) sc[ch] = 0.f;
        float sd = 0.f, sgx = 0.f, sgy = 0.f, sax = 0.f, say = 0.f, sxx = 0.f, sxy = 0.f, syy = 0.f, sa = 0.f;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const float w = lds32(row_w + 4u * j), a = lds32(row_a + 4u * j);
            const float4 dpx = lds128(dp + 16u * j);
            const float ddp = lds32(dd + 4u * j);
            const float dx = gx - (x0f + (float)(j & 7));
            const float dy = gy - (y0f + (float)(j >> 3));
            const float c4[4] = {dpx.x, dpx.y, dpx.z, dpx.w};
#pragma unroll
            for (int ch = 0; ch < C; ++ch) sc[ch] = fmaf(w, c4[ch], sc[ch]);
            sd = fmaf(w, ddp, sd);
            const float adx = a * dx, ady = a * dy;
            sxx = fmaf(adx, dx, sxx);
            sxy = fmaf(adx, dy, sxy);
            syy = fmaf(ady, dy, syy);
            const float ex = fmaf(adx, qxx, ady * qxy);    // a (Q d).x
            const float ey = fmaf(ady, qyy, adx * qxy);    // a (Q d).y
            sgx += ex;
            sgy += ey;
            sax += fabsf(ex);
            say += fabsf(ey);
            sa += a;
        }
#pragma unroll
        for (int ch = 0; ch < C; ++ch) sc[ch] += __shfl_xor_sync(FULL, sc[ch], 16);
        sd += __shfl_xor_sync(FULL, sd, 16);
        sgx += __shfl_xor_sync(FULL, sgx, 16);
        sgy += __shfl_xor_sync(FULL, sgy, 16);
        sax += __shfl_xor_sync(FULL, sax, 16);
        say += __shfl_xor_sync(FULL, say, 16);
        sxx += __shfl_xor_sync(FULL, sxx, 16);
        sxy += __shfl_xor_sync(FULL, sxy, 16);
        syy += __shfl_xor_sync(FULL, syy, 16);
        sa += __shfl_xor_sync(FULL, sa, 16);
        if (half == 0 && k < n) {
            const float o = gb.y;
            const uint32_t gid = (uint32_t)__float_as_int(gb.z);
#pragma unroll
            for (int ch = 0; ch < C; ++ch) atomicAdd(dL_dcolor + gid * C + ch, sc[ch]);
            atomicAdd(dL_ddepth + gid, sd);
            const float kx = o * ddelx_dx, ky = o * ddely_dy, kc = -0.5f * o;
            atomicAdd(dL_dmean2D + gid * 3u, -kx * sgx);
            atomicAdd(dL_dmean2D + gid * 3u + 1u, -ky * sgy);
            atomicAdd(dL_dabsmean2D + gid * 3u, kx * sax);
            atomicAdd(dL_dabsmean2D + gid * 3u + 1u, ky * say);
            atomicAdd(dL_dconic + gid * 4u, kc * sxx);
            atomicAdd(dL_dconic + gid * 4u + 1u, kc * sxy);
            atomicAdd(dL_dconic + gid * 4u + 3u, kc * syy);
            atomicAdd(dL_dopacity + gid, sa);
        }
        __syncwarp();
    };

    int rows = 0;

    // Pixel phase over the `cnt` records parked in the warp's buffer; `top_pos` is the list position
    // of the chunk's entry 0 (entries run towards the front of the list).
    // alpha of the k-th parked record for this lane's pixel, exactly as the forward evaluated it
    auto evaluate = [&](int k, int top_pos, float4& xyd, float4& co, float& G, float& alpha, bool& blend) {
        const uint32_t ra = rec_a + (uint32_t)k * (uint32_t)sizeof(PairRecord);
        const int pos = top_pos - (int)lds_u8(lane_a + k);
        xyd = lds128(ra);
        co = lds128(ra + 16);
        const float dx = __fsub_rn(xyd.x, pixf_x);
        const float dy = __fsub_rn(xyd.y, pixf_y);
        const float t1 = __fmul_rn(__fmul_rn(dy, co.z), dy);
        const float t3 = __fmul_rn(__fmul_rn(dx, co.y), dy);
        const float s = __fmaf_rn(dx, __fmul_rn(dx, co.x), t1);
        const float power = __fmaf_rn(s, -0.5f, -t3);
        G = exp_like_reference(power);
        alpha = fminf(0.99f, __fmul_rn(co.w, G));
        blend = (pos < last_contributor) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
    };
    // the pixel's recurrences for one record some pixel of the patch blends; fills one row
    auto apply = [&](int k, const float4& xyd, const float4& co, float G, float alpha, bool blend) {
        const uint32_t ra = rec_a + (uint32_t)k * (uint32_t)sizeof(PairRecord);
        float w = 0.f, a = 0.f;
        if (blend) {
            // one reciprocal serves T / (1 - alpha) and T_final / (1 - alpha); 1 - alpha is in
            // [0.01, 1], so MUFU.RCP plus one Newton step is within an ulp of the reference's division
            const float om = 1.f - alpha;
            float inv;
            asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(om));
            inv = __fmaf_rn(__fmaf_rn(-om, inv, 1.f), inv, inv);
            T = T * inv;
            w = alpha * T;
            // d = dL/dpix . (colour, depth, 1) of this Gaussian; channels >= 3 are masked out
            const float4 f = lds128(ra + 32);
            float d = fmaf(xyd.z, dL_dpixel_depth, dL_dpixel_alpha);
            d = fmaf(f.x, dL_dpixel[0], d);
            d = fmaf(f.y, dL_dpixel[1], d);
            d = fmaf(f.z, dL_dpixel[2], d);
            const float dL_dalpha = fmaf(T, d, -(behind * inv));
            behind = fmaf(w, d, behind);
            a = G * dL_dalpha;
        }
        s_w[warp][rows * BWD_STRIDE + lane] = w;
        s_a[warp][rows * BWD_STRIDE + lane] = a;
        if (lane == 0) {
            s_grec[warp][rows][0] = make_float4(xyd.x, xyd.y, co.x, co.y);
            s_grec[warp][rows][1] = make_float4(co.z, co.w, xyd.w, 0.f);
        }
        if (++rows == BWD_ROWS) {
            gaussian_phase(rows);
            rows = 0;
        }
    };

    // Pixel phase over the `cnt` records parked in the warp's buffer; `top_pos` is the list position
    // of the chunk's entry 0 (entries run towards the front of the list).  Two records are evaluated
    // together (their alphas do not depend on each other), then applied in list order.
    auto pixel_phase = [&](int cnt, int top_pos, bool final) {
        int k = 0;
        for (; k + 1 < cnt; k += 2) {
            float4 xa, ca, xb, cb;
            float Ga, aa, Gb, ab;
            bool ba, bb;
            evaluate(k, top_pos, xa, ca, Ga, aa, ba);
            evaluate(k + 1, top_pos, xb, cb, Gb, ab, bb);
            const bool any_a = __any_sync(FULL, ba), any_b = __any_sync(FULL, bb);
            if (any_a) apply(k, xa, ca, Ga, aa, ba);
            if (any_b) apply(k + 1, xb, cb, Gb, ab, bb);
        }
        if (k < cnt) {
            float4 xa, ca;
            float Ga, aa;
            bool ba;
            evaluate(k, top_pos, xa, ca, Ga, aa, ba);
            if (__any_sync(FULL, ba)) apply(k, xa, ca, Ga, aa, ba);
        }
        if (final && rows > 0) {         // after the last record: whatever rows are left
            gaussian_phase(rows);
            rows = 0;
        }
    };

    int parked = 0, parked_top = 0;
    const int start = lo + warp_last - 1;                 // global index of the first entry to visit
    uint32_t m_next = (start - lane >= lo) ? masks[start - lane] : 0u;

    for (int t = start;; t -= BLEND_CHUNK) {
        const bool more = t >= lo;                        // one extra pass drains the last parked chunk
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

        pixel_phase(parked, parked_top, !more);
        if (!more) break;

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
}

int launch_blend_backward(int P, int C, const ViewParams& vp, const uint2* ranges, const uint32_t* tile_order,
                          const PairRecord* records, const uint8_t* masks, const float* bg,
                          const float* out_alpha, const uint32_t* n_contrib, const float* dL_dpix,
                          const float* dL_dpix_depth, const float* dL_dpix_alpha, float* dL_dmean2D,
                          float* dL_dabsmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                          size_t color_view_stride, float* dL_ddepth, cudaStream_t stream)
{
    const int tiles = vp.tiles * vp.views;
    if (C == 4)
        launch_chained(blend_backward_kernel<4>, dim3(tiles * BLEND_SUBS), dim3(BLEND_THREADS), 0, stream,
            ranges, tile_order, records, masks, vp.W, vp.H, vp.grid_x, vp.tiles, P, bg, out_alpha, n_contrib, dL_dpix,
            dL_dpix_depth, dL_dpix_alpha, dL_dmean2D, dL_dabsmean2D, dL_dconic, dL_dopacity, dL_dcolor,
            color_view_stride, dL_ddepth);
    else if (C == 3)
        launch_chained(blend_backward_kernel<3>, dim3(tiles * BLEND_SUBS), dim3(BLEND_THREADS), 0, stream,
            ranges, tile_order, records, masks, vp.W, vp.H, vp.grid_x, vp.tiles, P, bg, out_alpha, n_contrib, dL_dpix,
            dL_dpix_depth, dL_dpix_alpha, dL_dmean2D, dL_dabsmean2D, dL_dconic, dL_dopacity, dL_dcolor,
            color_view_stride, dL_ddepth);
    else {
        set_error("blend_backward: unsupported channel count %d (3 or 4)", C);
        return -2;
    }
    count_launch();
    return 0;
}

}  // namespace sl
