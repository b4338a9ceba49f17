// splatloc_b200/csrc/blend_fwd.cu
//
// K6  forward alpha blending (reference: renderCUDA, cuda_rasterizer/forward.cu:844-963).
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
// Structure (see blend_common.cuh): one thread per pixel, a warp per 8x4 patch,
// each warp streaming the tile's PairRecords on its own.  Per 32-entry chunk a
// warp (1) ballots the mask bytes it prefetched, (2) issues the record loads of
// the survivors, (3) blends the survivors of the PREVIOUS chunk out of its
// private shared buffer while those loads fly, (4) parks the new records in the
// buffer.  A warp leaves as soon as its 32 pixels are saturated.  CTAs (half
// tiles) are launched longest list first.
#include "kernels.h"
#include "blend_common.cuh"

namespace sl {

template <int C, bool TOUCH>
__global__ void __launch_bounds__(BLEND_THREADS)
blend_forward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order,
                     const PairRecord* __restrict__ records, const uint8_t* __restrict__ masks,
                     int W, int H, int grid_x, int tiles_per_view, int P, const float* __restrict__ bg,
                     float* __restrict__ out_color, float* __restrict__ out_depth,
                     float* __restrict__ out_alpha, uint32_t* __restrict__ n_contrib,
                     int* __restrict__ n_touched)
{
    chain_wait();        // launched with launch_chained(): nothing before the previous stage is complete
    chain_release();
    __shared__ PairRecord s_rec[BLEND_WARPS][BLEND_CHUNK];
    __shared__ uint8_t s_lane[BLEND_WARPS][BLEND_CHUNK];   // chunk-relative index of each parked record

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = tile_order[blockIdx.x / BLEND_SUBS];          // global (view-major) tile number
    const int view = tile / tiles_per_view;
    const int patch = (blockIdx.x % BLEND_SUBS) * BLEND_WARPS + warp;
    const PatchGeom pg = patch_geom(tile - view * tiles_per_view, grid_x, patch, lane);
    const int px = pg.px, py = pg.py;
    const bool inside = px < W && py < H;
    const uint32_t pix_id = (uint32_t)W * py + px;
    const float pixf_x = (float)px, pixf_y = (float)py;
    const uint32_t rec_a = smem_addr(&s_rec[warp][0]), lane_a = smem_addr(&s_lane[warp][0]);
    const unsigned lt = (1u << lane) - 1u;

    const uint2 range = ranges[tile];
    const int lo = range.x, hi = range.y;
    {   // this view's images
        const size_t N = (size_t)W * H;
        out_color += (size_t)view * C * N;
        out_depth += (size_t)view * N;
        out_alpha += (size_t)view * N;
        n_contrib += (size_t)view * N;
        if (TOUCH) n_touched += (size_t)view * P;
    }

    bool done = !inside;
    float T = 1.0f;
    uint32_t last_contributor = 0;
    float acc[C];
#pragma unroll
    for (int ch = 0; ch < C; ++ch) acc[ch] = 0.f;
    float D = 0.f;

    // Blend the `cnt` records parked in the warp's buffer; `first_pos` is the 1-based list
    // position of the chunk's entry 0.  Returns true when all 32 pixels are finished.
    auto blend_parked = [&](int cnt, uint32_t first_pos) -> bool {
        for (int k0 = 0; k0 < cnt; k0 += 4) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int k = k0 + u;
                if (k < cnt && !done) {
                    const uint32_t ra = rec_a + (uint32_t)k * (uint32_t)sizeof(PairRecord);
                    const float4 xyd = lds128(ra);
                    const float4 co = lds128(ra + 16);
                    const float dx = __fsub_rn(xyd.x, pixf_x);
                    const float dy = __fsub_rn(xyd.y, pixf_y);
                    const float t1 = __fmul_rn(__fmul_rn(dy, co.z), dy);
                    const float t3 = __fmul_rn(__fmul_rn(dx, co.y), dy);
                    const float s = __fmaf_rn(dx, __fmul_rn(dx, co.x), t1);
                    const float power = __fmaf_rn(s, -0.5f, -t3);
                    bool touched = false;
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
                            last_contributor = first_pos + lds_u8(lane_a + k);
                            if (test_T < 0.0001f) done = true;
                            touched = test_T > 0.5f;
                        }
                    }
                    if (TOUCH) {
                        // extension (MonoGS-style n_touched): pixels in which this Gaussian is blended
                        // while more than half of the light still passes behind it
                        const unsigned tb = __ballot_sync(__activemask(), touched);
                        if (touched && (tb & ((1u << (threadIdx.x & 31)) - 1u)) == 0)
                            atomicAdd(&n_touched[__float_as_int(xyd.w)], __popc(tb));
                    }
                }
            }
            if (__all_sync(0xffffffffu, done)) return true;
        }
        return false;
    };

    bool finished = __all_sync(0xffffffffu, done);
    int parked = 0;            // records waiting in the buffer
    uint32_t parked_pos = 0;   // 1-based list position of their chunk's entry 0
    uint32_t m_next = (!finished && lo + lane < hi) ? masks[lo + lane] : 0u;

    for (int c = lo; c < hi && !finished; c += BLEND_CHUNK) {
        // (1) survivors of this chunk, (2) their records in flight
        const uint32_t m = m_next;
        const bool bit = (c + lane < hi) && ((m >> patch) & 1u);
        float4 r0 = make_float4(0.f, 0.f, 0.f, 0.f), r1 = r0, r2 = r0;
        if (bit) {
            const float4* src = reinterpret_cast<const float4*>(records + (c + lane));
            r0 = __ldg(src);
            r1 = __ldg(src + 1);
            r2 = __ldg(src + 2);
        }
        m_next = (c + BLEND_CHUNK + lane < hi) ? masks[c + BLEND_CHUNK + lane] : 0u;

        // (3) blend what the previous step parked
        finished = blend_parked(parked, parked_pos);
        if (finished) break;

        // (4) park this chunk's survivors, in list order
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
        parked_pos = (uint32_t)(c - lo) + 1u;
    }
    if (!finished) blend_parked(parked, parked_pos);

    if (inside) {
        out_alpha[pix_id] = __fsub_rn(1.0f, T);
        n_contrib[pix_id] = last_contributor;
#pragma unroll
        for (int ch = 0; ch < C; ++ch)
            out_color[(size_t)ch * H * W + pix_id] = __fmaf_rn(bg[ch], T, acc[ch]);
        out_depth[pix_id] = D;
    }
}

// ---------------------------------------------------------------------------------------------
// A/B variant (sl_set_record_stream(1)): the tile's records are staged in shared memory by the
// bulk-copy engine -- one `cp.async.bulk.shared.global` of a whole 32-record chunk (1536 B) per
// warp and stage, completion on an mbarrier, two stages -- instead of per-lane 16-byte loads of the
// mask survivors.  The copy engine moves every record of the chunk, survivors or not (~3x the
// bytes on the bench scene), but costs one instruction and no registers in flight; survivors are
// then blended straight out of the staged chunk in list order.  Same arithmetic, same results.
// Kept for the comparison profiles/README.md reports; the per-lane version is the default because
// it measured faster.
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@!p bra WAIT_%=;\n\t}"
        ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

template <int C, bool TOUCH>
__global__ void __launch_bounds__(BLEND_THREADS)
blend_forward_bulk_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order,
                          const PairRecord* __restrict__ records, const uint8_t* __restrict__ masks,
                          int W, int H, int grid_x, int tiles_per_view, int P, const float* __restrict__ bg,
                          float* __restrict__ out_color, float* __restrict__ out_depth,
                          float* __restrict__ out_alpha, uint32_t* __restrict__ n_contrib,
                          int* __restrict__ n_touched)
{
    chain_wait();        // launched with launch_chained(): nothing before the previous stage is complete
    chain_release();
    constexpr int STAGES = 2;
    __shared__ __align__(128) PairRecord s_stage[BLEND_WARPS][STAGES][BLEND_CHUNK];
    __shared__ __align__(8) unsigned long long s_bar[BLEND_WARPS][STAGES];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = tile_order[blockIdx.x / BLEND_SUBS];
    const int view = tile / tiles_per_view;
    const int patch = (blockIdx.x % BLEND_SUBS) * BLEND_WARPS + warp;
    const PatchGeom pg = patch_geom(tile - view * tiles_per_view, grid_x, patch, lane);
    const int px = pg.px, py = pg.py;
    const bool inside = px < W && py < H;
    const uint32_t pix_id = (uint32_t)W * py + px;
    const float pixf_x = (float)px, pixf_y = (float)py;
    const uint2 range = ranges[tile];
    const int lo = range.x, hi = range.y;
    {
        const size_t N = (size_t)W * H;
        out_color += (size_t)view * C * N;
        out_depth += (size_t)view * N;
        out_alpha += (size_t)view * N;
        n_contrib += (size_t)view * N;
        if (TOUCH) n_touched += (size_t)view * P;
    }
    const uint32_t stage_a = smem_addr(&s_stage[warp][0][0]), bar_a = smem_addr(&s_bar[warp][0]);
    constexpr uint32_t STAGE_BYTES = BLEND_CHUNK * sizeof(PairRecord);
    if (lane == 0) {
        for (int st = 0; st < STAGES; ++st) mbar_init(bar_a + 8u * st, 1u);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    auto issue = [&](int chunk) {                 // lane 0: one bulk copy of chunk `chunk` into its stage
        const int c = lo + chunk * BLEND_CHUNK;
        if (lane == 0 && c < hi) {
            const uint32_t bytes = (uint32_t)min(BLEND_CHUNK, hi - c) * (uint32_t)sizeof(PairRecord);
            const uint32_t st = (uint32_t)chunk % STAGES;
            mbar_expect_tx(bar_a + 8u * st, bytes);
            bulk_load(stage_a + st * STAGE_BYTES, records + c, bytes, bar_a + 8u * st);
        }
    };

    bool done = !inside;
    float T = 1.0f;
    uint32_t last_contributor = 0;
    float acc[C];
#pragma unroll
    for (int ch = 0; ch < C; ++ch) acc[ch] = 0.f;
    float D = 0.f;

    bool finished = __all_sync(0xffffffffu, done);
    const int n_chunks = (hi - lo + BLEND_CHUNK - 1) / BLEND_CHUNK;
    int issued = 0, consumed = 0;
    if (!finished && lo < hi) {
        issue(0);
        issue(1);
        issued = min(STAGES, n_chunks);
        uint32_t m_next = (lo + lane < hi) ? masks[lo + lane] : 0u;
        for (int chunk = 0, c = lo; c < hi && !finished; ++chunk, c += BLEND_CHUNK) {
            const uint32_t m = m_next;
            m_next = (c + BLEND_CHUNK + lane < hi) ? masks[c + BLEND_CHUNK + lane] : 0u;
            unsigned b = __ballot_sync(0xffffffffu, (c + lane < hi) && ((m >> patch) & 1u));
            const uint32_t st = (uint32_t)chunk % STAGES;
            mbar_wait(bar_a + 8u * st, ((uint32_t)chunk / STAGES) & 1u);
            const uint32_t base = stage_a + st * STAGE_BYTES;
            while (b) {
                const int k = __ffs(b) - 1;
                b &= b - 1;
                if (!done) {
                    const uint32_t ra = base + (uint32_t)k * (uint32_t)sizeof(PairRecord);
                    const float4 xyd = lds128(ra);
                    const float4 co = lds128(ra + 16);
                    const float dx = __fsub_rn(xyd.x, pixf_x);
                    const float dy = __fsub_rn(xyd.y, pixf_y);
                    const float t1 = __fmul_rn(__fmul_rn(dy, co.z), dy);
                    const float t3 = __fmul_rn(__fmul_rn(dx, co.y), dy);
                    const float s = __fmaf_rn(dx, __fmul_rn(dx, co.x), t1);
                    const float power = __fmaf_rn(s, -0.5f, -t3);
                    bool touched = false;
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
                            last_contributor = (uint32_t)(c - lo) + 1u + (uint32_t)k;
                            if (test_T < 0.0001f) done = true;
                            touched = test_T > 0.5f;
                        }
                    }
                    if (TOUCH) {
                        const unsigned tb = __ballot_sync(__activemask(), touched);
                        if (touched && (tb & ((1u << lane) - 1u)) == 0)
                            atomicAdd(&n_touched[__float_as_int(xyd.w)], __popc(tb));
                    }
                }
            }
            consumed = chunk + 1;
            finished = __all_sync(0xffffffffu, done);     // (also: every lane is done reading this stage)
            if (!finished && chunk + STAGES < n_chunks) {
                issue(chunk + STAGES);
                issued = chunk + STAGES + 1;
            }
        }
    }
    // a bulk copy still in flight must land before the CTA gives its shared memory back
    for (int chunk = consumed; chunk < issued; ++chunk)
        mbar_wait(bar_a + 8u * ((uint32_t)chunk % STAGES), ((uint32_t)chunk / STAGES) & 1u);

    if (inside) {
        out_alpha[pix_id] = __fsub_rn(1.0f, T);
        n_contrib[pix_id] = last_contributor;
#pragma unroll
        for (int ch = 0; ch < C; ++ch)
            out_color[(size_t)ch * H * W + pix_id] = __fmaf_rn(bg[ch], T, acc[ch]);
        out_depth[pix_id] = D;
    }
}

int launch_blend_forward(int P, int C, const ViewParams& vp, const uint2* ranges, const uint32_t* tile_order,
                         const PairRecord* records, const uint8_t* masks, const float* bg, float* out_color,
                         float* out_depth, float* out_alpha, uint32_t* n_contrib, int* n_touched,
                         int record_stream, cudaStream_t stream)
{
    const int tiles = vp.tiles * vp.views;
#define SL_LAUNCH_FWD(KK, CC, TT)                                                                                \
    launch_chained(KK<CC, TT>, dim3(tiles * BLEND_SUBS), dim3(BLEND_THREADS), 0, stream,                         \
        ranges, tile_order, records, masks, vp.W, vp.H, vp.grid_x, vp.tiles, P, bg, out_color, out_depth, out_alpha, \
        n_contrib, n_touched)
    if (C != 3 && C != 4) {
        set_error("blend_forward: unsupported channel count %d (3 or 4)", C);
        return -2;
    }
    if (record_stream == 1) {
        if (C == 4) { if (n_touched) SL_LAUNCH_FWD(blend_forward_bulk_kernel, 4, true); else SL_LAUNCH_FWD(blend_forward_bulk_kernel, 4, false); }
        else        { if (n_touched) SL_LAUNCH_FWD(blend_forward_bulk_kernel, 3, true); else SL_LAUNCH_FWD(blend_forward_bulk_kernel, 3, false); }
    } else {
        if (C == 4) { if (n_touched) SL_LAUNCH_FWD(blend_forward_kernel, 4, true); else SL_LAUNCH_FWD(blend_forward_kernel, 4, false); }
        else        { if (n_touched) SL_LAUNCH_FWD(blend_forward_kernel, 3, true); else SL_LAUNCH_FWD(blend_forward_kernel, 3, false); }
    }
#undef SL_LAUNCH_FWD
    count_launch();
    return 0;
}

}  // namespace sl
