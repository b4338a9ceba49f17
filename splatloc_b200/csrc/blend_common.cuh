// splatloc_b200/csrc/blend_common.cuh -- pieces shared by the forward and backward blend kernels.
//
// Work decomposition.  A warp owns an 8x4 pixel patch of a 16x16 tile and walks the tile's sorted
// list on its own: no shared staging between warps, no CTA barrier.  The list arrives as an array
// of 48-byte PairRecords (state.cuh) plus one mask byte per pair, both written once per frame by
// the binning stage (binning.cu).  Bit w of the mask says whether the Gaussian can pass the
// reference's alpha test anywhere in patch w:
//     alpha = min(0.99, o * exp(power)) >= 1/255   <=>   d^T Q d <= 2 ln(255 o),  Q = conic
// The mask byte holds the axis-aligned bounding box test of that ellipse, half-widths
//     hx = sqrt(2 ln(255 o) * Qyy / det Q),  hy = sqrt(2 ln(255 o) * Qxx / det Q);
// refined, for the patches the box touches, by the exact minimum of d^T Q d over the patch
// rectangle (patch_mask) -- everything padded by 0.2-0.4 % + 0.05 px and >= 2e-5 in power
// units, far more than any float32 / ex2.approx error.  A warp reads 32 mask bytes per step, ballots the survivors, fetches only their records
// into a warp-private shared buffer (while it blends the previous chunk) and evaluates those with
// the reference's exact arithmetic.  The test is conservative and list positions are kept, so
// colour, depth, alpha and n_contrib stay bit-identical -- the reference would have `continue`d on
// the dropped entries (forward.cu:925-931).  On the config-2 scene ~64 % of (patch, Gaussian)
// evaluations are dropped (tools/scene_stats.py).
#pragma once
#include "common.cuh"
#include "state.cuh"

namespace sl {

constexpr int BLEND_CHUNK = 32;            // list entries examined per warp step
// A CTA is BLEND_WARPS of a tile's 8 patches (a half tile by default).  Warps never synchronise
// with each other; the CTA size only sets the granularity at which work is handed to the SMs.
#ifndef SL_BLEND_CTA_WARPS
#define SL_BLEND_CTA_WARPS 4
#endif
constexpr int BLEND_WARPS = SL_BLEND_CTA_WARPS;
constexpr int BLEND_THREADS = BLEND_WARPS * 32;
constexpr int BLEND_SUBS = 8 / BLEND_WARPS;   // CTAs per tile
static_assert(BLEND_WARPS == 1 || BLEND_WARPS == 2 || BLEND_WARPS == 4 || BLEND_WARPS == 8, "patches per CTA");

struct PatchGeom {
    int px, py;          // this thread's pixel
    float wx0, wx1;      // this warp's patch bounds in pixel centres (inclusive)
    float wy0, wy1;
};

// `warp` is the patch index inside the tile, 0..7 (two columns x four rows of 8x4 pixels).
__device__ __forceinline__ PatchGeom patch_geom(int tile, int grid_x, int warp, int lane)
{
    const int tile_x = tile % grid_x, tile_y = tile / grid_x;
    const int x0 = tile_x * SL_TILE_X + (warp & 1) * 8;
    const int y0 = tile_y * SL_TILE_Y + (warp >> 1) * 4;
    PatchGeom g;
    g.px = x0 + (lane & 7);
    g.py = y0 + (lane >> 3);
    g.wx0 = (float)x0; g.wx1 = (float)(x0 + 7);
    g.wy0 = (float)y0; g.wy1 = (float)(y0 + 3);
    return g;
}

// Exact refinement of the reach mask: can d^T Q d drop to `limit` (2 ln(255 o), padded) anywhere on
// a patch rectangle (padded by 0.05 px)?  The form is convex with its minimum at the Gaussian's
// centre, so if the centre is outside the rectangle the minimum lies on an edge facing the centre (at
// most two), where it is the clamped vertex of a 1-D parabola.
// Bit w of the result: the Gaussian may reach a pixel of warp w's 8x4 patch.
__device__ __forceinline__ uint32_t patch_mask(float cx, float cy, const float4& co, int tile_x0, int tile_y0)
{
    // o <= 1/255 can never pass alpha >= 1/255 (exp(power) <= 1 wherever power <= 0)
    // +2e-5: absolute pad in `power` units (a 1e-5 relative margin on alpha), which also covers the
    // rounding of 255 * o and of logf when o is within float precision of 1/255
    const float t2 = 2.0f * logf(255.0f * co.w) + 2e-5f;
    if (!(t2 > 0.f)) return 0u;
    const float det = co.x * co.z - co.y * co.y;
    if (!(det > 0.f)) return 0xffu;   // not a proper ellipse: let the exact test decide
    const float inv = __fdividef(t2, det);
    const float hx = sqrtf(inv * co.z) * 1.002f + 0.05f;
    const float hy = sqrtf(inv * co.x) * 1.002f + 0.05f;
    if (!(hx == hx) || !(hy == hy)) return 0xffu;
    const float lx = cx - hx - (float)tile_x0, rx = cx + hx - (float)tile_x0;   // box in tile coordinates
    const float ly = cy - hy - (float)tile_y0, ry = cy + hy - (float)tile_y0;
    // columns: patch x-ranges [0,7] and [8,15]; rows: y-ranges [0,3], [4,7], [8,11], [12,15]
    uint32_t cols = 0u, rows = 0u;
    if (rx >= 0.f && lx <= 7.f) cols |= 1u;
    if (rx >= 8.f && lx <= 15.f) cols |= 2u;
    if (ry >= 0.f && ly <= 3.f) rows |= 1u;
    if (ry >= 4.f && ly <= 7.f) rows |= 2u;
    if (ry >= 8.f && ly <= 11.f) rows |= 4u;
    if (ry >= 12.f && ly <= 15.f) rows |= 8u;
    // warp w = 2 * row + col
    uint32_t m = 0u;
#pragma unroll
    for (int r = 0; r < 4; ++r)
        if (rows & (1u << r)) m |= cols << (2 * r);

    if (m == 0u) return 0u;

    // Exact test for all 8 patches in straight-line code: the per-column and
    // per-row terms are shared, and there is no data-dependent loop for the warp to diverge on.
    const float limit = t2 * 1.004f + 1e-4f;
    const float ux = cx - (float)tile_x0, uy = cy - (float)tile_y0;      // centre in tile coordinates
    const float rcx = __frcp_rn(co.x), rcz = __frcp_rn(co.z);
    const float two_b = 2.0f * co.y;
    float cx0[2], cx1[2], ue[2], tx[2], ax[2];
    bool in_x[2];
#pragma unroll
    for (int c = 0; c < 2; ++c) {
        cx0[c] = (float)(8 * c) - 0.05f - ux;             // patch column relative to the centre
        cx1[c] = cx0[c] + 7.1f;
        in_x[c] = cx0[c] <= 0.f && cx1[c] >= 0.f;
        ue[c] = cx0[c] > 0.f ? cx0[c] : cx1[c];           // the vertical edge facing the centre
        tx[c] = -co.y * ue[c] * rcz;                      // vertex of the parabola along that edge
        ax[c] = co.x * ue[c] * ue[c];
    }
    uint32_t exact = 0u;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const float y0 = (float)(4 * r) - 0.05f - uy, y1 = y0 + 3.1f;
        const bool in_y = y0 <= 0.f && y1 >= 0.f;
        const float ve = y0 > 0.f ? y0 : y1;              // the horizontal edge facing the centre
        const float ty = -co.y * ve * rcx;
        const float cy2 = co.z * ve * ve;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const float v = fminf(fmaxf(tx[c], y0), y1);
            const float best_x = in_x[c] ? 3.4e38f : ax[c] + (two_b * ue[c] + co.z * v) * v;
            const float u = fminf(fmaxf(ty, cx0[c]), cx1[c]);
            const float best_y = in_y ? 3.4e38f : cy2 + (two_b * ve + co.x * u) * u;
            const bool reached = (in_x[c] && in_y) || !(fminf(best_x, best_y) > limit);
            exact |= (reached ? 1u : 0u) << (2 * r + c);
        }
    }
    return m & exact;
}

// The feature channels of Gaussian `id` as the blend kernels read them.  `kp` != nullptr: raw-parameter
// mode, `features` holds f_dc (P, 3) and the colour is its degree-0 SH evaluation (common.cuh), with the
// key-point score as the 4th channel.
template <int C>
__device__ __forceinline__ float4 gather_features(const float* __restrict__ features, const float* __restrict__ kp,
                                                  bool raw, uint32_t id)
{
    float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
    if (raw) {
        f.x = act_sh0_colour(features[id * 3 + 0]);
        f.y = act_sh0_colour(features[id * 3 + 1]);
        f.z = act_sh0_colour(features[id * 3 + 2]);
        if (C == 4) f.w = kp[id];
    } else if (C == 4) {
        f = reinterpret_cast<const float4*>(features)[id];
    } else {
        f.x = features[id * C + 0];
        if (C > 1) f.y = features[id * C + 1];
        if (C > 2) f.z = features[id * C + 2];
    }
    return f;
}

// Shared-space (32-bit) addressing done by hand: with generic pointers the compiler re-derives the
// CTA's shared window base (S2R SR_CgaCtaId + shifts) inside the inner loop.
__device__ __forceinline__ uint32_t smem_addr(const void* p)
{
    uint32_t a = (uint32_t)__cvta_generic_to_shared(p);
    asm volatile("mov.u32 %0, %0;" : "+r"(a));   // opaque: keeps the address in a register instead of
    return a;                                     // letting the compiler rematerialise it per use
}
__device__ __forceinline__ float4 lds128(uint32_t a)
{
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t a)
{
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}

// exp(x) exactly as the reference's __expf computes it for every x that can still pass the alpha
// test: __expf(x) = ex2.approx(x * log2e) with a rescaling guard that only acts below 2^-126, where
// alpha = o * exp(x) is far under 1/255 anyway; .ftz drops the guard (3 instructions per evaluation).
__device__ __forceinline__ float exp_like_reference(float x)
{
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(__fmul_rn(x, 1.4426950216293334961f)));
    return r;
}

}  // namespace sl
