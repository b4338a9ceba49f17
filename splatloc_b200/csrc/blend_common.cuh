// splatloc_b200/csrc/blend_common.cuh -- pieces shared by the forward and backward blend kernels.
//
// Work decomposition.  A CTA owns one 16x16 tile, a warp an 8x4 pixel patch of it.  The tile's
// list is consumed in batches of 256 (tile, Gaussian) entries staged in shared memory.  While
// staging, each thread also computes for its Gaussian the axis-aligned bounding box of the region
// where it can pass the reference's alpha test at all:
//     alpha = min(0.99, o * exp(power)) >= 1/255   <=>   d^T Q d <= 2 ln(255 o),  Q = conic
// whose box has half-widths  hx = sqrt(2 ln(255 o) * Qyy / det Q),  hy = sqrt(2 ln(255 o) * Qxx / det Q)
// (padded by 0.2 % + 0.05 px, far more than any float32 / ex2.approx error).  Entries whose box
// misses a warp's patch are dropped from that warp's private, order-preserving index list, so the
// warp's inner loop only visits Gaussians that can contribute to one of its 32 pixels.  The test is
// conservative, every surviving entry is still evaluated with the reference's exact arithmetic and
// list positions are kept, so colour, depth, alpha and n_contrib stay bit-identical -- the
// reference would have `continue`d on the dropped ones (forward.cu:925-931).  On the config-2
// scene ~75 % of (patch, Gaussian) evaluations are dropped (tools/scene_stats.py).
#pragma once
#include "common.cuh"

namespace sl {

constexpr int BLEND_BATCH = 256;           // entries per staged batch
// A CTA owns BLEND_WARPS of a tile's 8 patches (a half tile by default): finer CTAs balance the
// SMs better at the end of the launch and keep fewer warps waiting at each batch barrier, at the
// price of staging every list 8 / BLEND_WARPS times.
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
    return m;
}

// One staged list entry: 48 bytes, read with three 16-byte shared loads off one base address.
struct __align__(16) Staged {
    float4 xyd;   // pixel-space x, y, view depth, Gaussian id (bit pattern)
    float4 co;    // conic xx, xy, yy, opacity
    float4 col;   // up to 4 feature channels
};

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

// Order-preserving compaction of the staged batch into this warp's private list.
// keep(j) decides per staged entry; returns the number of entries kept.
template <typename Keep>
__device__ __forceinline__ int compact_batch(uint8_t* __restrict__ list, int n, int lane, Keep keep)
{
    int cnt = 0;
    const unsigned lt = (1u << lane) - 1u;
#pragma unroll 1
    for (int c = 0; c * 32 < n; ++c) {
        const int j = c * 32 + lane;
        const bool bit = (j < n) && keep(j);
        const unsigned b = __ballot_sync(0xffffffffu, bit);
        if (bit) list[cnt + __popc(b & lt)] = (uint8_t)j;
        cnt += __popc(b);
    }
    __syncwarp();
    return cnt;
}

}  // namespace sl
