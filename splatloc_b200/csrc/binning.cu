// splatloc_b200/csrc/binning.cu
//
// K3  emit one (tile | depth) key and Gaussian id per overlapped tile
//     (reference: duplicateWithKeys, cuda_rasterizer/rasterizer_impl.cu:70-132).
//     Pairs are written at the Gaussian's prefix-sum offset in row-major tile
//     order, so the emission order -- and with it the tie-break of the stable
//     sort -- is the reference's.
// K5  per-tile [first, last+1) ranges in the sorted list (reference: identifyTileRanges,
//     rasterizer_impl.cu:137-159), folded into the record gather below.
// These two kernels serve the radix-sort fallback path; the default path is tile_binning.cu.
#include "kernels.h"
#include "blend_common.cuh"

namespace sl {

__global__ void __launch_bounds__(256)
emit_pairs_kernel(int P, const float2* __restrict__ means2D, const float* __restrict__ depths,
                  const uint32_t* __restrict__ offsets, const int* __restrict__ radii,
                  const uint8_t* __restrict__ tile_mask, int grid_x, int grid_y, int tiles_per_view,
                  uint64_t* __restrict__ keys, uint32_t* __restrict__ vals)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    // blockIdx.y: the view.  The prefix sum runs over all (view, Gaussian) entries, view-major.
    const size_t v = (size_t)blockIdx.y * P + idx;
    const int r = radii[v];
    if (r <= 0) return;

    uint32_t off = (v == 0) ? 0 : offsets[v - 1];
    const uint32_t end = offsets[v];
    if (off == end) return;  // visible radius but nothing to emit (masked out)
    const float2 xy = means2D[v];
    uint2 rmin, rmax;
    tile_rect(xy.x, xy.y, r, grid_x, grid_y, rmin, rmax);
    const uint64_t depth_bits = __float_as_uint(depths[v]);
    const uint64_t tile0 = (uint64_t)blockIdx.y * tiles_per_view;

    for (uint32_t y = rmin.y; y < rmax.y; y++) {
        for (uint32_t x = rmin.x; x < rmax.x; x++) {
            uint64_t key = y * grid_x + x;
            if (tile_mask != nullptr && !tile_mask[key]) continue;
            key += tile0;
            key <<= 32;
            key |= depth_bits;
            keys[off] = key;
            vals[off] = idx;
            off++;
        }
    }
}

// One thread per sorted pair: gather the Gaussian's screen-space data into a PairRecord, compute
// which 8x4 patches of the pair's tile it can reach (blend_common.cuh), and mark tile boundaries
// (the reference's identifyTileRanges).  After this kernel the blend kernels never touch the
// per-Gaussian arrays or point_list again: they stream `records` / `masks` linearly.
template <int C>
__global__ void __launch_bounds__(256)
build_records_kernel(int R, int P, const uint64_t* __restrict__ keys, const uint32_t* __restrict__ point_list,
                     const float2* __restrict__ means2D, const float* __restrict__ depths,
                     const float4* __restrict__ conic_opacity, const float* __restrict__ features,
                     size_t feature_view_stride, const float* __restrict__ kp, int raw, int grid_x, int tiles_per_view,
                     PairRecord* __restrict__ records, uint8_t* __restrict__ masks, uint2* __restrict__ ranges)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    const uint32_t tile = (uint32_t)(keys[i] >> 32);
    if (i == 0) {
        ranges[tile].x = 0;
    } else {
        const uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
        if (tile != prev) {
            ranges[prev].y = i;
            ranges[tile].x = i;
        }
    }
    if (i == R - 1) ranges[tile].y = R;

    const uint32_t view = tile / (uint32_t)tiles_per_view, local_tile = tile - view * (uint32_t)tiles_per_view;
    const int id = point_list[i];
    const size_t v = (size_t)view * P + id;
    const float2 xy = means2D[v];
    const float4 co = conic_opacity[v];
    features += view * feature_view_stride;
    const float4 f = gather_features<C>(features, kp, raw != 0, (uint32_t)id);
    PairRecord rec;
    rec.xyd = make_float4(xy.x, xy.y, depths[v], __int_as_float(id));
    rec.co = co;
    rec.col = f;
    records[i] = rec;
    const int tile_x0 = (local_tile % grid_x) * SL_TILE_X, tile_y0 = (local_tile / grid_x) * SL_TILE_Y;
    masks[i] = (uint8_t)patch_mask(xy.x, xy.y, co, tile_x0, tile_y0);
}

int launch_build_records(int R, int P, int C, const uint64_t* sorted_keys, const uint32_t* point_list,
                         const GeomState& g, const float* features, size_t feature_view_stride, const RawParams* rawp,
                         const ViewParams& vp, PairRecord* records, uint8_t* masks, uint2* ranges,
                         cudaStream_t stream)
{
    const float* kp = rawp != nullptr ? rawp->kp_score : nullptr;
    const int raw = rawp != nullptr ? 1 : 0;
    SL_CUDA(cudaMemsetAsync(ranges, 0, (size_t)vp.tiles * vp.views * sizeof(uint2), stream), "ranges memset");
    if (R > 0) {
        if (C == 4)
            build_records_kernel<4><<<ceil_div(R, 256), 256, 0, stream>>>(R, P, sorted_keys, point_list, g.means2D,
                g.depths, g.conic_opacity, features, feature_view_stride, kp, raw, vp.grid_x, vp.tiles, records, masks, ranges);
        else
            build_records_kernel<3><<<ceil_div(R, 256), 256, 0, stream>>>(R, P, sorted_keys, point_list, g.means2D,
                g.depths, g.conic_opacity, features, feature_view_stride, kp, raw, vp.grid_x, vp.tiles, records, masks, ranges);
        count_launch();
    }
    return 0;
}

int launch_emit_pairs(int P, const GeomState& g, const int* radii, const uint8_t* tile_mask,
                      const ViewParams& vp, uint64_t* keys, uint32_t* vals, cudaStream_t stream)
{
    const dim3 grid(ceil_div(P, 256), vp.views);
    emit_pairs_kernel<<<grid, 256, 0, stream>>>(P, g.means2D, g.depths, g.point_offsets, radii, tile_mask, vp.grid_x,
                                                vp.grid_y, vp.tiles, keys, vals); count_launch();
    return 0;
}

}  // namespace sl
