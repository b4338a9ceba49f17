// oracle/ref_shim.cu -- TEST INFRASTRUCTURE, not product code.
//
// A plain C ABI around the *unmodified* reference rasterizer so that the GPU
// parity tests and bench.py's reference arm can call it without torch/pybind.
// It is compiled together with the reference's own translation units, read in
// place from /root/reference (see oracle/build_ref.sh); nothing from the
// reference is copied into this repository.  The functions forward one-to-one
// to CudaRasterizer::Rasterizer::{forward,backward,markVisible}
// (submodules/diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h:71-76,
// 134-195) and SimpleKNN::knn (submodules/simple-knn/simple_knn.h:16-20).
#include <cstddef>
#include <cstdint>
#include <functional>
#include <cuda_runtime.h>
#ifdef REF_SHIM_RASTER
#include "cuda_rasterizer/config.h"
#include "cuda_rasterizer/rasterizer.h"
#endif
#ifdef REF_SHIM_KNN
#include "simple_knn.h"
#endif

typedef char* (*ref_alloc_fn)(void* ctx, size_t bytes);

extern "C" {

#ifdef REF_SHIM_RASTER
int ref_num_channels() { return NUM_CHANNELS; }

int ref_forward(ref_alloc_fn geom_alloc, void* geom_ctx,
                ref_alloc_fn binning_alloc, void* binning_ctx,
                ref_alloc_fn img_alloc, void* img_ctx,
                int P, int D, int M, const float* background, int width, int height,
                const float* means3D, const float* shs, const float* colors_precomp,
                const float* opacities, const float* scales, float scale_modifier,
                const float* rotations, const float* cov3D_precomp, const bool* tile_mask,
                const float* viewmatrix, const float* projmatrix, const float* cam_pos,
                float tan_fovx, float tan_fovy, int prefiltered,
                float* out_color, float* out_depth, float* out_alpha, int* radii, int debug)
{
    try {
        return CudaRasterizer::Rasterizer::forward(
            [=](size_t n) { return geom_alloc(geom_ctx, n); },
            [=](size_t n) { return binning_alloc(binning_ctx, n); },
            [=](size_t n) { return img_alloc(img_ctx, n); },
            P, D, M, background, width, height, means3D, shs, colors_precomp, opacities,
            scales, scale_modifier, rotations, cov3D_precomp, tile_mask, viewmatrix,
            projmatrix, cam_pos, tan_fovx, tan_fovy, prefiltered != 0,
            out_color, out_depth, out_alpha, radii, debug != 0);
    } catch (...) {
        return -1;
    }
}

int ref_backward(int P, int D, int M, int R, const float* background, int width, int height,
                 const float* means3D, const float* shs, const float* colors_precomp,
                 const float* scales, float scale_modifier, const float* rotations,
                 const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                 const float* campos, float tan_fovx, float tan_fovy, const int* radii,
                 char* geom_buffer, char* binning_buffer, char* img_buffer,
                 const float* accum_alphas, const float* dL_dpix, const float* dL_dpix_depth,
                 const float* dL_dpix_alpha, float* dL_dmean2D, float* dL_dabsmean2D,
                 float* dL_dconic, float* dL_dopacity, float* dL_dcolor, float* dL_ddepth,
                 float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale,
                 float* dL_drot, int debug)
{
    try {
        CudaRasterizer::Rasterizer::backward(
            P, D, M, R, background, width, height, means3D, shs, colors_precomp, scales,
            scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos,
            tan_fovx, tan_fovy, radii, geom_buffer, binning_buffer, img_buffer, accum_alphas,
            dL_dpix, dL_dpix_depth, dL_dpix_alpha, dL_dmean2D, dL_dabsmean2D, dL_dconic,
            dL_dopacity, dL_dcolor, dL_ddepth, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale,
            dL_drot, debug != 0);
        return 0;
    } catch (...) {
        return -1;
    }
}

void ref_mark_visible(int P, float* means3D, float* viewmatrix, float* projmatrix, bool* present)
{
    CudaRasterizer::Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix, present);
}
#endif  // REF_SHIM_RASTER

#ifdef REF_SHIM_KNN
void ref_knn(int P, float* points, float* mean_dists)
{
    SimpleKNN::knn(P, (float3*)points, mean_dists);
}
#endif

}  // extern "C"
