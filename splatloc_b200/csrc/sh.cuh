// splatloc_b200/csrc/sh.cuh
//
// Real spherical harmonics up to degree 3: colour evaluation and its backward.
// Same basis, constants and sign conventions as the reference
// (cuda_rasterizer/auxiliary.h:24-43 constants; forward.cu:20-71 evaluation;
// backward.cu:20-139 gradients), written per basis function instead of per
// GLM vector expression.  SH coefficients are laid out (P, M, 3).
#pragma once
#include "common.cuh"

namespace sl {

__device__ constexpr float kSH0 = 0.28209479177387814f;
__device__ constexpr float kSH1 = 0.4886025119029199f;
__device__ constexpr float kSH2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                      -1.0925484305920792f, 0.5462742152960396f};
__device__ constexpr float kSH3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                      0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                      -0.5900435899266435f};

// Basis values b[0..n) for unit direction (x, y, z); returns number of coefficients used.
__device__ __forceinline__ int sh_basis(int deg, float x, float y, float z, float* b)
{
    b[0] = kSH0;
    if (deg < 1) return 1;
    b[1] = -kSH1 * y;
    b[2] = kSH1 * z;
    b[3] = -kSH1 * x;
    if (deg < 2) return 4;
    float xx = x * x, yy = y * y, zz = z * z;
    float xy = x * y, yz = y * z, xz = x * z;
    b[4] = kSH2[0] * xy;
    b[5] = kSH2[1] * yz;
    b[6] = kSH2[2] * (2.0f * zz - xx - yy);
    b[7] = kSH2[3] * xz;
    b[8] = kSH2[4] * (xx - yy);
    if (deg < 3) return 9;
    b[9] = kSH3[0] * y * (3.0f * xx - yy);
    b[10] = kSH3[1] * xy * z;
    b[11] = kSH3[2] * y * (4.0f * zz - xx - yy);
    b[12] = kSH3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy);
    b[13] = kSH3[4] * x * (4.0f * zz - xx - yy);
    b[14] = kSH3[5] * z * (xx - yy);
    b[15] = kSH3[6] * x * (xx - 3.0f * yy);
    return 16;
}

// colour = max(sum_k b_k * sh_k + 0.5, 0); records which channels were clamped.
__device__ __forceinline__ f3 sh_to_rgb(int idx, int deg, int max_coeffs, const f3& pos, const f3& campos,
                                        const float* __restrict__ shs, uint8_t* __restrict__ clamped)
{
    f3 dir = {pos.x - campos.x, pos.y - campos.y, pos.z - campos.z};
    float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
    dir.x = dir.x / len;
    dir.y = dir.y / len;
    dir.z = dir.z / len;

    float b[16];
    int n = sh_basis(deg, dir.x, dir.y, dir.z, b);
    const float* sh = shs + (size_t)idx * max_coeffs * 3;
    float r = 0.f, g = 0.f, bl = 0.f;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        if (k < n) {
            r += b[k] * sh[3 * k + 0];
            g += b[k] * sh[3 * k + 1];
            bl += b[k] * sh[3 * k + 2];
        }
    }
    r += 0.5f;
    g += 0.5f;
    bl += 0.5f;
    clamped[3 * idx + 0] = (r < 0);
    clamped[3 * idx + 1] = (g < 0);
    clamped[3 * idx + 2] = (bl < 0);
    return {fmaxf(r, 0.0f), fmaxf(g, 0.0f), fmaxf(bl, 0.0f)};
}

// Backward of sh_to_rgb: writes dL/dsh (P, M, 3) and returns dL/dmean through the
// (normalised) view direction.
__device__ __forceinline__ f3 sh_to_rgb_backward(int idx, int deg, int max_coeffs, const f3& pos,
                                                 const f3& campos, const float* __restrict__ shs,
                                                 const uint8_t* __restrict__ clamped, f3 dL_dRGB,
                                                 float* __restrict__ dL_dshs, bool accumulate = false)
{
    f3 d0 = {pos.x - campos.x, pos.y - campos.y, pos.z - campos.z};
    float len = sqrtf(d0.x * d0.x + d0.y * d0.y + d0.z * d0.z);
    float x = d0.x / len, y = d0.y / len, z = d0.z / len;

    dL_dRGB.x *= clamped[3 * idx + 0] ? 0 : 1;
    dL_dRGB.y *= clamped[3 * idx + 1] ? 0 : 1;
    dL_dRGB.z *= clamped[3 * idx + 2] ? 0 : 1;

    float b[16];
    int n = sh_basis(deg, x, y, z, b);
    const float* sh = shs + (size_t)idx * max_coeffs * 3;
    float* dsh = dL_dshs + (size_t)idx * max_coeffs * 3;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        if (k < n && dL_dshs != nullptr) {   // nullptr: only the direction gradient is wanted
            // `accumulate`: a later view of a multi-view call adds to what an earlier one stored
            dsh[3 * k + 0] = b[k] * dL_dRGB.x + (accumulate ? dsh[3 * k + 0] : 0.f);
            dsh[3 * k + 1] = b[k] * dL_dRGB.y + (accumulate ? dsh[3 * k + 1] : 0.f);
            dsh[3 * k + 2] = b[k] * dL_dRGB.z + (accumulate ? dsh[3 * k + 2] : 0.f);
        }
    }

    // d basis_k / d(x,y,z), then contracted with sh_k (per channel) and dL/dRGB.
    float dbx[16], dby[16], dbz[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) dbx[k] = dby[k] = dbz[k] = 0.f;
    if (deg > 0) {
        dby[1] = -kSH1;
        dbz[2] = kSH1;
        dbx[3] = -kSH1;
        if (deg > 1) {
            float xx = x * x, yy = y * y, zz = z * z;
            float xy = x * y, yz = y * z, xz = x * z;
            dbx[4] = kSH2[0] * y;             dby[4] = kSH2[0] * x;
            dby[5] = kSH2[1] * z;             dbz[5] = kSH2[1] * y;
            dbx[6] = kSH2[2] * 2.f * -x;      dby[6] = kSH2[2] * 2.f * -y;   dbz[6] = kSH2[2] * 2.f * 2.f * z;
            dbx[7] = kSH2[3] * z;             dbz[7] = kSH2[3] * x;
            dbx[8] = kSH2[4] * 2.f * x;       dby[8] = kSH2[4] * 2.f * -y;
            if (deg > 2) {
                dbx[9] = kSH3[0] * 3.f * 2.f * xy;          dby[9] = kSH3[0] * 3.f * (xx - yy);
                dbx[10] = kSH3[1] * yz;                     dby[10] = kSH3[1] * xz;                 dbz[10] = kSH3[1] * xy;
                dbx[11] = kSH3[2] * -2.f * xy;              dby[11] = kSH3[2] * (-3.f * yy + 4.f * zz - xx);  dbz[11] = kSH3[2] * 4.f * 2.f * yz;
                dbx[12] = kSH3[3] * -3.f * 2.f * xz;        dby[12] = kSH3[3] * -3.f * 2.f * yz;    dbz[12] = kSH3[3] * 3.f * (2.f * zz - xx - yy);
                dbx[13] = kSH3[4] * (-3.f * xx + 4.f * zz - yy);  dby[13] = kSH3[4] * -2.f * xy;    dbz[13] = kSH3[4] * 4.f * 2.f * xz;
                dbx[14] = kSH3[5] * 2.f * xz;               dby[14] = kSH3[5] * -2.f * yz;          dbz[14] = kSH3[5] * (xx - yy);
                dbx[15] = kSH3[6] * 3.f * (xx - yy);        dby[15] = kSH3[6] * -3.f * 2.f * xy;
            }
        }
    }
    f3 dL_ddir = {0.f, 0.f, 0.f};
#pragma unroll
    for (int k = 1; k < 16; ++k) {
        if (k < n) {
            float s = sh[3 * k + 0] * dL_dRGB.x + sh[3 * k + 1] * dL_dRGB.y + sh[3 * k + 2] * dL_dRGB.z;
            dL_ddir.x += dbx[k] * s;
            dL_ddir.y += dby[k] * s;
            dL_ddir.z += dbz[k] * s;
        }
    }

    // Jacobian of v / |v| (auxiliary.h:113-123).
    float sum2 = d0.x * d0.x + d0.y * d0.y + d0.z * d0.z;
    float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    f3 g;
    g.x = ((+sum2 - d0.x * d0.x) * dL_ddir.x - d0.y * d0.x * dL_ddir.y - d0.z * d0.x * dL_ddir.z) * invsum32;
    g.y = (-d0.x * d0.y * dL_ddir.x + (sum2 - d0.y * d0.y) * dL_ddir.y - d0.z * d0.y * dL_ddir.z) * invsum32;
    g.z = (-d0.x * d0.z * dL_ddir.x - d0.y * d0.z * dL_ddir.y + (sum2 - d0.z * d0.z) * dL_ddir.z) * invsum32;
    return g;
}

}  // namespace sl
