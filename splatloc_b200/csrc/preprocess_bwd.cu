// splatloc_b200/csrc/preprocess_bwd.cu
//
// K8+K9 fused per-Gaussian backward.  The reference runs two kernels
// (computeCov2DCUDA, cuda_rasterizer/backward.cu:534-664, then preprocessCUDA,
// :949-1015), the second re-reading and += -ing what the first stored; here one
// thread carries a Gaussian through both and stores each gradient once:
//   conic grads -> 2D covariance -> 3D covariance (6) and camera-space mean,
//   + screen-space mean grads through the projection,
//   + depth grads through the view matrix (row 2),
//   + optional SH backward (colour -> SH coefficients and view direction),
//   + optional 3D covariance -> scale / raw quaternion (no normalisation
//     Jacobian, as in the reference, backward.cu:730).
// Only Gaussians with radii > 0 in some view are touched; all other rows keep the
// caller's zeros.  HBM-bound: ~100 B in per (visible Gaussian, view), 36..64 B out
// per Gaussian.  With several views per call (sl_backward_batch) the gradients of
// a Gaussian are summed over the views before they are stored.
#include "kernels.h"
#include "sh.cuh"

namespace sl {

// One Gaussian in one view where it is visible.  The pointers are the view's slices of the per-view
// arrays.  Its dL/dmean3D and dL/dcov3D are stored into (first == true) or added to the running sums
// over the views, `g_mean_acc` / `g_cov_acc`; dL/dsh likewise goes to / adds to memory.  `pose` (when
// wanted) receives this Gaussian's contribution to the gradient of the loss w.r.t. a left
// perturbation Exp(rho, theta) of this view's world-to-camera transform.
__device__ __forceinline__ void
per_gaussian_backward(int idx, int D, int M, int C,
                           const float* __restrict__ means3D, const int* __restrict__ radii,
                           const float* __restrict__ shs, const uint8_t* __restrict__ clamped,
                           const float* __restrict__ scales, const float* __restrict__ rotations,
                           const float* __restrict__ cov3Ds, const float* __restrict__ view,
                           const float* __restrict__ proj, const float* __restrict__ campos,
                           ViewParams vp,
                           const float* __restrict__ dL_dmean2D,  // (P,3)
                           const float* __restrict__ dL_dconics,  // (P,4)
                           const float* __restrict__ dL_dcolor,   // (P,C)
                           const float* __restrict__ dL_ddepth,   // (P,1)
                           float* __restrict__ dL_dsh,            // (P,M,3)
                           bool first, f3& g_mean_acc, float (&g_cov_acc)[6],
                           bool want_pose, float (&pose)[6])
{
    const float* cov3D = cov3Ds + 6 * idx;
    const f3 mean = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
    const float h_x = vp.focal_x, h_y = vp.focal_y;

    // ---------------------------------------------------------------- conic -> cov2D -> cov3D, t
    const f3 dL_dconic = {dL_dconics[4 * idx], dL_dconics[4 * idx + 1], dL_dconics[4 * idx + 3]};
    f3 t = xform_4x3(mean, view);
    const f3 t_cam = t;   // camera-frame position before the covariance clamp

    const float limx = 1.3f * vp.tan_fovx;
    const float limy = 1.3f * vp.tan_fovy;
    const float txtz = t.x / t.z;
    const float tytz = t.y / t.z;
    t.x = min(limx, max(-limx, txtz)) * t.z;
    t.y = min(limy, max(-limy, tytz)) * t.z;

    const float x_grad_mul = txtz < -limx || txtz > limx ? 0 : 1;
    const float y_grad_mul = tytz < -limy || tytz > limy ? 0 : 1;

    mat3 J = mat3_cols(h_x / t.z, 0.0f, -(h_x * t.x) / (t.z * t.z),
                       0.0f, h_y / t.z, -(h_y * t.y) / (t.z * t.z),
                       0, 0, 0);
    mat3 Wm = mat3_cols(view[0], view[4], view[8],
                        view[1], view[5], view[9],
                        view[2], view[6], view[10]);
    mat3 Vrk = mat3_cols(cov3D[0], cov3D[1], cov3D[2],
                         cov3D[1], cov3D[3], cov3D[4],
                         cov3D[2], cov3D[4], cov3D[5]);
    mat3 T = mul(Wm, J);
    mat3 cov2D = mul(mul(transpose(T), transpose(Vrk)), T);

    const float a = cov2D.c[0][0] + 0.3f;
    const float b = cov2D.c[0][1];
    const float c = cov2D.c[1][1] + 0.3f;

    const float denom = a * c - b * b;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);

    float g_cov[6];
    if (denom2inv != 0) {
        dL_da = denom2inv * (-c * c * dL_dconic.x + 2 * b * c * dL_dconic.y + (denom - a * c) * dL_dconic.z);
        dL_dc = denom2inv * (-a * a * dL_dconic.z + 2 * a * b * dL_dconic.y + (denom - a * c) * dL_dconic.x);
        dL_db = denom2inv * 2 * (b * c * dL_dconic.x - (denom + 2 * b * b) * dL_dconic.y + a * b * dL_dconic.z);

        g_cov[0] = (T.c[0][0] * T.c[0][0] * dL_da + T.c[0][0] * T.c[1][0] * dL_db + T.c[1][0] * T.c[1][0] * dL_dc);
        g_cov[3] = (T.c[0][1] * T.c[0][1] * dL_da + T.c[0][1] * T.c[1][1] * dL_db + T.c[1][1] * T.c[1][1] * dL_dc);
        g_cov[5] = (T.c[0][2] * T.c[0][2] * dL_da + T.c[0][2] * T.c[1][2] * dL_db + T.c[1][2] * T.c[1][2] * dL_dc);
        g_cov[1] = 2 * T.c[0][0] * T.c[0][1] * dL_da + (T.c[0][0] * T.c[1][1] + T.c[0][1] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][1] * dL_dc;
        g_cov[2] = 2 * T.c[0][0] * T.c[0][2] * dL_da + (T.c[0][0] * T.c[1][2] + T.c[0][2] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][2] * dL_dc;
        g_cov[4] = 2 * T.c[0][2] * T.c[0][1] * dL_da + (T.c[0][1] * T.c[1][2] + T.c[0][2] * T.c[1][1]) * dL_db + 2 * T.c[1][1] * T.c[1][2] * dL_dc;
    } else {
#pragma unroll
        for (int i = 0; i < 6; i++) g_cov[i] = 0;
    }
#pragma unroll
    for (int i = 0; i < 6; i++) g_cov_acc[i] = first ? g_cov[i] : g_cov_acc[i] + g_cov[i];

    // gradients w.r.t. the upper 2x3 of T (cov2D = T^T Vrk^T T)
    const float dL_dT00 = 2 * (T.c[0][0] * Vrk.c[0][0] + T.c[0][1] * Vrk.c[0][1] + T.c[0][2] * Vrk.c[0][2]) * dL_da +
                          (T.c[1][0] * Vrk.c[0][0] + T.c[1][1] * Vrk.c[0][1] + T.c[1][2] * Vrk.c[0][2]) * dL_db;
    const float dL_dT01 = 2 * (T.c[0][0] * Vrk.c[1][0] + T.c[0][1] * Vrk.c[1][1] + T.c[0][2] * Vrk.c[1][2]) * dL_da +
                          (T.c[1][0] * Vrk.c[1][0] + T.c[1][1] * Vrk.c[1][1] + T.c[1][2] * Vrk.c[1][2]) * dL_db;
    const float dL_dT02 = 2 * (T.c[0][0] * Vrk.c[2][0] + T.c[0][1] * Vrk.c[2][1] + T.c[0][2] * Vrk.c[2][2]) * dL_da +
                          (T.c[1][0] * Vrk.c[2][0] + T.c[1][1] * Vrk.c[2][1] + T.c[1][2] * Vrk.c[2][2]) * dL_db;
    const float dL_dT10 = 2 * (T.c[1][0] * Vrk.c[0][0] + T.c[1][1] * Vrk.c[0][1] + T.c[1][2] * Vrk.c[0][2]) * dL_dc +
                          (T.c[0][0] * Vrk.c[0][0] + T.c[0][1] * Vrk.c[0][1] + T.c[0][2] * Vrk.c[0][2]) * dL_db;
    const float dL_dT11 = 2 * (T.c[1][0] * Vrk.c[1][0] + T.c[1][1] * Vrk.c[1][1] + T.c[1][2] * Vrk.c[1][2]) * dL_dc +
                          (T.c[0][0] * Vrk.c[1][0] + T.c[0][1] * Vrk.c[1][1] + T.c[0][2] * Vrk.c[1][2]) * dL_db;
    const float dL_dT12 = 2 * (T.c[1][0] * Vrk.c[2][0] + T.c[1][1] * Vrk.c[2][1] + T.c[1][2] * Vrk.c[2][2]) * dL_dc +
                          (T.c[0][0] * Vrk.c[2][0] + T.c[0][1] * Vrk.c[2][1] + T.c[0][2] * Vrk.c[2][2]) * dL_db;

    // T = W * J
    const float dL_dJ00 = Wm.c[0][0] * dL_dT00 + Wm.c[0][1] * dL_dT01 + Wm.c[0][2] * dL_dT02;
    const float dL_dJ02 = Wm.c[2][0] * dL_dT00 + Wm.c[2][1] * dL_dT01 + Wm.c[2][2] * dL_dT02;
    const float dL_dJ11 = Wm.c[1][0] * dL_dT10 + Wm.c[1][1] * dL_dT11 + Wm.c[1][2] * dL_dT12;
    const float dL_dJ12 = Wm.c[2][0] * dL_dT10 + Wm.c[2][1] * dL_dT11 + Wm.c[2][2] * dL_dT12;

    const float tz = 1.f / t.z;
    const float tz2 = tz * tz;
    const float tz3 = tz2 * tz;

    const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
    const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
    const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 +
                         (2 * h_y * t.y) * tz3 * dL_dJ12;

    // back through t = view * mean (rotation part, transposed)
    f3 g_mean;
    g_mean.x = view[0] * dL_dtx + view[1] * dL_dty + view[2] * dL_dtz;
    g_mean.y = view[4] * dL_dtx + view[5] * dL_dty + view[6] * dL_dtz;
    g_mean.z = view[8] * dL_dtx + view[9] * dL_dty + view[10] * dL_dtz;

    // ---------------------------------------------------------------- screen-space mean -> mean
    {
        float4 m_hom = xform_4x4(mean, proj);
        float m_w = 1.0f / (m_hom.w + 0.0000001f);
        const float gx = dL_dmean2D[3 * idx], gy = dL_dmean2D[3 * idx + 1];
        float mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * m_w * m_w;
        float mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * m_w * m_w;
        f3 g;
        g.x = (proj[0] * m_w - proj[3] * mul1) * gx + (proj[1] * m_w - proj[3] * mul2) * gy;
        g.y = (proj[4] * m_w - proj[7] * mul1) * gx + (proj[5] * m_w - proj[7] * mul2) * gy;
        g.z = (proj[8] * m_w - proj[11] * mul1) * gx + (proj[9] * m_w - proj[11] * mul2) * gy;
        g_mean.x += g.x;
        g_mean.y += g.y;
        g_mean.z += g.z;
    }

    // ---------------------------------------------------------------- depth -> mean (view row 2)
    {
        const float gd = dL_ddepth[idx];
        float mul3 = view[2] * mean.x + view[6] * mean.y + view[10] * mean.z + view[14];
        f3 g;
        g.x = (view[2] - view[3] * mul3) * gd;
        g.y = (view[6] - view[7] * mul3) * gd;
        g.z = (view[10] - view[11] * mul3) * gd;
        g_mean.x += g.x;
        g_mean.y += g.y;
        g_mean.z += g.z;
    }

    // ---------------------------------------------------------------- camera pose (extension)
    // W2C' = Exp(rho, theta) W2C moves every camera-frame point by rho + theta x t and rotates the
    // rows of the world-to-camera rotation R (the W of the EWA projection, T = W J).  With g_t the
    // gradient w.r.t. the camera-frame mean (covariance path directly, projection and depth paths
    // rotated from the world frame) and G = dL/dR from dL/dT:
    //     dL/drho = g_t (+ R g_SH),   dL/dtheta = t x g_t + sum_r R[:, r] x G[:, r].
    if (want_pose) {
        // world-frame pieces so far minus the covariance piece = projection + depth paths
        const float wx = g_mean.x - (view[0] * dL_dtx + view[1] * dL_dty + view[2] * dL_dtz);
        const float wy = g_mean.y - (view[4] * dL_dtx + view[5] * dL_dty + view[6] * dL_dtz);
        const float wz = g_mean.z - (view[8] * dL_dtx + view[9] * dL_dty + view[10] * dL_dtz);
        const float gtx = dL_dtx + view[0] * wx + view[4] * wy + view[8] * wz;
        const float gty = dL_dty + view[1] * wx + view[5] * wy + view[9] * wz;
        const float gtz = dL_dtz + view[2] * wx + view[6] * wy + view[10] * wz;
        pose[0] = gtx;
        pose[1] = gty;
        pose[2] = gtz;
        pose[3] = t_cam.y * gtz - t_cam.z * gty;
        pose[4] = t_cam.z * gtx - t_cam.x * gtz;
        pose[5] = t_cam.x * gty - t_cam.y * gtx;
        const float dT0[3] = {dL_dT00, dL_dT01, dL_dT02}, dT1[3] = {dL_dT10, dL_dT11, dL_dT12};
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const float G0 = dT0[r] * J.c[0][0], G1 = dT1[r] * J.c[1][1];
            const float G2 = dT0[r] * J.c[0][2] + dT1[r] * J.c[1][2];
            const float R0 = view[4 * r + 0], R1 = view[4 * r + 1], R2 = view[4 * r + 2];   // column r of R
            pose[3] += R1 * G2 - R2 * G1;
            pose[4] += R2 * G0 - R0 * G2;
            pose[5] += R0 * G1 - R1 * G0;
        }
    }

    // ---------------------------------------------------------------- colour -> SH (+ view direction)
    if (shs != nullptr) {
        f3 cp = {campos[0], campos[1], campos[2]};
        f3 dL_dRGB = {dL_dcolor[idx * C + 0], dL_dcolor[idx * C + 1], dL_dcolor[idx * C + 2]};
        f3 g = sh_to_rgb_backward(idx, D, M, mean, cp, shs, clamped, dL_dRGB, dL_dsh, !first);
        g_mean.x += g.x;
        g_mean.y += g.y;
        g_mean.z += g.z;
        if (want_pose) {   // the camera centre moves by -R^T rho: dL/drho += R g_SH
            pose[0] += view[0] * g.x + view[4] * g.y + view[8] * g.z;
            pose[1] += view[1] * g.x + view[5] * g.y + view[9] * g.z;
            pose[2] += view[2] * g.x + view[6] * g.y + view[10] * g.z;
        }
    }
    g_mean_acc.x = first ? g_mean.x : g_mean_acc.x + g_mean.x;
    g_mean_acc.y = first ? g_mean.y : g_mean_acc.y + g_mean.y;
    g_mean_acc.z = first ? g_mean.z : g_mean_acc.z + g_mean.z;
}

// dL/dcov3D (summed over the views) -> dL/dscale, dL/drotation.  The map from the covariance gradient
// is linear and view-independent, so it runs once per Gaussian however many views there are.
// `s_act`, `q`: the activated scale and rotation.  Returns the gradients w.r.t. (scale_modifier * s_act)
// -- the modifier's factor is not applied, as in the reference (backward.cu:668-731) -- and w.r.t. q.
__device__ __forceinline__ void
cov3d_to_scale_rot_backward(float scale_modifier, const f3 s_act, const float4 q, const float (&g_cov)[6],
                            f3& d_scale, float4& d_rot)
{
    {
        const float r = q.x, x = q.y, y = q.z, z = q.w;
        mat3 R = mat3_cols(
            1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
            2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
            2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
        const f3 s = {scale_modifier * s_act.x, scale_modifier * s_act.y, scale_modifier * s_act.z};
        mat3 S = mat3_cols(1.0f, 0.0f, 0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.0f, 1.0f);
        S.c[0][0] = s.x;
        S.c[1][1] = s.y;
        S.c[2][2] = s.z;
        mat3 Mx = mul(S, R);

        mat3 dL_dSigma = mat3_cols(
            g_cov[0], 0.5f * g_cov[1], 0.5f * g_cov[2],
            0.5f * g_cov[1], g_cov[3], 0.5f * g_cov[4],
            0.5f * g_cov[2], 0.5f * g_cov[4], g_cov[5]);

        // dSigma/dM = 2 M
        mat3 M2;
#pragma unroll
        for (int cc = 0; cc < 3; ++cc)
#pragma unroll
            for (int rr = 0; rr < 3; ++rr) M2.c[cc][rr] = 2.0f * Mx.c[cc][rr];
        mat3 dL_dM = mul(M2, dL_dSigma);

        mat3 Rt = transpose(R);
        mat3 dL_dMt = transpose(dL_dM);

        d_scale.x = Rt.c[0][0] * dL_dMt.c[0][0] + Rt.c[0][1] * dL_dMt.c[0][1] + Rt.c[0][2] * dL_dMt.c[0][2];
        d_scale.y = Rt.c[1][0] * dL_dMt.c[1][0] + Rt.c[1][1] * dL_dMt.c[1][1] + Rt.c[1][2] * dL_dMt.c[1][2];
        d_scale.z = Rt.c[2][0] * dL_dMt.c[2][0] + Rt.c[2][1] * dL_dMt.c[2][1] + Rt.c[2][2] * dL_dMt.c[2][2];

#pragma unroll
        for (int rr = 0; rr < 3; ++rr) {
            dL_dMt.c[0][rr] *= s.x;
            dL_dMt.c[1][rr] *= s.y;
            dL_dMt.c[2][rr] *= s.z;
        }

        float4 dq;
        dq.x = 2 * z * (dL_dMt.c[0][1] - dL_dMt.c[1][0]) + 2 * y * (dL_dMt.c[2][0] - dL_dMt.c[0][2]) + 2 * x * (dL_dMt.c[1][2] - dL_dMt.c[2][1]);
        dq.y = 2 * y * (dL_dMt.c[1][0] + dL_dMt.c[0][1]) + 2 * z * (dL_dMt.c[2][0] + dL_dMt.c[0][2]) + 2 * r * (dL_dMt.c[1][2] - dL_dMt.c[2][1]) - 4 * x * (dL_dMt.c[2][2] + dL_dMt.c[1][1]);
        dq.z = 2 * x * (dL_dMt.c[1][0] + dL_dMt.c[0][1]) + 2 * r * (dL_dMt.c[2][0] - dL_dMt.c[0][2]) + 2 * z * (dL_dMt.c[1][2] + dL_dMt.c[2][1]) - 4 * y * (dL_dMt.c[2][2] + dL_dMt.c[0][0]);
        dq.w = 2 * r * (dL_dMt.c[0][1] - dL_dMt.c[1][0]) + 2 * x * (dL_dMt.c[2][0] + dL_dMt.c[0][2]) + 2 * y * (dL_dMt.c[1][2] + dL_dMt.c[2][1]) - 4 * z * (dL_dMt.c[1][1] + dL_dMt.c[0][0]);
        d_rot = dq;
    }
}

// One thread per Gaussian, walking the views of the call: the Gaussian's gradients are summed over the
// views in registers and stored once (rows that no view sees keep the caller's zeros), the camera-pose
// gradient of every view is reduced over the CTA and leaves with one atomic per CTA, view and value.
// One view: 4 CTAs per SM (64 registers; the kernel waits on scattered loads, measured 27 us against 30 at
// 3 CTAs / 80 registers).  Several views: the running sums get 80 registers (at 64 they spill 256 bytes).
#ifndef SL_PBWD_MULTI_CTAS
#define SL_PBWD_MULTI_CTAS 3
#endif
template <bool MULTI, bool RAW>
__global__ void __launch_bounds__(256, MULTI ? SL_PBWD_MULTI_CTAS : 4)
preprocess_backward_kernel(int P, int D, int M, int C,
                           const float* __restrict__ means3D, const int* __restrict__ radii,
                           const float* __restrict__ shs, const uint8_t* __restrict__ clamped,
                           const float* __restrict__ scales, const float* __restrict__ rotations,
                           const float* __restrict__ cov3Ds, const float* __restrict__ viewmatrices,
                           const float* __restrict__ projmatrices, const float* __restrict__ cam_positions,
                           ViewParams vp,
                           const float* __restrict__ dL_dmean2D, const float* __restrict__ dL_dconics,
                           float* __restrict__ dL_dmeans, const float* __restrict__ dL_dcolor,
                           size_t color_view_stride, const float* __restrict__ dL_ddepth,
                           float* __restrict__ dL_dcov, float* __restrict__ dL_dsh, float* __restrict__ dL_dscale,
                           float* __restrict__ dL_drot, float* dL_dtau,
                           int scale_dim, const float* __restrict__ raw_opacities, const float* __restrict__ features_dc,
                           float* __restrict__ dL_dopacity, float* __restrict__ dL_dfdc, float* __restrict__ dL_dkp,
                           float* push_mc)
{
    chain_wait();        // launched with launch_chained(): nothing before the previous stage is complete
    chain_release();
    __shared__ float s_pose[8][6];
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool want_pose = dL_dtau != nullptr;
    f3 g_mean = {0.f, 0.f, 0.f};
    float g_cov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    bool seen = false;
    int n_seen = 0;                                  // views in which this Gaussian was visible
    const int n_views = MULTI ? vp.views : 1;
    for (int v = 0; v < n_views; ++v) {
        const size_t vP = (size_t)v * P;
        float pose[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (idx < P && radii[vP + idx] > 0) {
            per_gaussian_backward(idx, D, M, C, means3D, radii, shs, clamped + 3 * vP, scales, rotations, cov3Ds,
                                  viewmatrices + 16 * v, projmatrices + 16 * v,
                                  cam_positions != nullptr ? cam_positions + 3 * v : nullptr, vp,
                                  dL_dmean2D + 3 * vP, dL_dconics + 4 * vP, dL_dcolor + v * color_view_stride,
                                  dL_ddepth + vP, dL_dsh, MULTI ? !seen : true, g_mean, g_cov, want_pose, pose);
            seen = true;
            ++n_seen;
        }
        if (want_pose) {
            // 6 floats per view: warp shuffle, then shared memory across the 8 warps
#pragma unroll
            for (int k = 0; k < 6; ++k) {
                float x = pose[k];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
                if (lane == 0) s_pose[warp][k] = x;
            }
            __syncthreads();
            if (threadIdx.x < 6) {
                float x = 0.f;
#pragma unroll
                for (int w = 0; w < 8; ++w) x += s_pose[w][threadIdx.x];
                if (x != 0.f) atomicAdd(&dL_dtau[6 * v + threadIdx.x], x);
            }
            __syncthreads();
        }
    }
    if (dL_dmeans == nullptr) return;            // only the pose gradients were wanted
    if (push_mc == nullptr && !seen) return;     // (with the exchange every lane stays for the staged push)
    if (seen) {
    dL_dmeans[3 * idx + 0] = g_mean.x;
    dL_dmeans[3 * idx + 1] = g_mean.y;
    dL_dmeans[3 * idx + 2] = g_mean.z;
#pragma unroll
    for (int i = 0; i < 6; i++) dL_dcov[6 * idx + i] = g_cov[i];
    if (scales != nullptr) {
        float4 q = reinterpret_cast<const float4*>(rotations)[idx];
        f3 s, ds;
        float4 dq;
        if (RAW) {
            // the optimiser's raw parameters: activate as the forward did (common.cuh) ...
            if (scale_dim == 1) {
                s.x = s.y = s.z = expf(scales[idx]);
            } else {
                s = {expf(scales[3 * idx]), expf(scales[3 * idx + 1]), expf(scales[3 * idx + 2])};
            }
            const float4 q_raw = q;
            q = act_normalize(q_raw);
            cov3d_to_scale_rot_backward(vp.scale_modifier, s, q, g_cov, ds, dq);
            // ... and take the gradients back through the activations: exp' = exp, and for y = x / |x|
            // dx = (dy - y (y . dy)) / |x|
            ds.x *= s.x; ds.y *= s.y; ds.z *= s.z;
            if (scale_dim == 1) {
                dL_dscale[idx] = ds.x + ds.y + ds.z;       // render() repeats the one scale three times
            } else {
                dL_dscale[3 * idx + 0] = ds.x; dL_dscale[3 * idx + 1] = ds.y; dL_dscale[3 * idx + 2] = ds.z;
            }
            const float n = fmaxf(sqrtf(q_raw.x * q_raw.x + q_raw.y * q_raw.y + q_raw.z * q_raw.z + q_raw.w * q_raw.w), 1e-12f);
            const float yd = q.x * dq.x + q.y * dq.y + q.z * dq.z + q.w * dq.w;
            reinterpret_cast<float4*>(dL_drot)[idx] = make_float4((dq.x - q.x * yd) / n, (dq.y - q.y * yd) / n,
                                                                  (dq.z - q.z * yd) / n, (dq.w - q.w * yd) / n);
        } else {
            s = {scales[3 * idx], scales[3 * idx + 1], scales[3 * idx + 2]};
            cov3d_to_scale_rot_backward(vp.scale_modifier, s, q, g_cov, ds, dq);
            dL_dscale[3 * idx + 0] = ds.x; dL_dscale[3 * idx + 1] = ds.y; dL_dscale[3 * idx + 2] = ds.z;
            reinterpret_cast<float4*>(dL_drot)[idx] = dq;
        }
    }
    if (RAW) {
        // sigmoid' = o (1 - o), in place; colour = max(C0 f_dc + 0.5, 0) (PyTorch's clamp passes the
        // gradient where the input is >= the bound); the 4th channel is the key-point score itself
        const float o = act_sigmoid(raw_opacities[idx]);
        dL_dopacity[idx] = dL_dopacity[idx] * (1.0f - o) * o;
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            const float pre = __fadd_rn(__fmul_rn(0.28209479177387814f, features_dc[3 * idx + ch]), 0.5f);
            dL_dfdc[3 * idx + ch] = pre >= 0.0f ? 0.28209479177387814f * dL_dcolor[idx * C + ch] : 0.0f;
        }
        if (C == 4 && dL_dkp != nullptr) dL_dkp[idx] = dL_dcolor[idx * C + 3];
    }
    }   // seen
    // Multi-GPU training step: instead of a separate all-reduce of the gradient block afterwards, the
    // finished gradients of every Gaussian some view of THIS GPU saw go straight into every GPU's
    // exchange buffer through the switch (common.cuh).  Rows no GPU saw are never touched: the traffic
    // is the visible fraction of the block, and it leaves while the rest of the kernel is still
    // computing.  The rows are staged through shared memory so that one warp instruction adds whole
    // 64-byte rows (lanes 4j .. 4j+3: the four quarters of one row; 16-byte accesses scattered over
    // 32 rows per instruction measured 4x slower on the link).
    if (push_mc != nullptr && scales != nullptr) {
        extern __shared__ __align__(16) float s_rows[];           // [256][EXCHANGE_ROW] + [256] flags
        float* mine = s_rows + (size_t)threadIdx.x * EXCHANGE_ROW;
        uint32_t* s_seen = reinterpret_cast<uint32_t*>(s_rows + 256 * EXCHANGE_ROW);
        s_seen[threadIdx.x] = seen ? 1u : 0u;
        if (seen) {
            const float4 rq = reinterpret_cast<const float4*>(dL_drot)[idx];
            float sx, sy = 0.f, sz = 0.f;
            if (RAW && scale_dim == 1) {
                sx = dL_dscale[idx];
            } else {
                sx = dL_dscale[3 * idx]; sy = dL_dscale[3 * idx + 1]; sz = dL_dscale[3 * idx + 2];
            }
            float c0, c1, c2, c3 = 0.f;
            if (RAW) {
                c0 = dL_dfdc[3 * idx]; c1 = dL_dfdc[3 * idx + 1]; c2 = dL_dfdc[3 * idx + 2];
                if (C == 4 && dL_dkp != nullptr) c3 = dL_dkp[idx];
            } else {
                c0 = dL_dcolor[idx * C]; c1 = dL_dcolor[idx * C + 1]; c2 = dL_dcolor[idx * C + 2];
                if (C == 4) c3 = dL_dcolor[idx * C + 3];
            }
            reinterpret_cast<float4*>(mine)[0] = make_float4(g_mean.x, g_mean.y, g_mean.z, dL_dopacity[idx]);
            reinterpret_cast<float4*>(mine)[1] = make_float4(sx, sy, sz, (float)n_seen);
            reinterpret_cast<float4*>(mine)[2] = rq;
            reinterpret_cast<float4*>(mine)[3] = make_float4(c0, c1, c2, c3);
        }
        __syncwarp();
        const int warp_row0 = threadIdx.x & ~31;                   // first row of this warp in the stage
        const size_t gauss0 = (size_t)blockIdx.x * blockDim.x + warp_row0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int r = 8 * k + (lane >> 2), q = lane & 3;       // row within the warp, quarter of the row
            if (s_seen[warp_row0 + r]) {
                const float4 v = reinterpret_cast<const float4*>(s_rows + (size_t)(warp_row0 + r) * EXCHANGE_ROW)[q];
                multimem_red_add_v4(push_mc + (gauss0 + r) * EXCHANGE_ROW + 4 * q, v.x, v.y, v.z, v.w);
            }
        }
    }
}

int launch_preprocess_backward(int P, int D, int M, int C, const float* means3D, const int* radii,
                               const float* shs, const uint8_t* clamped, const float* scales,
                               const float* rotations, const float* cov3D, const float* viewmatrices,
                               const float* projmatrices, const float* cam_positions, const ViewParams& vp,
                               const float* dL_dmean2D, const float* dL_dconic, float* dL_dmean3D,
                               const float* dL_dcolor, size_t color_view_stride, const float* dL_ddepth,
                               float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot, float* dL_dtau,
                               const RawParams* raw, const float* raw_opacities, const float* features_dc,
                               float* dL_dopacity, float* dL_dfdc, float* dL_dkp, float* push_mc, cudaStream_t stream)
{
    const int scale_dim = raw != nullptr ? raw->scale_dim : 3;
    const size_t smem = push_mc != nullptr ? (size_t)256 * (EXCHANGE_ROW + 1) * sizeof(float) : 0;
#define SL_LAUNCH_PBWD(MM, RR)                                                                                   \
    launch_chained(preprocess_backward_kernel<MM, RR>, dim3(ceil_div(P, 256)), dim3(256), smem, stream,          \
        P, D, M, C, means3D, radii, shs, clamped, scales, rotations, cov3D, viewmatrices, projmatrices,         \
        cam_positions, vp, dL_dmean2D, dL_dconic, dL_dmean3D, dL_dcolor, color_view_stride, dL_ddepth, dL_dcov3D, \
        dL_dsh, dL_dscale, dL_drot, dL_dtau, scale_dim, raw_opacities, features_dc, dL_dopacity, dL_dfdc, dL_dkp, \
        push_mc)
    if (raw != nullptr) {
        if (vp.views > 1) SL_LAUNCH_PBWD(true, true); else SL_LAUNCH_PBWD(false, true);
    } else {
        if (vp.views > 1) SL_LAUNCH_PBWD(true, false); else SL_LAUNCH_PBWD(false, false);
    }
#undef SL_LAUNCH_PBWD
    count_launch();
    return 0;
}

}  // namespace sl
