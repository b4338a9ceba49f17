// splatloc_b200/csrc/ssim_loss.cu
//
// SURVEY 8(f) row 2, SSIM variant: the photometric loss of SplatLoc's colour-refinement loop
// (train_gaussians.py:283-286), the step immediately after the rasterizer's forward in that loop:
//     loss = (1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image, gt))
//   ssim / _ssim      gaussian_splatting/utils/loss_utils.py:61-102   (11 x 11 Gaussian window, sigma 1.5,
//                     zero padding, C1 = 0.01^2, C2 = 0.03^2, mean over all elements)
//   l1_loss           gaussian_splatting/utils/loss_utils.py:  mean |image - gt|
// The reference runs five depth-wise 11 x 11 convolutions, ~15 elementwise kernels and the autograd
// mirror of all of them per iteration.  Here one kernel reads a 26 x 26 halo tile of both images per
// 16 x 16 output tile into shared memory, filters {x, y, x^2, y^2, xy} separably (the window is an outer
// product, loss_utils.py:48-58), forms the SSIM map and the three partial-derivative maps a backward needs,
// and reduces SSIM and L1 sums deterministically (one partial per CTA, summed in a fixed order by the last CTA to
// finish); a second kernel filters the derivative maps with the
// same window and writes d loss / d image.
//
// With  mu1 = w*x, mu2 = w*y, s11 = w*x^2, s22 = w*y^2, s12 = w*xy  (w* = window filter),
//   A = 2 mu1 mu2 + C1,  B = 2 (s12 - mu1 mu2) + C2,  D = mu1^2 + mu2^2 + C1,  E = (s11 - mu1^2) + (s22 - mu2^2) + C2,
//   m = A B / (D E),
//   dm/ds11 = -A B / (D E^2),  dm/ds12 = 2 A / (D E),
//   dm/dmu1 = 2 mu2 B / (D E) - 2 mu1 A B / (D^2 E) - 2 mu1 dm/ds11 - mu2 dm/ds12,
//   d(sum m)/dx(p) = sum_q w(q - p) [ dm/dmu1(q) + 2 x(p) dm/ds11(q) + y(p) dm/ds12(q) ].
//
// HBM: forward reads 2 and writes 3 floats per element, backward reads 5 and writes 1: 44 B per element.
#include "kernels.h"

namespace sl {

namespace {

constexpr int SS_T = 16;             // output tile edge
constexpr int SS_R = 5;              // window radius (11 taps)
constexpr int SS_H = SS_T + 2 * SS_R;   // halo tile edge: 26
constexpr float SS_C1 = 0.01f * 0.01f, SS_C2 = 0.03f * 0.03f;

struct Window11 { float w[11]; };

__device__ __forceinline__ float load_or_zero(const float* __restrict__ img, int x, int y, int W, int H)
{
    return (x >= 0 && x < W && y >= 0 && y < H) ? img[(size_t)y * W + x] : 0.0f;
}

// grid: (tiles_x, tiles_y, C); block 16 x 16
__global__ void __launch_bounds__(SS_T * SS_T)
ssim_forward_kernel(int W, int H, const float* __restrict__ img1, const float* __restrict__ img2, Window11 win,
                    float* __restrict__ maps, double* __restrict__ partial, unsigned int* ticket, double inv_n,
                    float* __restrict__ out)
{
    chain_wait();
    chain_release();
    __shared__ float s_x[SS_H][SS_H + 1], s_y[SS_H][SS_H + 1];
    __shared__ float s_h[5][SS_H][SS_T + 1];
    __shared__ double s_red[SS_T * SS_T / 32][2];
    __shared__ double s_fin[SS_T * SS_T][2];
    __shared__ bool s_last;
    const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * SS_T + tx;
    const int x0 = blockIdx.x * SS_T, y0 = blockIdx.y * SS_T, c = blockIdx.z;
    const size_t plane = (size_t)W * H;
    const float* a = img1 + c * plane;
    const float* b = img2 + c * plane;

    for (int i = tid; i < SS_H * SS_H; i += SS_T * SS_T) {
        const int ly = i / SS_H, lx = i - ly * SS_H;
        s_x[ly][lx] = load_or_zero(a, x0 + lx - SS_R, y0 + ly - SS_R, W, H);
        s_y[ly][lx] = load_or_zero(b, x0 + lx - SS_R, y0 + ly - SS_R, W, H);
    }
    __syncthreads();
    // horizontal pass: 26 rows x 16 columns of {x, y, x^2, y^2, xy}
    for (int i = tid; i < SS_H * SS_T; i += SS_T * SS_T) {
        const int ly = i / SS_T, lx = i - ly * SS_T;
        float hx = 0.f, hy = 0.f, hxx = 0.f, hyy = 0.f, hxy = 0.f;
#pragma unroll
        for (int k = 0; k < 11; ++k) {
            const float w = win.w[k], xv = s_x[ly][lx + k], yv = s_y[ly][lx + k];
            hx = fmaf(w, xv, hx);
            hy = fmaf(w, yv, hy);
            hxx = fmaf(w, xv * xv, hxx);
            hyy = fmaf(w, yv * yv, hyy);
            hxy = fmaf(w, xv * yv, hxy);
        }
        s_h[0][ly][lx] = hx; s_h[1][ly][lx] = hy; s_h[2][ly][lx] = hxx; s_h[3][ly][lx] = hyy; s_h[4][ly][lx] = hxy;
    }
    __syncthreads();
    float mu1 = 0.f, mu2 = 0.f, s11 = 0.f, s22 = 0.f, s12 = 0.f;
#pragma unroll
    for (int k = 0; k < 11; ++k) {
        const float w = win.w[k];
        mu1 = fmaf(w, s_h[0][ty + k][tx], mu1);
        mu2 = fmaf(w, s_h[1][ty + k][tx], mu2);
        s11 = fmaf(w, s_h[2][ty + k][tx], s11);
        s22 = fmaf(w, s_h[3][ty + k][tx], s22);
        s12 = fmaf(w, s_h[4][ty + k][tx], s12);
    }
    const int px = x0 + tx, py = y0 + ty;
    const bool inside = px < W && py < H;
    double m_d = 0.0, l1_d = 0.0;
    if (inside) {
        const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
        const float sig1 = s11 - mu1_sq, sig2 = s22 - mu2_sq, sig12 = s12 - mu12;
        const float A = 2.0f * mu12 + SS_C1, B = 2.0f * sig12 + SS_C2;
        const float D = mu1_sq + mu2_sq + SS_C1, E = sig1 + sig2 + SS_C2;
        const float inv_DE = 1.0f / (D * E);
        const float m = A * B * inv_DE;
        const float dm_ds11 = -m / E;
        const float dm_ds12 = 2.0f * A * inv_DE;
        const float dm_dmu1 = 2.0f * mu2 * B * inv_DE - 2.0f * mu1 * m / D - 2.0f * mu1 * dm_ds11 - mu2 * dm_ds12;
        const size_t o = c * plane + (size_t)py * W + px;
        const size_t CHW = (size_t)gridDim.z * plane;
        maps[o] = dm_dmu1;
        maps[CHW + o] = dm_ds11;
        maps[2 * CHW + o] = dm_ds12;
        m_d = (double)m;
        l1_d = (double)fabsf(s_x[ty + SS_R][tx + SS_R] - s_y[ty + SS_R][tx + SS_R]);
    }
    // deterministic sums: warp shuffle, shared memory over the 8 warps, one partial per CTA
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m_d += __shfl_xor_sync(0xffffffffu, m_d, o);
        l1_d += __shfl_xor_sync(0xffffffffu, l1_d, o);
    }
    if ((tid & 31) == 0) { s_red[tid >> 5][0] = m_d; s_red[tid >> 5][1] = l1_d; }
    __syncthreads();
    if (tid == 0) {
        double sm = 0.0, sl = 0.0;
#pragma unroll
        for (int w = 0; w < SS_T * SS_T / 32; ++w) { sm += s_red[w][0]; sl += s_red[w][1]; }
        const size_t blk = ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        partial[2 * blk] = sm;
        partial[2 * blk + 1] = sl;
        __threadfence();
        const unsigned int n_blocks = gridDim.x * gridDim.y * gridDim.z;
        s_last = atomicAdd(ticket, 1u) == n_blocks - 1;
    }
    __syncthreads();
    if (!s_last) return;
    // the last CTA to finish sums the partials in a fixed order (thread-strided, then a tree): the means
    // do not depend on which CTA that was
    __threadfence();
    const int n_partials = (int)(gridDim.x * gridDim.y * gridDim.z);
    double sm = 0.0, sl = 0.0;
    for (int i = tid; i < n_partials; i += SS_T * SS_T) { sm += partial[2 * i]; sl += partial[2 * i + 1]; }
    s_fin[tid][0] = sm; s_fin[tid][1] = sl;
    __syncthreads();
    for (int o = SS_T * SS_T / 2; o > 0; o >>= 1) {
        if (tid < o) { s_fin[tid][0] += s_fin[tid + o][0]; s_fin[tid][1] += s_fin[tid + o][1]; }
        __syncthreads();
    }
    if (tid == 0) {
        out[0] = (float)(s_fin[0][0] * inv_n);
        out[1] = (float)(s_fin[0][1] * inv_n);
        *ticket = 0u;                                   // ready for the next call
    }
}

// d loss / d img1 = upstream * ( scale_ssim * d(sum m)/dx + scale_l1 * sign(x - y) )
__global__ void __launch_bounds__(SS_T * SS_T)
ssim_backward_kernel(int W, int H, const float* __restrict__ img1, const float* __restrict__ img2, Window11 win,
                     const float* __restrict__ maps, const float* __restrict__ upstream, float scale_ssim,
                     float scale_l1, float* __restrict__ dL_dimg1)
{
    chain_wait();
    chain_release();
    __shared__ float s_m[3][SS_H][SS_H + 1];
    __shared__ float s_h[3][SS_H][SS_T + 1];
    const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * SS_T + tx;
    const int x0 = blockIdx.x * SS_T, y0 = blockIdx.y * SS_T, c = blockIdx.z;
    const size_t plane = (size_t)W * H, CHW = (size_t)gridDim.z * plane;

    for (int i = tid; i < SS_H * SS_H; i += SS_T * SS_T) {
        const int ly = i / SS_H, lx = i - ly * SS_H;
#pragma unroll
        for (int q = 0; q < 3; ++q)
            s_m[q][ly][lx] = load_or_zero(maps + q * CHW + c * plane, x0 + lx - SS_R, y0 + ly - SS_R, W, H);
    }
    __syncthreads();
    for (int i = tid; i < SS_H * SS_T; i += SS_T * SS_T) {
        const int ly = i / SS_T, lx = i - ly * SS_T;
        float h0 = 0.f, h1 = 0.f, h2 = 0.f;
#pragma unroll
        for (int k = 0; k < 11; ++k) {
            const float w = win.w[k];
            h0 = fmaf(w, s_m[0][ly][lx + k], h0);
            h1 = fmaf(w, s_m[1][ly][lx + k], h1);
            h2 = fmaf(w, s_m[2][ly][lx + k], h2);
        }
        s_h[0][ly][lx] = h0; s_h[1][ly][lx] = h1; s_h[2][ly][lx] = h2;
    }
    __syncthreads();
    const int px = x0 + tx, py = y0 + ty;
    if (px >= W || py >= H) return;
    float g0 = 0.f, g1 = 0.f, g2 = 0.f;
#pragma unroll
    for (int k = 0; k < 11; ++k) {
        const float w = win.w[k];
        g0 = fmaf(w, s_h[0][ty + k][tx], g0);
        g1 = fmaf(w, s_h[1][ty + k][tx], g1);
        g2 = fmaf(w, s_h[2][ty + k][tx], g2);
    }
    const size_t o = c * plane + (size_t)py * W + px;
    const float x = img1[o], y = img2[o];
    const float d = x - y;
    const float sgn = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
    const float up = upstream != nullptr ? upstream[0] : 1.0f;
    dL_dimg1[o] = up * (scale_ssim * (g0 + 2.0f * x * g1 + y * g2) + scale_l1 * sgn);
}

}  // namespace

size_t ssim_workspace_bytes(int W, int H, int C)
{
    return 16 + (size_t)2 * sizeof(double) * ceil_div(W, SS_T) * ceil_div(H, SS_T) * C;    // ticket, then the partials
}

int launch_ssim_forward(int W, int H, int C, const float* img1, const float* img2, const float* window11, float* maps,
                        float* out, char* workspace, cudaStream_t stream)
{
    Window11 win;
    for (int k = 0; k < 11; ++k) win.w[k] = window11[k];
    const dim3 grid(ceil_div(W, SS_T), ceil_div(H, SS_T), C), block(SS_T, SS_T);
    unsigned int* ticket = reinterpret_cast<unsigned int*>(workspace);
    double* partial = reinterpret_cast<double*>(workspace + 16);
    launch_chained(ssim_forward_kernel, grid, block, 0, stream, W, H, img1, img2, win, maps, partial, ticket,
                   1.0 / ((double)W * H * C), out);
    count_launch();
    return 0;
}

int launch_ssim_backward(int W, int H, int C, const float* img1, const float* img2, const float* window11,
                         const float* maps, const float* upstream, float scale_ssim, float scale_l1, float* dL_dimg1,
                         cudaStream_t stream)
{
    Window11 win;
    for (int k = 0; k < 11; ++k) win.w[k] = window11[k];
    const dim3 grid(ceil_div(W, SS_T), ceil_div(H, SS_T), C), block(SS_T, SS_T);
    launch_chained(ssim_backward_kernel, grid, block, 0, stream, W, H, img1, img2, win, maps, upstream, scale_ssim,
                   scale_l1, dL_dimg1);
    count_launch();
    return 0;
}

}  // namespace sl
