// splatloc_b200/csrc/optimizer.cu
//
// SURVEY 8(f) row 4, second half: what SplatLoc's mapping loop does to the Gaussian store between
// two rasterizer calls (reference: gaussian_splatting/scene/gaussian_model.py), as stream-ordered
// kernels over the store's arrays instead of dozens of PyTorch indexing / cat / optimizer-state ops:
//
//   adam_kernel           torch.optim.Adam(l, lr=0.0, eps=1e-15).step() over the eight parameter
//                         groups of training_setup (:250-300) in ONE launch (blockIdx.y = group);
//   densify_stats_kernel  add_densification_stats (:677-679) + the max_radii2D update of
//                         train_gaussians.py:241-244, for all views of a window at once;
//   densify_plan_kernel   the selection masks of densify_and_clone (:640-647), densify_and_split
//                         (:602-611) and the final prune of densify_and_prune (:661-675);
//   densify_apply_kernel  clone / split / prune + the optimizer-state surgery of
//                         cat_tensors_to_optimizer (:529-551) and _prune_optimizer (:493-510) as
//                         ONE compaction: every surviving or new row is written once at its final
//                         position (prefix sums of the plan's flags give the positions).
//
// Row order of the result is the reference's: surviving original rows in ascending order, then
// the clones (ascending source row), then the first child of every split Gaussian (ascending
// source row), then the second children.  New rows get zero Adam moments, kept rows keep theirs;
// the densification statistics of the new store are all zero (densification_postfix :583-585).
// All HBM-bound: ~28 B per parameter element for Adam, ~12 B per moved float for the compaction.
#include "kernels.h"

namespace sl {

// ------------------------------------------------------------------ fused Adam
__global__ void adam_count_kernel(int* step_counter) { *step_counter += 1; }

__global__ void __launch_bounds__(256)
adam_kernel(AdamArgs a, int P, const float* __restrict__ marker, float marker_threshold)
{
    const AdamGroup g = a.group[blockIdx.y];
    float step_size = g.step_size, bc2_sqrt = a.bias_correction2_sqrt;
    if (a.step_counter != nullptr) {
        // bias corrections from the device-side step count (uniform over the grid)
        __shared__ float s_step_size, s_bc2;
        if (threadIdx.x == 0) {
            const double t = (double)*a.step_counter;
            s_step_size = (float)((double)g.step_size / (1.0 - pow(a.beta1_d, t)));
            s_bc2 = (float)sqrt(1.0 - pow(a.beta2_d, t));
        }
        __syncthreads();
        step_size = s_step_size;
        bc2_sqrt = s_bc2;
    }
    const size_t n = (size_t)P * g.width;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        float grad = g.grad[i];
        // train_gaussians.py:232-234: key primitives (marker > 0.005) get a zero position gradient
        // (the moments still move them, as in the reference)
        if (g.freeze_by_marker && marker != nullptr && marker[i / g.width] > marker_threshold) grad = 0.f;
        float m = g.exp_avg[i], v = g.exp_avg_sq[i];
        m = fmaf(a.one_minus_beta1, grad - m, m);                 // exp_avg.lerp_(grad, 1 - beta1)
        v = fmaf(a.one_minus_beta2 * grad, grad, v * a.beta2);    // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
        const float denom = sqrtf(v) / bc2_sqrt + a.eps;
        g.param[i] = g.param[i] - step_size * (m / denom);      // param.addcdiv_(exp_avg, denom, value=-step_size)
        g.exp_avg[i] = m;
        g.exp_avg_sq[i] = v;
    }
}

int launch_adam(const AdamArgs& args, int P, const float* marker, float marker_threshold, int* step_counter,
                cudaStream_t stream)
{
    if (P <= 0 || args.n_groups <= 0) return 0;
    if (step_counter != nullptr) { adam_count_kernel<<<1, 1, 0, stream>>>(step_counter); count_launch(); }
    int widest = 1;
    for (int k = 0; k < args.n_groups; ++k) widest = args.group[k].width > widest ? args.group[k].width : widest;
    const size_t n = (size_t)P * widest;
    const int blocks = (int)((n + 256 * 4 - 1) / (256 * 4));
    const dim3 grid(blocks < 148 * 16 ? blocks : 148 * 16, args.n_groups);
    adam_kernel<<<grid, 256, 0, stream>>>(args, P, marker, marker_threshold); count_launch();
    return 0;
}

// ------------------------------------------------------------------ densification statistics
__global__ void __launch_bounds__(256)
densify_stats_kernel(int P, int n_views, const float* __restrict__ means2D_grad, const int* __restrict__ radii,
                     float* __restrict__ xyz_gradient_accum, float* __restrict__ denom, float* __restrict__ max_radii2D)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    float acc = 0.f, cnt = 0.f, mr = max_radii2D[i];
    for (int v = 0; v < n_views; ++v) {
        const size_t k = (size_t)v * P + i;
        const int r = radii[k];
        if (r > 0) {
            const float gx = means2D_grad[3 * k], gy = means2D_grad[3 * k + 1];
            acc += sqrtf(gx * gx + gy * gy);
            cnt += 1.f;
            mr = fmaxf(mr, (float)r);
        }
    }
    if (cnt > 0.f) {
        xyz_gradient_accum[i] += acc;
        denom[i] += cnt;
        max_radii2D[i] = mr;
    }
}

int launch_densify_stats(int P, int n_views, const float* means2D_grad, const int* radii, float* xyz_gradient_accum,
                         float* denom, float* max_radii2D, cudaStream_t stream)
{
    if (P <= 0) return 0;
    densify_stats_kernel<<<ceil_div(P, 256), 256, 0, stream>>>(P, n_views, means2D_grad, radii, xyz_gradient_accum, denom,
                                                               max_radii2D);
    count_launch();
    return 0;
}

// ------------------------------------------------------------------ densify: plan
// flags[0..3][P] (u32): the row survives as itself | is cloned | is split and its children survive |
// is split (children or not: ranks the noise rows).  Their inclusive prefix sums give the positions.
__global__ void __launch_bounds__(256)
densify_plan_kernel(DensifyParams dp, int P, const float* __restrict__ xyz_gradient_accum,
                    const float* __restrict__ denom, const float* __restrict__ raw_scaling,
                    const float* __restrict__ raw_opacity, const float* __restrict__ marker, uint32_t* __restrict__ flags)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    float grad = xyz_gradient_accum[i] / denom[i];
    if (grad != grad) grad = 0.f;                                   // grads[grads.isnan()] = 0.0
    float s_max = expf(raw_scaling[(size_t)i * dp.scale_dim]);
    for (int k = 1; k < dp.scale_dim; ++k) s_max = fmaxf(s_max, expf(raw_scaling[(size_t)i * dp.scale_dim + k]));
    const bool hot = grad >= dp.grad_threshold;
    const bool clone = hot && s_max <= dp.dense_extent;             // densify_and_clone :642-647
    const bool split = hot && s_max > dp.dense_extent;              // densify_and_split :606-611
    // final prune (:661-675).  max_radii2D was reset to zero by densification_postfix before the
    // reference looks at it, so its "too big on screen" test never fires; the world-size test does.
    const float opacity = act_sigmoid(raw_opacity[i]);
    bool prune_self = opacity < dp.min_opacity, prune_child = prune_self;
    if (dp.use_size_test) {
        prune_self = prune_self || s_max > 0.1f * dp.extent;
        prune_child = prune_child || s_max / (0.8f * 2) > 0.1f * dp.extent;
    }
    if (dp.primitive_reg && marker != nullptr) {                    // key primitives are never pruned
        const bool plain = marker[i] <= 0.005f;
        prune_self = prune_self && plain;
        prune_child = prune_child && plain;
    }
    flags[i] = (!split && !prune_self) ? 1u : 0u;
    flags[(size_t)P + i] = (clone && !prune_self) ? 1u : 0u;
    flags[2 * (size_t)P + i] = (split && !prune_child) ? 1u : 0u;
    flags[3 * (size_t)P + i] = split ? 1u : 0u;
}

__global__ void densify_totals_kernel(int P, const uint32_t* __restrict__ incl, uint32_t* __restrict__ totals)
{
    if (threadIdx.x < 4) totals[threadIdx.x] = incl[(size_t)threadIdx.x * P + P - 1];
}

size_t densify_scratch_bytes(int P)
{
    const size_t scan_blocks = ((size_t)(P > 0 ? P : 1) + SCAN_ITEMS_PER_BLOCK - 1) / SCAN_ITEMS_PER_BLOCK;
    return align_up(8 * (size_t)(P > 0 ? P : 1) * sizeof(uint32_t), 256) + align_up((scan_blocks + 1) * 8, 256) + 512;
}

// scratch: flags [4P] | inclusive scans [4P] | scan descriptors | ticket
int launch_densify_plan(const DensifyParams& dp, int P, const float* xyz_gradient_accum, const float* denom,
                        const float* raw_scaling, const float* raw_opacity, const float* marker, char* scratch,
                        uint32_t* totals, cudaStream_t stream)
{
    if (P <= 0) return 0;
    uint32_t* flags = reinterpret_cast<uint32_t*>(scratch);
    uint32_t* incl = flags + 4 * (size_t)P;
    char* p = scratch + align_up(8 * (size_t)P * sizeof(uint32_t), 256);
    unsigned long long* desc = reinterpret_cast<unsigned long long*>(p);
    const size_t scan_blocks = ((size_t)P + SCAN_ITEMS_PER_BLOCK - 1) / SCAN_ITEMS_PER_BLOCK;
    uint32_t* ticket = reinterpret_cast<uint32_t*>(p + align_up((scan_blocks + 1) * 8, 256));
    densify_plan_kernel<<<ceil_div(P, 256), 256, 0, stream>>>(dp, P, xyz_gradient_accum, denom, raw_scaling, raw_opacity,
                                                              marker, flags);
    count_launch();
    for (int k = 0; k < 4; ++k) {
        const int rc = launch_inclusive_scan(flags + (size_t)k * P, incl + (size_t)k * P, P, desc, ticket, stream);
        if (rc) return rc;
    }
    densify_totals_kernel<<<1, 32, 0, stream>>>(P, incl, totals); count_launch();
    return 0;
}

// ------------------------------------------------------------------ densify: apply
__device__ __forceinline__ void copy_row(const float* __restrict__ src, float* __restrict__ dst, int w)
{
    for (int c = 0; c < w; ++c) dst[c] = src[c];
}
__device__ __forceinline__ void zero_row(float* __restrict__ dst, int w)
{
    for (int c = 0; c < w; ++c) dst[c] = 0.f;
}

__global__ void __launch_bounds__(256)
densify_apply_kernel(DensifyArrays arr, int P, int scale_dim, const uint32_t* __restrict__ flags,
                     const uint32_t* __restrict__ incl, const uint32_t* __restrict__ totals,
                     const float* __restrict__ noise /* [2 * n_split, 3] standard normal */)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const uint32_t nA = totals[0], nB = totals[1], nC = totals[2], nS = totals[3];
    const bool keepA = flags[i] != 0, keepB = flags[(size_t)P + i] != 0, keepC = flags[2 * (size_t)P + i] != 0;
    if (!keepA && !keepB && !keepC) return;
    const uint32_t posA = incl[i] - 1, posB = nA + incl[(size_t)P + i] - 1;
    const uint32_t posC0 = nA + nB + incl[2 * (size_t)P + i] - 1, posC1 = posC0 + nC;
    const uint32_t rankS = incl[3 * (size_t)P + i] - 1;            // row of this Gaussian in the noise block

    // children of a split: position = R(q) (s * z) + xyz, scaling = log(s / (0.8 * 2))   (:613-617)
    float child_xyz[2][3], child_scaling[3];
    if (keepC) {
        const float* q4 = arr.group[arr.rotation_group].param_in + 4 * (size_t)i;
        const float qn = sqrtf(q4[0] * q4[0] + q4[1] * q4[1] + q4[2] * q4[2] + q4[3] * q4[3]);
        const float r = q4[0] / qn, x = q4[1] / qn, y = q4[2] / qn, z = q4[3] / qn;
        const float R[3][3] = {{1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y)},
                               {2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x)},
                               {2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)}};
        const float* sc = arr.group[arr.scaling_group].param_in + (size_t)scale_dim * i;
        float s[3];
        for (int c = 0; c < 3; ++c) s[c] = expf(sc[scale_dim == 1 ? 0 : c]);
        for (int c = 0; c < scale_dim; ++c) child_scaling[c] = logf(s[c] / (0.8f * 2));
        const float* xyz = arr.group[arr.xyz_group].param_in + 3 * (size_t)i;
        for (int k = 0; k < 2; ++k) {
            const float* zz = noise + 3 * ((size_t)k * nS + rankS);
            const float sx = s[0] * zz[0], sy = s[1] * zz[1], sz = s[2] * zz[2];
            for (int c = 0; c < 3; ++c) child_xyz[k][c] = R[c][0] * sx + R[c][1] * sy + R[c][2] * sz + xyz[c];
        }
    }

    for (int gi = 0; gi < arr.n_groups; ++gi) {
        const DensifyGroup g = arr.group[gi];
        const int w = g.width;
        if (w == 0) continue;
        const float* src = g.param_in + (size_t)w * i;
        if (keepA) {
            copy_row(src, g.param_out + (size_t)w * posA, w);
            if (g.exp_avg_in != nullptr) {
                copy_row(g.exp_avg_in + (size_t)w * i, g.exp_avg_out + (size_t)w * posA, w);
                copy_row(g.exp_avg_sq_in + (size_t)w * i, g.exp_avg_sq_out + (size_t)w * posA, w);
            }
        }
        if (keepB) {
            copy_row(src, g.param_out + (size_t)w * posB, w);
            if (g.exp_avg_out != nullptr) {
                zero_row(g.exp_avg_out + (size_t)w * posB, w);
                zero_row(g.exp_avg_sq_out + (size_t)w * posB, w);
            }
        }
        if (keepC) {
            for (int k = 0; k < 2; ++k) {
                const uint32_t pos = k == 0 ? posC0 : posC1;
                float* dst = g.param_out + (size_t)w * pos;
                if (gi == arr.xyz_group) {
                    for (int c = 0; c < 3; ++c) dst[c] = child_xyz[k][c];
                } else if (gi == arr.scaling_group) {
                    for (int c = 0; c < w; ++c) dst[c] = child_scaling[c];
                } else {
                    copy_row(src, dst, w);
                }
                if (g.exp_avg_out != nullptr) {
                    zero_row(g.exp_avg_out + (size_t)w * pos, w);
                    zero_row(g.exp_avg_sq_out + (size_t)w * pos, w);
                }
            }
        }
    }
}

int launch_densify_apply(const DensifyArrays& arr, int P, int scale_dim, const char* scratch, const uint32_t* totals,
                         const float* noise, cudaStream_t stream)
{
    if (P <= 0) return 0;
    const uint32_t* flags = reinterpret_cast<const uint32_t*>(scratch);
    const uint32_t* incl = flags + 4 * (size_t)P;
    densify_apply_kernel<<<ceil_div(P, 256), 256, 0, stream>>>(arr, P, scale_dim, flags, incl, totals, noise);
    count_launch();
    return 0;
}

}  // namespace sl
