// splatloc_b200/csrc/common.cuh
//
// Shared device/host helpers for the B200 (sm_100a) Gaussian-splat rasterizer.
// No torch, no GLM, no CUB: everything the kernels need lives here.
//
// Arithmetic notes.  Integer outputs of the path (radii, tile counts, sort
// keys, tile ranges, n_contrib) must be bit-identical to the reference, and
// they are functions of FP32 intermediates.  The helpers below therefore keep
// the *term order* of the reference's expressions (its matrix helpers are in
// cuda_rasterizer/auxiliary.h:62-103 and the vendored GLM's operator*,
// third_party/glm/glm/detail/type_mat3x3.inl:486-520) so that nvcc's default
// fused-multiply-add contraction produces the same roundings.
#pragma once
#include <utility>
#include <cstdint>
#include <cstddef>
#include <cuda_runtime.h>

#define SL_TILE_X 16
#define SL_TILE_Y 16
#define SL_TILE_PIX (SL_TILE_X * SL_TILE_Y)

namespace sl {

// ---------------------------------------------------------------- error state
// Thread-local last-error text for the C ABI (sl_last_error()).
void set_error(const char* fmt, ...);
const char* last_error();

// Returns 0 on success; on failure records the CUDA error text and returns
// the negative error code used by the C ABI.
int check_cuda(cudaError_t e, const char* what);

#define SL_CUDA(call, what)                         \
    do {                                            \
        int _rc = ::sl::check_cuda((call), (what)); \
        if (_rc != 0) return _rc;                   \
    } while (0)

// After a launch: always pick up launch-configuration errors; with `debug`
// also synchronise so that execution errors surface at the stage that caused
// them (reference: CHECK_CUDA, cuda_rasterizer/auxiliary.h:274-283).
int check_stage(bool debug, cudaStream_t stream, const char* what);

#define SL_STAGE(debug, stream, what)                          \
    do {                                                       \
        int _rc = ::sl::check_stage((debug), (stream), (what)); \
        if (_rc != 0) return _rc;                              \
    } while (0)

// Counts this library's own kernel launches (read back through sl_launch_count()).
void count_launch(int n = 1);

// ---- chained launches (programmatic dependent launch) ------------------------------------------
// The stages of a step form one dependent chain of short kernels; between two of them the GPU would
// drain, take the next launch and ramp up again.  Launched through launch_chained() a kernel's CTAs may
// become resident while the previous kernel of the stream is still running its last CTAs; they block
// in chain_wait() -- the first statement of every such kernel, before ANY global access -- until that
// kernel has finished and its writes are visible, and let the next kernel's CTAs in behind them with
// chain_release().  Off (sl_set_chained_launches(0)) the same kernels are launched plainly and both
// calls are no-ops.
bool chained_launches_enabled();
#ifdef __CUDACC__
__device__ __forceinline__ void chain_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void chain_release() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_chained(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                  Args&&... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = chained_launches_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}
#endif

// --------------------------------------------------------------- small maths
struct f3 { float x, y, z; };

// Column-major 3x3, c[col][row]; product term order matches GLM so that the
// compiler's fma contraction rounds the same way.
struct mat3 {
    float c[3][3];
};

__device__ __forceinline__ mat3 mat3_cols(float a00, float a01, float a02,
                                          float a10, float a11, float a12,
                                          float a20, float a21, float a22)
{
    mat3 m;
    m.c[0][0] = a00; m.c[0][1] = a01; m.c[0][2] = a02;
    m.c[1][0] = a10; m.c[1][1] = a11; m.c[1][2] = a12;
    m.c[2][0] = a20; m.c[2][1] = a21; m.c[2][2] = a22;
    return m;
}

__device__ __forceinline__ mat3 mul(const mat3& a, const mat3& b)
{
    mat3 r;
#pragma unroll
    for (int col = 0; col < 3; ++col)
#pragma unroll
        for (int row = 0; row < 3; ++row)
            r.c[col][row] = a.c[0][row] * b.c[col][0] + a.c[1][row] * b.c[col][1] + a.c[2][row] * b.c[col][2];
    return r;
}

__device__ __forceinline__ mat3 transpose(const mat3& a)
{
    mat3 r;
#pragma unroll
    for (int col = 0; col < 3; ++col)
#pragma unroll
        for (int row = 0; row < 3; ++row)
            r.c[col][row] = a.c[row][col];
    return r;
}

// p' = M p for the 4x4 matrices handed to the op: element m[4*c + r] is the
// mathematical M[r][c] (auxiliary.h:62-81).
__device__ __forceinline__ f3 xform_4x3(const f3& p, const float* __restrict__ m)
{
    f3 t;
    t.x = m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12];
    t.y = m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13];
    t.z = m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14];
    return t;
}

__device__ __forceinline__ float4 xform_4x4(const f3& p, const float* __restrict__ m)
{
    float4 t;
    t.x = m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12];
    t.y = m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13];
    t.z = m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14];
    t.w = m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15];
    return t;
}

// NDC -> pixel; evaluated in double like the reference (auxiliary.h:45-48:
// the literals 1.0 / 0.5 are doubles).
__device__ __forceinline__ float ndc_to_pix(float v, int S)
{
    return ((v + 1.0) * S - 1.0) * 0.5;
}

// Tile rectangle of a splat (auxiliary.h:50-60); double arithmetic, truncation
// toward zero, clamped to the grid.
__device__ __forceinline__ void tile_rect(float px, float py, int radius, int grid_x, int grid_y,
                                          uint2& rmin, uint2& rmax)
{
    rmin.x = min(grid_x, max(0, (int)((px - radius - 0.5) / SL_TILE_X)));
    rmin.y = min(grid_y, max(0, (int)((py - radius - 0.5) / SL_TILE_Y)));
    rmax.x = min(grid_x, max(0, (int)((px + radius + SL_TILE_X - 1 + 0.5) / SL_TILE_X)));
    rmax.y = min(grid_y, max(0, (int)((py + radius + SL_TILE_Y - 1 + 0.5) / SL_TILE_Y)));
}

// Frustum test (auxiliary.h:148-168): NDC z in (-1.01, 1.01), x,y in (-1.2, 1.2).
__device__ __forceinline__ bool in_frustum_ndc(const f3& p, const float* __restrict__ proj)
{
    float4 h = xform_4x4(p, proj);
    float w = 1.0f / (h.w + 0.0000001f);
    float x = h.x * w, y = h.y * w, z = h.z * w;
    const float pad = 0.01f, xy_pad = 0.2f;
    return (z > -1 - pad) && (z < 1 + pad) && (x > -1 - xy_pad) && (x < 1. + xy_pad) &&
           (y > -1 - xy_pad) && (y < 1. + xy_pad);
}

// ---- raw-parameter mode: the activations SplatLoc's render() applies in PyTorch
// before it calls the rasterizer (gaussian_model.py:75-110 getters, gaussian_renderer/__init__.py:85-102),
// evaluated inside the kernels.  Written to round like the PyTorch float32 kernels they replace:
//   opacity  = sigmoid(x)            = 1 / (1 + exp(-x))
//   scale    = exp(x)
//   rotation = x / max(|x|, 1e-12)   (torch.nn.functional.normalize)
//   colour   = max(C0 * f_dc + 0.5, 0), a separate multiply and add (eval_sh at degree 0, sh_utils.py:73;
//              then clamp_min(. + 0.5, 0))
struct RawParams {
    const float* kp_score;   // [P] appended as the 4th feature channel (render()'s torch.cat), or NULL
    int scale_dim;           // 3, or 1 for isotropic Gaussians (render() repeats the one scale)
};

__device__ __forceinline__ float act_sigmoid(float x) { return 1.0f / (1.0f + expf(-x)); }
__device__ __forceinline__ float act_sh0_colour(float f)
{
    return fmaxf(__fadd_rn(__fmul_rn(0.28209479177387814f, f), 0.5f), 0.0f);
}
__device__ __forceinline__ float4 act_normalize(float4 q)
{
    const float n = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
    const float d = fmaxf(n, 1e-12f);
    return make_float4(q.x / d, q.y / d, q.z / d, q.w / d);
}

// ---- multi-GPU gradient exchange fused into the per-Gaussian backward (csrc/preprocess_bwd.cu) ----
// One row of EXCHANGE_ROW floats per Gaussian: {dL/dmean xyz, dL/dopacity | dL/dscale xyz, seen |
// dL/drot wxyz | dL/dcolour c0..c3}; "seen" counts the views (over all GPUs) in which the Gaussian was
// visible: every push adds the number of its own views that saw it.  `mc` is a MULTICAST address (NVSwitch, NVLS) that maps the same row of every GPU's exchange
// buffer: one multimem.red adds the four floats into all of them inside the switch.
constexpr int EXCHANGE_ROW = 16;
__device__ __forceinline__ void multimem_red_add_v4(float* mc, float a, float b, float c, float d)
{
    asm volatile("multimem.red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};"
                 ::"l"(mc), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__host__ __device__ __forceinline__ size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

}  // namespace sl
