// Shared device helpers for the B200 (sm_100a) Gaussian-splatting path.
//
// The arithmetic recipes here are the ones that decide bit-exact parity with the
// reference rasterizer (SURVEY.md §8 a-bis): point transforms as left-to-right
// mul/add chains (reference: cuda_rasterizer/auxiliary.h:58-97), ndc2Pix in FP64
// (auxiliary.h:41-44), tile rectangle with truncating casts (auxiliary.h:46-56),
// column-major 3x3 products evaluated per element as a0*b0 + a1*b1 + a2*b2.
// Compiled with nvcc defaults (-fmad=true, no fast-math) exactly like the reference.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#define GS_TILE_X 16
#define GS_TILE_Y 16
#define GS_TILE_PIX (GS_TILE_X * GS_TILE_Y)

#define GS_DEV __device__ __forceinline__

// Error codes returned through the C ABI (see include/gs_b200.h).
#define GS_OK 0
#define GS_ERR_ARG (-1)
#define GS_ERR_WORKSPACE (-2)
#define GS_ERR_LIMIT (-3)

#define GS_CUDA_TRY(expr)                                  \
    do {                                                   \
        cudaError_t _e = (expr);                           \
        if (_e != cudaSuccess) return (int)_e;             \
    } while (0)

#define GS_LAUNCH_CHECK()                                  \
    do {                                                   \
        cudaError_t _e = cudaGetLastError();               \
        if (_e != cudaSuccess) return (int)_e;             \
    } while (0)

namespace gs {

// ------------------------------------------------------------------------------------------------
// Programmatic dependent launch (sm_90+). A step is a chain of ~25 short dependent kernels on one stream; between two of
// them the stream normally drains, flushes and only then starts launching the next grid (~2 us each). A kernel launched
// with launch_pdl() may have its CTAs scheduled as soon as every CTA of its predecessor has executed pdl_trigger() (i.e. the
// predecessor is fully resident); they park at pdl_wait(), which returns once the predecessor grid has COMPLETED and its
// writes are visible. Rule for every kernel on the chain: `pdl_trigger(); pdl_wait();` are its first statements, before
// any global-memory access. Both are no-ops for a grid launched the ordinary way or preceded by a memset / memcpy.
// Used for scenes of >= 1M Gaussians (pdl_hint); GS_PDL=0 / 1 forces it off / on.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
bool pdl_enabled();               // api.cu
void pdl_hint(long long gaussians);   // api.cu: called at the head of every entry point that launches a chain
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// Spherical-harmonics basis constants (real SH, degrees 0..3).
__device__ constexpr float kSH0 = 0.28209479177387814f;
__device__ constexpr float kSH1 = 0.4886025119029199f;
__device__ constexpr float kSH2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                     -1.0925484305920792f, 0.5462742152960396f};
__device__ constexpr float kSH3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                     0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                                     -0.5900435899266435f};

// Real SH basis b[0..(deg+1)^2) at the unit direction (x, y, z), every operation rounded on its own (no FMA contraction):
// the per-Gaussian backward (preprocess_bwd.cu) and the exchange's rebuild kernels (dp_exchange.cu) must produce the same
// bits from the same direction whatever the compiler does around them.
__device__ __forceinline__ void sh_basis_rn(int deg, float x, float y, float z, float* __restrict__ b) {
    b[0] = kSH0;
    if (deg < 1) return;
    b[1] = __fmul_rn(-kSH1, y);
    b[2] = __fmul_rn(kSH1, z);
    b[3] = __fmul_rn(-kSH1, x);
    if (deg < 2) return;
    const float xx = __fmul_rn(x, x), yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
    const float xy = __fmul_rn(x, y), yz = __fmul_rn(y, z), xz = __fmul_rn(x, z);
    const float xx_yy = __fsub_rn(xx, yy);
    b[4] = __fmul_rn(kSH2[0], xy);
    b[5] = __fmul_rn(kSH2[1], yz);
    b[6] = __fmul_rn(kSH2[2], __fsub_rn(__fsub_rn(__fmul_rn(2.f, zz), xx), yy));
    b[7] = __fmul_rn(kSH2[3], xz);
    b[8] = __fmul_rn(kSH2[4], xx_yy);
    if (deg < 3) return;
    const float four_zz_xx_yy = __fsub_rn(__fsub_rn(__fmul_rn(4.f, zz), xx), yy);
    b[9] = __fmul_rn(__fmul_rn(kSH3[0], y), __fsub_rn(__fmul_rn(3.f, xx), yy));
    b[10] = __fmul_rn(__fmul_rn(kSH3[1], xy), z);
    b[11] = __fmul_rn(__fmul_rn(kSH3[2], y), four_zz_xx_yy);
    b[12] = __fmul_rn(__fmul_rn(kSH3[3], z), __fsub_rn(__fsub_rn(__fmul_rn(2.f, zz), __fmul_rn(3.f, xx)), __fmul_rn(3.f, yy)));
    b[13] = __fmul_rn(__fmul_rn(kSH3[4], x), four_zz_xx_yy);
    b[14] = __fmul_rn(__fmul_rn(kSH3[5], z), xx_yy);
    b[15] = __fmul_rn(__fmul_rn(kSH3[6], x), __fsub_rn(xx, __fmul_rn(3.f, yy)));
}


struct V3 {
    float x, y, z;
    GS_DEV V3() {}
    GS_DEV V3(float a, float b, float c) : x(a), y(b), z(c) {}
    GS_DEV float& operator[](int i) { return (&x)[i]; }
    GS_DEV const float& operator[](int i) const { return (&x)[i]; }
};
GS_DEV V3 operator+(const V3& a, const V3& b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
GS_DEV V3 operator-(const V3& a, const V3& b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
GS_DEV V3 operator*(float s, const V3& v) { return V3(s * v.x, s * v.y, s * v.z); }
GS_DEV V3 operator*(const V3& v, float s) { return V3(v.x * s, v.y * s, v.z * s); }
GS_DEV V3 operator/(const V3& v, float s) { return V3(v.x / s, v.y / s, v.z / s); }
GS_DEV float dot3(const V3& a, const V3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// |v|^2 with the contraction nvcc gives dot3(v, v) in a straight-line context (y*y, then x and z fused), spelled out:
// the view direction must come out bit-identical in every kernel that rebuilds an SH basis from it (preprocess
// backward, dp_exchange's sh_grad_combine), and the compiler's own contraction depends on block placement.
GS_DEV float norm2_fixed(const V3& v) { return __fmaf_rn(v.z, v.z, __fmaf_rn(v.x, v.x, __fmul_rn(v.y, v.y))); }

// Column-major 3x3: c[j] is column j, c[j][i] is row i of column j.
struct M3 {
    V3 c[3];
    GS_DEV M3() {}
    GS_DEV M3(float x0, float y0, float z0, float x1, float y1, float z1, float x2, float y2, float z2) {
        c[0] = V3(x0, y0, z0); c[1] = V3(x1, y1, z1); c[2] = V3(x2, y2, z2);
    }
    GS_DEV V3& operator[](int j) { return c[j]; }
    GS_DEV const V3& operator[](int j) const { return c[j]; }
};
GS_DEV M3 operator*(const M3& a, const M3& b) {
    M3 r;
#pragma unroll
    for (int j = 0; j < 3; ++j)
#pragma unroll
        for (int i = 0; i < 3; ++i)
            r[j][i] = a[0][i] * b[j][0] + a[1][i] * b[j][1] + a[2][i] * b[j][2];
    return r;
}
GS_DEV M3 operator*(float s, const M3& m) {
    M3 r; r[0] = m[0] * s; r[1] = m[1] * s; r[2] = m[2] * s; return r;
}
GS_DEV M3 transpose(const M3& m) {
    return M3(m[0][0], m[1][0], m[2][0], m[0][1], m[1][1], m[2][1], m[0][2], m[1][2], m[2][2]);
}

// matrix = 16 floats, element [4*i+j] of the transposed matrix (column-major of the math matrix).
GS_DEV float3 xform_point_4x3(const float3& p, const float* __restrict__ m) {
    float3 t = {
        m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
        m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
        m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14],
    };
    return t;
}
GS_DEV float4 xform_point_4x4(const float3& p, const float* __restrict__ m) {
    float4 t = {
        m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
        m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
        m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14],
        m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15]};
    return t;
}
GS_DEV float3 xform_vec_4x3_T(const float3& p, const float* __restrict__ m) {
    float3 t = {
        m[0] * p.x + m[1] * p.y + m[2] * p.z,
        m[4] * p.x + m[5] * p.y + m[6] * p.z,
        m[8] * p.x + m[9] * p.y + m[10] * p.z,
    };
    return t;
}

// NDC -> pixel; the literals are doubles on purpose (the reference evaluates this in FP64).
GS_DEV float ndc_to_pix(float v, int S) { return ((v + 1.0) * S - 1.0) * 0.5; }

GS_DEV void tile_rect(const float2 p, int max_radius, uint2& rect_min, uint2& rect_max, const uint2 grid) {
    rect_min = {min(grid.x, (unsigned)max((int)0, (int)((p.x - max_radius) / GS_TILE_X))),
                min(grid.y, (unsigned)max((int)0, (int)((p.y - max_radius) / GS_TILE_Y)))};
    rect_max = {min(grid.x, (unsigned)max((int)0, (int)((p.x + max_radius + GS_TILE_X - 1) / GS_TILE_X))),
                min(grid.y, (unsigned)max((int)0, (int)((p.y + max_radius + GS_TILE_Y - 1) / GS_TILE_Y)))};
}

GS_DEV float3 dnormvdv3(float3 v, float3 dv) {
    float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
    float invsum32 = 1.0f / sqrt(sum2 * sum2 * sum2);
    float3 r;
    r.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
    r.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
    r.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
    return r;
}

// Quaternion i of a [P][4] array: one 16-byte access when the base is 16-byte aligned, four scalar ones otherwise
// (a contiguous view with an odd storage offset, or an arena slice at an odd Gaussian count).
GS_DEV float4 load_quat(const float* __restrict__ q, size_t i, bool aligned) {
    if (aligned) return reinterpret_cast<const float4*>(q)[i];
    return make_float4(q[4 * i], q[4 * i + 1], q[4 * i + 2], q[4 * i + 3]);
}
GS_DEV void store_quat(float* __restrict__ q, size_t i, bool aligned, const float4 v) {
    if (aligned) { reinterpret_cast<float4*>(q)[i] = v; return; }
    q[4 * i] = v.x; q[4 * i + 1] = v.y; q[4 * i + 2] = v.z; q[4 * i + 3] = v.w;
}
static inline bool is_aligned16(const void* p) { return (reinterpret_cast<size_t>(p) & 15) == 0; }

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace gs
