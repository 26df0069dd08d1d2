// Fused parameter activations and their regularisers.
//
// The reference turns the raw parameters into rasterizer inputs with torch ops on every iteration
// (scene/gaussian_model_static.py:43-56 setup_functions: scaling = exp, opacity = sigmoid, rotation = F.normalize;
// properties :126-146, called from gaussian_renderer/__init__.py:53-60) and evaluates them a second time for the two
// regularisers of the MCMC loss (train_static.py:99-100: opacity_reg * |opacity|.mean() + scale_reg * |scaling|.mean()):
// about 12 elementwise / reduction kernels forward and as many backward, each a full pass over P.
// Here: ONE forward kernel (64 B per Gaussian) that also produces the two regulariser sums, and ONE backward kernel that
// folds the regularisers' gradients into the chain rule:
//   d raw_opacity  = (g_o + c_o sign(o)) o (1 - o)          o = 1 / (1 + exp(-raw))
//   d raw_scaling  = (g_s + c_s sign(s)) s                   s = exp(raw)
//   d raw_rotation = (g_q - qh (qh . g_q)) / max(|q|, 1e-12) qh = q / max(|q|, 1e-12)
#include "gs_common.cuh"
#include "kernels.cuh"

namespace gs {

constexpr int kActThreads = 256;

GS_DEV float sigmoidf_ref(float x) { return 1.0f / (1.0f + expf(-x)); }

__global__ void __launch_bounds__(kActThreads) activate_fwd_kernel(int P, const float* __restrict__ raw_opacity,
                                                                   const float* __restrict__ raw_scaling,
                                                                   const float* __restrict__ raw_rotation,
                                                                   float* __restrict__ opacity, float* __restrict__ scaling,
                                                                   float* __restrict__ rotation, float* __restrict__ partial) {
    __shared__ float s_red[2][kActThreads / 32];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float abs_o = 0.f, abs_s = 0.f;
    if (i < P) {
        const float o = sigmoidf_ref(raw_opacity[i]);
        opacity[i] = o;
        abs_o = fabsf(o);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float s = expf(raw_scaling[3 * (size_t)i + k]);
            scaling[3 * (size_t)i + k] = s;
            abs_s += fabsf(s);
        }
        const float4 q = reinterpret_cast<const float4*>(raw_rotation)[i];
        const float n = fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);
        reinterpret_cast<float4*>(rotation)[i] = make_float4(q.x / n, q.y / n, q.z / n, q.w / n);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        abs_o += __shfl_xor_sync(0xffffffffu, abs_o, o);
        abs_s += __shfl_xor_sync(0xffffffffu, abs_s, o);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { s_red[0][warp] = abs_o; s_red[1][warp] = abs_s; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float a = 0.f, b = 0.f;
        for (int w = 0; w < kActThreads / 32; ++w) { a += s_red[0][w]; b += s_red[1][w]; }
        partial[2 * (size_t)blockIdx.x] = a;
        partial[2 * (size_t)blockIdx.x + 1] = b;
    }
}

__global__ void __launch_bounds__(kActThreads) activate_bwd_kernel(int P, const float* __restrict__ raw_opacity,
                                                                   const float* __restrict__ raw_scaling,
                                                                   const float* __restrict__ raw_rotation,
                                                                   const float* __restrict__ g_opacity,   // may be null
                                                                   const float* __restrict__ g_scaling,   // may be null
                                                                   const float* __restrict__ g_rotation,  // may be null
                                                                   float c_opacity, float c_scaling,
                                                                   const float* __restrict__ reg_grad,     // device scalar or null
                                                                   float* __restrict__ d_raw_opacity, float* __restrict__ d_raw_scaling,
                                                                   float* __restrict__ d_raw_rotation) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const float up = reg_grad ? __ldg(reg_grad) : 0.f;
    const float co = c_opacity * up, cs = c_scaling * up;
    const float o = sigmoidf_ref(raw_opacity[i]);
    const float go = (g_opacity ? g_opacity[i] : 0.f) + (o > 0.f ? co : 0.f);
    d_raw_opacity[i] = go * o * (1.f - o);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const size_t e = 3 * (size_t)i + k;
        const float s = expf(raw_scaling[e]);
        const float gs = (g_scaling ? g_scaling[e] : 0.f) + (s > 0.f ? cs : 0.f);
        d_raw_scaling[e] = gs * s;
    }
    const float4 q = reinterpret_cast<const float4*>(raw_rotation)[i];
    float4 d = make_float4(0.f, 0.f, 0.f, 0.f);
    if (g_rotation) {
        const float4 g = reinterpret_cast<const float4*>(g_rotation)[i];
        const float norm = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
        const float n = fmaxf(norm, 1e-12f);
        const float inv = 1.f / n;
        if (norm > 1e-12f) {
            const float4 h = make_float4(q.x * inv, q.y * inv, q.z * inv, q.w * inv);
            const float dot = h.x * g.x + h.y * g.y + h.z * g.z + h.w * g.w;
            d = make_float4((g.x - h.x * dot) * inv, (g.y - h.y * dot) * inv, (g.z - h.z * dot) * inv, (g.w - h.w * dot) * inv);
        } else {
            d = make_float4(g.x * inv, g.y * inv, g.z * inv, g.w * inv);   // clamped norm: a plain scale
        }
    }
    reinterpret_cast<float4*>(d_raw_rotation)[i] = d;
}

int launch_activate_fwd(int P, const float* raw_o, const float* raw_s, const float* raw_q, float* o, float* s, float* q,
                        float* partial, float* out_means, cudaStream_t stream) {
    if (P <= 0) return GS_OK;
    const int blocks = (P + kActThreads - 1) / kActThreads;
    activate_fwd_kernel<<<blocks, kActThreads, 0, stream>>>(P, raw_o, raw_s, raw_q, o, s, q, partial);
    GS_LAUNCH_CHECK();
    return launch_pair_reduce(partial, blocks, 1.0 / (double)P, 1.0 / (3.0 * (double)P), out_means, stream);
}

int launch_activate_bwd(int P, const float* raw_o, const float* raw_s, const float* raw_q, const float* g_o, const float* g_s,
                        const float* g_q, float c_o, float c_s, const float* reg_grad, float* d_o, float* d_s, float* d_q,
                        cudaStream_t stream) {
    if (P <= 0) return GS_OK;
    const int blocks = (P + kActThreads - 1) / kActThreads;
    activate_bwd_kernel<<<blocks, kActThreads, 0, stream>>>(P, raw_o, raw_s, raw_q, g_o, g_s, g_q, c_o, c_s, reg_grad, d_o, d_s, d_q);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

size_t activate_partial_count(int P) { return (size_t)((P + kActThreads - 1) / kActThreads); }

}  // namespace gs
