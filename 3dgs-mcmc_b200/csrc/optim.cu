// Multi-tensor Adam step for the Gaussian parameters: ONE launch updates all of the model's
// parameter tensors (xyz, f_dc, f_rest, opacity, scaling, rotation: 59 floats per Gaussian) where the reference's
// torch.optim.Adam (scene/gaussian_model_static.py:191-200: per-group lr, betas (0.9, 0.999), eps 1e-15, no weight
// decay, no amsgrad; stepped at train_static.py:129-130) runs about ten foreach kernels over them.
//
// HBM bound: 16 B read (grad, param, exp_avg, exp_avg_sq) + 12 B written per float = 28 B x 59 x P per step
// (4.96 GB at 3M Gaussians). The arithmetic follows torch's own op sequence so results agree to rounding:
//   exp_avg    = lerp(exp_avg, grad, 1 - beta1)            = fma(1 - beta1, grad - exp_avg, exp_avg)
//   exp_avg_sq = exp_avg_sq * beta2 + (1 - beta2) * grad * grad
//   denom      = sqrt(exp_avg_sq) / sqrt(1 - beta2^step) + eps
//   param      = param - (lr / (1 - beta1^step)) * (exp_avg / denom)
// with the bias corrections computed on the host in double, as torch does for non-capturable steps.
#include "gs_common.cuh"
#include "kernels.cuh"

namespace gs {

constexpr int kAdamThreads = 256;
constexpr int kAdamVecPerThread = 4;                                   // float4 per thread per chunk
constexpr int kAdamChunk = kAdamThreads * kAdamVecPerThread * 4;       // 4096 floats per chunk

GS_DEV void adam_one(float& p, float g, float& m, float& v, float w1, float beta2, float w2,
                     float bc2_sqrt, float eps, float neg_step_size) {
    m = fmaf(w1, g - m, m);
    v = fmaf(w2 * g, g, v * beta2);
    const float denom = sqrtf(v) / bc2_sqrt + eps;
    p = fmaf(neg_step_size, m / denom, p);
}

__global__ void __launch_bounds__(kAdamThreads) adam_step_kernel(const __grid_constant__ AdamLaunch L) {
    // chunk -> tensor (at most 8 tensors: linear search over the prefix table)
    const unsigned long long chunk = blockIdx.x;
    int t = 0;
#pragma unroll
    for (int i = 1; i < 8; ++i)
        if (i < L.n_tensors && chunk >= L.first_chunk[i]) t = i;
    const AdamTensor T = L.t[t];
    const unsigned long long base = (chunk - L.first_chunk[t]) * (unsigned long long)kAdamChunk;
    const unsigned long long n = T.n;
    const unsigned long long left = n - base;
    if (T.vec_ok && left >= (unsigned long long)kAdamChunk) {
        float4 g[kAdamVecPerThread], p[kAdamVecPerThread], m[kAdamVecPerThread], v[kAdamVecPerThread];
#pragma unroll
        for (int k = 0; k < kAdamVecPerThread; ++k) {
            const unsigned long long e = base / 4 + (unsigned long long)k * kAdamThreads + threadIdx.x;
            g[k] = __ldcs(reinterpret_cast<const float4*>(T.grad) + e);
            p[k] = __ldcs(reinterpret_cast<const float4*>(T.param) + e);
            m[k] = __ldcs(reinterpret_cast<const float4*>(T.m) + e);
            v[k] = __ldcs(reinterpret_cast<const float4*>(T.v) + e);
        }
#pragma unroll
        for (int k = 0; k < kAdamVecPerThread; ++k) {
            const unsigned long long e = base / 4 + (unsigned long long)k * kAdamThreads + threadIdx.x;
            adam_one(p[k].x, g[k].x, m[k].x, v[k].x, L.w1, L.beta2, L.w2, T.bc2_sqrt, L.eps, T.neg_step_size);
            adam_one(p[k].y, g[k].y, m[k].y, v[k].y, L.w1, L.beta2, L.w2, T.bc2_sqrt, L.eps, T.neg_step_size);
            adam_one(p[k].z, g[k].z, m[k].z, v[k].z, L.w1, L.beta2, L.w2, T.bc2_sqrt, L.eps, T.neg_step_size);
            adam_one(p[k].w, g[k].w, m[k].w, v[k].w, L.w1, L.beta2, L.w2, T.bc2_sqrt, L.eps, T.neg_step_size);
            __stcs(reinterpret_cast<float4*>(T.param) + e, p[k]);
            __stcs(reinterpret_cast<float4*>(T.m) + e, m[k]);
            __stcs(reinterpret_cast<float4*>(T.v) + e, v[k]);
        }
    } else {
        const unsigned long long end = left < (unsigned long long)kAdamChunk ? n : base + kAdamChunk;
        for (unsigned long long e = base + threadIdx.x; e < end; e += kAdamThreads) {
            float p = T.param[e], m = T.m[e], v = T.v[e];
            adam_one(p, T.grad[e], m, v, L.w1, L.beta2, L.w2, T.bc2_sqrt, L.eps, T.neg_step_size);
            T.param[e] = p; T.m[e] = m; T.v[e] = v;
        }
    }
}

int launch_adam_step(AdamLaunch& L, cudaStream_t stream) {
    unsigned long long chunks = 0;
    for (int i = 0; i < L.n_tensors; ++i) {
        L.first_chunk[i] = chunks;
        chunks += (L.t[i].n + kAdamChunk - 1) / kAdamChunk;
        const size_t a = reinterpret_cast<size_t>(L.t[i].param) | reinterpret_cast<size_t>(L.t[i].grad) |
                         reinterpret_cast<size_t>(L.t[i].m) | reinterpret_cast<size_t>(L.t[i].v);
        L.t[i].vec_ok = (a & 15) == 0;
    }
    if (chunks == 0) return GS_OK;
    if (chunks > 0x7fffffffull) return GS_ERR_ARG;
    adam_step_kernel<<<(unsigned)chunks, kAdamThreads, 0, stream>>>(L);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

}  // namespace gs
