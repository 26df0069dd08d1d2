// MCMC update kernels for sm_100a: compute_relocation (Eq. 9 of "3D Gaussian Splatting as Markov
// Chain Monte Carlo") and the fused add-noise perturbation.
//
// compute_relocation follows the reference kernel's arithmetic exactly
// (cuda_rasterizer/utils.cu:6-36): o' = 1 - powf(1 - o, 1/N) in FP32; each term
// (+-1/sqrt(k+1)) * o'^(k+1) evaluated in FP64 with the accurate pow and rounded to FP32; the sum
// accumulated in FP32 in the reference's (i outer, k inner) order. The reference recomputes the term
// inside the double loop (N(N+1)/2 double pows); term k does not depend on i, so this kernel computes
// each of the N terms once, parks them in shared memory and replays the same accumulation order —
// identical bits, O(N) pows.
//
// add_noise fuses the ~15 torch launches of train_static.py:132-140 (+ build_rotation /
// build_scaling_rotation, utils/general_utils.py:78-110) into one pass:
//   xyz += (R diag(s)) (R diag(s))^T (n * sigmoid(100 ((1 - sigmoid(o_raw)) - 0.995)) * noise_lr * xyz_lr)
// with s = exp(raw scale), R from the normalised quaternion. Normal deviates either come from the
// caller (torch.randn_like, for parity with the reference's RNG stream) or from an in-kernel
// counter-based generator.
#include "gs_common.cuh"
#include "kernels.cuh"

namespace gs {

constexpr int kRelocThreads = 128;
constexpr int kRelocMaxN = 64;

__global__ void __launch_bounds__(kRelocThreads) compute_relocation_kernel(int P, const float* __restrict__ opacity_old,
                                                                          const float* __restrict__ scale_old,
                                                                          const int* __restrict__ N,
                                                                          const float* __restrict__ binoms, int n_max,
                                                                          float* __restrict__ opacity_new,
                                                                          float* __restrict__ scale_new) {
    __shared__ float s_term[kRelocMaxN][kRelocThreads];  // term k of thread t at [k][t]: conflict-free
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const int t = threadIdx.x;
    const int N_idx = N[idx];
    const float o_old = opacity_old[idx];
    const float o_new = 1.0f - powf(1.0f - o_old, 1.0f / N_idx);
    opacity_new[idx] = o_new;

    float denom_sum = 0.0f;
    if (N_idx <= kRelocMaxN) {
        for (int i = 1; i <= N_idx; ++i) {
            const int kn = i - 1;  // the new term of this row
            // pow(-1, kn) is exactly +-1 for an integer kn: the sign is taken directly instead of through the general FP64 pow
            s_term[kn][t] = (float)((((kn & 1) ? -1.0 : 1.0) / sqrt((double)(kn + 1))) * pow((double)o_new, (double)(kn + 1)));
            const float* __restrict__ row = binoms + (size_t)(i - 1) * n_max;
            for (int k = 0; k <= kn; ++k) denom_sum += (row[k] * s_term[k][t]);
        }
    } else {
        // Out-of-contract N (the caller clamps to [1, 50]); direct evaluation.
        for (int i = 1; i <= N_idx; ++i)
            for (int k = 0; k <= (i - 1); ++k) {
                const float term = (float)((pow(-1.0, (double)k) / sqrt((double)(k + 1))) * pow((double)o_new, (double)(k + 1)));
                denom_sum += (binoms[(size_t)(i - 1) * n_max + k] * term);
            }
    }
    const float coeff = (o_old / denom_sum);
#pragma unroll
    for (int i = 0; i < 3; ++i) scale_new[idx * 3 + i] = coeff * scale_old[idx * 3 + i];
}

int launch_compute_relocation(int P, const float* opacity_old, const float* scale_old, const int* N, const float* binoms,
                              int n_max, float* opacity_new, float* scale_new, cudaStream_t stream) {
    if (P <= 0) return GS_OK;
    compute_relocation_kernel<<<(P + kRelocThreads - 1) / kRelocThreads, kRelocThreads, 0, stream>>>(
        P, opacity_old, scale_old, N, binoms, n_max, opacity_new, scale_new);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

// ---- counter-based normal generator (Philox-4x32-10 + Box-Muller), used when no noise tensor is given ----
GS_DEV uint4 philox4x32_10(uint4 ctr, uint2 key) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += 0x9E3779B9u;
        key.y += 0xBB67AE85u;
    }
    return ctr;
}
GS_DEV float2 box_muller(uint32_t a, uint32_t b) {
    const float u1 = ((float)a + 1.0f) * 2.3283064365386963e-10f;  // (0, 1]
    const float u2 = (float)b * 2.3283064365386963e-10f;
    const float rad = sqrtf(-2.0f * logf(u1));
    float s, c;
    sincosf(6.283185307179586f * u2, &s, &c);
    return make_float2(rad * c, rad * s);
}

__global__ void __launch_bounds__(256) add_noise_kernel(int P, float* __restrict__ xyz, const float* __restrict__ scaling_raw,
                                                        const float* __restrict__ rotation_raw,
                                                        const float* __restrict__ opacity_raw,
                                                        const float* __restrict__ noise /* [P,3] or null */, float lr_scale,
                                                        unsigned long long seed, unsigned long long offset, bool rot_aligned) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const size_t i = (size_t)idx;
    float3 n;
    if (noise) {
        n = make_float3(noise[3 * i], noise[3 * i + 1], noise[3 * i + 2]);
    } else {
        const uint4 r = philox4x32_10(make_uint4((uint32_t)idx, (uint32_t)offset, (uint32_t)(offset >> 32), 0u),
                                      make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
        const float2 a = box_muller(r.x, r.y), b = box_muller(r.z, r.w);
        n = make_float3(a.x, a.y, b.x);
    }
    // gate = sigmoid(100 * ((1 - sigmoid(o)) - 0.995)), written the way the reference composes it.
    const float opacity = 1.0f / (1.0f + expf(-opacity_raw[idx]));
    const float gate = 1.0f / (1.0f + expf(-100.0f * ((1.0f - opacity) - 0.995f)));
    const float gs_ = gate * lr_scale;
    n.x *= gs_; n.y *= gs_; n.z *= gs_;

    const float4 qr = load_quat(rotation_raw, i, rot_aligned);
    const float norm = sqrtf(qr.x * qr.x + qr.y * qr.y + qr.z * qr.z + qr.w * qr.w);
    const float r = qr.x / norm, x = qr.y / norm, y = qr.z / norm, z = qr.w / norm;
    const float s0 = expf(scaling_raw[3 * i]), s1 = expf(scaling_raw[3 * i + 1]), s2 = expf(scaling_raw[3 * i + 2]);
    // L = R * diag(s), row-major R as in build_rotation.
    const float R00 = 1.f - 2.f * (y * y + z * z), R01 = 2.f * (x * y - r * z), R02 = 2.f * (x * z + r * y);
    const float R10 = 2.f * (x * y + r * z), R11 = 1.f - 2.f * (x * x + z * z), R12 = 2.f * (y * z - r * x);
    const float R20 = 2.f * (x * z - r * y), R21 = 2.f * (y * z + r * x), R22 = 1.f - 2.f * (x * x + y * y);
    const float L00 = R00 * s0, L01 = R01 * s1, L02 = R02 * s2;
    const float L10 = R10 * s0, L11 = R11 * s1, L12 = R12 * s2;
    const float L20 = R20 * s0, L21 = R21 * s1, L22 = R22 * s2;
    // Sigma = L L^T, then Sigma * n.
    const float S00 = L00 * L00 + L01 * L01 + L02 * L02, S01 = L00 * L10 + L01 * L11 + L02 * L12;
    const float S02 = L00 * L20 + L01 * L21 + L02 * L22, S11 = L10 * L10 + L11 * L11 + L12 * L12;
    const float S12 = L10 * L20 + L11 * L21 + L12 * L22, S22 = L20 * L20 + L21 * L21 + L22 * L22;
    xyz[3 * i] += S00 * n.x + S01 * n.y + S02 * n.z;
    xyz[3 * i + 1] += S01 * n.x + S11 * n.y + S12 * n.z;
    xyz[3 * i + 2] += S02 * n.x + S12 * n.y + S22 * n.z;
}

int launch_add_noise(int P, float* xyz, const float* scaling_raw, const float* rotation_raw, const float* opacity_raw,
                     const float* noise, float lr_scale, unsigned long long seed, unsigned long long offset,
                     cudaStream_t stream) {
    if (P <= 0) return GS_OK;
    add_noise_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, xyz, scaling_raw, rotation_raw, opacity_raw, noise, lr_scale, seed, offset,
                                                   is_aligned16(rotation_raw));
    GS_LAUNCH_CHECK();
    return GS_OK;
}

}  // namespace gs
