// Fused photometric loss of the training step:
//     loss = (1 - lambda) * mean |x - y|  +  lambda * (1 - mean ssim_map(x, y))
// as train_static.py:94-97 composes it from utils/loss_utils.py:21-22 (l1_loss) and :40-67 (ssim / _ssim: 11x11
// Gaussian window, sigma 1.5, zero padding 5, per-channel (groups = C), C1 = 0.01^2, C2 = 0.03^2, mean over all
// elements). The reference runs five conv2d calls plus ~20 elementwise kernels forward and their autograd mirror
// backward; here the forward is one tiled kernel (separable window, both images staged once in shared memory) that
// also emits the three derivative maps, and the backward is one tiled kernel that convolves those maps and adds the
// L1 term. x is the render (gets the gradient), y the ground truth (none).
//
// With mu = w*x, nu = w*y, e11 = w*x^2, e22 = w*y^2, e12 = w*xy (w* = windowed sum),
//   A1 = 2 mu nu + C1, A2 = 2 (e12 - mu nu) + C2, B1 = mu^2 + nu^2 + C1, B2 = (e11 - mu^2) + (e22 - nu^2) + C2,
//   S = A1 A2 / (B1 B2),
//   dS/de11 = -S / B2,  dS/de12 = 2 A1 / (B1 B2),
//   dS/dmu  = 2 nu (A2 - A1) / (B1 B2) - 2 mu S (1/B1 - 1/B2)          (total derivative, e11/e12 held fixed)
// and  d(sum_q S_q)/dx_p = sum_q w(p - q) [ dS/dmu_q + 2 x_p dS/de11_q + y_p dS/de12_q ].
#include <mutex>

#include "gs_common.cuh"
#include "kernels.cuh"

namespace gs {

constexpr int kLossTile = 32;                          // output tile (pixels, square)
constexpr int kLossHalo = 5;
constexpr int kLossSpan = kLossTile + 2 * kLossHalo;   // 42
constexpr int kLossThreads = 256;
constexpr int kLossBlock = 4;                          // outputs per thread along the filtered axis (register blocking:
                                                       // 14 shared-memory loads feed 4 x 11 taps)
constexpr int kLossGroups = kLossTile / kLossBlock;    // 8

GS_DEV float block_sum_256(float v, float* s_red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) s_red[warp] = v;
    __syncthreads();
    float t = 0.f;
    if (threadIdx.x < kLossThreads / 32) t = s_red[threadIdx.x];
    if (warp == 0) {
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    }
    __syncthreads();
    return t;   // valid in thread 0
}

__global__ void __launch_bounds__(kLossThreads) ssim_l1_fwd_kernel(const __grid_constant__ LossParams p) {
    // odd row strides: the row-blocked horizontal pass and the column-per-lane vertical pass are both conflict free
    __shared__ float s_x[kLossSpan][kLossSpan + 1];
    __shared__ float s_y[kLossSpan][kLossSpan + 1];
    __shared__ float s_h[5][kLossSpan][kLossTile + 1];
    __shared__ float s_red[kLossThreads / 32];
    const int x0 = blockIdx.x * kLossTile, y0 = blockIdx.y * kLossTile;
    const size_t plane = (size_t)blockIdx.z * p.H * p.W;
    const float* __restrict__ X = p.img + plane;
    const float* __restrict__ Y = p.gt + plane;
    // ---- stage both images with their halo (zero padding outside the image, F.conv2d padding=5) ----
    for (int i = threadIdx.x; i < kLossSpan * kLossSpan; i += kLossThreads) {
        const int r = i / kLossSpan, c = i - r * kLossSpan;
        const int gy = y0 + r - kLossHalo, gx = x0 + c - kLossHalo;
        float a = 0.f, b = 0.f;
        if (gy >= 0 && gy < p.H && gx >= 0 && gx < p.W) {
            a = X[(size_t)gy * p.W + gx];
            b = Y[(size_t)gy * p.W + gx];
        }
        s_x[r][c] = a;
        s_y[r][c] = b;
    }
    __syncthreads();
    // ---- horizontal pass: 42 rows x 8 groups of 4 columns x 5 maps ----
    for (int i = threadIdx.x; i < kLossSpan * kLossGroups; i += kLossThreads) {
        const int r = i / kLossGroups, c0 = (i - r * kLossGroups) * kLossBlock;
        float a[kLossBlock + 10], b[kLossBlock + 10], aa[kLossBlock + 10], bb[kLossBlock + 10], ab[kLossBlock + 10];
#pragma unroll
        for (int j = 0; j < kLossBlock + 10; ++j) {
            a[j] = s_x[r][c0 + j];
            b[j] = s_y[r][c0 + j];
            aa[j] = a[j] * a[j]; bb[j] = b[j] * b[j]; ab[j] = a[j] * b[j];
        }
#pragma unroll
        for (int o = 0; o < kLossBlock; ++o) {
            float m1 = 0.f, m2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
            for (int k = 0; k < 11; ++k) {
                const float w = p.win[k];
                m1 = fmaf(w, a[o + k], m1);
                m2 = fmaf(w, b[o + k], m2);
                e11 = fmaf(w, aa[o + k], e11);
                e22 = fmaf(w, bb[o + k], e22);
                e12 = fmaf(w, ab[o + k], e12);
            }
            s_h[0][r][c0 + o] = m1; s_h[1][r][c0 + o] = m2; s_h[2][r][c0 + o] = e11; s_h[3][r][c0 + o] = e22;
            s_h[4][r][c0 + o] = e12;
        }
    }
    __syncthreads();
    // ---- vertical pass: lane = column, each thread 4 rows; then SSIM and its partial derivatives per pixel ----
    const int c = threadIdx.x % kLossTile, r0 = (threadIdx.x / kLossTile) * kLossBlock;
    float v[5][kLossBlock];
#pragma unroll
    for (int m = 0; m < 5; ++m) {
        float q[kLossBlock + 10];
#pragma unroll
        for (int j = 0; j < kLossBlock + 10; ++j) q[j] = s_h[m][r0 + j][c];
#pragma unroll
        for (int o = 0; o < kLossBlock; ++o) {
            float t = 0.f;
#pragma unroll
            for (int k = 0; k < 11; ++k) t = fmaf(p.win[k], q[o + k], t);
            v[m][o] = t;
        }
    }
    const int gx = x0 + c;
    float ssim_acc = 0.f, l1_acc = 0.f;
#pragma unroll
    for (int o = 0; o < kLossBlock; ++o) {
        const int gy = y0 + r0 + o;
        if (gx < p.W && gy < p.H) {
            const float mu = v[0][o], nu = v[1][o], e11 = v[2][o], e22 = v[3][o], e12 = v[4][o];
            const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
            const float mu_sq = mu * mu, nu_sq = nu * nu, munu = mu * nu;
            const float s1 = e11 - mu_sq, s2 = e22 - nu_sq, s12 = e12 - munu;
            const float A1 = 2.f * munu + C1, A2 = 2.f * s12 + C2, B1 = mu_sq + nu_sq + C1, B2 = s1 + s2 + C2;
            const float inv_B1 = 1.f / B1, inv_B2 = 1.f / B2;
            const float inv_B12 = inv_B1 * inv_B2;
            const float ssim = (A1 * A2) * inv_B12;
            const size_t idx = plane + (size_t)gy * p.W + gx;
            p.d_mu[idx] = 2.f * nu * (A2 - A1) * inv_B12 - 2.f * mu * ssim * (inv_B1 - inv_B2);
            p.d_e11[idx] = -ssim * inv_B2;
            p.d_e12[idx] = 2.f * A1 * inv_B12;
            ssim_acc += ssim;
            l1_acc += fabsf(s_x[r0 + o + kLossHalo][c + kLossHalo] - s_y[r0 + o + kLossHalo][c + kLossHalo]);
        }
    }
    // ---- per-block partial sums (summed in a fixed order by the reduce kernel: the loss value is deterministic) ----
    const float l1_sum = block_sum_256(l1_acc, s_red);
    const float ssim_sum = block_sum_256(ssim_acc, s_red);
    if (threadIdx.x == 0) {
        const size_t bl = ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        p.partial[2 * bl] = l1_sum;
        p.partial[2 * bl + 1] = ssim_sum;
    }
}

__global__ void __launch_bounds__(1024) loss_reduce_kernel(const float* __restrict__ partial, int nblocks, double inv_a,
                                                           double inv_b, float* __restrict__ out) {
    __shared__ double s_a[32], s_b[32];
    double a = 0.0, b = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += 1024) {
        a += (double)partial[2 * i];
        b += (double)partial[2 * i + 1];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, o);
        b += __shfl_xor_sync(0xffffffffu, b, o);
    }
    if ((threadIdx.x & 31) == 0) { s_a[threadIdx.x >> 5] = a; s_b[threadIdx.x >> 5] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double ta = 0.0, tb = 0.0;
        for (int w = 0; w < 32; ++w) { ta += s_a[w]; tb += s_b[w]; }
        out[0] = (float)(ta * inv_a);   // loss: mean |x - y|
        out[1] = (float)(tb * inv_b);   // loss: mean ssim_map
    }
}

__global__ void __launch_bounds__(kLossThreads) ssim_l1_bwd_kernel(const __grid_constant__ LossParams p, float c_l1, float c_ssim,
                                                                   const float* __restrict__ grad_scale) {
    // dL/dimg = (upstream gradient of the scalar loss) x (c_l1 * sign(x - y) + c_ssim * d(sum S)/dx)
    const float up = grad_scale ? __ldg(grad_scale) : 1.f;
    const float w_l1 = c_l1 * up, w_ssim = c_ssim * up;
    __shared__ float s_d[3][kLossSpan][kLossSpan + 1];
    __shared__ float s_h[3][kLossSpan][kLossTile + 1];
    const int x0 = blockIdx.x * kLossTile, y0 = blockIdx.y * kLossTile;
    const size_t plane = (size_t)blockIdx.z * p.H * p.W;
    for (int i = threadIdx.x; i < kLossSpan * kLossSpan; i += kLossThreads) {
        const int r = i / kLossSpan, c = i - r * kLossSpan;
        const int gy = y0 + r - kLossHalo, gx = x0 + c - kLossHalo;
        float a = 0.f, b = 0.f, d = 0.f;
        if (gy >= 0 && gy < p.H && gx >= 0 && gx < p.W) {   // pixels outside the image have no ssim_map entry
            const size_t o = plane + (size_t)gy * p.W + gx;
            a = p.d_mu[o]; b = p.d_e11[o]; d = p.d_e12[o];
        }
        s_d[0][r][c] = a; s_d[1][r][c] = b; s_d[2][r][c] = d;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kLossSpan * kLossGroups; i += kLossThreads) {
        const int r = i / kLossGroups, c0 = (i - r * kLossGroups) * kLossBlock;
#pragma unroll
        for (int m = 0; m < 3; ++m) {
            float q[kLossBlock + 10];
#pragma unroll
            for (int j = 0; j < kLossBlock + 10; ++j) q[j] = s_d[m][r][c0 + j];
#pragma unroll
            for (int o = 0; o < kLossBlock; ++o) {
                float t = 0.f;
#pragma unroll
                for (int k = 0; k < 11; ++k) t = fmaf(p.win[k], q[o + k], t);
                s_h[m][r][c0 + o] = t;
            }
        }
    }
    __syncthreads();
    const int c = threadIdx.x % kLossTile, r0 = (threadIdx.x / kLossTile) * kLossBlock;
    float v[3][kLossBlock];
#pragma unroll
    for (int m = 0; m < 3; ++m) {
        float q[kLossBlock + 10];
#pragma unroll
        for (int j = 0; j < kLossBlock + 10; ++j) q[j] = s_h[m][r0 + j][c];
#pragma unroll
        for (int o = 0; o < kLossBlock; ++o) {
            float t = 0.f;
#pragma unroll
            for (int k = 0; k < 11; ++k) t = fmaf(p.win[k], q[o + k], t);
            v[m][o] = t;
        }
    }
    const int gx = x0 + c;
#pragma unroll
    for (int o = 0; o < kLossBlock; ++o) {
        const int gy = y0 + r0 + o;
        if (gx < p.W && gy < p.H) {
            const size_t idx = plane + (size_t)gy * p.W + gx;
            const float x = p.img[idx], y = p.gt[idx];
            const float diff = x - y;
            const float sgn = diff > 0.f ? 1.f : (diff < 0.f ? -1.f : 0.f);   // torch: d|t|/dt = sign(t), 0 at 0
            p.dL_dimg[idx] = w_ssim * (v[0][o] + 2.f * x * v[1][o] + y * v[2][o]) + w_l1 * sgn;
        }
    }
}

int launch_loss_forward(const LossParams& p, float* out_means, cudaStream_t stream) {
    const dim3 grid((p.W + kLossTile - 1) / kLossTile, (p.H + kLossTile - 1) / kLossTile, p.C);
    ssim_l1_fwd_kernel<<<grid, kLossThreads, 0, stream>>>(p);
    GS_LAUNCH_CHECK();
    const int nblocks = (int)(grid.x * grid.y * grid.z);
    const double inv_n = 1.0 / ((double)p.C * p.H * p.W);
    return launch_pair_reduce(p.partial, nblocks, inv_n, inv_n, out_means, stream);
}

// out[0] = inv_a * sum_i partial[2i], out[1] = inv_b * sum_i partial[2i+1], summed in a fixed order in double.
int launch_pair_reduce(const float* partial, int n, double inv_a, double inv_b, float* out, cudaStream_t stream) {
    loss_reduce_kernel<<<1, 1024, 0, stream>>>(partial, n, inv_a, inv_b, out);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

int launch_loss_backward(const LossParams& p, float c_l1, float c_ssim, const float* grad_scale, cudaStream_t stream) {
    const dim3 grid((p.W + kLossTile - 1) / kLossTile, (p.H + kLossTile - 1) / kLossTile, p.C);
    ssim_l1_bwd_kernel<<<grid, kLossThreads, 0, stream>>>(p, c_l1, c_ssim, grad_scale);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

size_t loss_partial_count(int C, int H, int W) {
    return (size_t)((W + kLossTile - 1) / kLossTile) * ((H + kLossTile - 1) / kLossTile) * C;
}

// ------------------------------------------------------------------------------------------------------------------
// Plain L1 (utils/loss_utils.py:21-22: torch.abs(network_output - gt).mean()) for callers that use it without the SSIM
// term: one streaming pass + the fixed-order reduction above forward, one streaming pass backward
// (grad * sign(x - y) / n, sign(0) = 0 as torch's abs backward), instead of torch's ~8 elementwise / reduction launches.
// ------------------------------------------------------------------------------------------------------------------
constexpr int kL1Threads = 256;
constexpr int kL1Chunk = kL1Threads * 16;   // floats per block iteration (four float4 per thread)
constexpr int kL1MaxBlocks = 2048;

int l1_block_count(size_t n) {
    const size_t chunks = (n + kL1Chunk - 1) / kL1Chunk;
    return (int)(chunks < (size_t)kL1MaxBlocks ? (chunks ? chunks : 1) : (size_t)kL1MaxBlocks);
}

template <bool kVec>
__global__ void __launch_bounds__(kL1Threads) l1_fwd_kernel(size_t n, const float* __restrict__ x, const float* __restrict__ y,
                                                            float* __restrict__ partial) {
    __shared__ float s_w[kL1Threads / 32];
    float acc = 0.f;
    // block b takes chunks b, b + gridDim.x, ...: the assignment, and so the summation order, is fixed
    for (size_t base = (size_t)blockIdx.x * kL1Chunk; base < n; base += (size_t)gridDim.x * kL1Chunk) {
        if (kVec && base + kL1Chunk <= n) {
            float4 a[4], b[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const size_t e = base / 4 + (size_t)k * kL1Threads + threadIdx.x;
                a[k] = __ldcs(reinterpret_cast<const float4*>(x) + e);
                b[k] = __ldcs(reinterpret_cast<const float4*>(y) + e);
            }
#pragma unroll
            for (int k = 0; k < 4; ++k)
                acc += (fabsf(a[k].x - b[k].x) + fabsf(a[k].y - b[k].y)) + (fabsf(a[k].z - b[k].z) + fabsf(a[k].w - b[k].w));
        } else {
            const size_t end = base + kL1Chunk < n ? base + kL1Chunk : n;
            for (size_t e = base + threadIdx.x; e < end; e += kL1Threads) acc += fabsf(x[e] - y[e]);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < kL1Threads / 32; ++w) t += s_w[w];
        partial[2 * blockIdx.x] = t;
        partial[2 * blockIdx.x + 1] = 0.f;
    }
}

GS_DEV float l1_sign(float d) { return d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f); }

template <bool kVec>
__global__ void __launch_bounds__(kL1Threads) l1_bwd_kernel(size_t n, const float* __restrict__ x, const float* __restrict__ y,
                                                            const float* __restrict__ grad_scale, float inv_n,
                                                            float* __restrict__ dL_dx) {
    const float g = (grad_scale ? __ldg(grad_scale) : 1.f) * inv_n;
    const size_t base = (size_t)blockIdx.x * kL1Chunk;
    if (kVec && base + kL1Chunk <= n) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const size_t e = base / 4 + (size_t)k * kL1Threads + threadIdx.x;
            const float4 a = __ldcs(reinterpret_cast<const float4*>(x) + e);
            const float4 b = __ldcs(reinterpret_cast<const float4*>(y) + e);
            reinterpret_cast<float4*>(dL_dx)[e] =
                make_float4(g * l1_sign(a.x - b.x), g * l1_sign(a.y - b.y), g * l1_sign(a.z - b.z), g * l1_sign(a.w - b.w));
        }
    } else {
        const size_t end = base + kL1Chunk < n ? base + kL1Chunk : n;
        for (size_t e = base + threadIdx.x; e < end; e += kL1Threads) dL_dx[e] = g * l1_sign(x[e] - y[e]);
    }
}

int launch_l1_forward(size_t n, const float* x, const float* y, float* partial, float* out_mean, cudaStream_t stream) {
    const int nb = l1_block_count(n);
    const bool vec = ((reinterpret_cast<size_t>(x) | reinterpret_cast<size_t>(y)) & 15) == 0;
    if (vec) l1_fwd_kernel<true><<<nb, kL1Threads, 0, stream>>>(n, x, y, partial);
    else l1_fwd_kernel<false><<<nb, kL1Threads, 0, stream>>>(n, x, y, partial);
    GS_LAUNCH_CHECK();
    return launch_pair_reduce(partial, nb, 1.0 / (double)n, 0.0, out_mean, stream);
}

int launch_l1_backward(size_t n, const float* x, const float* y, const float* grad_scale, float* dL_dx, cudaStream_t stream) {
    const size_t chunks = (n + kL1Chunk - 1) / kL1Chunk;
    if (chunks > 0x7fffffffull) return GS_ERR_ARG;
    const bool vec = ((reinterpret_cast<size_t>(x) | reinterpret_cast<size_t>(y) | reinterpret_cast<size_t>(dL_dx)) & 15) == 0;
    const float inv_n = (float)(1.0 / (double)n);
    if (vec) l1_bwd_kernel<true><<<(unsigned)chunks, kL1Threads, 0, stream>>>(n, x, y, grad_scale, inv_n, dL_dx);
    else l1_bwd_kernel<false><<<(unsigned)chunks, kL1Threads, 0, stream>>>(n, x, y, grad_scale, inv_n, dL_dx);
    GS_LAUNCH_CHECK();
    return GS_OK;
}


// ------------------------------------------------------------------------------------------------------------------
// gs_l1_loss_forward_static: the L1 forward for callers that have no way to hand this library a scratch buffer.
//
// gs_l1_loss_forward needs a caller-allocated workspace for its per-block partial sums. Some hosts of this library
// cannot produce one: a plugin running inside an engine that owns the device allocator and does not expose it, a
// real-time loop in which cudaMalloc's implicit device synchronisation is forbidden, or a process whose memory
// budget was carved up front and that must not grow. For them the partial sums live in the array below, which is
// part of the library image: each GPU that loads the library carries these 16 KB until the process exits, and no
// call ever allocates or frees anything.
//
// The price is a size limit and shared state. The array holds one pair of floats per block and one block per 4096
// elements, so n is limited to kL1StaticMaxN = 2048 * 4096 = 8 388 608 floats (a 3 x 1080 x 1920 image is 6 220 800);
// larger inputs are refused with GS_ERR_LIMIT. Up to that size the launch shape, the summation order and therefore
// the result are the same as gs_l1_loss_forward's, bit for bit. Every call on a device uses the same array: two calls
// whose kernels can overlap on that GPU (different streams, no dependency) race on it, so keep the calls ordered.
// The array keeps whatever the previous call left; that is harmless because each call writes every slot it later
// reads (slots [0, 2 * blocks)) and nothing is accumulated into it, so there is nothing that has to be zeroed.
// ------------------------------------------------------------------------------------------------------------------
constexpr size_t kL1StaticMaxN = (size_t)kL1MaxBlocks * kL1Chunk;
static __device__ __align__(256) float g_l1_static_partial[2 * kL1MaxBlocks];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_l1_static_mutex;
float* g_l1_static_ptr[kMaxDevices] = {};   // device address of g_l1_static_partial on each device, once resolved

// In host code `g_l1_static_partial` is only a handle for the runtime, not a pointer a kernel could use; the real
// address differs per device and is looked up once per device, with that device current.
int l1_static_scratch(float** out) {
    int dev = -1;
    GS_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return GS_ERR_ARG;
    std::lock_guard<std::mutex> lock(g_l1_static_mutex);
    if (!g_l1_static_ptr[dev]) {
        float* ptr = nullptr;
        GS_CUDA_TRY(cudaGetSymbolAddress((void**)&ptr, g_l1_static_partial));
        if (!ptr) return (int)cudaErrorInvalidSymbol;
        g_l1_static_ptr[dev] = ptr;
    }
    *out = g_l1_static_ptr[dev];
    return GS_OK;
}
}  // namespace

}  // namespace gs

// Declared in include/gs_b200.h. `stream` must belong to the current device. The first call on each device resolves
// the scratch address (cudaGetSymbolAddress, which may load the module); make that first call outside any stream capture.
extern "C" int gs_l1_loss_forward_static(void* stream_, size_t n, const float* x, const float* y, float* out_mean) {
    if (n == 0 || !x || !y || !out_mean) return GS_ERR_ARG;
    if (n > gs::kL1StaticMaxN) return GS_ERR_LIMIT;
    float* partial = nullptr;
    const int rc = gs::l1_static_scratch(&partial);
    if (rc) return rc;
    return gs::launch_l1_forward(n, x, y, partial, out_mean, (cudaStream_t)stream_);
}
