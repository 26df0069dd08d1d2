// View-sharded data parallelism: the kernels of the inter-GPU gradient exchange.
//
// The reference is single-GPU; the north star adds one exchange step per training step: every rank renders its own
// camera view against replicated Gaussians and the per-Gaussian parameter gradients are summed over the ranks.
// A plain all-reduce moves 59 floats per Gaussian (708 MB at 3M), 48 of them the SH gradient. That slab is rank one
// per view: dL/dsh[k][c] = basis_k(dir) * dL/dRGB[c] (cuda_rasterizer/backward.cu:20-139), and dir depends only on
// the Gaussian's position (replicated) and the camera position (3 floats per view). So the ranks exchange the
// clamp-masked dL/dRGB (3 floats per Gaussian per view, all-gather) and every rank rebuilds the summed SH gradient
// locally; only the remaining 11 floats per Gaussian (means 3, opacity 1, scales 3, rotations 4) are all-reduced.
//
//   color_grad_push_kernel   masks dL/dRGB with the forward's clamp bits and stores it into slot `rank` of the
//                            [V][P][3] gather buffer of EVERY rank, through peer pointers or one NVSwitch
//                            multicast store (multimem.st): producer fused with its all-gather, no waits inside.
//   sh_grad_combine_kernel   dL/dsh[i] = sum_v basis(dir(i, cam_v)) (x) g_v[i], views added in rank order with
//                            separately rounded products, i.e. bit-identical to summing the per-view dL/dsh
//                            tensors in that order, and bit-identical on every rank.
//   slab_all_reduce_kernel   two-shot all-reduce of a symmetric-memory slab over NVLink: each rank reduces its
//                            1/N slice (multimem.ld_reduce in the switch, or peer loads added in rank order) and
//                            broadcasts it (multimem.st / peer stores). Cross-rank ordering comes from barriers
//                            the host enqueues before and after (no spinning inside these kernels).
#include <cstdlib>

#include "gs_common.cuh"
#include "kernels.cuh"
#include "sh_stage.cuh"

namespace gs {

GS_DEV void st_multicast_v4(float* addr, float4 v) {
    asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}
GS_DEV void st_multicast_f32(float* addr, float v) {
    asm volatile("multimem.st.weak.global.f32 [%0], %1;" ::"l"(addr), "f"(v) : "memory");
}
GS_DEV float4 ld_reduce_multicast_v4(const float* addr) {
    float4 v;
    asm volatile("multimem.ld_reduce.weak.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(addr)
                 : "memory");
    return v;
}

// ---------------------------------------------------------------------------------------------------------------
// masked colour gradient -> every rank's gather buffer
// ---------------------------------------------------------------------------------------------------------------
constexpr int kPushThreads = 256;

// Thread t produces floats [4t, 4t+4) of the flat [3P] colour-gradient array (they belong to one or two Gaussians), so
// every warp store is 512 contiguous bytes: NVLink / multicast stores leave as full packets.
GS_DEV float masked_channel(int f, int P, const int* __restrict__ radii, const uint8_t* __restrict__ clamped,
                            const float4* __restrict__ acc) {
    const int i = f / 3, ch = f - 3 * i;
    if (i >= P || radii[i] <= 0) return 0.f;   // culled Gaussians have no gradient
    // clamped channels pass no gradient (backward.cu:31-34)
    if ((clamped[i] >> ch) & 1) return 0.f;
    const float4 a = acc[i];
    return ch == 0 ? a.x : (ch == 1 ? a.y : a.z);
}

template <bool kMulticast>
__global__ void __launch_bounds__(kPushThreads) color_grad_push_kernel(int P, const int* __restrict__ radii,
                                                                       const uint8_t* __restrict__ clamped,
                                                                       const float4* __restrict__ acc_color_opacity,
                                                                       const __grid_constant__ DpPeers dst, size_t slot_offset) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int first = 4 * t;
    const int total = 3 * P;
    if (first >= total) return;
    float4 v;
    v.x = masked_channel(first, P, radii, clamped, acc_color_opacity);
    v.y = masked_channel(first + 1, P, radii, clamped, acc_color_opacity);
    v.z = masked_channel(first + 2, P, radii, clamped, acc_color_opacity);
    v.w = masked_channel(first + 3, P, radii, clamped, acc_color_opacity);
    const bool full = first + 3 < total;
#pragma unroll 1
    for (int d = 0; d < dst.n; ++d) {
        float* base = dst.ptr[d] + slot_offset + first;
        if (full) {
            if (kMulticast) st_multicast_v4(base, v);
            else *reinterpret_cast<float4*>(base) = v;
        } else {
            const float out[4] = {v.x, v.y, v.z, v.w};
            for (int k = 0; k < total - first; ++k) {
                if (kMulticast) st_multicast_f32(base + k, out[k]);
                else base[k] = out[k];
            }
        }
    }
}

int launch_color_grad_push(int P, const int* radii, const uint8_t* clamped, const float4* acc, const DpPeers& dst,
                           size_t slot_offset, bool multicast, cudaStream_t stream) {
    if (P <= 0) return GS_OK;
    const int threads_needed = (3 * P + 3) / 4;
    const int blocks = (threads_needed + kPushThreads - 1) / kPushThreads;
    if (multicast) color_grad_push_kernel<true><<<blocks, kPushThreads, 0, stream>>>(P, radii, clamped, acc, dst, slot_offset);
    else color_grad_push_kernel<false><<<blocks, kPushThreads, 0, stream>>>(P, radii, clamped, acc, dst, slot_offset);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// summed SH gradient from the gathered colour gradients
// ---------------------------------------------------------------------------------------------------------------
// Same expressions as sh_backward's second phase (preprocess_bwd.cu), so one view reproduces the per-view gradient
// bit for bit.
GS_DEV void sh_basis(int deg, float x, float y, float z, float* b) {
    const float xx = x * x, yy = y * y, zz = z * z;
    const float xy = x * y, yz = y * z, xz = x * z;
    b[0] = kSH0;
    if (deg > 0) {
        b[1] = -kSH1 * y;
        b[2] = kSH1 * z;
        b[3] = -kSH1 * x;
        if (deg > 1) {
            b[4] = kSH2[0] * xy;
            b[5] = kSH2[1] * yz;
            b[6] = kSH2[2] * (2.f * zz - xx - yy);
            b[7] = kSH2[3] * xz;
            b[8] = kSH2[4] * (xx - yy);
            if (deg > 2) {
                b[9] = kSH3[0] * y * (3.f * xx - yy);
                b[10] = kSH3[1] * xy * z;
                b[11] = kSH3[2] * y * (4.f * zz - xx - yy);
                b[12] = kSH3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy);
                b[13] = kSH3[4] * x * (4.f * zz - xx - yy);
                b[14] = kSH3[5] * z * (xx - yy);
                b[15] = kSH3[6] * x * (xx - 3.f * yy);
            }
        }
    }
}

constexpr int kCombineThreads = 128;

template <int kDeg>
__global__ void __launch_bounds__(kCombineThreads) sh_grad_combine_kernel(int P, int M, int V,
                                                                          const float* __restrict__ means3D,
                                                                          const float* __restrict__ campos,   // [V][3]
                                                                          const float* __restrict__ gathered, // [V][slot_stride]
                                                                          size_t slot_stride, float* __restrict__ dL_dsh) {
    constexpr int kCoef = (kDeg + 1) * (kDeg + 1);
    constexpr int kRowPitch4 = 13;                                            // float4 per staged row (12 + 1 pad)
    __shared__ __align__(16) float s_rows[kCombineThreads / 32][32 * 4 * kRowPitch4];
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool valid = idx < P;
    const size_t i = (size_t)(valid ? idx : 0);
    const int Mrow = 3 * M;
    const int row_stride = Mrow | 1;
    const int warp_first = idx - lane;
    const int rows_valid = min(32, P - warp_first);
    float acc[3 * kCoef];
#pragma unroll
    for (int k = 0; k < 3 * kCoef; ++k) acc[k] = 0.f;
    bool any = false;
    if (valid) {
        const V3 pos(means3D[3 * i], means3D[3 * i + 1], means3D[3 * i + 2]);
        // views in groups of kViewBatch: all colour triples of a group are requested before the first is used
        constexpr int kViewBatch = 8;
#pragma unroll 1
        for (int v0 = 0; v0 < V; v0 += kViewBatch) {
            float gq[kViewBatch][3];
#pragma unroll
            for (int u = 0; u < kViewBatch; ++u) {
                const int v = v0 + u;
                if (v < V) {
                    const float* g = gathered + (size_t)v * slot_stride + 3 * i;
                    gq[u][0] = __ldcs(g); gq[u][1] = __ldcs(g + 1); gq[u][2] = __ldcs(g + 2);
                } else {
                    gq[u][0] = gq[u][1] = gq[u][2] = 0.f;
                }
            }
#pragma unroll
            for (int u = 0; u < kViewBatch; ++u) {
                const int v = v0 + u;
                const float gr = gq[u][0], gg = gq[u][1], gb = gq[u][2];
                if (v >= V || (gr == 0.f && gg == 0.f && gb == 0.f)) continue;   // culled in this view: adds nothing
                any = true;
                const V3 cam(campos[3 * v], campos[3 * v + 1], campos[3 * v + 2]);
                const V3 dir_orig = pos - cam;
                const V3 dir = dir_orig / sqrt(norm2_fixed(dir_orig));
                float b[16];
                sh_basis(kDeg, dir.x, dir.y, dir.z, b);
#pragma unroll
                for (int k = 0; k < kCoef; ++k) {
                    acc[3 * k + 0] = __fadd_rn(acc[3 * k + 0], __fmul_rn(b[k], gr));
                    acc[3 * k + 1] = __fadd_rn(acc[3 * k + 1], __fmul_rn(b[k], gg));
                    acc[3 * k + 2] = __fadd_rn(acc[3 * k + 2], __fmul_rn(b[k], gb));
                }
            }
        }
    }
    // ---- rows out ----
    if (kCoef == 16 && Mrow == 48 && (reinterpret_cast<size_t>(dL_dsh) & 15) == 0) {
        // degree 3, M = 16 (the training configuration): 16-byte traffic on both sides of the tile. Each thread parks its
        // 192-byte row in shared memory with twelve STS.128 (row pitch 208 B: conflict free), then the warp streams its
        // 6 KB of rows out as twelve fully coalesced 512-byte stores. (Per-thread 16-byte global stores cost one LSU
        // pass per row and instruction, 384 per warp; the scalar staging ~400 instructions per warp.)
        float4* tile4 = reinterpret_cast<float4*>(&s_rows[warp][0]);
        if (valid) {
#pragma unroll
            for (int q = 0; q < 12; ++q) {
                const int b0 = 4 * q;
                tile4[lane * kRowPitch4 + q] = make_float4(acc[b0 < 3 * kCoef ? b0 : 0], acc[b0 + 1 < 3 * kCoef ? b0 + 1 : 0],
                                                           acc[b0 + 2 < 3 * kCoef ? b0 + 2 : 0], acc[b0 + 3 < 3 * kCoef ? b0 + 3 : 0]);
            }
        }
        __syncwarp();
        if (rows_valid > 0) {
            float4* dst = reinterpret_cast<float4*>(dL_dsh + (size_t)warp_first * 48);
            const int chunks = rows_valid * 12;
#pragma unroll
            for (int it = 0; it < 12; ++it) {
                const int c = it * 32 + lane;
                if (c < chunks) {
                    const int r = c / 12;
                    __stcs(dst + c, tile4[r * kRowPitch4 + (c - 12 * r)]);
                }
            }
        }
        return;
    }
    const uint32_t row_mask = __ballot_sync(0xffffffffu, any);
    float* tile = &s_rows[warp][0];
    if (any) {
        float* row = tile + lane * row_stride;
#pragma unroll
        for (int k = 0; k < 3 * kCoef; ++k) row[k] = acc[k];
    }
    __syncwarp();
    if (rows_valid > 0) {
        float* dst = dL_dsh + (size_t)warp_first * Mrow;
        if (Mrow == 48) stage_sh_rows<false, 48>(tile, row_stride, dst, Mrow, rows_valid, row_mask, 3 * kCoef, lane);
        else stage_sh_rows<false, 0>(tile, row_stride, dst, Mrow, rows_valid, row_mask, 3 * kCoef, lane);
    }
}

int launch_sh_grad_combine(int P, int D, int M, int V, const float* means3D, const float* campos, const float* gathered,
                           size_t slot_stride, float* dL_dsh, cudaStream_t stream) {
    if (P <= 0) return GS_OK;
    if (3 * M > kShMaxRow || M < (D + 1) * (D + 1)) return GS_ERR_ARG;
    const int blocks = (P + kCombineThreads - 1) / kCombineThreads;
    switch (D) {
        case 0: sh_grad_combine_kernel<0><<<blocks, kCombineThreads, 0, stream>>>(P, M, V, means3D, campos, gathered, slot_stride, dL_dsh); break;
        case 1: sh_grad_combine_kernel<1><<<blocks, kCombineThreads, 0, stream>>>(P, M, V, means3D, campos, gathered, slot_stride, dL_dsh); break;
        case 2: sh_grad_combine_kernel<2><<<blocks, kCombineThreads, 0, stream>>>(P, M, V, means3D, campos, gathered, slot_stride, dL_dsh); break;
        case 3: sh_grad_combine_kernel<3><<<blocks, kCombineThreads, 0, stream>>>(P, M, V, means3D, campos, gathered, slot_stride, dL_dsh); break;
        default: return GS_ERR_ARG;
    }
    GS_LAUNCH_CHECK();
    return GS_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// two-shot all-reduce of a symmetric slab over NVLink (barriers before and after are the caller's)
// ---------------------------------------------------------------------------------------------------------------
constexpr int kArThreads = 512;

constexpr int kArUnroll = 4;

template <bool kMulticast>
__global__ void __launch_bounds__(kArThreads) slab_all_reduce_kernel(const __grid_constant__ DpPeers peers, size_t first_vec, size_t n_vec) {
    // this rank owns float4 elements [first_vec, first_vec + n_vec) of the slab; kArUnroll independent 16-byte
    // reductions per thread are in flight before the first store (NVLink round trips are microseconds)
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t e0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e0 < n_vec; e0 += stride * kArUnroll) {
        float4 s[kArUnroll];
#pragma unroll
        for (int u = 0; u < kArUnroll; ++u) {
            const size_t e = e0 + (size_t)u * stride;
            if (e >= n_vec) break;
            const size_t off = 4 * (first_vec + e);
            if (kMulticast) {
                s[u] = ld_reduce_multicast_v4(peers.ptr[0] + off);
            } else {
                s[u] = __ldcs(reinterpret_cast<const float4*>(peers.ptr[0] + off));
            }
        }
        if (!kMulticast) {
#pragma unroll 1
            for (int r = 1; r < peers.n; ++r) {   // rank order: the sum is the same bits on every rank
                float4 v[kArUnroll];
#pragma unroll
                for (int u = 0; u < kArUnroll; ++u) {
                    const size_t e = e0 + (size_t)u * stride;
                    if (e >= n_vec) break;
                    v[u] = __ldcs(reinterpret_cast<const float4*>(peers.ptr[r] + 4 * (first_vec + e)));
                }
#pragma unroll
                for (int u = 0; u < kArUnroll; ++u) {
                    const size_t e = e0 + (size_t)u * stride;
                    if (e >= n_vec) break;
                    s[u].x += v[u].x; s[u].y += v[u].y; s[u].z += v[u].z; s[u].w += v[u].w;
                }
            }
        }
#pragma unroll
        for (int u = 0; u < kArUnroll; ++u) {
            const size_t e = e0 + (size_t)u * stride;
            if (e >= n_vec) break;
            const size_t off = 4 * (first_vec + e);
            if (kMulticast) {
                st_multicast_v4(peers.ptr[0] + off, s[u]);
            } else {
                for (int r = 0; r < peers.n; ++r) *reinterpret_cast<float4*>(peers.ptr[r] + off) = s[u];
            }
        }
    }
}

int launch_slab_all_reduce(const DpPeers& peers, bool multicast, size_t n_floats, int rank, int world, int sm_count,
                           cudaStream_t stream) {
    if (n_floats == 0 || world <= 1) return GS_OK;
    if (n_floats % 4 != 0 || rank < 0 || rank >= world) return GS_ERR_ARG;
    const size_t n_vec_total = n_floats / 4;
    const size_t per = (n_vec_total + world - 1) / world;
    const size_t first = per * rank;
    if (first >= n_vec_total) return GS_OK;
    const size_t n_vec = (first + per <= n_vec_total) ? per : n_vec_total - first;
    size_t blocks = (n_vec + (size_t)kArThreads * kArUnroll - 1) / ((size_t)kArThreads * kArUnroll);
    // NVLink bound: a fraction of the SMs keeps enough 16-byte requests in flight, and the rest stay free for the SH
    // rebuild that runs beside this kernel. GS_DP_AR_BLOCKS overrides the block count (tuning).
    static const int env_blocks = [] { const char* e = getenv("GS_DP_AR_BLOCKS"); return e ? atoi(e) : 0; }();
    const size_t cap = env_blocks > 0 ? (size_t)env_blocks : (size_t)(sm_count > 0 ? sm_count : 148) / 2;
    if (blocks > cap) blocks = cap;
    if (multicast) slab_all_reduce_kernel<true><<<(unsigned)blocks, kArThreads, 0, stream>>>(peers, first, n_vec);
    else slab_all_reduce_kernel<false><<<(unsigned)blocks, kArThreads, 0, stream>>>(peers, first, n_vec);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

}  // namespace gs
