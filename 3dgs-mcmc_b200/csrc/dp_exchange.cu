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
//   color_grad_push_sparse_kernel  the same, but only the non-zero triples travel (packed per 256-Gaussian chunk
//                            behind a bit mask), from a bounded number of CTAs; the default layout.
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
// sparse layout: only the non-zero colour gradients travel
// ---------------------------------------------------------------------------------------------------------------
// A view's colour gradient is exactly zero for every Gaussian that was culled, and for every visible one that no pixel
// blended (bwd_zero clears the accumulator; only blended splats add to it) - at the bench scene 75 % are visible and fewer
// contribute. The dense layout ships those zeros over NVLink and the rebuild reads them back.
//
// Sparse slot: one record of kChunkFloats floats per chunk of 256 consecutive Gaussians,
//   words [0, 16)    8 x (mask, before): bit l of mask w <=> Gaussian 256c + 32w + l has a non-zero masked gradient;
//                    before = number of set bits in the chunk's earlier mask words
//   floats [16, ...) the chunk's non-zero triples, packed in index order
// so the address of every record is fixed (no allocation, no atomics, a deterministic image of the slot) and the bytes
// stored are 64 + 12 x (non-zero Gaussians of the chunk), as one contiguous 16-byte aligned run per chunk. The rebuild
// finds Gaussian i's triple in view v with one 8-byte header load per warp and a popcount.
constexpr int kChunk = 256;
constexpr int kChunkHeader = 16;                       // floats (= u32 words)
constexpr int kChunkFloats = kChunkHeader + 3 * kChunk;

size_t dp_sparse_slot_floats(int P) { return (size_t)kChunkFloats * (size_t)((P + kChunk - 1) / kChunk); }

template <bool kMulticast>
__global__ void __launch_bounds__(kChunk) color_grad_push_sparse_kernel(int P, const int* __restrict__ radii,
                                                                        const uint8_t* __restrict__ clamped,
                                                                        const float4* __restrict__ acc_color_opacity,
                                                                        const __grid_constant__ DpPeers dst, size_t slot_offset) {
    __shared__ __align__(16) uint32_t s_rec[kChunkFloats];
    __shared__ int s_count[kChunk / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nchunks = (P + kChunk - 1) / kChunk;
    float* s_tri = reinterpret_cast<float*>(s_rec) + kChunkHeader;

    auto fetch = [&](int c, float& r, float& g, float& b) {
        const int i = c * kChunk + tid;
        r = g = b = 0.f;
        if (c < nchunks && i < P) {
            const int rad = radii[i];
            const uint32_t cl = clamped[i];
            const float4 a = __ldcs(acc_color_opacity + i);
            if (rad > 0) {                                          // culled Gaussians have no gradient
                r = (cl & 1u) ? 0.f : a.x;                          // clamped channels pass none (backward.cu:31-34)
                g = (cl & 2u) ? 0.f : a.y;
                b = (cl & 4u) ? 0.f : a.z;
            }
        }
    };
    float r, g, b;
    fetch(blockIdx.x, r, g, b);
    // a bounded number of CTAs walks the chunks (the HBM-bound per-Gaussian backward runs beside this kernel); the next
    // chunk's loads are in flight while this one is packed and stored
#pragma unroll 1
    for (int c = blockIdx.x; c < nchunks; c += gridDim.x) {
        float nr, ng, nb;
        fetch(c + gridDim.x, nr, ng, nb);
        const bool nz = (r != 0.f) || (g != 0.f) || (b != 0.f);
        const uint32_t m = __ballot_sync(0xffffffffu, nz);
        if (lane == 0) s_count[warp] = __popc(m);
        __syncthreads();
        int before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < kChunk / 32; ++w) {
            const int n = s_count[w];
            before += w < warp ? n : 0;
            total += n;
        }
        if (lane == 0) {
            s_rec[2 * warp] = m;
            s_rec[2 * warp + 1] = (uint32_t)before;
        }
        if (nz) {
            float* t = s_tri + 3 * (before + __popc(m & ((1u << lane) - 1u)));
            t[0] = r; t[1] = g; t[2] = b;
        }
        __syncthreads();
        const int n4 = kChunkHeader / 4 + ((3 * total + 3) >> 2);   // 16-byte pieces of the run: header + packed triples
        for (int q = tid; q < n4; q += kChunk) {
            const float4 v = reinterpret_cast<const float4*>(s_rec)[q];
#pragma unroll 1
            for (int d = 0; d < dst.n; ++d) {
                float* out = dst.ptr[d] + slot_offset + (size_t)c * kChunkFloats + 4 * q;
                if (kMulticast) st_multicast_v4(out, v);
                else *reinterpret_cast<float4*>(out) = v;
            }
        }
        __syncthreads();
        r = nr; g = ng; b = nb;
    }
}

int launch_color_grad_push_sparse(int P, const int* radii, const uint8_t* clamped, const float4* acc, const DpPeers& dst,
                                  size_t slot_offset, bool multicast, int sm_count, cudaStream_t stream) {
    if (P <= 0) return GS_OK;
    const int nchunks = (P + kChunk - 1) / kChunk;
    // two CTAs per SM by default: a quarter of the thread slots, enough loads in flight for the 21 B/Gaussian this kernel
    // reads, and the per-Gaussian backward beside it keeps most of every SM. GS_DP_PUSH_BLOCKS overrides (tuning).
    static const int env_blocks = [] { const char* e = getenv("GS_DP_PUSH_BLOCKS"); return e ? atoi(e) : 0; }();
    int blocks = env_blocks > 0 ? env_blocks : 2 * (sm_count > 0 ? sm_count : 148);
    if (blocks > nchunks) blocks = nchunks;
    if (multicast) color_grad_push_sparse_kernel<true><<<blocks, kChunk, 0, stream>>>(P, radii, clamped, acc, dst, slot_offset);
    else color_grad_push_sparse_kernel<false><<<blocks, kChunk, 0, stream>>>(P, radii, clamped, acc, dst, slot_offset);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// summed SH gradient from the gathered colour gradients
// ---------------------------------------------------------------------------------------------------------------
// Same expressions as sh_backward's second phase (preprocess_bwd.cu), so one view reproduces the per-view gradient
// bit for bit.
GS_DEV void sh_basis(int deg, float x, float y, float z, float* b) { sh_basis_rn(deg, x, y, z, b); }

constexpr int kCombineThreads = 128;

template <int kDeg, bool kSparse>
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
            if (kSparse) {
                // sparse slots: the warp's (mask, before) header word of every view first (one 8-byte broadcast load each),
                // then the triples of the views this Gaussian is present in
                const size_t rec = (size_t)(idx >> 8) * kChunkFloats;
                const int w = (idx >> 5) & 7;
                uint2 hd[kViewBatch];
#pragma unroll
                for (int u = 0; u < kViewBatch; ++u) {
                    const int v = v0 + u;
                    hd[u] = v < V ? __ldg(reinterpret_cast<const uint2*>(gathered + (size_t)v * slot_stride + rec) + w) : make_uint2(0u, 0u);
                }
#pragma unroll
                for (int u = 0; u < kViewBatch; ++u) {
                    gq[u][0] = gq[u][1] = gq[u][2] = 0.f;
                    if ((hd[u].x >> lane) & 1u) {
                        const float* g = gathered + (size_t)(v0 + u) * slot_stride + rec + kChunkHeader +
                                         3 * (hd[u].y + __popc(hd[u].x & ((1u << lane) - 1u)));
                        gq[u][0] = __ldcs(g); gq[u][1] = __ldcs(g + 1); gq[u][2] = __ldcs(g + 2);
                    }
                }
            } else {
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

template <bool kSparse>
static int launch_sh_grad_combine_t(int P, int D, int M, int V, const float* means3D, const float* campos, const float* gathered,
                                    size_t slot_stride, float* dL_dsh, cudaStream_t stream) {
    const int blocks = (P + kCombineThreads - 1) / kCombineThreads;
    switch (D) {
        case 0: sh_grad_combine_kernel<0, kSparse><<<blocks, kCombineThreads, 0, stream>>>(P, M, V, means3D, campos, gathered, slot_stride, dL_dsh); break;
        case 1: sh_grad_combine_kernel<1, kSparse><<<blocks, kCombineThreads, 0, stream>>>(P, M, V, means3D, campos, gathered, slot_stride, dL_dsh); break;
        case 2: sh_grad_combine_kernel<2, kSparse><<<blocks, kCombineThreads, 0, stream>>>(P, M, V, means3D, campos, gathered, slot_stride, dL_dsh); break;
        case 3: sh_grad_combine_kernel<3, kSparse><<<blocks, kCombineThreads, 0, stream>>>(P, M, V, means3D, campos, gathered, slot_stride, dL_dsh); break;
        default: return GS_ERR_ARG;
    }
    GS_LAUNCH_CHECK();
    return GS_OK;
}

int launch_sh_grad_combine(int P, int D, int M, int V, const float* means3D, const float* campos, const float* gathered,
                           size_t slot_stride, bool sparse, float* dL_dsh, cudaStream_t stream) {
    if (P <= 0) return GS_OK;
    if (3 * M > kShMaxRow || M < (D + 1) * (D + 1)) return GS_ERR_ARG;
    return sparse ? launch_sh_grad_combine_t<true>(P, D, M, V, means3D, campos, gathered, slot_stride, dL_dsh, stream)
                  : launch_sh_grad_combine_t<false>(P, D, M, V, means3D, campos, gathered, slot_stride, dL_dsh, stream);
}

// ---------------------------------------------------------------------------------------------------------------
// "records" layout: the whole per-Gaussian gradient travels sparsely, no all-reduce
// ---------------------------------------------------------------------------------------------------------------
// A view's gradient w.r.t. a Gaussian's mean / opacity / scale / rotation is exactly zero wherever its colour gradient is
// (culled, or never blended): at the bench scene 7.5 % of the Gaussians per view carry a gradient at all. So the
// per-Gaussian backward (preprocess_bwd.cu, push epilogue) stores, per chunk of 128 Gaussians, a bit mask and the 14-float
// records of the non-zero ones into slot `rank` of every rank's gather buffer, and this kernel - the same on every rank -
// adds the views in rank order: the four small gradients directly, the SH gradient as basis(dir_v) (x) dL/dRGB_v.
// The outputs must hold zeros on entry (only rows that some view touched are written). One warp per chunk: the V x 4 header
// words go to shared memory, the union of the masks is compacted so that lane t takes the t-th Gaussian with a gradient.
constexpr int kRcWarps = 4;
constexpr int kRcMaxViews = 32;

template <int kDeg>
__global__ void __launch_bounds__(32 * kRcWarps) grad_record_combine_kernel(int P, int M, int V, const float* __restrict__ means3D,
                                                                           const float* __restrict__ campos,     // [V][3]
                                                                           const float* __restrict__ gathered,   // [V][slot_stride]
                                                                           size_t slot_stride, float* __restrict__ dL_dmeans3D,
                                                                           float* __restrict__ dL_dopacity, float* __restrict__ dL_dscales,
                                                                           float* __restrict__ dL_drots, float* __restrict__ dL_dsh,
                                                                           float* __restrict__ dL_dsh_rest) {
    constexpr int kCoef = (kDeg + 1) * (kDeg + 1);
    constexpr int kWords = kRecChunk / 32;
    __shared__ uint2 s_hdr[kRcWarps][kRcMaxViews][kWords];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int chunk = blockIdx.x * kRcWarps + warp;
    const int nchunks = (P + kRecChunk - 1) / kRecChunk;
    if (chunk >= nchunks) return;
    const size_t rec = (size_t)chunk * kRecFloats;
    // headers in; union of the masks per word
    uint32_t u_mine = 0;                                   // lane l: OR over the views it loaded of word (l % kWords)
    for (int e = lane; e < V * kWords; e += 32) {
        const int v = e / kWords, w = e - v * kWords;
        const uint2 h = __ldcs(reinterpret_cast<const uint2*>(gathered + (size_t)v * slot_stride + rec) + w);
        s_hdr[warp][v][w] = h;
        u_mine |= h.x;
    }
#pragma unroll
    for (int d = kWords; d < 32; d <<= 1) u_mine |= __shfl_xor_sync(0xffffffffu, u_mine, d);   // lanes l, l + 4, ... share a word
    uint32_t U[kWords];
    int first[kWords + 1];
    first[0] = 0;
#pragma unroll
    for (int w = 0; w < kWords; ++w) {
        U[w] = __shfl_sync(0xffffffffu, u_mine, w);
        first[w + 1] = first[w] + __popc(U[w]);
    }
    __syncwarp();
    const int total = first[kWords];
#pragma unroll 1
    for (int t = lane; t < total; t += 32) {
        int w = 0;
#pragma unroll
        for (int k = 1; k < kWords; ++k) w += t >= first[k] ? 1 : 0;
        uint32_t uw = U[0];
        int f0 = first[0];
#pragma unroll
        for (int k = 1; k < kWords; ++k) {
            if (w == k) { uw = U[k]; f0 = first[k]; }
        }
        const int bit = __fns(uw, 0, t - f0 + 1);          // position of this lane's set bit in the union word
        const size_t i = (size_t)chunk * kRecChunk + 32 * w + bit;
        const V3 pos(means3D[3 * i], means3D[3 * i + 1], means3D[3 * i + 2]);
        float acc[3 * kCoef];
#pragma unroll
        for (int k = 0; k < 3 * kCoef; ++k) acc[k] = 0.f;
        float sum[11];
#pragma unroll
        for (int k = 0; k < 11; ++k) sum[k] = 0.f;
        // views in batches: all records of a batch are requested before the first is used
        constexpr int kBatch = 4;
#pragma unroll 1
        for (int v0 = 0; v0 < V; v0 += kBatch) {
            float2 r[kBatch][kRecWidth / 2];
            bool have[kBatch];
#pragma unroll
            for (int u = 0; u < kBatch; ++u) {
                const int v = v0 + u;
                have[u] = false;
                if (v < V) {
                    const uint2 h = s_hdr[warp][v][w];
                    if ((h.x >> bit) & 1u) {
                        have[u] = true;
                        const float2* src = reinterpret_cast<const float2*>(
                            gathered + (size_t)v * slot_stride + rec + kRecHeader + kRecWidth * (h.y + __popc(h.x & ((1u << bit) - 1u))));
#pragma unroll
                        for (int k = 0; k < kRecWidth / 2; ++k) r[u][k] = __ldcs(src + k);
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < kBatch; ++u) {
                if (!have[u]) continue;
                const int v = v0 + u;
                const float f[kRecWidth] = {r[u][0].x, r[u][0].y, r[u][1].x, r[u][1].y, r[u][2].x, r[u][2].y, r[u][3].x,
                                            r[u][3].y, r[u][4].x, r[u][4].y, r[u][5].x, r[u][5].y, r[u][6].x, r[u][6].y};
#pragma unroll
                for (int k = 0; k < 11; ++k) sum[k] = __fadd_rn(sum[k], f[3 + k]);
                const float gr = f[0], gg = f[1], gb = f[2];
                if (gr == 0.f && gg == 0.f && gb == 0.f) continue;
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
        dL_dmeans3D[3 * i] = sum[0]; dL_dmeans3D[3 * i + 1] = sum[1]; dL_dmeans3D[3 * i + 2] = sum[2];
        dL_dopacity[i] = sum[3];
        dL_dscales[3 * i] = sum[4]; dL_dscales[3 * i + 1] = sum[5]; dL_dscales[3 * i + 2] = sum[6];
        dL_drots[4 * i] = sum[7]; dL_drots[4 * i + 1] = sum[8]; dL_drots[4 * i + 2] = sum[9]; dL_drots[4 * i + 3] = sum[10];
        if (dL_dsh_rest) {
            // split layout: dL_dsh is the [P][1][3] gradient of features_dc, dL_dsh_rest the [P][M-1][3] one of features_rest
#pragma unroll
            for (int k = 0; k < 3; ++k) dL_dsh[3 * i + k] = acc[k];
            float* rrow = dL_dsh_rest + i * (size_t)(3 * (M - 1));
#pragma unroll
            for (int k = 3; k < 3 * kCoef; ++k) rrow[k - 3] = acc[k];
            continue;
        }
        float* row = dL_dsh + i * (size_t)(3 * M);
        if ((3 * kCoef) % 4 == 0 && ((3 * M) % 4) == 0 && (reinterpret_cast<size_t>(dL_dsh) & 15) == 0) {
#pragma unroll
            for (int q = 0; q < 3 * kCoef / 4; ++q)
                reinterpret_cast<float4*>(row)[q] = make_float4(acc[4 * q], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3]);
        } else {
#pragma unroll
            for (int k = 0; k < 3 * kCoef; ++k) row[k] = acc[k];
        }
    }
}

int launch_grad_record_combine(int P, int D, int M, int V, const float* means3D, const float* campos, const float* gathered,
                               size_t slot_stride, float* dL_dmeans3D, float* dL_dopacity, float* dL_dscales, float* dL_drots,
                               float* dL_dsh, float* dL_dsh_rest, cudaStream_t stream) {
    if (P <= 0) return GS_OK;
    if (M < (D + 1) * (D + 1) || V < 1 || V > kRcMaxViews) return GS_ERR_ARG;
    const int nchunks = (P + kRecChunk - 1) / kRecChunk;
    const int blocks = (nchunks + kRcWarps - 1) / kRcWarps;
#define GS_RC_LAUNCH(DEG)                                                                                                        \
    grad_record_combine_kernel<DEG><<<blocks, 32 * kRcWarps, 0, stream>>>(P, M, V, means3D, campos, gathered, slot_stride,       \
                                                                          dL_dmeans3D, dL_dopacity, dL_dscales, dL_drots, dL_dsh,  \
                                                                          dL_dsh_rest)
    switch (D) {
        case 0: GS_RC_LAUNCH(0); break;
        case 1: GS_RC_LAUNCH(1); break;
        case 2: GS_RC_LAUNCH(2); break;
        case 3: GS_RC_LAUNCH(3); break;
        default: return GS_ERR_ARG;
    }
#undef GS_RC_LAUNCH
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
