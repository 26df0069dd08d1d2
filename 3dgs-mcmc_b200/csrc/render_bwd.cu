// Backward tile renderer for sm_100a: back-to-front replay of the blend with warp-aggregated gradient
// reductions.
//
// Same CTA / warp / pixel mapping and the same staging as the forward (bulk-copied instance slices,
// cp.async-gathered 48-byte records, see render_common.cuh), streamed in reverse and only up to the
// tile's last blended position (tile_max_contrib, recorded by the forward) instead of the end of the
// tile's list. Per (pixel, splat) pair the recomputation of G and alpha and every gradient expression
// follow the reference (cuda_rasterizer/backward.cu:399-557). The reference issues 9 scalar float
// atomics per contributing pair; here each warp culls splats that cannot touch its 8x4 patch, reduces the
// nine per-pair terms over its 32 pixels with a fold tree (14 shuffles) and issues one predicated
// red.global.add.f32 from nine lanes per (warp, splat) — and nothing at all when no lane of the warp
// blended the splat. Summation order differs from the reference (as it does between two runs of the
// reference), hence tolerance parity on gradients.
#include "bulk_copy.cuh"
#include "gs_common.cuh"
#include "kernels.cuh"
#include "render_common.cuh"
#include "workspace.cuh"

namespace gs {

GS_DEV void red_add_f32(float* addr, const float v) {
    asm volatile("red.global.add.f32 [%0], %1;" ::"l"(addr), "f"(v) : "memory");
}

__global__ void __launch_bounds__(GS_TILE_PIX) render_bwd_kernel(const RenderBwdParams p) {
    __shared__ __align__(128) SplatRec s_rec[kRecStages][kChunk];
    __shared__ __align__(16) uint64_t s_inst[kInstStages][kChunk + 2];
    __shared__ __align__(8) uint64_t s_inst_full[kInstStages];
    __shared__ __align__(8) uint64_t s_rec_full[kRecStages];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x;
    const int tile_x = tile % p.tiles_x, tile_y = tile / p.tiles_x;
    const uint32_t px = tile_x * GS_TILE_X + (warp & 1) * 8 + (lane & 7);
    const uint32_t py = tile_y * GS_TILE_Y + (warp >> 1) * 4 + (lane >> 3);
    const bool inside = px < (uint32_t)p.W && py < (uint32_t)p.H;
    const uint32_t pix_id = (uint32_t)p.W * py + px;
    const float2 pixf = {(float)px, (float)py};
    const float patch_x0 = (float)(tile_x * GS_TILE_X + (warp & 1) * 8), patch_y0 = (float)(tile_y * GS_TILE_Y + (warp >> 1) * 4);

    const uint2 range = p.ranges[tile];
    const int n_total = min((int)(range.y - range.x), (int)p.tile_max_contrib[tile]);
    const int nchunks = (n_total + kChunk - 1) / kChunk;
    if (nchunks == 0) return;

    if (tid == 0) stager_init(s_inst_full, s_rec_full);
    __syncthreads();
    // Stream chunk k is list chunk c = nchunks-1-k.
    Stager st(p.inst, p.rec, range.x, n_total, nchunks, /*reversed=*/true, &s_inst[0][0], s_inst_full, &s_rec[0][0], s_rec_full);
    st.prologue(tid);

    const size_t plane = (size_t)p.H * p.W;
    const float T_final = inside ? p.final_T[pix_id] : 0.f;
    float T = T_final;
    const int last_contributor = inside ? (int)p.n_contrib[pix_id] : 0;
    const int warp_last = __reduce_max_sync(0xffffffffu, last_contributor);
    float dL_dpixel[3] = {0.f, 0.f, 0.f};
    if (inside) {
        dL_dpixel[0] = p.dL_dpix[0 * plane + pix_id];
        dL_dpixel[1] = p.dL_dpix[1 * plane + pix_id];
        dL_dpixel[2] = p.dL_dpix[2 * plane + pix_id];
    }
    float bg_dot_dpixel = 0.f;
#pragma unroll
    for (int i = 0; i < 3; ++i) bg_dot_dpixel += p.bg[i] * dL_dpixel[i];

    float accum_rec[3] = {0.f, 0.f, 0.f};
    float last_alpha = 0.f;
    float last_color[3] = {0.f, 0.f, 0.f};
    // Gradient of pixel coordinates w.r.t. normalised screen coordinates.
    const float ddelx_dx = 0.5 * p.W;
    const float ddely_dy = 0.5 * p.H;
    const bool b4 = (lane & 16) != 0, b3 = (lane & 8) != 0, b2 = (lane & 4) != 0;

    for (int k = 0; k < nchunks; ++k) {
        if (k > 0) __syncthreads();  // every warp is done with chunk k-1: its stages can be refilled
        st.advance(tid, k);
        const int s = k % kRecStages;
        const int c = nchunks - 1 - k;
        const int n = min(kChunk, n_total - c * kChunk);
        mbar_wait(&s_rec_full[s], (uint32_t)((k / kRecStages) & 1));
        const float4* __restrict__ rec = reinterpret_cast<const float4*>(&s_rec[s][0]);
        // Positions >= warp_last were blended by no pixel of this warp.
        const int jstart = min(n - 1, warp_last - 1 - c * kChunk);
        // Lane-parallel patch cull (see render_common.cuh), 32 list positions at a time, walked back to front.
        for (int base = jstart & ~31; base >= 0; base -= 32) {
            const int jt = base + lane;
            bool may = false;
            if (jt <= jstart) {
                const float4 t0 = rec[3 * jt], t1 = rec[3 * jt + 1];
                may = patch_may_contribute(t0.x, t0.y, t0.z, t0.w, t1.x, t1.z, patch_x0, patch_y0);
            }
            uint32_t m = __ballot_sync(0xffffffffu, may);
            while (m) {
                const int b = 31 - __clz(m);
                m &= ~(1u << b);
                const int j = base + b;
                const int pos = c * kChunk + j;  // 0-based position in the tile list
                const float4 q0 = rec[3 * j];       // x, y, A, B
                const float4 q1 = rec[3 * j + 1];   // C, opacity, thr, id
                const float2 d = {q0.x - pixf.x, q0.y - pixf.y};
                const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
                bool act = pos < last_contributor && !(power > 0.0f) && !(power < q1.z);
                if (!__any_sync(0xffffffffu, act)) continue;
                const float G = exp(power);
                const float alpha = min(0.99f, q1.y * G);
                act = act && !(alpha < 1.0f / 255.0f);
                if (!__any_sync(0xffffffffu, act)) continue;

                float v[8];
                float v8 = 0.f;
#pragma unroll
                for (int i = 0; i < 8; ++i) v[i] = 0.f;
                if (act) {
                    const float4 q2 = rec[3 * j + 2];  // r, g, b
                    const float col[3] = {q2.x, q2.y, q2.z};
                    // 1 - alpha is in [0.01, 1]: the approximate reciprocal (1 ulp) stands in for the reference's two
                    // divisions; the difference is far inside the gradient tolerance.
                    const float inv_1ma = __fdividef(1.0f, 1.f - alpha);
                    T = T * inv_1ma;
                    const float dchannel_dcolor = alpha * T;
                    float dL_dalpha = 0.0f;
#pragma unroll
                    for (int ch = 0; ch < 3; ++ch) {
                        const float cc = col[ch];
                        accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                        last_color[ch] = cc;
                        const float dL_dchannel = dL_dpixel[ch];
                        dL_dalpha += (cc - accum_rec[ch]) * dL_dchannel;
                        v[ch] = dchannel_dcolor * dL_dchannel;
                    }
                    dL_dalpha *= T;
                    last_alpha = alpha;
                    dL_dalpha += (-T_final * inv_1ma) * bg_dot_dpixel;

                    const float dL_dG = q1.y * dL_dalpha;
                    const float gdx = G * d.x;
                    const float gdy = G * d.y;
                    const float dG_ddelx = -gdx * q0.z - gdy * q0.w;
                    const float dG_ddely = -gdy * q1.x - gdx * q0.w;
                    v[3] = G * dL_dalpha;                    // opacity
                    v[4] = dL_dG * dG_ddelx * ddelx_dx;      // mean2D.x
                    v[5] = dL_dG * dG_ddely * ddely_dy;      // mean2D.y
                    v[6] = -0.5f * gdx * d.x * dL_dG;        // conic.x
                    v[7] = -0.5f * gdx * d.y * dL_dG;        // conic.y
                    v8 = -0.5f * gdy * d.y * dL_dG;          // conic.w
                }
                // Fold tree: 8 values -> 4 -> 2 -> 1 per lane over lane bits 4, 3, 2 (7 shuffles), then two plain
                // exchanges over bits 1, 0: lane 4k ends with the warp total of v[k]. v8 takes a 5-step butterfly.
                float r4[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float send = b4 ? v[i] : v[i + 4];
                    const float keep = b4 ? v[i + 4] : v[i];
                    r4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
                }
                float r2[2];
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const float send = b3 ? r4[i] : r4[i + 2];
                    const float keep = b3 ? r4[i + 2] : r4[i];
                    r2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
                }
                float r1;
                {
                    const float send = b2 ? r2[0] : r2[1];
                    const float keep = b2 ? r2[1] : r2[0];
                    r1 = keep + __shfl_xor_sync(0xffffffffu, send, 4);
                }
                r1 += __shfl_xor_sync(0xffffffffu, r1, 2);
                r1 += __shfl_xor_sync(0xffffffffu, r1, 1);
#pragma unroll
                for (int msk = 16; msk >= 1; msk >>= 1) v8 += __shfl_xor_sync(0xffffffffu, v8, msk);
                const int id = __float_as_int(q1.w);
                // lane 4k adds v[k] (k < 4: colour.rgb, opacity; k >= 4: mean2D.xy, conic.xy); lane 1 adds conic.w
                if ((lane & 3) == 0) {
                    float* dst = reinterpret_cast<float*>((lane & 16 ? p.acc_mean_conic : p.acc_color_opacity) + id) + ((lane >> 2) & 3);
                    red_add_f32(dst, r1);
                } else if (lane == 1) {
                    red_add_f32(p.acc_conic_w + id, v8);
                }
            }
        }
    }
    st.drain(tid);  // every copy was consumed (no early exit); this only retires the cp.async groups
}

int launch_render_bwd(const RenderBwdParams& p, int tiles, cudaStream_t stream) {
    if (tiles <= 0) return GS_OK;
    render_bwd_kernel<<<tiles, GS_TILE_PIX, 0, stream>>>(p);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

}  // namespace gs
