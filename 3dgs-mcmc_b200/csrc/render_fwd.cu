// Forward tile renderer for sm_100a: per-tile front-to-back alpha blending.
//
// One CTA per 16x16 tile; each warp owns an 8x4 pixel patch (compact footprint -> whole-warp rejects
// are frequent). The tile's slice of the sorted instance list is streamed into shared memory by 1-D bulk
// async copies (cp.async.bulk + mbarrier, 4-deep ring, one issuing thread); the 48-byte render records
// those instances name are then gathered straight into a 3-deep shared-memory ring with 16-byte
// cp.async copies (no registers, completion tracked by an mbarrier), two chunks ahead of the blending,
// so no indirect load sits on the blending critical path and no tile-ordered copy of the records is ever
// written to HBM (the reference gathers with plain loads inside the loop, forward.cu:318-321).
// Per 32 records each lane runs a conservative patch cull for one record (render_common.cuh); only the
// survivors are evaluated per pixel. Every blend decision and every arithmetic expression that reaches
// the output is the reference's (cuda_rasterizer/forward.cu:261-377), so the image, final_T and n_contrib
// match it bit for bit. is_used flags are collected per stage in shared memory and written once per
// (tile, splat) instead of once per (pixel, splat).
#include "bulk_copy.cuh"
#include "gs_common.cuh"
#include "kernels.cuh"
#include "render_common.cuh"
#include "workspace.cuh"

namespace gs {

__global__ void __launch_bounds__(GS_TILE_PIX) render_fwd_kernel(const RenderFwdParams p) {
    __shared__ __align__(128) SplatRec s_rec[kRecStages][kChunk];
    __shared__ __align__(16) uint32_t s_inst[kInstStages][kChunk + kInstPad];
    __shared__ __align__(8) uint64_t s_inst_full[kInstStages];
    __shared__ __align__(8) uint64_t s_rec_full[kRecStages];
    __shared__ int s_used[kRecStages][kChunk];  // id+1 of splats blended by some pixel, 0 otherwise
    __shared__ uint32_t s_maxc;

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
    const int n_total = (int)(range.y - range.x);
    const int nchunks = (n_total + kChunk - 1) / kChunk;

    for (int i = tid; i < kRecStages * kChunk; i += GS_TILE_PIX) (&s_used[0][0])[i] = 0;
    if (tid == 0) {
        s_maxc = 0;
        stager_init<kChunk>(s_inst_full, s_rec_full);
    }
    __syncthreads();

    // List chunk c covers positions [c*kChunk, c*kChunk + n) of the tile's list.
    Stager<kChunk> st(p.ids, p.rec, range.x, n_total, nchunks, /*reversed=*/false, &s_inst[0][0], s_inst_full, &s_rec[0][0], s_rec_full);
    st.prologue(tid);

    auto flush_used = [&](int c) {
        const int s = c % kRecStages;
        if (tid < kChunk) {
            const int u = s_used[s][tid];
            if (u) {
                p.is_used[u - 1] = 1;
                s_used[s][tid] = 0;
            }
        }
    };

    float T = 1.0f;
    uint32_t last_contributor = 0;
    float C[3] = {0.f, 0.f, 0.f};
    bool done = !inside;

    int c = 0;
    for (; c < nchunks; ++c) {
        const int ndone = __syncthreads_count(done);
        if (c > 0) flush_used(c - 1);
        if (ndone == GS_TILE_PIX) break;
        st.advance(tid, c);

        const int s = c % kRecStages;
        const int n = min(kChunk, n_total - c * kChunk);
        mbar_wait(&s_rec_full[s], (uint32_t)((c / kRecStages) & 1));
        // Lane-parallel patch cull: lane l tests splat base+l against this warp's 8x4 patch; only the survivors
        // are evaluated per pixel. Positions keep their list index, so n_contrib is unchanged.
        if (!__all_sync(0xffffffffu, done)) {
            const float4* __restrict__ rec = reinterpret_cast<const float4*>(&s_rec[s][0]);
            for (int base = 0; base < n; base += 32) {
                if (__all_sync(0xffffffffu, done)) break;
                const int jt = base + lane;
                bool may = false;
                if (jt < n) {
                    const float4 t0 = rec[3 * jt], t1 = rec[3 * jt + 1];
                    may = patch_may_contribute(t0.x, t0.y, t0.z, t0.w, t1.x, t1.z, patch_x0, patch_y0);
                }
                uint32_t m = __ballot_sync(0xffffffffu, may);
                while (m) {
                    const int j = base + __ffs(m) - 1;
                    m &= m - 1;
                    if (done) continue;
                    const float4 q0 = rec[3 * j];      // x, y, A, B
                    const float4 q1 = rec[3 * j + 1];  // C, opacity, thr, id
                    const float2 d = {q0.x - pixf.x, q0.y - pixf.y};
                    const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
                    if (power > 0.0f) continue;
                    if (power < q1.z) continue;  // cannot reach alpha >= 1/255 (see preprocess_one)
                    const float alpha = min(0.99f, q1.y * exp(power));
                    if (alpha < 1.0f / 255.0f) continue;
                    const float test_T = T * (1 - alpha);
                    if (test_T < 0.0001f) {
                        done = true;
                        continue;
                    }
                    const float4 q2 = rec[3 * j + 2];  // r, g, b
                    C[0] += q2.x * alpha * T;
                    C[1] += q2.y * alpha * T;
                    C[2] += q2.z * alpha * T;
                    T = test_T;
                    last_contributor = (uint32_t)(c * kChunk + j + 1);
                    s_used[s][j] = __float_as_int(q1.w) + 1;
                }
            }
        }
    }
    __syncthreads();
    if (c == nchunks && nchunks > 0) flush_used(nchunks - 1);
    // Copies that were issued but never consumed (early exit) must land before the CTA retires.
    st.drain(tid);

    const uint32_t wmax = __reduce_max_sync(0xffffffffu, last_contributor);
    if (lane == 0 && wmax) atomicMax(&s_maxc, wmax);
    if (inside) {
        p.final_T[pix_id] = T;
        p.n_contrib[pix_id] = last_contributor;
        const size_t plane = (size_t)p.H * p.W;
        p.out_color[0 * plane + pix_id] = C[0] + T * p.bg[0];
        p.out_color[1 * plane + pix_id] = C[1] + T * p.bg[1];
        p.out_color[2 * plane + pix_id] = C[2] + T * p.bg[2];
    }
    __syncthreads();
    if (tid == 0) p.tile_max_contrib[tile] = s_maxc;
}

int launch_render_fwd(const RenderFwdParams& p, int tiles, cudaStream_t stream) {
    if (tiles <= 0) return GS_OK;
    render_fwd_kernel<<<tiles, GS_TILE_PIX, 0, stream>>>(p);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

}  // namespace gs
