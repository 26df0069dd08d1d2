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

#include <cstdlib>

namespace gs {

// CH records per staged chunk, RS record stages, GA gather-ahead. kTwoLevel: two-level cull (see below).
//
// Two-level cull. Most of a tile's list never reaches its pixels: the reference puts a Gaussian into every tile its
// 3-sigma bounding SQUARE touches, but only 42 % of those instances come within alpha >= 1/255 of ANY pixel of the tile,
// and the forward - unlike the backward, which stops at the last blended position - walks the list until its pixels
// saturate, which most never do. Testing every record against every warp's patch (8 tests per record) was 58 % of this
// kernel's instructions. Level 1: thread r tests record r of the chunk against the whole 16x16 tile, once per CTA
// (box_may_contribute<15, 15>: the same conservative test, wider box); the ballots give a survivor bitmap of the chunk.
// Level 2: every warp compacts the bitmap into a list of record indices (list order preserved) and runs its patch test
// only on those. Nothing that can contribute is ever dropped, so image / final_T / n_contrib / is_used are unchanged.
template <int CH, int RS, int GA, bool kTwoLevel, bool kCount = false>
__global__ void __launch_bounds__(GS_TILE_PIX, 6) render_fwd_kernel(const RenderFwdParams p) {
    pdl_trigger();
    pdl_wait();
    constexpr int kWords = CH / 32;              // survivor bitmap words per chunk
    __shared__ __align__(128) SplatRec s_rec[RS][CH];
    __shared__ __align__(16) uint32_t s_inst[kInstStages][CH + kInstPad];
    __shared__ __align__(8) uint64_t s_inst_full[kInstStages];
    __shared__ __align__(8) uint64_t s_rec_full[RS];
    __shared__ int s_used[RS][CH];  // id of splats blended by some pixel, -1 otherwise
    __shared__ uint32_t s_mask[kTwoLevel ? kWords : 1];
    __shared__ uint8_t s_cidx[kTwoLevel ? GS_TILE_PIX / 32 : 1][kTwoLevel ? CH : 1];   // per warp: compacted record indices
    __shared__ uint32_t s_maxc;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x;
    const int tile_x = tile % p.tiles_x, tile_y = tile / p.tiles_x;
    const uint32_t px = tile_x * GS_TILE_X + (warp & 1) * 8 + (lane & 7);
    const uint32_t py = tile_y * GS_TILE_Y + (warp >> 1) * 4 + (lane >> 3);
    const bool inside = px < (uint32_t)p.W && py < (uint32_t)p.H;
    const uint32_t pix_id = (uint32_t)p.W * py + px;
    const float2 pixf = {(float)px, (float)py};
    const float tile_x0 = (float)(tile_x * GS_TILE_X), tile_y0 = (float)(tile_y * GS_TILE_Y);
    const float patch_x0 = tile_x0 + (float)((warp & 1) * 8), patch_y0 = tile_y0 + (float)((warp >> 1) * 4);

    const uint2 range = p.ranges[tile];
    const int n_total = (int)(range.y - range.x);
    const int nchunks = (n_total + CH - 1) / CH;

    for (int i = tid; i < RS * CH; i += GS_TILE_PIX) (&s_used[0][0])[i] = -1;
    if (tid == 0) {
        s_maxc = 0;
        stager_init<CH, RS>(s_inst_full, s_rec_full);
    }
    __syncthreads();

    // List chunk c covers positions [c*CH, c*CH + n) of the tile's list.
    Stager<CH, RS, GA> st(p.ids, p.rec, range.x, n_total, nchunks, /*reversed=*/false, &s_inst[0][0], s_inst_full, &s_rec[0][0], s_rec_full);
    st.prologue(tid);

    auto flush_used = [&](int c) {
        const int s = c % RS;
        for (int i = tid; i < CH; i += GS_TILE_PIX) {
            const int u = s_used[s][i];
            if (u >= 0) {
                p.is_used[u] = 1;
                s_used[s][i] = -1;
            }
        }
    };

    // kCount (measurement build of the same kernel, gs_debug_forward_work): per-thread tallies, summed into p.counters at the end:
    // [0] warp-chunks walked, [1] records patch-tested (x 1 lane each), [2] (warp, record) pairs evaluated per pixel,
    // [3] (pixel, record) blends, [4] records that survived the tile-level test (per warp-chunk)
    unsigned long long cnt[5] = {0, 0, 0, 0, 0};
    float T = 1.0f;
    uint32_t last_contributor = 0;
    float C[3] = {0.f, 0.f, 0.f};
    // `gate` is the pixel's "done" state folded into the alpha-threshold test: -inf while the pixel still blends (the test
    // is then exactly the record's own `power < thr`), +inf once it has saturated (every record fails the test).
    float gate = inside ? -INFINITY : INFINITY;
#define done (gate > 0.0f)
    const uint32_t lt_mask = (1u << lane) - 1u;

    // Blend record j of the chunk into this lane's pixel (the reference's expression tree, forward.cu:330-358).
    auto blend = [&](const float4* __restrict__ rec, const int s, const int c, const int j) {
        const float4 q0 = rec[3 * j];      // x, y, A, B
        const float4 q1 = rec[3 * j + 1];  // C, opacity, thr, id
        const float2 d = {q0.x - pixf.x, q0.y - pixf.y};
        const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
        if (power > 0.0f) return;
        if (power < fmaxf(q1.z, gate)) return;  // cannot reach alpha >= 1/255 (see preprocess_fwd_kernel), or pixel done
        const float alpha = min(0.99f, q1.y * exp(power));
        if (alpha < 1.0f / 255.0f) return;
        const float test_T = T * (1 - alpha);
        if (test_T < 0.0001f) {
            asm volatile("mov.f32 %0, 0f7f800000;" : "=f"(gate));   // +inf; volatile: not to be hoisted and undone on the blend path
            return;
        }
        const float4 q2 = rec[3 * j + 2];  // r, g, b
        C[0] += q2.x * alpha * T;
        C[1] += q2.y * alpha * T;
        C[2] += q2.z * alpha * T;
        T = test_T;
        last_contributor = (uint32_t)(c * CH + j + 1);
        s_used[s][j] = __float_as_int(q1.w);
        if (kCount) ++cnt[3];
    };

    int c = 0;
    for (; c < nchunks; ++c) {
        const int ndone = __syncthreads_count(done);
        if (c > 0) flush_used(c - 1);
        if (ndone == GS_TILE_PIX) break;
        st.advance(tid, c);

        const int s = c % RS;
        const int n = min(CH, n_total - c * CH);
        mbar_wait(&s_rec_full[s], (uint32_t)((c / RS) & 1));
        const float4* __restrict__ rec = reinterpret_cast<const float4*>(&s_rec[s][0]);
        if (kTwoLevel) {
            // ---- level 1: one tile-level test per record, once per CTA ----
            if (tid < CH) {
                bool may = false;
                if (tid < n) {
                    const float4 t0 = rec[3 * tid], t1 = rec[3 * tid + 1];
                    may = box_may_contribute<GS_TILE_X - 1, GS_TILE_Y - 1>(t0.x, t0.y, t0.z, t0.w, t1.x, t1.z, tile_x0, tile_y0);
                }
                const uint32_t m = __ballot_sync(0xffffffffu, may);
                if (lane == 0) s_mask[warp] = m;
            }
            __syncthreads();
            // ---- level 2: per warp, compact the tile-level survivors (list order kept), patch-test those ----
            if (!__all_sync(0xffffffffu, done)) {
                uint8_t* const cidx = s_cidx[warp];
                int total = 0;
#pragma unroll
                for (int w = 0; w < kWords; ++w) {
                    const uint32_t m = s_mask[w];
                    if ((m >> lane) & 1u) cidx[total + __popc(m & lt_mask)] = (uint8_t)(32 * w + lane);
                    total += __popc(m);
                }
                __syncwarp();
                if (kCount && lane == 0) { ++cnt[0]; cnt[4] += (unsigned)total; }
                for (int base = 0; base < total; base += 32) {
                    if (__all_sync(0xffffffffu, done)) break;
                    const int k = base + lane;
                    if (kCount && k < total) ++cnt[1];
                    bool may = false;
                    int jt = 0;
                    if (k < total) {
                        jt = cidx[k];
                        const float4 t0 = rec[3 * jt], t1 = rec[3 * jt + 1];
                        may = patch_may_contribute(t0.x, t0.y, t0.z, t0.w, t1.x, t1.z, patch_x0, patch_y0);
                    }
                    uint32_t m = __ballot_sync(0xffffffffu, may);
                    if (kCount && lane == 0) cnt[2] += (unsigned)__popc(m);
                    while (m) {
                        const int b = __ffs(m) - 1;
                        m &= m - 1;
                        blend(rec, s, c, __shfl_sync(0xffffffffu, jt, b));
                    }
                }
                __syncwarp();   // cidx is rewritten for the next chunk only after every lane has read it
            }
        } else {
            // Lane-parallel patch cull: lane l tests splat base+l against this warp's 8x4 patch; only the survivors
            // are evaluated per pixel. Positions keep their list index, so n_contrib is unchanged.
            if (!__all_sync(0xffffffffu, done)) {
                if (kCount && lane == 0) { ++cnt[0]; cnt[4] += (unsigned)n; }
                for (int base = 0; base < n; base += 32) {
                    if (__all_sync(0xffffffffu, done)) break;
                    const int jt = base + lane;
                    bool may = false;
                    if (kCount && jt < n) ++cnt[1];
                    if (jt < n) {
                        const float4 t0 = rec[3 * jt], t1 = rec[3 * jt + 1];
                        may = patch_may_contribute(t0.x, t0.y, t0.z, t0.w, t1.x, t1.z, patch_x0, patch_y0);
                    }
                    uint32_t m = __ballot_sync(0xffffffffu, may);
                    if (kCount && lane == 0) cnt[2] += (unsigned)__popc(m);
                    while (m) {
                        const int j = base + __ffs(m) - 1;
                        m &= m - 1;
                        blend(rec, s, c, j);
                    }
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
    if (kCount && p.counters) {
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            const unsigned long long v = cnt[k];
            if (v) atomicAdd(p.counters + k, v);
        }
    }
}

#undef done

int launch_render_fwd(const RenderFwdParams& p, int tiles, cudaStream_t stream) {
    if (tiles <= 0) return GS_OK;
    // GS_FWD_VARIANT (measurement): 0 = one-level cull (round 1), 2 = two-level with 128-record chunks, default = two-level
    // with 256-record chunks (one level-1 test per thread)
    static const int variant = [] { const char* e = getenv("GS_FWD_VARIANT"); return e ? atoi(e) : 1; }();
    if (p.counters) {   // measurement: the same kernels with tallies
        if (variant == 0) render_fwd_kernel<128, 3, 2, false, true><<<tiles, GS_TILE_PIX, 0, stream>>>(p);
        else render_fwd_kernel<256, 2, 1, true, true><<<tiles, GS_TILE_PIX, 0, stream>>>(p);
        GS_LAUNCH_CHECK();
        return GS_OK;
    }
    switch (variant) {
        case 0: render_fwd_kernel<128, 3, 2, false><<<tiles, GS_TILE_PIX, 0, stream>>>(p); break;
        case 2: render_fwd_kernel<128, 3, 2, true><<<tiles, GS_TILE_PIX, 0, stream>>>(p); break;
        default: launch_pdl(render_fwd_kernel<256, 2, 1, true, false>, dim3(tiles), dim3(GS_TILE_PIX), 0, stream, p); break;
    }
    GS_LAUNCH_CHECK();
    return GS_OK;
}

}  // namespace gs
