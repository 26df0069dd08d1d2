// Backward tile renderer for sm_100a: back-to-front replay of the blend with deferred, moment-based
// gradient reductions.
//
// Same CTA / warp / pixel mapping and the same staging as the forward (bulk-copied instance slices,
// cp.async-gathered 48-byte records, see render_common.cuh), streamed in reverse and only up to the
// tile's last blended position (tile_max_contrib, recorded by the forward) instead of the end of the
// tile's list. Per (pixel, splat) pair the recomputation of G and alpha and the expression for dL/dalpha
// follow the reference (cuda_rasterizer/backward.cu:399-557).
//
// The reference issues 9 scalar float atomics per contributing pair (backward.cu:523-554). All nine sums
// over the pixels of a warp's 8x4 patch are linear in just two per-pixel scalars,
//     u = alpha * T                     (colour:   dL/dc_ch  = sum_pix u * dL/dpix_ch)
//     w = G * dL/dalpha                 (opacity:  dL/do     = sum_pix w)
// because with d = (X - px, Y - py), px = patch_x0 + lx, py = patch_y0 + ly (lx 0..7, ly 0..3 are lane constants)
//     dL/dmean2D, dL/dconic  are  o * {A, B, C}-weighted combinations of  sum w*dx, sum w*dy, sum w*dx^2,
//     sum w*dx*dy, sum w*dy^2,  and those follow from the six moments  sum w * {1, lx, ly, lx^2, lx*ly, ly^2}.
// So each warp appends (u, w) of every splat it blended as one 32-float row pair to a small shared-memory
// buffer (two STS per splat instead of a 14-shuffle fold tree) and, every ROWS splats, reduces the rows
// as a tiny dense product against lane-constant coefficients: lane (row r, column group g) sums its
// columns with immediates for lx / ly, the groups combine with one or two butterfly steps, and the row's
// owner issues three reductions (two 16-byte `red.global.add.v4.f32`, one scalar) for the whole patch.
// Summation order differs from the reference (as it does between two runs of the reference), hence
// tolerance parity on gradients.
#include "bulk_copy.cuh"
#include "gs_common.cuh"
#include "kernels.cuh"
#include "render_common.cuh"
#include "workspace.cuh"

#include <atomic>
#include <cstdlib>

namespace gs {

GS_DEV void red_add_f32(float* addr, const float v) {
    asm volatile("red.global.add.f32 [%0], %1;" ::"l"(addr), "f"(v) : "memory");
}
GS_DEV uint32_t bfind_u32(const uint32_t v) {   // index of the most significant set bit (v != 0): one FLO
    uint32_t r;
    asm("bfind.u32 %0, %1;" : "=r"(r) : "r"(v));
    return r;
}
GS_DEV void red_add_f32x4(float4* addr, const float a, const float b, const float c, const float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

constexpr int kBwdWarps = GS_TILE_PIX / 32;
// Pending (u, w) rows live at a pitch of 36 floats: the row-wise appends (32 consecutive floats) and the row-per-lane
// float4 reads of the flush (8 lanes per LDS.128 phase, rows r..r+7: banks 4 r + 4 q mod 32) are both conflict free, and the
// append address is a running pointer (one add per splat) instead of a swizzle recomputed per splat.
constexpr int kRowPitch = 36;

// Dynamic shared-memory layout of render_bwd_kernel<CH, ROWS>.
template <int CH, int ROWS, int RS = kRecStages>
struct BwdSmem {
    static constexpr size_t kRec = 0;
    static constexpr size_t kInst = kRec + sizeof(SplatRec) * RS * CH;
    static constexpr size_t kInstFull = kInst + sizeof(uint32_t) * kInstStages * (CH + kInstPad);
    static constexpr size_t kRecFull = kInstFull + sizeof(uint64_t) * kInstStages;
    static constexpr size_t kRows = (kRecFull + sizeof(uint64_t) * RS + 127) / 128 * 128;
    static constexpr size_t kDpix = kRows + sizeof(float) * kBwdWarps * 2 * ROWS * kRowPitch;
    static constexpr size_t kBytes = kDpix + sizeof(float4) * kBwdWarps * 32;
};

// Reduce the pending rows of one warp and add them to the per-Gaussian accumulators.
// U, W: [ROWS][kRowPitch] floats, element (row, lane) at row * kRowPitch + lane.
template <int ROWS>
GS_DEV void flush_rows(const RenderBwdParams& p, const float* __restrict__ U, const float* __restrict__ W,
                                        const float4* __restrict__ dpix, const int nrows, const int myid, const int lane,
                                        const float patch_x0, const float patch_y0) {
    constexpr int GROUPS = 32 / ROWS;        // column groups per row (2 or 4)
    constexpr int COLS = 32 / GROUPS;        // columns per lane (16 or 8)
    const int r = lane & (ROWS - 1), g = lane / ROWS;
    const bool valid = r < nrows;
    float4 m0 = {0.f, 0.f, 0.f, 0.f}, m1 = {0.f, 0.f, 0.f, 0.f};
    if (valid) {  // x, y, A, B | C, opacity, thr, id of the row's splat (L2-resident: this CTA gathered it moments ago)
        const float4* rp = reinterpret_cast<const float4*>(p.rec + myid);
        m0 = __ldg(rp);
        m1 = __ldg(rp + 1);
    }
    const float4* Ur = reinterpret_cast<const float4*>(U + r * kRowPitch);
    const float4* Wr = reinterpret_cast<const float4*>(W + r * kRowPitch);
    const float4* dp = dpix + COLS * g;
    float cr = 0.f, cg = 0.f, cb = 0.f;
    float wv[COLS];
#pragma unroll
    for (int q = 0; q < COLS / 4; ++q) {
        const int chunk = (COLS / 4) * g + q;
        const float4 u4 = Ur[chunk], w4 = Wr[chunk];
        wv[4 * q + 0] = w4.x; wv[4 * q + 1] = w4.y; wv[4 * q + 2] = w4.z; wv[4 * q + 3] = w4.w;
        const float uu[4] = {u4.x, u4.y, u4.z, u4.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float4 d = dp[4 * q + e];
            cr = fmaf(uu[e], d.x, cr);
            cg = fmaf(uu[e], d.y, cg);
            cb = fmaf(uu[e], d.z, cb);
        }
    }
    // Moments of w over this lane's columns; column c of the warp is pixel (lx, ly) = (c & 7, c >> 3).
    float S0, Sx, Sy, Sxx, Sxy, Syy;
    if constexpr (ROWS == 16) {
        // columns 16g .. 16g+15: ly = 2g + (k >> 3), lx = k & 7
        float c8[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) c8[k] = wv[k] + wv[8 + k];
        S0 = c8[0]; Sx = 0.f; Sxx = 0.f;
        float Sy1 = wv[8], Sxy1 = 0.f;
#pragma unroll
        for (int k = 1; k < 8; ++k) {
            S0 += c8[k];
            Sx = fmaf((float)k, c8[k], Sx);
            Sxx = fmaf((float)(k * k), c8[k], Sxx);
            Sy1 += wv[8 + k];
            Sxy1 = fmaf((float)k, wv[8 + k], Sxy1);
        }
        const float hf = (float)(2 * g);  // ly = ly' + hf with ly' in {0, 1}: ly^2 = ly' + 2 hf ly' + hf^2
        Sy = fmaf(hf, S0, Sy1);
        Sxy = fmaf(hf, Sx, Sxy1);
        Syy = fmaf(hf, fmaf(hf, S0, 2.f * Sy1), Sy1);
    } else {
        // columns 8g .. 8g+7: ly = g, lx = k
        S0 = wv[0]; Sx = 0.f; Sxx = 0.f;
#pragma unroll
        for (int k = 1; k < 8; ++k) {
            S0 += wv[k];
            Sx = fmaf((float)k, wv[k], Sx);
            Sxx = fmaf((float)(k * k), wv[k], Sxx);
        }
        const float gf = (float)g;
        Sy = gf * S0;
        Sxy = gf * Sx;
        Syy = gf * Sy;
    }
#pragma unroll
    for (int msk = 16; msk >= ROWS; msk >>= 1) {
        cr += __shfl_xor_sync(0xffffffffu, cr, msk);
        cg += __shfl_xor_sync(0xffffffffu, cg, msk);
        cb += __shfl_xor_sync(0xffffffffu, cb, msk);
        S0 += __shfl_xor_sync(0xffffffffu, S0, msk);
        Sx += __shfl_xor_sync(0xffffffffu, Sx, msk);
        Sy += __shfl_xor_sync(0xffffffffu, Sy, msk);
        Sxx += __shfl_xor_sync(0xffffffffu, Sxx, msk);
        Sxy += __shfl_xor_sync(0xffffffffu, Sxy, msk);
        Syy += __shfl_xor_sync(0xffffffffu, Syy, msk);
    }
    if (valid) {
        if (g == 0) {
            red_add_f32x4(p.acc_color_opacity + myid, cr, cg, cb, S0);
        } else {
            // d = (Xp - lx, Yp - ly): sum w dx = Xp S0 - Sx, sum w dx^2 = Xp (Xp S0 - Sx) - (Xp Sx - Sxx), ...
            const float Xp = m0.x - patch_x0, Yp = m0.y - patch_y0;
            const float A = m0.z, B = m0.w, C = m1.x, o = m1.y;
            const float Mx = fmaf(Xp, S0, -Sx), My = fmaf(Yp, S0, -Sy);
            if (g == 1) {
                const float Mxx = fmaf(Xp, Mx, -fmaf(Xp, Sx, -Sxx));
                const float Mxy = fmaf(Xp, My, -fmaf(Yp, Sx, -Sxy));
                // Gradient of pixel coordinates w.r.t. normalised screen coordinates (reference backward.cu:460-461).
                const float ddelx_dx = 0.5f * p.W, ddely_dy = 0.5f * p.H;
                const float ho = -0.5f * o;
                red_add_f32x4(p.acc_mean_conic + myid, o * (-A * Mx - B * My) * ddelx_dx, o * (-C * My - B * Mx) * ddely_dy,
                              ho * Mxx, ho * Mxy);
                if constexpr (GROUPS == 2) {
                    const float Myy = fmaf(Yp, My, -fmaf(Yp, Sy, -Syy));
                    red_add_f32(p.acc_conic_w + myid, ho * Myy);
                }
            } else if (GROUPS > 2 && g == 2) {
                const float Myy = fmaf(Yp, My, -fmaf(Yp, Sy, -Syy));
                red_add_f32(p.acc_conic_w + myid, -0.5f * o * Myy);
            }
        }
    }
    __syncwarp();  // the rows may be overwritten once every lane has read them
}

template <int CH, int ROWS, int MIN_CTAS, int RS = kRecStages, int GA = kGatherAhead>
__global__ void __launch_bounds__(GS_TILE_PIX, MIN_CTAS) render_bwd_kernel(const RenderBwdParams p) {
    using L = BwdSmem<CH, ROWS, RS>;
    extern __shared__ __align__(128) unsigned char smem[];
    SplatRec* s_rec = reinterpret_cast<SplatRec*>(smem + L::kRec);
    uint32_t* s_inst = reinterpret_cast<uint32_t*>(smem + L::kInst);
    uint64_t* s_inst_full = reinterpret_cast<uint64_t*>(smem + L::kInstFull);
    uint64_t* s_rec_full = reinterpret_cast<uint64_t*>(smem + L::kRecFull);

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
    const int nchunks = (n_total + CH - 1) / CH;
    if (nchunks == 0) return;

    if (tid == 0) stager_init<CH, RS>(s_inst_full, s_rec_full);
    __syncthreads();
    // Stream chunk k is list chunk c = nchunks-1-k.
    Stager<CH, RS, GA> st(p.ids, p.rec, range.x, n_total, nchunks, /*reversed=*/true, s_inst, s_inst_full, s_rec, s_rec_full);
    st.prologue(tid);

    float* const U = reinterpret_cast<float*>(smem + L::kRows) + warp * (2 * ROWS * kRowPitch);
    float* const W = U + ROWS * kRowPitch;
    float4* const dpix = reinterpret_cast<float4*>(smem + L::kDpix) + warp * 32;

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
    dpix[lane] = make_float4(dL_dpixel[0], dL_dpixel[1], dL_dpixel[2], 0.f);
    float bg_dot_dpixel = 0.f;
#pragma unroll
    for (int i = 0; i < 3; ++i) bg_dot_dpixel += p.bg[i] * dL_dpixel[i];

    // The reference carries accum_rec[ch] / last_color[ch] / last_alpha per channel (backward.cu:487-499) and forms
    // dL/dalpha = sum_ch (c_ch - accum_rec_ch) dL/dpix_ch. Only that dot product is ever used, and it obeys a scalar recurrence:
    // with S = sum_ch accum_rec_ch dL/dpix_ch and cdot = sum_ch c_ch dL/dpix_ch, the next splat's S is S + alpha (cdot - S).
    // Same value up to summation order (gradients are tolerance-gated), 14 instead of 26 instructions per blended pair.
    float S = 0.f;
    const float nTb = -T_final * bg_dot_dpixel;
    int nrows = 0;   // pending rows of this warp (uniform)
    int myid = 0;    // Gaussian index of pending row (lane & (ROWS-1))
    float* urow = U + lane;   // append slot of the next row
    __syncwarp();

    for (int k = 0; k < nchunks; ++k) {
        if (k > 0) __syncthreads();  // every warp is done with chunk k-1: its stages can be refilled
        st.advance(tid, k);
        const int s = k % RS;
        const int c = nchunks - 1 - k;
        const int n = min(CH, n_total - c * CH);
        mbar_wait(&s_rec_full[s], (uint32_t)((k / RS) & 1));
        const float4* __restrict__ rec = reinterpret_cast<const float4*>(s_rec + s * CH);
        // Positions >= warp_last were blended by no pixel of this warp.
        const int jstart = min(n - 1, warp_last - 1 - c * CH);
        const int lane_last = last_contributor - c * CH;   // this pixel blended chunk positions j < lane_last
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
                const int b = (int)bfind_u32(m);
                m ^= 1u << b;
                const int j = base + b;
                const float4 q0 = rec[3 * j];       // x, y, A, B
                const float4 q1 = rec[3 * j + 1];   // C, opacity, thr, id
                const float2 d = {q0.x - pixf.x, q0.y - pixf.y};
                const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
                // The same expf as the forward, on purpose: the replay T <- T / (1 - alpha) only telescopes back to the
                // forward's transmittances when alpha is reproduced bit for bit. With ex2.approx (7 instructions fewer
                // per pair) the gradients drifted by ~3e-5 relative and dL/dopacity, a sum with heavy cancellation, by
                // 1.1e-3 against the reference (tools/diag_grad_noise.py) - outside the 1e-3 tolerance.
                const float G = exp(power);
                const float alpha = min(0.99f, q1.y * G);
                const bool act = j < lane_last && !(power > 0.0f) && !(power < q1.z) && !(alpha < 1.0f / 255.0f);
                // No warp-wide early-outs here: 99 % of the splats that pass the patch cull are blended by some lane.

                float u = 0.f, w = 0.f;
                if (act) {
                    const float4 q2 = rec[3 * j + 2];  // r, g, b
                    const float cdot = fmaf(q2.z, dL_dpixel[2], fmaf(q2.y, dL_dpixel[1], q2.x * dL_dpixel[0]));
                    // 1 - alpha is in [0.01, 1]: the approximate reciprocal (1 ulp) stands in for the reference's two
                    // divisions; the difference is far inside the gradient tolerance.
                    const float inv_1ma = rcp_approx(1.f - alpha);
                    T = T * inv_1ma;
                    u = alpha * T;
                    const float t = cdot - S;
                    S = fmaf(alpha, t, S);
                    w = G * fmaf(nTb, inv_1ma, t * T);
                }
                // Append the splat's row: lane l holds pixel l's (u, w).
                urow[0] = u;
                urow[ROWS * kRowPitch] = w;
                urow += kRowPitch;
                if ((lane & (ROWS - 1)) == nrows) myid = __float_as_int(q1.w);
                if (++nrows == ROWS) {
                    __syncwarp();
                    flush_rows<ROWS>(p, U, W, dpix, ROWS, myid, lane, patch_x0, patch_y0);
                    nrows = 0;
                    urow = U + lane;
                }
            }
        }
    }
    if (nrows > 0) {
        __syncwarp();
        flush_rows<ROWS>(p, U, W, dpix, nrows, myid, lane, patch_x0, patch_y0);
    }
    st.drain(tid);  // every copy was consumed (no early exit); this only retires the cp.async groups
}

template <int CH, int ROWS, int MIN_CTAS, int RS = kRecStages, int GA = kGatherAhead>
static int launch_variant(const RenderBwdParams& p, int tiles, cudaStream_t stream) {
    auto kern = render_bwd_kernel<CH, ROWS, MIN_CTAS, RS, GA>;
    constexpr size_t bytes = BwdSmem<CH, ROWS, RS>::kBytes;
    // The opt-in to more than 48 KB of dynamic shared memory is per function and per device and sticks: set it the first time
    // this instantiation runs on a device, not on every step (small scenes are bound by the host's launch rate).
    static std::atomic<unsigned long long> configured{0};
    int dev = 0;
    GS_CUDA_TRY(cudaGetDevice(&dev));
    const unsigned long long bit = 1ull << (dev & 63);
    if (dev >= 64 || !(configured.load(std::memory_order_acquire) & bit)) {
        GS_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
        if (dev < 64) configured.fetch_or(bit, std::memory_order_release);
    }
    kern<<<tiles, GS_TILE_PIX, bytes, stream>>>(p);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

// Measurement only (gs_debug_render_work): per tile, the number of (warp patch, splat) positions the backward walks
// (every list position below the patch's last blended one - the "pre-cull" count the roofline's evaluated_pairs is built
// from) and how many of them survive the conservative patch cull and are actually evaluated per pixel (x 32 lanes).
__global__ void __launch_bounds__(GS_TILE_PIX) render_work_kernel(const RenderBwdParams p, uint32_t* __restrict__ walked,
                                                                  uint32_t* __restrict__ survivors) {
    __shared__ uint32_t s_cnt[2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x;
    const int tile_x = tile % p.tiles_x, tile_y = tile / p.tiles_x;
    const uint32_t px = tile_x * GS_TILE_X + (warp & 1) * 8 + (lane & 7);
    const uint32_t py = tile_y * GS_TILE_Y + (warp >> 1) * 4 + (lane >> 3);
    const bool inside = px < (uint32_t)p.W && py < (uint32_t)p.H;
    const float patch_x0 = (float)(tile_x * GS_TILE_X + (warp & 1) * 8), patch_y0 = (float)(tile_y * GS_TILE_Y + (warp >> 1) * 4);
    if (tid < 2) s_cnt[tid] = 0;
    __syncthreads();
    const uint2 range = p.ranges[tile];
    const int last = inside ? (int)p.n_contrib[(size_t)p.W * py + px] : 0;
    const int warp_last = min((int)(range.y - range.x), __reduce_max_sync(0xffffffffu, last));
    uint32_t surv = 0;
    for (int base = 0; base < warp_last; base += 32) {
        const int j = base + lane;
        bool may = false;
        if (j < warp_last) {
            const float4* r = reinterpret_cast<const float4*>(p.rec + p.ids[range.x + j]);
            const float4 t0 = __ldg(r), t1 = __ldg(r + 1);
            may = patch_may_contribute(t0.x, t0.y, t0.z, t0.w, t1.x, t1.z, patch_x0, patch_y0);
        }
        surv += (uint32_t)__popc(__ballot_sync(0xffffffffu, may));
    }
    if (lane == 0) {
        atomicAdd(&s_cnt[0], (uint32_t)warp_last);
        atomicAdd(&s_cnt[1], surv);
    }
    __syncthreads();
    if (tid == 0) { walked[tile] = s_cnt[0]; survivors[tile] = s_cnt[1]; }
}

int launch_render_work(const RenderBwdParams& p, int tiles, uint32_t* walked, uint32_t* survivors, cudaStream_t stream) {
    if (tiles <= 0) return GS_OK;
    render_work_kernel<<<tiles, GS_TILE_PIX, 0, stream>>>(p, walked, survivors);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

int launch_render_bwd(const RenderBwdParams& p, int tiles, cudaStream_t stream) {
    if (tiles <= 0) return GS_OK;
    // GS_BWD_VARIANT (measurement). Default since round 2: 128-record chunks, two record stages (55 KB of shared memory), 64
    // registers -> 4 CTAs per SM. After the per-pair instruction cuts the kernel no longer saturates the issue slots at 3 CTAs
    // (72.6 % issue utilisation, few eligible warps): 1.182 -> 1.138 ms. 0 = three record stages at 80 registers / 3 CTAs
    // (round 1), 1 = 64-record chunks at 4 CTAs (1.166), 2 = 8-row flushes at 4 CTAs (1.187), 3 = 256-record chunks (1.182).
    static const int variant = [] { const char* e = getenv("GS_BWD_VARIANT"); return e ? atoi(e) : 4; }();
    switch (variant) {
        case 0: return launch_variant<128, 16, 3>(p, tiles, stream);
        case 1: return launch_variant<64, 16, 4>(p, tiles, stream);
        case 2: return launch_variant<128, 8, 4>(p, tiles, stream);
        case 3: return launch_variant<256, 16, 3, 2, 1>(p, tiles, stream);
        default: return launch_variant<128, 16, 4, 2, 1>(p, tiles, stream);
    }
}

}  // namespace gs
