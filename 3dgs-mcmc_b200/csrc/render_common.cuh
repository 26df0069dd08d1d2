// Shared pieces of the forward / backward tile renderers.
#pragma once
#include "bulk_copy.cuh"
#include "gs_common.cuh"
#include "workspace.cuh"

namespace gs {

// Each warp of a 16x16-tile CTA owns an 8x4 pixel patch.
constexpr float kPatchW = 7.0f;  // pixel-centre span of the patch in x (8 pixels)
constexpr float kPatchH = 3.0f;  // and in y (4 pixels)

// Conservative patch-level cull. The reference evaluates, per (pixel, splat),
//     power = -0.5 (A dx^2 + C dy^2) - B dx dy,   d = splat_xy - pixel      (cuda_rasterizer/forward.cu:330-338)
// and drops the pair when power > 0 or opacity * exp(power) < 1/255. `thr` (ranges_and_pack_kernel) is a
// lower bound on the power of any pair that survives the alpha test. This returns false only when NO pixel of
// the patch whose first pixel is (X0, Y0) can reach power >= thr: it minimises the quadratic form
// q(d) = A dx^2 + 2 B dx dy + C dy^2 over the patch's d-box exactly (interior, else the best of the four edges)
// and compares -0.5 q_min, widened by a rounding allowance proportional to the largest term either evaluation can
// see, with thr. Anything odd (non-positive-definite conic, NaN) answers true and takes the exact per-pixel path,
// so no blend decision ever changes.
GS_DEV float rcp_approx(const float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// box_may_contribute is the same test for any axis-aligned box of pixel centres [X0, X0 + BW] x [Y0, Y0 + BH]: the warp's
// 8x4 patch (BW = 7, BH = 3) or the CTA's whole 16x16 tile (BW = BH = 15, the first level of the forward's two-level cull).
template <int BW, int BH>
GS_DEV bool box_may_contribute(const float mx, const float my, const float A, const float B, const float C,
                               const float thr, const float X0, const float Y0) {
    if (!(A > 1e-30f) || !(C > 1e-30f)) return true;   // (also keeps the bare reciprocals below in range)
    const float hx = (mx - X0) + 1e-3f, lx = (mx - X0) - ((float)BW + 1e-3f);
    const float hy = (my - Y0) + 1e-3f, ly = (my - Y0) - ((float)BH + 1e-3f);
    const bool inside = (lx <= 0.0f) && (hx >= 0.0f) && (ly <= 0.0f) && (hy >= 0.0f);
    // The clamp points only have to be near the edge minima (q is stationary there: an error d in t costs C d^2,
    // far below the allowance), so bare rcp.approx (no range-scaling guards) is enough.
    const float nBrC = -B * rcp_approx(C), nBrA = -B * rcp_approx(A);
    const float B2 = 2.0f * B;
    // vertical edges dx = e: best dy = clamp(-B e / C)
    float t, q;
    t = fminf(fmaxf(nBrC * lx, ly), hy);
    float qmin = A * lx * lx + t * (B2 * lx + C * t);
    t = fminf(fmaxf(nBrC * hx, ly), hy);
    q = A * hx * hx + t * (B2 * hx + C * t);
    qmin = fminf(qmin, q);
    // horizontal edges dy = e: best dx = clamp(-B e / A)
    t = fminf(fmaxf(nBrA * ly, lx), hx);
    q = C * ly * ly + t * (B2 * ly + A * t);
    qmin = fminf(qmin, q);
    t = fminf(fmaxf(nBrA * hy, lx), hx);
    q = C * hy * hy + t * (B2 * hy + A * t);
    qmin = fminf(qmin, q);
    const float Dx = fmaxf(fabsf(lx), fabsf(hx)), Dy = fmaxf(fabsf(ly), fabsf(hy));
    const float mag = A * Dx * Dx + C * Dy * Dy + fabsf(B2) * Dx * Dy;
    const float pmax = inside ? 0.0f : -0.5f * qmin;
    return !(pmax + 4e-6f * mag + 1e-4f < thr);
}

GS_DEV bool patch_may_contribute(const float mx, const float my, const float A, const float B, const float C,
                                 const float thr, const float X0, const float Y0) {
    return box_may_contribute<7, 3>(mx, my, A, B, C, thr, X0, Y0);
}

// ------------------------------------------------------------------------------------------------
// Shared-memory staging of one tile's splats, used by both renderers.
//   id ring     : kInstStages x (CH + 8) u32 Gaussian ids, filled by 1-D bulk copies (one issuing thread);
//   record ring : kRecStages x CH SplatRec, filled by 16-byte cp.async gathers (threads 0..CH-1,
//                 one record each), two stream chunks ahead of the consumer.
// "Stream chunk" k is list chunk k for the forward and nchunks-1-k for the backward (reversed).
// Barrier protocol: chunk k's ids complete phase (k / kInstStages) & 1 of inst_full[k % kInstStages]
// (armed with the byte count by the issuing thread); its records complete phase (k / kRecStages) & 1 of
// rec_full[k % kRecStages] after CH thread arrivals, each triggered when that thread's copies have landed.
// A stage is refilled only after a CTA-wide barrier that follows its last reader, so phases never alias.
// ------------------------------------------------------------------------------------------------
constexpr int kChunk = 128;
constexpr int kRecStages = 3;
constexpr int kInstStages = 4;
constexpr int kGatherAhead = 2;
constexpr int kInstPad = 8;   // ring slots per stage beyond CH (16-byte alignment slack on both ends)

template <int CH = kChunk, int RS = kRecStages>
GS_DEV void stager_init(uint64_t* inst_full, uint64_t* rec_full) {
#pragma unroll
    for (int s = 0; s < kInstStages; ++s) mbar_init(&inst_full[s], 1);
#pragma unroll
    for (int s = 0; s < RS; ++s) mbar_init(&rec_full[s], CH);
    mbar_fence_init();
}

// CH = records per chunk, RS = record ring depth, GA = chunks the record gather runs ahead of the consumer (GA < RS).
template <int CH = kChunk, int RS = kRecStages, int GA = kGatherAhead>
struct Stager {
    static_assert(GA >= 1 && GA < RS, "the stage being refilled must not be the one being read");
    static_assert(CH % 4 == 0 && CH <= GS_TILE_PIX, "chunk starts must keep the 16-byte phase of the tile's first id");
    const uint32_t* __restrict__ ids;
    const SplatRec* __restrict__ rec;
    uint32_t first;
    int n_total, nchunks;
    bool reversed;
    uint32_t* s_inst;
    uint64_t* inst_full;
    SplatRec* s_rec;
    uint64_t* rec_full;
    int mis;          // first & 3: entries between the previous 16-byte boundary and the slice start
    int inst_issued;  // meaningful in thread 0
    int gathered;     // uniform

    GS_DEV Stager(const uint32_t* ids_, const SplatRec* rec_, uint32_t first_, int n_total_, int nchunks_, bool reversed_,
                  uint32_t* s_inst_, uint64_t* inst_full_, SplatRec* s_rec_, uint64_t* rec_full_)
        : ids(ids_), rec(rec_), first(first_), n_total(n_total_), nchunks(nchunks_), reversed(reversed_), s_inst(s_inst_),
          inst_full(inst_full_), s_rec(s_rec_), rec_full(rec_full_), mis((int)(first_ & 3u)), inst_issued(0), gathered(0) {}

    GS_DEV int list_chunk(int k) const { return reversed ? nchunks - 1 - k : k; }
    GS_DEV int chunk_n(int lc) const { return min(CH, n_total - lc * CH); }

    // Thread 0: bulk-copy the id slice of stream chunk k. Bulk copies move 16-byte units from 16-byte aligned
    // addresses, so the copy starts up to three entries early and may run up to three entries past the slice's
    // end (the id array is padded for that).
    GS_DEV void issue_inst(int k) {
        const int lc = list_chunk(k), n = chunk_n(lc), s = k % kInstStages;
        const size_t start = (size_t)first + (size_t)lc * CH;
        const uint32_t bytes = (uint32_t)((n + mis + 3) & ~3) * 4u;
        mbar_arrive_expect_tx(&inst_full[s], bytes);
        bulk_copy_g2s(s_inst + s * (CH + kInstPad), ids + (start - mis), bytes, &inst_full[s]);
    }

    // Threads 0..CH-1: gather the records of stream chunk k (its ids must have been issued).
    GS_DEV void gather(int tid, int k) {
        if (tid < CH) {
            const int lc = list_chunk(k), n = chunk_n(lc), si = k % kInstStages, sr = k % RS;
            mbar_wait(&inst_full[si], (uint32_t)((k / kInstStages) & 1));
            if (tid < n) {
                const uint32_t id = s_inst[si * (CH + kInstPad) + mis + tid];
                const float4* src = reinterpret_cast<const float4*>(rec + id);
                float4* dst = reinterpret_cast<float4*>(s_rec + sr * CH + tid);
                cp_async_16(dst, src);
                cp_async_16(dst + 1, src + 1);
                cp_async_16(dst + 2, src + 2);
            }
            cp_async_mbar_arrive_noinc(&rec_full[sr]);
        }
    }

    GS_DEV void prologue(int tid) {
        if (tid == 0)
            for (; inst_issued < min(kInstStages, nchunks); ++inst_issued) issue_inst(inst_issued);
        for (; gathered < min(GA, nchunks); ++gathered) gather(tid, gathered);
    }

    // Call at iteration c (consumer about to read stream chunk c), after the CTA-wide barrier at the loop top.
    GS_DEV void advance(int tid, int c) {
        if (c + GA < nchunks) {
            gather(tid, c + GA);
            gathered = c + GA + 1;
        }
        if (tid == 0 && c + kInstStages < nchunks) {
            issue_inst(c + kInstStages);
            inst_issued = c + kInstStages + 1;
        }
    }

    // Before the CTA retires: every asynchronous copy aimed at its shared memory must have landed.
    GS_DEV void drain(int tid) {
        cp_async_wait_all();
        if (tid == 0)
            for (int k = gathered; k < inst_issued; ++k) mbar_wait(&inst_full[k % kInstStages], (uint32_t)((k / kInstStages) & 1));
    }
};

}  // namespace gs
