// Coalesced staging of spherical-harmonics rows between global and shared memory, shared by the forward and
// backward per-Gaussian kernels.
#pragma once
#include "bulk_copy.cuh"
#include "gs_common.cuh"

namespace gs {

constexpr int kShMaxRow = 48;  // floats per SH row that the shared staging supports (M <= 16)

// Coalesced staging of the warp's 32 SH rows (row r = Gaussian warp_first + r) between global memory and a
// [32][row_stride] shared tile whose odd stride makes both the warp-wide copies and the per-thread row walks
// bank-conflict free. A thread reading / writing its own 192-byte row straight from global memory touches 32
// different lines per warp instruction; this way every warp instruction moves whole 128-byte lines.
// kMrow > 0: compile-time row length (fully unrolled, loads issued as cp.async so that all of them are in flight
// at once); kMrow == 0: run-time row length.
template <bool kLoad, int kMrow>
GS_DEV void stage_sh_rows(float* tile, int row_stride, float* gmem /* first row of the warp */, int Mrow_rt, int rows_valid,
                          uint32_t row_mask, int ncoef3, int lane) {
    const int Mrow = kMrow > 0 ? kMrow : Mrow_rt;
    if (kMrow > 0) {
#pragma unroll 12
        for (int it = 0; it < kMrow; ++it) {
            const int f = it * 32 + lane;
            const int row = f / (kMrow > 0 ? kMrow : 1), col = f - row * Mrow;
            if (row < rows_valid) {
                const bool on = (row_mask >> row) & 1u;
                if (kLoad) {
                    if (on && (ncoef3 <= 0 || col < ncoef3))
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(tile + row * row_stride + col)),
                                     "l"(gmem + (size_t)row * Mrow + col)
                                     : "memory");
                } else {
                    gmem[(size_t)row * Mrow + col] = (on && col < ncoef3) ? tile[row * row_stride + col] : 0.f;
                }
            }
        }
        return;  // loads: the caller waits (cp_async_wait_all) before reading the tile
    }
    int row = 0, col = lane;
    while (col >= Mrow) { col -= Mrow; ++row; }
    for (int it = 0; it < Mrow; ++it) {
        if (row < rows_valid) {
            const bool on = (row_mask >> row) & 1u;
            if (kLoad) {
                if (on && (ncoef3 <= 0 || col < ncoef3)) tile[row * row_stride + col] = gmem[(size_t)row * Mrow + col];
            } else {
                // rows of culled Gaussians and coefficients above the active degree are zero
                gmem[(size_t)row * Mrow + col] = (on && col < ncoef3) ? tile[row * row_stride + col] : 0.f;
            }
        }
        col += 32;
        while (col >= Mrow) { col -= Mrow; ++row; }
    }
}

// ---- 16-byte variant for degree-3-sized rows (48 floats = twelve 16-byte chunks, row base 16-byte aligned) ----
// Tile layout [32][kShPitch4] float4: the 208-byte pitch makes a thread's STS.128 / LDS.128 on its own row conflict
// free (13 r + q is distinct mod 8 over any 8 consecutive lanes) while the warp-wide sweeps move twelve fully
// coalesced 512-byte lines instead of forty-eight 128-byte ones: a quarter of the copy instructions and index math.
constexpr int kShPitch4 = 13;
constexpr int kShTileFloats = 32 * 4 * kShPitch4;   // per warp; also covers the scalar layout (32 x 49)

// Starts the copies of the rows in row_mask (only the chunks holding the first ncoef3 floats); the caller waits with
// cp_async_wait_all() + __syncwarp() before reading.
GS_DEV void stage_sh_rows16_load(float4* tile4, const float* gmem /* first row of the warp */, int rows_valid,
                                 uint32_t row_mask, int ncoef3, int lane) {
    const int nchunk = (ncoef3 + 3) >> 2;
    const float4* src = reinterpret_cast<const float4*>(gmem);
#pragma unroll
    for (int it = 0; it < 12; ++it) {
        const int c = it * 32 + lane;
        const int r = c / 12, q = c - 12 * r;
        if (r < rows_valid && ((row_mask >> r) & 1u) && q < nchunk)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(tile4 + r * kShPitch4 + q)), "l"(src + c)
                         : "memory");
    }
}

// The thread's own row into registers (every index static, so the array never touches local memory).
GS_DEV void sh_row_to_regs(const float4* tile4, int lane, float* c48) {
#pragma unroll
    for (int q = 0; q < 12; ++q) {
        const float4 t = tile4[lane * kShPitch4 + q];
        c48[4 * q] = t.x; c48[4 * q + 1] = t.y; c48[4 * q + 2] = t.z; c48[4 * q + 3] = t.w;
    }
}

// Rows parked by their threads (tile4[lane * kShPitch4 + q]) streamed out as coalesced 512-byte warp stores.
GS_DEV void stage_sh_rows16_store(const float4* tile4, float* gmem /* first row of the warp */, int rows_valid, int lane) {
    float4* dst = reinterpret_cast<float4*>(gmem);
    const int chunks = rows_valid * 12;
#pragma unroll
    for (int it = 0; it < 12; ++it) {
        const int c = it * 32 + lane;
        if (c < chunks) {
            const int r = c / 12;
            __stcs(dst + c, tile4[r * kShPitch4 + (c - 12 * r)]);
        }
    }
}

// ---- split SH tensors: features_dc [P][1][3] and features_rest [P][15][3] (scene/gaussian_model_static.py:139-144) ----
// The model keeps the two apart and the reference concatenates them every iteration. Their rows (12 B and 180 B) are not
// 16-byte aligned, but a warp's 32 rows are contiguous in each tensor and 16-byte sized (384 B = 24 chunks, 5760 B = 360
// chunks), so they are copied as they lie into two linear tiles - dc at floats [0, 96), rest at [96, 1536) of the warp's
// tile - with cp.async.16, and a thread walks its rows with scalar LDS / STS at strides 3 and 45 floats (both odd: conflict
// free). The same instruction count as the 16-byte path of the concatenated layout, no copy of the 576 MB tensor.
constexpr int kShSplitDc = 96;            // floats of the dc tile (32 rows x 3)
constexpr int kShSplitRestRow = 45;
static_assert(kShSplitDc + 32 * kShSplitRestRow <= kShTileFloats, "split tiles fit the warp's staging tile");

// Starts the copies of every 16-byte chunk that holds part of a row in row_mask; the caller waits with
// cp_async_wait_all() + __syncwarp(). A chunk that would run past the last row of the tensor is copied float by float.
GS_DEV void stage_sh_split_load(float* tile, const float* dc /* first row of the warp */, const float* rest, int rows_valid,
                                uint32_t row_mask, int lane) {
#pragma unroll
    for (int it = 0; it < 12; ++it) {
        const int c = it * 32 + lane;                     // chunk of the warp's 24 + 360
        const bool is_dc = c < 24;
        const int f0 = 4 * (is_dc ? c : c - 24);          // first float of the chunk within its tensor block
        const int rowlen = is_dc ? 3 : kShSplitRestRow;
        const int r0 = f0 / rowlen, r1 = (f0 + 3) / rowlen;
        if (r0 >= rows_valid) continue;
        const bool wanted = ((row_mask >> r0) & 1u) || (r1 < 32 && ((row_mask >> r1) & 1u));
        if (!wanted) continue;
        const float* src = (is_dc ? dc : rest) + f0;
        float* dst = tile + (is_dc ? 0 : kShSplitDc) + f0;
        const int floats_left = rows_valid * rowlen - f0;
        if (floats_left >= 4) {
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
        } else {
            for (int k = 0; k < floats_left; ++k)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst + k)), "l"(src + k) : "memory");
        }
    }
}

// The thread's own rows from the split tiles into registers, in the concatenated order (dc first).
GS_DEV void sh_split_row_to_regs(const float* tile, int lane, float* c48) {
#pragma unroll
    for (int k = 0; k < 3; ++k) c48[k] = tile[3 * lane + k];
#pragma unroll
    for (int k = 0; k < kShSplitRestRow; ++k) c48[3 + k] = tile[kShSplitDc + kShSplitRestRow * lane + k];
}

// Gradient rows parked in the split tiles (thread `lane`: dc at tile[3 lane ...], rest at tile[96 + 45 lane ...]) streamed out
// to the two gradient tensors as coalesced 16-byte stores; the ragged end of the last warp goes float by float.
GS_DEV void stage_sh_split_store(const float* tile, float* dc /* first row of the warp */, float* rest, int rows_valid, int lane) {
#pragma unroll
    for (int it = 0; it < 12; ++it) {
        const int c = it * 32 + lane;
        const bool is_dc = c < 24;
        const int f0 = 4 * (is_dc ? c : c - 24);
        const int rowlen = is_dc ? 3 : kShSplitRestRow;
        const int floats_left = rows_valid * rowlen - f0;
        if (floats_left <= 0) continue;
        const float* src = tile + (is_dc ? 0 : kShSplitDc) + f0;
        float* dst = (is_dc ? dc : rest) + f0;
        if (floats_left >= 4) {
            __stcs(reinterpret_cast<float4*>(dst), *reinterpret_cast<const float4*>(src));
        } else {
            for (int k = 0; k < floats_left; ++k) dst[k] = src[k];
        }
    }
}

}  // namespace gs
