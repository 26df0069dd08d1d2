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

}  // namespace gs
