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

}  // namespace gs
