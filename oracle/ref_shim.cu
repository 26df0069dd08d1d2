// C-ABI shim around the UNMODIFIED reference CUDA rasterizer.
//
// TEST INFRASTRUCTURE ONLY (oracle/). This file is compiled together with the
// reference's own sources *where they lie* under
//   /root/reference/submodules/diff-gaussian-rasterization/cuda_rasterizer/
// (see oracle/Makefile) into oracle/_ref/libref_dgr.so. It replaces only the
// reference's torch glue (rasterize_points.cu:36-248), which needs a torch C++
// build; the library API it calls is the reference's public one:
//   CudaRasterizer::Rasterizer::forward/backward/markVisible (rasterizer.h:20-87)
//   UTILS::ComputeRelocation                                  (utils.h:8-19)
// The three byte buffers the reference grows through std::function callbacks
// (rasterize_points.cu:28-34) are grow-only cudaMalloc blocks owned by the shim;
// ref_fetch() decodes them with the reference's own fromChunk()
// (rasterizer_impl.cu:155-194) so tests can compare keys / order / ranges bit for bit.
//
// Only tests/, __graft_entry__.smoke() and bench.py's baseline legs may load the
// resulting library. Nothing under 3dgs-mcmc_b200/ does.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <stdexcept>

#include <cuda_runtime.h>

#include "cuda_rasterizer/config.h"
#include "cuda_rasterizer/rasterizer.h"
#include "cuda_rasterizer/rasterizer_impl.h"
#include "cuda_rasterizer/utils.h"

namespace {

struct GrowBuf {
    char* ptr = nullptr;
    size_t cap = 0;
    size_t used = 0;
    char* get(size_t n) {
        if (n > cap) {
            if (ptr) cudaFree(ptr);
            size_t want = n + n / 4 + 4096;
            if (cudaMalloc(&ptr, want) != cudaSuccess) {
                ptr = nullptr; cap = 0;
                throw std::runtime_error("ref_shim: cudaMalloc failed");
            }
            cap = want;
        }
        used = n;
        return ptr;
    }
};

GrowBuf g_geom, g_bin, g_img;
int g_P = 0, g_R = 0, g_W = 0, g_H = 0;

}  // namespace

extern "C" {

int ref_forward(int P, int D, int M, const float* bg, int W, int H,
                const float* means3D, const float* shs, const float* colors_precomp,
                const float* opacities, const float* scales, float scale_modifier,
                const float* rotations, const float* cov3D_precomp,
                const float* viewmatrix, const float* projmatrix, const float* campos,
                float tanfovx, float tanfovy, int prefiltered,
                float* out_color, int* is_used, int* radii, int debug)
{
    try {
        std::function<char*(size_t)> geomF = [](size_t n) { return g_geom.get(n); };
        std::function<char*(size_t)> binF = [](size_t n) { return g_bin.get(n); };
        std::function<char*(size_t)> imgF = [](size_t n) { return g_img.get(n); };
        int R = CudaRasterizer::Rasterizer::forward(
            geomF, binF, imgF, P, D, M, bg, W, H, means3D, shs, colors_precomp, opacities,
            scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos,
            tanfovx, tanfovy, prefiltered != 0, out_color, is_used, radii, debug != 0);
        g_P = P; g_R = R; g_W = W; g_H = H;
        return R;
    } catch (const std::exception& e) {
        fprintf(stderr, "ref_forward: %s\n", e.what());
        return -1;
    }
}

int ref_backward(int P, int D, int M, int R, const float* bg, int W, int H,
                 const float* means3D, const float* shs, const float* colors_precomp,
                 const float* scales, float scale_modifier, const float* rotations,
                 const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                 const float* campos, float tanfovx, float tanfovy, const int* radii,
                 const float* dL_dpix, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                 float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                 float* dL_dscale, float* dL_drot, int debug)
{
    try {
        CudaRasterizer::Rasterizer::backward(
            P, D, M, R, bg, W, H, means3D, shs, colors_precomp, scales, scale_modifier, rotations,
            cov3D_precomp, viewmatrix, projmatrix, campos, tanfovx, tanfovy, radii,
            g_geom.ptr, g_bin.ptr, g_img.ptr, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity,
            dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, debug != 0);
        return 0;
    } catch (const std::exception& e) {
        fprintf(stderr, "ref_backward: %s\n", e.what());
        return -1;
    }
}

int ref_mark_visible(int P, float* means3D, float* viewmatrix, float* projmatrix, bool* present)
{
    CudaRasterizer::Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix, present);
    return (int)cudaGetLastError();
}

int ref_compute_relocation(int P, float* opacity_old, float* scale_old, int* N, float* binoms,
                           int n_max, float* opacity_new, float* scale_new)
{
    UTILS::ComputeRelocation(P, opacity_old, scale_old, N, binoms, n_max, opacity_new, scale_new);
    return (int)cudaGetLastError();
}

// Copy one decoded state array of the last ref_forward() into dst (device memory).
// which: 0 depths f32[P] | 1 clamped bool[3P] | 2 internal_radii i32[P] | 3 means2D f32[2P]
//        4 cov3D f32[6P] | 5 conic_opacity f32[4P] | 6 rgb f32[3P] | 7 tiles_touched u32[P]
//        8 point_offsets u32[P] | 10 point_list u32[R] | 11 point_list_unsorted u32[R]
//        12 point_list_keys u64[R] | 13 point_list_keys_unsorted u64[R]
//        20 accum_alpha f32[WH] | 21 n_contrib u32[WH] | 22 ranges u32[2*tiles]
// Returns the number of bytes of the array (copied only if <= max_bytes), or -1.
long long ref_fetch(int which, void* dst, size_t max_bytes)
{
    using namespace CudaRasterizer;
    const void* src = nullptr;
    size_t n = 0;
    const size_t P = (size_t)g_P, R = (size_t)g_R, N = (size_t)g_W * (size_t)g_H;
    const size_t tiles = (size_t)((g_W + BLOCK_X - 1) / BLOCK_X) * (size_t)((g_H + BLOCK_Y - 1) / BLOCK_Y);
    if (which < 10) {
        if (!g_geom.ptr) return -1;
        char* c = g_geom.ptr;
        GeometryState s = GeometryState::fromChunk(c, P);
        switch (which) {
            case 0: src = s.depths; n = P * 4; break;
            case 1: src = s.clamped; n = P * 3; break;
            case 2: src = s.internal_radii; n = P * 4; break;
            case 3: src = s.means2D; n = P * 8; break;
            case 4: src = s.cov3D; n = P * 24; break;
            case 5: src = s.conic_opacity; n = P * 16; break;
            case 6: src = s.rgb; n = P * 12; break;
            case 7: src = s.tiles_touched; n = P * 4; break;
            case 8: src = s.point_offsets; n = P * 4; break;
            default: return -1;
        }
    } else if (which < 20) {
        if (!g_bin.ptr) return -1;
        char* c = g_bin.ptr;
        BinningState s = BinningState::fromChunk(c, R);
        switch (which) {
            case 10: src = s.point_list; n = R * 4; break;
            case 11: src = s.point_list_unsorted; n = R * 4; break;
            case 12: src = s.point_list_keys; n = R * 8; break;
            case 13: src = s.point_list_keys_unsorted; n = R * 8; break;
            default: return -1;
        }
    } else {
        if (!g_img.ptr) return -1;
        char* c = g_img.ptr;
        ImageState s = ImageState::fromChunk(c, N);
        switch (which) {
            case 20: src = s.accum_alpha; n = N * 4; break;
            case 21: src = s.n_contrib; n = N * 4; break;
            case 22: src = s.ranges; n = tiles * 8; break;
            default: return -1;
        }
    }
    if (dst && n <= max_bytes && n > 0) {
        if (cudaMemcpy(dst, src, n, cudaMemcpyDeviceToDevice) != cudaSuccess) return -1;
    }
    return (long long)n;
}

void ref_release()
{
    if (g_geom.ptr) cudaFree(g_geom.ptr);
    if (g_bin.ptr) cudaFree(g_bin.ptr);
    if (g_img.ptr) cudaFree(g_img.ptr);
    g_geom = GrowBuf(); g_bin = GrowBuf(); g_img = GrowBuf();
}

}  // extern "C"
