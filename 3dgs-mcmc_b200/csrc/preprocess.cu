// Per-Gaussian forward preprocessing for sm_100a: frustum cull, clip-space projection, 3D covariance,
// EWA 2D covariance, conic / radius / tile rectangle and degree-0..3 spherical-harmonics colour,
// fused in one HBM-bound pass, plus the stand-alone frustum test.
//
// Behaviour follows the reference's preprocessCUDA and helpers
// (cuda_rasterizer/forward.cu:155-256, :118-152, :74-113, :20-71; auxiliary.h:139-164) so that
// radii / means2D / depths / conics come out bit-identical; the structure is ours: one thread per
// Gaussian with 128-bit read-only loads of the SH block, output zero-fill of radii / is_used /
// tiles_touched folded in (no separate memsets), 3 clamp flags packed in one byte.
#include "gs_common.cuh"
#include "kernels.cuh"
#include "sh_stage.cuh"
#include "workspace.cuh"

namespace gs {

// Sigma = (S R)^T (S R), S = diag(mod * scale), R from the quaternion as given (not normalised).
GS_DEV void cov3d_from_scale_rot(const V3 scale, float mod, const float4 rot, float* cov3D) {
    const float sx = mod * scale.x, sy = mod * scale.y, sz = mod * scale.z;
    const float r = rot.x, x = rot.y, y = rot.z, z = rot.w;
    const M3 R(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
               2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
               2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
    M3 S(sx, 0.f, 0.f, 0.f, sy, 0.f, 0.f, 0.f, sz);
    const M3 Mx = S * R;
    const M3 Sigma = transpose(Mx) * Mx;
    cov3D[0] = Sigma[0][0];
    cov3D[1] = Sigma[0][1];
    cov3D[2] = Sigma[0][2];
    cov3D[3] = Sigma[1][1];
    cov3D[4] = Sigma[1][2];
    cov3D[5] = Sigma[2][2];
}

// EWA splatting: cov2D = (W J)^T Vrk^T (W J) + 0.3 I (upper 2x2).
GS_DEV float3 cov2d_ewa(const float3& mean, float focal_x, float focal_y, float tan_fovx, float tan_fovy,
                        const float* cov3D, const float* __restrict__ view) {
    float3 t = xform_point_4x3(mean, view);
    const float limx = 1.3f * tan_fovx;
    const float limy = 1.3f * tan_fovy;
    const float txtz = t.x / t.z;
    const float tytz = t.y / t.z;
    t.x = min(limx, max(-limx, txtz)) * t.z;
    t.y = min(limy, max(-limy, tytz)) * t.z;

    const M3 J(focal_x / t.z, 0.0f, -(focal_x * t.x) / (t.z * t.z),
               0.0f, focal_y / t.z, -(focal_y * t.y) / (t.z * t.z),
               0.f, 0.f, 0.f);
    const M3 Wm(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]);
    const M3 T = Wm * J;
    const M3 Vrk(cov3D[0], cov3D[1], cov3D[2], cov3D[1], cov3D[3], cov3D[4], cov3D[2], cov3D[4], cov3D[5]);
    M3 cov = transpose(T) * transpose(Vrk) * T;
    cov[0][0] += 0.3f;
    cov[1][1] += 0.3f;
    return {cov[0][0], cov[0][1], cov[1][1]};
}

// Evaluate SH colour. `sh` points at this Gaussian's [M][3] block.
// Not inlined on purpose: bit-exact parity of the colours depends on which multiplies the compiler contracts
// into FMAs, and that choice must not drift with edits to the caller.
// kVec: sh is the thread's row in the 16-byte staged tile (sh_stage.cuh); it is pulled into registers with twelve
// LDS.128 (chunks above the active degree hold stale shared memory and are never used).
// kSplit (with kVec): sh is the warp's pair of split tiles (features_dc / features_rest staged as they lie) and the row is
// gathered from them instead; the arithmetic below is the same expression tree for both layouts (the bit-identity is
// tested: tests/test_gpu_dp_exchange.py::test_split_sh_tensors_equal_the_concatenated_layout).
template <bool kVec, bool kSplit = false>
__device__ __noinline__ V3 sh_to_rgb(int deg, const V3 pos, const V3 campos, const float* __restrict__ sh, uint8_t& clamp_bits,
                                     int split_lane = 0) {
    V3 dir = pos - campos;
    dir = dir / sqrt(dot3(dir, dir));
    float c48[kVec ? 48 : 1];
    if (kVec) {
        if (kSplit) {
            sh_split_row_to_regs(sh, split_lane, c48);
        } else {
#pragma unroll
            for (int q = 0; q < 12; ++q) {
                const float4 t = reinterpret_cast<const float4*>(sh)[q];
                c48[kVec ? 4 * q : 0] = t.x; c48[kVec ? 4 * q + 1 : 0] = t.y; c48[kVec ? 4 * q + 2 : 0] = t.z; c48[kVec ? 4 * q + 3 : 0] = t.w;
            }
        }
    }
    auto C = [&](int k) {
        return kVec ? V3(c48[kVec ? 3 * k : 0], c48[kVec ? 3 * k + 1 : 0], c48[kVec ? 3 * k + 2 : 0])
                    : V3(sh[3 * k], sh[3 * k + 1], sh[3 * k + 2]);
    };

    // explicitly rounded: with the row in registers the compiler duplicates this product into the degree > 0 path, where
    // ptxas would fuse it with the subtraction that follows (the reference's build keeps a plain multiply here)
    const V3 c0 = C(0);
    V3 result(__fmul_rn(kSH0, c0.x), __fmul_rn(kSH0, c0.y), __fmul_rn(kSH0, c0.z));
    if (deg > 0) {
        const float x = dir.x, y = dir.y, z = dir.z;
        result = result - kSH1 * y * C(1) + kSH1 * z * C(2) - kSH1 * x * C(3);
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z;
            const float xy = x * y, yz = y * z, xz = x * z;
            result = result + kSH2[0] * xy * C(4) + kSH2[1] * yz * C(5) + kSH2[2] * (2.0f * zz - xx - yy) * C(6) +
                     kSH2[3] * xz * C(7) + kSH2[4] * (xx - yy) * C(8);
            if (deg > 2) {
                result = result + kSH3[0] * y * (3.0f * xx - yy) * C(9) + kSH3[1] * xy * z * C(10) +
                         kSH3[2] * y * (4.0f * zz - xx - yy) * C(11) +
                         kSH3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * C(12) +
                         kSH3[4] * x * (4.0f * zz - xx - yy) * C(13) + kSH3[5] * z * (xx - yy) * C(14) +
                         kSH3[6] * x * (xx - 3.0f * yy) * C(15);
            }
        }
    }
    result.x += 0.5f;
    result.y += 0.5f;
    result.z += 0.5f;
    clamp_bits = (uint8_t)((result.x < 0 ? 1 : 0) | (result.y < 0 ? 2 : 0) | (result.z < 0 ? 4 : 0));
    return V3(fmaxf(result.x, 0.0f), fmaxf(result.y, 0.0f), fmaxf(result.z, 0.0f));
}

constexpr int kPfThreads = 128;

// One thread per Gaussian, in two phases around the arrival of the SH rows:
//   phase 1  cull, projection, covariance, conic, radius, tile rectangle (everything that does not need colour);
//   phase 2  SH colour, per-Gaussian outputs and the 48-byte render record.
// The warp's SH rows are copied to shared memory with coalesced cp.async sweeps (sh_stage.cuh) issued right after the
// frustum test, so the copies overlap phase 1; a thread reading its own 192-byte row straight from global memory
// would touch 32 different lines per load instruction (the reference's access pattern, forward.cu:29).
// Every per-Gaussian output is written (defaults for culled ones, replacing three host-side fills).
template <bool kSplit>
__global__ void __launch_bounds__(kPfThreads) preprocess_fwd_kernel(const PreprocessParams p) {
    __shared__ __align__(16) float s_sh[kPfThreads / 32][kShTileFloats];
    __shared__ uint32_t s_tiles, s_vis;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { s_tiles = 0; s_vis = 0; }
    __syncthreads();
    const bool valid = idx < p.P;
    const size_t i = (size_t)(valid ? idx : 0);

    int out_radius = 0;
    uint32_t out_tiles = 0;
    uint2 out_rect = {0u, 0u};
    uint32_t depth_bits = 0xffffffffu;  // sentinel: culled Gaussians are dropped by the first depth-sort pass

    const float3 p_orig = {p.means3D[3 * i], p.means3D[3 * i + 1], p.means3D[3 * i + 2]};
    const float4 p_hom = xform_point_4x4(p_orig, p.projmatrix);
    const float p_w = 1.0f / (p_hom.w + 0.0000001f);
    const float3 p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};
    const float3 p_view = xform_point_4x3(p_orig, p.viewmatrix);

    const bool alive = valid && p_view.z > 0.2f;
    if (valid && !alive && p.prefiltered) {
        printf("Point is filtered although prefiltered is set. This shouldn't happen!");
        __trap();
    }

    // ---- phase 1 ----
    bool visible = false;
    float3 conic = {0.f, 0.f, 0.f};
    float2 point_image = {0.f, 0.f};
    float my_radius = 0.f;
    if (alive) {
        const float* cov3D;
        float cov_local[6];
        if (p.cov3D_precomp != nullptr) {
            cov3D = p.cov3D_precomp + 6 * i;
        } else {
            const V3 s(p.scales[3 * i], p.scales[3 * i + 1], p.scales[3 * i + 2]);
            const float4 q = load_quat(p.rotations, i, p.rot_aligned != 0);
            cov3d_from_scale_rot(s, p.scale_modifier, q, cov_local);
#pragma unroll
            for (int k = 0; k < 6; ++k) p.cov3D[6 * i + k] = cov_local[k];
            cov3D = cov_local;
        }
        const float3 cov = cov2d_ewa(p_orig, p.focal_x, p.focal_y, p.tan_fovx, p.tan_fovy, cov3D, p.viewmatrix);

        const float det = (cov.x * cov.z - cov.y * cov.y);
        if (det != 0.0f) {
            const float det_inv = 1.f / det;
            conic = {cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv};
            const float mid = 0.5f * (cov.x + cov.z);
            const float lambda1 = mid + sqrt(max(0.1f, mid * mid - det));
            const float lambda2 = mid - sqrt(max(0.1f, mid * mid - det));
            my_radius = ceil(3.f * sqrt(max(lambda1, lambda2)));
            point_image = {ndc_to_pix(p_proj.x, p.W), ndc_to_pix(p_proj.y, p.H)};
            uint2 rect_min, rect_max;
            tile_rect(point_image, (int)my_radius, rect_min, rect_max, p.grid);
            const uint32_t ntiles = (rect_max.x - rect_min.x) * (rect_max.y - rect_min.y);
            if (ntiles != 0) {
                visible = true;
                out_radius = (int)my_radius;
                out_tiles = ntiles;
                out_rect = make_uint2(rect_min.x | (rect_min.y << 16), (rect_max.x - rect_min.x) | ((rect_max.y - rect_min.y) << 16));
                depth_bits = __float_as_uint(p_view.z);
            }
        }
    }

    // ---- SH rows of the VISIBLE Gaussians only. Round 1 started these copies right after the frustum test, under phase 1; but
    //      ~20 % of the Gaussians in front of the camera touch no tile (off screen), the kernel is DRAM bound (70 % of peak),
    //      and their 192-byte rows were 13 % of its read traffic. The other resident warps cover the copy latency. ----
    const int Mrow = 3 * p.M;
    const bool use_sh = p.colors_precomp == nullptr;
    const bool staged = use_sh && Mrow <= kShMaxRow;
    const int row_stride = Mrow | 1;
    const int warp_first = idx - lane;
    const int rows_valid = min(32, p.P - warp_first);
    const uint32_t alive_mask = __ballot_sync(0xffffffffu, visible);   // rows to stage
    float* tile = &s_sh[warp][0];
    const int ncoef3 = 3 * (p.D + 1) * (p.D + 1);
    // degree-3-sized rows on a 16-byte aligned base (the training layout) move as 16-byte chunks
    const bool split = kSplit && use_sh;                   // shs = features_dc [P][1][3], shs_rest = features_rest [P][15][3]
    const bool rows16 = !split && staged && Mrow == 48 && (reinterpret_cast<size_t>(p.shs) & 15) == 0;
    if (split) {
        if (alive_mask)
            stage_sh_split_load(tile, p.shs + (size_t)warp_first * 3, p.shs_rest + (size_t)warp_first * kShSplitRestRow, rows_valid,
                                alive_mask, lane);
    } else if (staged && alive_mask) {
        float* src = const_cast<float*>(p.shs) + (size_t)warp_first * Mrow;
        if (rows16) stage_sh_rows16_load(reinterpret_cast<float4*>(tile), src, rows_valid, alive_mask, ncoef3, lane);
        else if (Mrow == 48) stage_sh_rows<true, 48>(tile, row_stride, src, Mrow, rows_valid, alive_mask, ncoef3, lane);
        else stage_sh_rows<true, 0>(tile, row_stride, src, Mrow, rows_valid, alive_mask, ncoef3, lane);
    }

    // ---- the SH rows have landed ----
    cp_async_wait_all();
    __syncwarp();

    // ---- phase 2 ----
    if (visible) {
        float3 col;
        if (use_sh) {
            uint8_t cb;
            const V3 campos(p.campos[0], p.campos[1], p.campos[2]);
            V3 c;
            if (kSplit) {
                c = sh_to_rgb<true, kSplit>(p.D, V3(p_orig.x, p_orig.y, p_orig.z), campos, tile, cb, lane);
            } else if (rows16) {
                c = sh_to_rgb<true>(p.D, V3(p_orig.x, p_orig.y, p_orig.z), campos, tile + lane * (4 * kShPitch4), cb);
            } else {
                const float* my_sh = staged ? tile + lane * row_stride : p.shs + i * Mrow;
                c = sh_to_rgb<false>(p.D, V3(p_orig.x, p_orig.y, p_orig.z), campos, my_sh, cb);
            }
            p.clamped[idx] = cb;
            col = make_float3(c.x, c.y, c.z);
        } else {
            col = make_float3(p.colors_precomp[3 * i], p.colors_precomp[3 * i + 1], p.colors_precomp[3 * i + 2]);
        }
        const float opacity = p.opacities[idx];
        // Render record. thr is a conservative reject threshold: power < thr  ==>  opacity * expf(power) < 1/255
        // (1e-3 slack in the exponent, far above the expf / logf / product rounding errors), so the renderers can
        // skip expf — and whole pixel patches — for such pairs without changing any decision.
        float thr;
        if (opacity > 0.0f) {
            const float lt = logf(1.0f / (255.0f * opacity));
            thr = lt - 0.001f - 1e-6f * fabsf(lt);
        } else {
            thr = opacity == opacity ? 1.0f : -INFINITY;  // non-positive opacity never passes; NaN takes the exact path
        }
        float4* dst = reinterpret_cast<float4*>(p.rec + idx);
        dst[0] = make_float4(point_image.x, point_image.y, conic.x, conic.y);
        dst[1] = make_float4(conic.z, opacity, thr, __int_as_float(idx));
        dst[2] = make_float4(col.x, col.y, col.z, p_view.z);
    }
    if (valid) {
        p.is_used[idx] = 0;
        p.radii[idx] = out_radius;
        p.rect[idx] = out_rect;
        p.depth_keys[idx] = ((uint64_t)depth_bits << 32) | (uint32_t)idx;   // index order: ties in depth stay in index order
    }
    // Instance count R = sum of tiles_touched: one atomic per warp into shared memory, one per block to global.
    const uint32_t wsum = __reduce_add_sync(0xffffffffu, out_tiles);
    if (lane == 0 && wsum) atomicAdd(&s_tiles, wsum);
    // num_rendered[2] counts the visible Gaussians: the depth sort drops the culled ones (sentinel keys) in its first pass
    // and runs its other passes, the scans and the emission table on the visible ones only.
    const uint32_t vis_mask = __ballot_sync(0xffffffffu, visible);
    if (lane == 0 && vis_mask) atomicAdd(&s_vis, (uint32_t)__popc(vis_mask));
    __syncthreads();
    if (threadIdx.x == 0) {
        if (s_tiles) atomicAdd(p.num_rendered, s_tiles);
        if (s_vis) atomicAdd(p.num_rendered + 2, s_vis);
    }
}

__global__ void __launch_bounds__(256) mark_visible_kernel(int P, const float* __restrict__ means3D,
                                                           const float* __restrict__ view, uint8_t* __restrict__ present) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float3 p_orig = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
    const float3 p_view = xform_point_4x3(p_orig, view);
    present[idx] = p_view.z > 0.2f ? 1 : 0;
}

int launch_preprocess_fwd(const PreprocessParams& p, cudaStream_t stream) {
    if (p.P <= 0) return GS_OK;
    if (p.shs_rest && !p.colors_precomp) preprocess_fwd_kernel<true><<<(p.P + kPfThreads - 1) / kPfThreads, kPfThreads, 0, stream>>>(p);
    else preprocess_fwd_kernel<false><<<(p.P + kPfThreads - 1) / kPfThreads, kPfThreads, 0, stream>>>(p);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

int launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t stream) {
    if (P <= 0) return GS_OK;
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, means3D, view, present);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

}  // namespace gs
