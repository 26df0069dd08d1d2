// Per-Gaussian backward for sm_100a: conic -> 2D covariance -> (3D covariance, view-space mean),
// screen-space mean -> 3D mean, colour -> SH coefficients and view direction, 3D covariance ->
// scale / quaternion, fused into ONE pass over the Gaussians.
//
// The reference runs two kernels here (computeCov2DCUDA then preprocessCUDA,
// cuda_rasterizer/backward.cu:144-274, :346-396 with helpers :20-139, :278-341) on gradient tensors
// that the torch glue zero-fills first (rasterize_points.cu:154-162, 300 B per Gaussian). This kernel
// reads the render-backward accumulators once and writes every output gradient for every Gaussian
// (zeros for culled ones and for SH coefficients above the active degree), so the outputs need no
// memset and dL_dmeans3D is never read back.
//
// The arithmetic is derived from the maths, in matrix form (tests/test_bwd_math.py restates it in float64 numpy and pins it
// against the CPU oracle of the reference's scalar expansions; gradients agree with the reference CUDA to rounding):
//   M = J Wv (2x3),  S = M Sigma M^T + 0.3 I,  conic K = S^-1
//   G_S = -adj(S) G_K adj(S) / (det^2 + 1e-7),  G_K = [[gA, gB], [gB, gC]]   (the tile backward accumulates half of d/dB)
//   G_Sigma = M^T G_S M  (dL_dcov3D: off-diagonals doubled),  G_M = 2 G_S M Sigma,  G_J = G_M Wv^T,  then through J to the mean
//   colour: s_k = <sh_k, g>,  dL/ddir = sum_k s_k grad B_k(dir),  dL/dv = (dL/ddir - d <d, dL/ddir>) / |v|
//   Sigma = A A^T, A = R diag(s):  G_A = 2 G_Sigma A,  dL/ds_j = <G_A[:, j], R[:, j]>,  G_R = G_A diag(s),
//   quaternion (r, u): a = axial(G_R - G_R^T),  dL/dr = 2 <u, a>,  dL/du = 2 [(G_R + G_R^T) u - 2 tr(G_R) u + r a]
// The three contributions to dL_dmeans3D are added in the order covariance path, projection path, view-direction path.
#include "bulk_copy.cuh"
#include "gs_common.cuh"
#include "kernels.cuh"
#include "sh_stage.cuh"
#include "workspace.cuh"

namespace gs {

GS_DEV void store3(float* __restrict__ dst, size_t idx, float a, float b, float c) {
    dst[3 * idx] = a;
    dst[3 * idx + 1] = b;
    dst[3 * idx + 2] = c;
}

// SH backward: returns dL/dmean through the view direction; basis_out[k] receives basis_k(dir) for k < (deg+1)^2 and
// dL_dRGB_out the clamp-masked colour gradient: dL/dsh[k][ch] = basis_out[k] * dL_dRGB_out[ch], stored by the caller.
// kVec: sh is the thread's row in the 16-byte staged tile (sh_stage.cuh), read with LDS.128; otherwise a row of floats in
// shared or global memory. |dir|^2 is spelled out (norm2_fixed): with two inlined instantiations the compiler merges their
// common prologue across blocks and its own FMA contraction of the dot product is lost.
// split_lane >= 0 (kVec only): sh is the warp's pair of split tiles (sh_stage.cuh) and the row is gathered from them.
template <bool kVec>
GS_DEV float3 sh_backward(int deg, const V3 pos, const V3 campos, const float* sh, uint8_t clamp_bits, V3 dL_dRGB,
                          float (&basis_out)[16], V3& dL_dRGB_out, int split_lane = -1) {
    const V3 dir_orig = pos - campos;
    const float norm = sqrt(norm2_fixed(dir_orig));
    const V3 dir = dir_orig / norm;
    const float inv_norm = 1.0f / norm;
    float c48[kVec ? 48 : 1];
    if (kVec) {
        if (split_lane >= 0) sh_split_row_to_regs(sh, split_lane, c48);
        else sh_row_to_regs(reinterpret_cast<const float4*>(sh), 0, c48);
    }
    auto C = [&](int k) {
        return kVec ? V3(c48[kVec ? 3 * k : 0], c48[kVec ? 3 * k + 1 : 0], c48[kVec ? 3 * k + 2 : 0])
                    : V3(sh[3 * k], sh[3 * k + 1], sh[3 * k + 2]);
    };
    // Clamped channels pass no gradient.
    dL_dRGB.x *= (clamp_bits & 1) ? 0.f : 1.f;
    dL_dRGB.y *= (clamp_bits & 2) ? 0.f : 1.f;
    dL_dRGB.z *= (clamp_bits & 4) ? 0.f : 1.f;

    const float x = dir.x, y = dir.y, z = dir.z;
    const float xx = x * x, yy = y * y, zz = z * z;
    const float xy = x * y, yz = y * z, xz = x * z;
    // ---- phase 1: the colour is sum_k B_k(dir) sh_k, so dL/ddir = sum_k <sh_k, g> grad B_k(dir): one dot product per
    //      coefficient, then the gradients of the basis polynomials (reads the coefficients) ----
    auto S = [&](int k) { return dot3(C(k), dL_dRGB); };
    V3 g_dir(0.f, 0.f, 0.f);
    if (deg > 0) {
        // B1 = -c1 y, B2 = c1 z, B3 = -c1 x
        g_dir = V3(-kSH1 * S(3), -kSH1 * S(1), kSH1 * S(2));
        if (deg > 1) {
            // B4 = a xy, B5 = b yz, B6 = c (2zz - xx - yy), B7 = d xz, B8 = e (xx - yy)
            const float s4 = kSH2[0] * S(4), s5 = kSH2[1] * S(5), s6 = kSH2[2] * S(6), s7 = kSH2[3] * S(7), s8 = kSH2[4] * S(8);
            g_dir.x += s4 * y - 2.f * s6 * x + s7 * z + 2.f * s8 * x;
            g_dir.y += s4 * x + s5 * z - 2.f * s6 * y - 2.f * s8 * y;
            g_dir.z += s5 * y + 4.f * s6 * z + s7 * x;
            if (deg > 2) {
                // B9 = a y (3xx - yy), B10 = b xyz, B11 = c y (4zz - xx - yy), B12 = d z (2zz - 3xx - 3yy),
                // B13 = e x (4zz - xx - yy), B14 = f z (xx - yy), B15 = g x (xx - 3yy)
                const float s9 = kSH3[0] * S(9), s10 = kSH3[1] * S(10), s11 = kSH3[2] * S(11), s12 = kSH3[3] * S(12);
                const float s13 = kSH3[4] * S(13), s14 = kSH3[5] * S(14), s15 = kSH3[6] * S(15);
                const float dxy = xx - yy;
                g_dir.x += 6.f * s9 * xy + s10 * yz - 2.f * s11 * xy - 6.f * s12 * xz + s13 * (4.f * zz - 3.f * xx - yy) +
                           2.f * s14 * xz + 3.f * s15 * dxy;
                g_dir.y += 3.f * s9 * dxy + s10 * xz + s11 * (4.f * zz - xx - 3.f * yy) - 6.f * s12 * yz - 2.f * s13 * xy -
                           2.f * s14 * yz - 6.f * s15 * xy;
                g_dir.z += s10 * xy + 8.f * s11 * yz + s12 * (6.f * zz - 3.f * xx - 3.f * yy) + 8.f * s13 * xz + s14 * dxy;
            }
        }
    }
    // through the normalisation dir = v / |v|: the component of the gradient along dir is projected out
    const float along = dot3(dir, g_dir);
    const float3 dmean = {(g_dir.x - dir.x * along) * inv_norm, (g_dir.y - dir.y * along) * inv_norm, (g_dir.z - dir.z * along) * inv_norm};
    // ---- phase 2: dL/dsh[k] = basis_k(dir) * dL/dRGB ----
    dL_dRGB_out = dL_dRGB;
    sh_basis_rn(deg, x, y, z, basis_out);
    return dmean;
}

// Sigma = A A^T with A = R diag(s), s = mod * scale, R the rotation polynomial of the quaternion as given (not normalised):
// gradient w.r.t. s (returned as is, like the reference: backward.cu:318-321) and the quaternion. dcov: the six-vector
// gradient (off-diagonal entries count both symmetric positions).
GS_DEV void cov3d_backward(const V3 scale, float mod, const float4 rot, const float* dcov, float* __restrict__ dL_dscale,
                           float4* __restrict__ dL_drot) {
    const float r = rot.x, x = rot.y, y = rot.z, z = rot.w;
    // rows of R
    const float R[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                           {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                           {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
    const float sv[3] = {mod * scale.x, mod * scale.y, mod * scale.z};
    // full symmetric gradient matrix
    const float G[3][3] = {{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]},
                           {0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
                           {0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}};
    // G_A = 2 G A, A[i][j] = R[i][j] s_j;  dL/ds_j = sum_i G_A[i][j] R[i][j];  G_R[i][j] = G_A[i][j] s_j
    float GR[3][3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        float ds = 0.f;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            const float ga = 2.f * sv[j] * (G[i][0] * R[0][j] + G[i][1] * R[1][j] + G[i][2] * R[2][j]);
            ds += ga * R[i][j];
            GR[i][j] = ga * sv[j];
        }
        dL_dscale[j] = ds;
    }
    // R is quadratic in (r, u): its derivative contracts G_R into the axial vector of its antisymmetric part and the
    // symmetric part applied to u
    const float ax = GR[2][1] - GR[1][2], ay = GR[0][2] - GR[2][0], az = GR[1][0] - GR[0][1];
    const float tr2 = 2.f * (GR[0][0] + GR[1][1] + GR[2][2]);
    const float sx = 2.f * GR[0][0] * x + (GR[0][1] + GR[1][0]) * y + (GR[0][2] + GR[2][0]) * z;
    const float sy = (GR[0][1] + GR[1][0]) * x + 2.f * GR[1][1] * y + (GR[1][2] + GR[2][1]) * z;
    const float sz = (GR[0][2] + GR[2][0]) * x + (GR[1][2] + GR[2][1]) * y + 2.f * GR[2][2] * z;
    float4 dq;
    dq.x = 2.f * (x * ax + y * ay + z * az);
    dq.y = 2.f * (sx - tr2 * x + r * ax);
    dq.z = 2.f * (sy - tr2 * y + r * ay);
    dq.w = 2.f * (sz - tr2 * z + r * az);
    *dL_drot = dq;  // gradient w.r.t. the quaternion as given; normalisation is the caller's (PyTorch's) job
}

constexpr int kPbThreads = 128;
static_assert(kPbThreads == kRecChunk, "one block of the per-Gaussian backward produces one gradient record chunk");
static_assert(kRecFloats <= (kPbThreads / 32) * kShTileFloats, "the record is staged in the (dead) SH tile");

__global__ void __launch_bounds__(kPbThreads) preprocess_bwd_kernel(const PreprocessBwdParams p) {
    pdl_trigger();
    pdl_wait();
    __shared__ __align__(16) float s_sh[kPbThreads / 32][kShTileFloats];
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool valid = idx < p.P;
    const size_t i = (size_t)(valid ? idx : 0);
    const bool visible = valid && p.radii[i] > 0;
    // ---- SH rows in: one coalesced sweep per warp ----
    const int Mrow = 3 * p.M;
    const bool staged = p.shs != nullptr && Mrow <= kShMaxRow;
    const int row_stride = Mrow | 1;
    const int warp_first = idx - lane;
    const int rows_valid = min(32, p.P - warp_first);
    const uint32_t vis_mask = __ballot_sync(0xffffffffu, visible);
    float* tile = &s_sh[warp][0];
    const bool write_dsh = p.dL_dsh != nullptr;   // false: the caller rebuilds dL_dsh from the colour gradients (dp_exchange.cu)
    const int ncoef = (p.D + 1) * (p.D + 1);
    // Degree-3-sized rows (M = 16, the training layout) on 16-byte aligned bases move as 16-byte chunks both ways:
    // cp.async.16 in, LDS.128 into registers, and the gradient rows (basis x masked colour gradient, formed in registers)
    // parked with STS.128 and streamed out as 512-byte warp stores.
    const bool split = p.shs != nullptr && p.shs_rest != nullptr;    // features_dc / features_rest as two tensors (M = 16)
    const bool rows16 = !split && staged && Mrow == 48 && (reinterpret_cast<size_t>(p.shs) & 15) == 0 &&
                        (!write_dsh || (reinterpret_cast<size_t>(p.dL_dsh) & 15) == 0);
    if (split) {
        if (rows_valid > 0 && vis_mask)
            stage_sh_split_load(tile, p.shs + (size_t)warp_first * 3, p.shs_rest + (size_t)warp_first * kShSplitRestRow, rows_valid,
                                vis_mask, lane);
    } else if (staged && rows_valid > 0) {
        float* src = const_cast<float*>(p.shs) + (size_t)warp_first * Mrow;
        if (rows16) stage_sh_rows16_load(reinterpret_cast<float4*>(tile), src, rows_valid, vis_mask, 3 * ncoef, lane);
        else if (Mrow == 48) stage_sh_rows<true, 48>(tile, row_stride, src, Mrow, rows_valid, vis_mask, 0, lane);
        else stage_sh_rows<true, 0>(tile, row_stride, src, Mrow, rows_valid, vis_mask, 0, lane);
        // not waited for yet: the covariance / projection part below runs while the rows arrive
    }

    float4 acc0 = make_float4(0.f, 0.f, 0.f, 0.f), acc1 = acc0;
    float acc2 = 0.f;
    if (visible) {
        acc0 = p.acc_color_opacity[idx];
        acc1 = p.acc_mean_conic[idx];
        acc2 = p.acc_conic_w[idx];
    }
    // Pass-through outputs.
    if (valid) {
        store3(p.dL_dmeans2D, i, acc1.x, acc1.y, 0.f);
        if (p.dL_dcolors) store3(p.dL_dcolors, i, acc0.x, acc0.y, acc0.z);
        if (p.dL_dopacity) p.dL_dopacity[idx] = acc0.w;
    }

    float dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float3 dmean = {0.f, 0.f, 0.f};
    float dscale[3] = {0.f, 0.f, 0.f};
    float4 drot = make_float4(0.f, 0.f, 0.f, 0.f);
    float* my_dsh = (p.shs && !split) ? (staged ? tile + lane * row_stride : (write_dsh ? p.dL_dsh + i * Mrow : nullptr)) : nullptr;
    const float* my_sh = p.shs ? (staged ? tile + lane * row_stride : p.shs + i * Mrow) : nullptr;

    if (visible) {
        const float3 mean = {p.means3D[3 * i], p.means3D[3 * i + 1], p.means3D[3 * i + 2]};
        const float* cov3D = p.cov3D + 6 * i;
        const float* view = p.viewmatrix;

        // ---- conic -> cov2D -> cov3D and view-space mean ----
        // rows of the world -> view rotation (view is the transposed 4x4 the rasterizer takes)
        const V3 w0(view[0], view[4], view[8]), w1(view[1], view[5], view[9]), w2(view[2], view[6], view[10]);
        float3 t = xform_point_4x3(mean, view);
        const float limx = 1.3f * p.tan_fovx, limy = 1.3f * p.tan_fovy;
        const float txtz = t.x / t.z, tytz = t.y / t.z;
        const bool x_free = !(txtz < -limx || txtz > limx), y_free = !(tytz < -limy || tytz > limy);   // clamped axes pass no gradient
        t.x = min(limx, max(-limx, txtz)) * t.z;
        t.y = min(limy, max(-limy, tytz)) * t.z;
        const float iz = 1.f / t.z, iz2 = iz * iz;
        // J = [[j00, 0, j02], [0, j11, j12]],  M = J Wv
        const float j00 = p.focal_x * iz, j02 = -p.focal_x * t.x * iz2, j11 = p.focal_y * iz, j12 = -p.focal_y * t.y * iz2;
        const V3 m0 = j00 * w0 + j02 * w2, m1 = j11 * w1 + j12 * w2;
        // Sigma (symmetric) times the rows of M
        const float c0 = cov3D[0], c1 = cov3D[1], c2 = cov3D[2], c3 = cov3D[3], c4 = cov3D[4], c5 = cov3D[5];
        auto sig = [&](const V3& v) { return V3(c0 * v.x + c1 * v.y + c2 * v.z, c1 * v.x + c3 * v.y + c4 * v.z, c2 * v.x + c4 * v.y + c5 * v.z); };
        const V3 ms0 = sig(m0), ms1 = sig(m1);
        const float a = dot3(m0, ms0) + 0.3f, b = dot3(m0, ms1), c = dot3(m1, ms1) + 0.3f;
        const float det = a * c - b * b;
        const float scale_s = -1.0f / (det * det + 0.0000001f);
        // G_S = -adj(S) G_K adj(S) / (det^2 + eps), adj(S) = [[c, -b], [-b, a]], G_K = [[gA, gB], [gB, gC]]
        const float gA = acc1.z, gB = acc1.w, gC = acc2;
        const float k0 = c * gA - b * gB, k1 = c * gB - b * gC;     // first row of adj(S) G_K
        const float k2 = a * gB - b * gA, k3 = a * gC - b * gB;     // second row
        const float s00 = scale_s * (k0 * c - k1 * b), s01 = scale_s * (k1 * a - k0 * b), s11 = scale_s * (k3 * a - k2 * b);
        // G_Sigma = M^T G_S M = s00 m0 m0^T + s01 (m0 m1^T + m1 m0^T) + s11 m1 m1^T
        const V3 u0 = s00 * m0 + s01 * m1, u1 = s01 * m0 + s11 * m1;   // rows of G_S M
        dcov[0] = m0.x * u0.x + m1.x * u1.x;
        dcov[3] = m0.y * u0.y + m1.y * u1.y;
        dcov[5] = m0.z * u0.z + m1.z * u1.z;
        dcov[1] = 2.f * (m0.x * u0.y + m1.x * u1.y);     // off-diagonal entries appear twice in the symmetric matrix
        dcov[2] = 2.f * (m0.x * u0.z + m1.x * u1.z);
        dcov[4] = 2.f * (m0.y * u0.z + m1.y * u1.z);
        // G_M = 2 G_S M Sigma, G_J = G_M Wv^T (only J's four non-constant entries are needed)
        const V3 gm0 = 2.f * (s00 * ms0 + s01 * ms1), gm1 = 2.f * (s01 * ms0 + s11 * ms1);
        const float gj00 = dot3(gm0, w0), gj02 = dot3(gm0, w2), gj11 = dot3(gm1, w1), gj12 = dot3(gm1, w2);
        // through J(t): j00 = fx / tz, j02 = -fx tx / tz^2 (and the y row alike)
        const float gx = x_free ? -p.focal_x * iz2 * gj02 : 0.f;
        const float gy = y_free ? -p.focal_y * iz2 * gj12 : 0.f;
        const float gz = -iz2 * (p.focal_x * gj00 + p.focal_y * gj11) + 2.f * iz2 * iz * (p.focal_x * t.x * gj02 + p.focal_y * t.y * gj12);
        const V3 dm = gx * w0 + gy * w1 + gz * w2;
        dmean = {dm.x, dm.y, dm.z};

        // ---- screen-space mean -> 3D mean through the perspective division: pix = P_row(axis) . p / (P_row(3) . p) ----
        const float* proj = p.projmatrix;
        const float4 hom = xform_point_4x4(mean, proj);
        const float iw = 1.0f / (hom.w + 0.0000001f);
        const float fx_ = acc1.x * iw, fy_ = acc1.y * iw;                 // gradient per unit of the homogeneous x / y
        const float fw = -(fx_ * hom.x + fy_ * hom.y) * iw;               // and of w
        dmean.x += proj[0] * fx_ + proj[1] * fy_ + proj[3] * fw;
        dmean.y += proj[4] * fx_ + proj[5] * fy_ + proj[7] * fw;
        dmean.z += proj[8] * fx_ + proj[9] * fy_ + proj[11] * fw;

    }
    cp_async_wait_all();
    __syncwarp();
    float basis[16];
    V3 g_rgb(0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < 16; ++k) basis[k] = 0.f;
    if (visible) {
        const float3 mean = {p.means3D[3 * i], p.means3D[3 * i + 1], p.means3D[3 * i + 2]};
        // ---- colour -> SH and view direction ----
        if (p.shs) {
            const V3 campos(p.campos[0], p.campos[1], p.campos[2]);
            const float3 dsh_mean =
                split  ? sh_backward<true>(p.D, V3(mean.x, mean.y, mean.z), campos, tile, p.clamped[idx],
                                           V3(acc0.x, acc0.y, acc0.z), basis, g_rgb, lane)
                : rows16 ? sh_backward<true>(p.D, V3(mean.x, mean.y, mean.z), campos, tile + lane * (4 * kShPitch4), p.clamped[idx],
                                           V3(acc0.x, acc0.y, acc0.z), basis, g_rgb)
                       : sh_backward<false>(p.D, V3(mean.x, mean.y, mean.z), campos, my_sh, p.clamped[idx],
                                            V3(acc0.x, acc0.y, acc0.z), basis, g_rgb);
            if (!rows16 && !split && my_dsh) {
                // every read of the row (shared tile or memory) is above: the gradient row may now overwrite it
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    if (k < ncoef) {
                        my_dsh[3 * k] = basis[k] * g_rgb.x;
                        my_dsh[3 * k + 1] = basis[k] * g_rgb.y;
                        my_dsh[3 * k + 2] = basis[k] * g_rgb.z;
                    }
                }
            }
            dmean.x += dsh_mean.x;
            dmean.y += dsh_mean.y;
            dmean.z += dsh_mean.z;
        }
        // ---- 3D covariance -> scale, rotation ----
        if (p.scales) {
            const V3 sc(p.scales[3 * i], p.scales[3 * i + 1], p.scales[3 * i + 2]);
            const float4 q = load_quat(p.rotations, i, p.rot_aligned != 0);
            cov3d_backward(sc, p.scale_modifier, q, dcov, dscale, &drot);
        }
    }

    if (p.push.n > 0) {
        // ---- view-sharded exchange: this block's chunk record goes straight to every rank's gather buffer ----
        // (the SH tile is dead here: every warp is past sh_backward once the block has met at the first barrier)
        __shared__ int s_cnt[kPbThreads / 32];
        uint32_t* rec = reinterpret_cast<uint32_t*>(&s_sh[0][0]);
        const float vals[kRecWidth] = {g_rgb.x, g_rgb.y, g_rgb.z, dmean.x, dmean.y, dmean.z, acc0.w,
                                       dscale[0], dscale[1], dscale[2], drot.x, drot.y, drot.z, drot.w};
        bool nz = false;
#pragma unroll
        for (int k = 0; k < kRecWidth; ++k) nz |= vals[k] != 0.f;
        nz &= visible;
        const uint32_t m = __ballot_sync(0xffffffffu, nz);
        if (lane == 0) s_cnt[warp] = __popc(m);
        __syncthreads();
        int before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < kPbThreads / 32; ++w) {
            const int n = s_cnt[w];
            before += w < warp ? n : 0;
            total += n;
        }
        if (lane == 0) {
            rec[2 * warp] = m;
            rec[2 * warp + 1] = (uint32_t)before;
        }
        if (nz) {
            float* r = reinterpret_cast<float*>(rec) + kRecHeader + kRecWidth * (before + __popc(m & ((1u << lane) - 1u)));
#pragma unroll
            for (int k = 0; k < kRecWidth; k += 2) *reinterpret_cast<float2*>(r + k) = make_float2(vals[k], vals[k + 1]);
        }
        __syncthreads();
        const int n4 = kRecHeader / 4 + ((kRecWidth * total + 3) >> 2);      // 16-byte pieces: header + packed records
        const size_t base = p.push_offset + (size_t)blockIdx.x * kRecFloats;
        for (int q = threadIdx.x; q < n4; q += kPbThreads) {
            const float4 v = reinterpret_cast<const float4*>(rec)[q];
#pragma unroll 1
            for (int d = 0; d < p.push.n; ++d) {
                float* out = p.push.ptr[d] + base + 4 * q;
                if (p.push_multicast)
                    asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(out), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
                else
                    *reinterpret_cast<float4*>(out) = v;
            }
        }
        return;
    }
    if (valid) {
        store3(p.dL_dmeans3D, i, dmean.x, dmean.y, dmean.z);
        if (p.dL_dcov3D) {
#pragma unroll
            for (int k = 0; k < 6; ++k) p.dL_dcov3D[6 * i + k] = dcov[k];
        }
        store3(p.dL_dscales, i, dscale[0], dscale[1], dscale[2]);
        store_quat(p.dL_drots, i, p.drot_aligned != 0, drot);
    }
    if (split) {
        if (write_dsh) {
            // every lane's reads of the tiles are done (sh_backward pulled the rows into registers): the gradient rows go
            // into the same places and leave as 16-byte chunks of the two gradient tensors
            __syncwarp();
            const float gc[3] = {g_rgb.x, g_rgb.y, g_rgb.z};
#pragma unroll
            for (int k = 0; k < 3; ++k) tile[3 * lane + k] = basis[0] * gc[k];
#pragma unroll
            for (int k = 0; k < kShSplitRestRow; ++k) tile[kShSplitDc + kShSplitRestRow * lane + k] = basis[1 + k / 3] * gc[k % 3];
            __syncwarp();
            if (rows_valid > 0)
                stage_sh_split_store(tile, p.dL_dsh + (size_t)warp_first * 3, p.dL_dsh_rest + (size_t)warp_first * kShSplitRestRow,
                                     rows_valid, lane);
        }
    } else if (rows16) {
        if (write_dsh) {
            float4* tile4 = reinterpret_cast<float4*>(tile);
            const float gc[3] = {g_rgb.x, g_rgb.y, g_rgb.z};
#pragma unroll
            for (int q = 0; q < 12; ++q) {
                float v[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int f = 4 * q + j;
                    v[j] = basis[f / 3] * gc[f % 3];      // basis is zero above the active degree and for culled Gaussians
                }
                tile4[lane * kShPitch4 + q] = make_float4(v[0], v[1], v[2], v[3]);
            }
            __syncwarp();
            if (rows_valid > 0) stage_sh_rows16_store(tile4, p.dL_dsh + (size_t)warp_first * 48, rows_valid, lane);
        }
    } else if (staged) {
        // ---- SH gradient rows out: coalesced, zeros for culled Gaussians and unused coefficients ----
        __syncwarp();
        if (rows_valid > 0 && write_dsh) {
            float* dst = p.dL_dsh + (size_t)warp_first * Mrow;
            if (Mrow == 48) stage_sh_rows<false, 48>(tile, row_stride, dst, Mrow, rows_valid, vis_mask, 3 * ncoef, lane);
            else stage_sh_rows<false, 0>(tile, row_stride, dst, Mrow, rows_valid, vis_mask, 3 * ncoef, lane);
        }
    } else if (my_dsh && valid) {
        // Zero the coefficients the SH backward did not write (all of them for culled Gaussians).
        const int first = visible ? 3 * ncoef : 0;
        for (int k = first; k < Mrow; ++k) my_dsh[k] = 0.f;
    }
}

int launch_preprocess_bwd(const PreprocessBwdParams& p, cudaStream_t stream) {
    if (p.P <= 0) return GS_OK;
    launch_pdl(preprocess_bwd_kernel, dim3((p.P + kPbThreads - 1) / kPbThreads), dim3(kPbThreads), 0, stream, p);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

}  // namespace gs
