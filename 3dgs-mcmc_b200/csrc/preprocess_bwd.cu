// Per-Gaussian backward for sm_100a: conic -> 2D covariance -> (3D covariance, view-space mean),
// screen-space mean -> 3D mean, colour -> SH coefficients and view direction, 3D covariance ->
// scale / quaternion, fused into ONE pass over the Gaussians.
//
// The reference runs two kernels here (computeCov2DCUDA then preprocessCUDA,
// cuda_rasterizer/backward.cu:144-274, :346-396 with helpers :20-139, :278-341) on gradient tensors
// that the torch glue zero-fills first (rasterize_points.cu:154-162, 300 B per Gaussian). This kernel
// reads the render-backward accumulators once and writes every output gradient for every Gaussian
// (zeros for culled ones and for SH coefficients above the active degree), so the outputs need no
// memset and dL_dmeans3D is never read back. Expressions follow the reference so results agree to
// rounding; the three contributions to dL_dmeans3D are added in the reference's order
// (covariance path, projection path, view-direction path).
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
template <bool kVec>
GS_DEV float3 sh_backward(int deg, const V3 pos, const V3 campos, const float* sh, uint8_t clamp_bits, V3 dL_dRGB,
                          float (&basis_out)[16], V3& dL_dRGB_out) {
    const V3 dir_orig = pos - campos;
    const V3 dir = dir_orig / sqrt(norm2_fixed(dir_orig));
    float c48[kVec ? 48 : 1];
    if (kVec) sh_row_to_regs(reinterpret_cast<const float4*>(sh), 0, c48);
    auto C = [&](int k) {
        return kVec ? V3(c48[kVec ? 3 * k : 0], c48[kVec ? 3 * k + 1 : 0], c48[kVec ? 3 * k + 2 : 0])
                    : V3(sh[3 * k], sh[3 * k + 1], sh[3 * k + 2]);
    };
    // Clamped channels pass no gradient.
    dL_dRGB.x *= (clamp_bits & 1) ? 0.f : 1.f;
    dL_dRGB.y *= (clamp_bits & 2) ? 0.f : 1.f;
    dL_dRGB.z *= (clamp_bits & 4) ? 0.f : 1.f;

    V3 dRGBdx(0.f, 0.f, 0.f), dRGBdy(0.f, 0.f, 0.f), dRGBdz(0.f, 0.f, 0.f);
    const float x = dir.x, y = dir.y, z = dir.z;
    const float xx = x * x, yy = y * y, zz = z * z;
    const float xy = x * y, yz = y * z, xz = x * z;
    // ---- phase 1: derivative of the colour w.r.t. the view direction (reads the coefficients) ----
    if (deg > 0) {
        dRGBdx = -kSH1 * C(3);
        dRGBdy = -kSH1 * C(1);
        dRGBdz = kSH1 * C(2);
        if (deg > 1) {
            dRGBdx = dRGBdx + (kSH2[0] * y * C(4) + kSH2[2] * 2.f * -x * C(6) + kSH2[3] * z * C(7) + kSH2[4] * 2.f * x * C(8));
            dRGBdy = dRGBdy + (kSH2[0] * x * C(4) + kSH2[1] * z * C(5) + kSH2[2] * 2.f * -y * C(6) + kSH2[4] * 2.f * -y * C(8));
            dRGBdz = dRGBdz + (kSH2[1] * y * C(5) + kSH2[2] * 2.f * 2.f * z * C(6) + kSH2[3] * x * C(7));
            if (deg > 2) {
                dRGBdx = dRGBdx + (kSH3[0] * C(9) * 3.f * 2.f * xy + kSH3[1] * C(10) * yz + kSH3[2] * C(11) * -2.f * xy +
                                   kSH3[3] * C(12) * -3.f * 2.f * xz + kSH3[4] * C(13) * (-3.f * xx + 4.f * zz - yy) +
                                   kSH3[5] * C(14) * 2.f * xz + kSH3[6] * C(15) * 3.f * (xx - yy));
                dRGBdy = dRGBdy + (kSH3[0] * C(9) * 3.f * (xx - yy) + kSH3[1] * C(10) * xz +
                                   kSH3[2] * C(11) * (-3.f * yy + 4.f * zz - xx) + kSH3[3] * C(12) * -3.f * 2.f * yz +
                                   kSH3[4] * C(13) * -2.f * xy + kSH3[5] * C(14) * -2.f * yz +
                                   kSH3[6] * C(15) * -3.f * 2.f * xy);
                dRGBdz = dRGBdz + (kSH3[1] * C(10) * xy + kSH3[2] * C(11) * 4.f * 2.f * yz +
                                   kSH3[3] * C(12) * 3.f * (2.f * zz - xx - yy) + kSH3[4] * C(13) * 4.f * 2.f * xz +
                                   kSH3[5] * C(14) * (xx - yy));
            }
        }
    }
    const float3 dL_ddir = {dot3(dRGBdx, dL_dRGB), dot3(dRGBdy, dL_dRGB), dot3(dRGBdz, dL_dRGB)};
    const float3 dmean = dnormvdv3(float3{dir_orig.x, dir_orig.y, dir_orig.z}, dL_ddir);
    // ---- phase 2: dL/dsh[k] = basis_k(dir) * dL/dRGB ----
    dL_dRGB_out = dL_dRGB;
    auto W = [&](int k, float basis) { basis_out[k] = basis; };
    W(0, kSH0);
    if (deg > 0) {
        W(1, -kSH1 * y);
        W(2, kSH1 * z);
        W(3, -kSH1 * x);
        if (deg > 1) {
            W(4, kSH2[0] * xy);
            W(5, kSH2[1] * yz);
            W(6, kSH2[2] * (2.f * zz - xx - yy));
            W(7, kSH2[3] * xz);
            W(8, kSH2[4] * (xx - yy));
            if (deg > 2) {
                W(9, kSH3[0] * y * (3.f * xx - yy));
                W(10, kSH3[1] * xy * z);
                W(11, kSH3[2] * y * (4.f * zz - xx - yy));
                W(12, kSH3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
                W(13, kSH3[4] * x * (4.f * zz - xx - yy));
                W(14, kSH3[5] * z * (xx - yy));
                W(15, kSH3[6] * x * (xx - 3.f * yy));
            }
        }
    }
    return dmean;
}

// Sigma = M^T M with M = S R: gradient w.r.t. scale and the (un-normalised) quaternion.
GS_DEV void cov3d_backward(const V3 scale, float mod, const float4 rot, const float* dL_dcov3D, float* __restrict__ dL_dscale,
                           float4* __restrict__ dL_drot) {
    const float r = rot.x, x = rot.y, y = rot.z, z = rot.w;
    const M3 R(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
               2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
               2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
    const V3 s = mod * scale;
    const M3 S(s.x, 0.f, 0.f, 0.f, s.y, 0.f, 0.f, 0.f, s.z);
    const M3 Mx = S * R;
    const M3 dL_dSigma(dL_dcov3D[0], 0.5f * dL_dcov3D[1], 0.5f * dL_dcov3D[2],
                       0.5f * dL_dcov3D[1], dL_dcov3D[3], 0.5f * dL_dcov3D[4],
                       0.5f * dL_dcov3D[2], 0.5f * dL_dcov3D[4], dL_dcov3D[5]);
    const M3 dL_dM = (2.0f * Mx) * dL_dSigma;
    const M3 Rt = transpose(R);
    M3 dL_dMt = transpose(dL_dM);

    dL_dscale[0] = dot3(Rt[0], dL_dMt[0]);
    dL_dscale[1] = dot3(Rt[1], dL_dMt[1]);
    dL_dscale[2] = dot3(Rt[2], dL_dMt[2]);

    dL_dMt[0] = dL_dMt[0] * s.x;
    dL_dMt[1] = dL_dMt[1] * s.y;
    dL_dMt[2] = dL_dMt[2] * s.z;

    float4 dq;
    dq.x = 2 * z * (dL_dMt[0][1] - dL_dMt[1][0]) + 2 * y * (dL_dMt[2][0] - dL_dMt[0][2]) + 2 * x * (dL_dMt[1][2] - dL_dMt[2][1]);
    dq.y = 2 * y * (dL_dMt[1][0] + dL_dMt[0][1]) + 2 * z * (dL_dMt[2][0] + dL_dMt[0][2]) + 2 * r * (dL_dMt[1][2] - dL_dMt[2][1]) -
           4 * x * (dL_dMt[2][2] + dL_dMt[1][1]);
    dq.z = 2 * x * (dL_dMt[1][0] + dL_dMt[0][1]) + 2 * r * (dL_dMt[2][0] - dL_dMt[0][2]) + 2 * z * (dL_dMt[1][2] + dL_dMt[2][1]) -
           4 * y * (dL_dMt[2][2] + dL_dMt[0][0]);
    dq.w = 2 * r * (dL_dMt[0][1] - dL_dMt[1][0]) + 2 * x * (dL_dMt[2][0] + dL_dMt[0][2]) + 2 * y * (dL_dMt[1][2] + dL_dMt[2][1]) -
           4 * z * (dL_dMt[1][1] + dL_dMt[0][0]);
    *dL_drot = dq;  // gradient w.r.t. the quaternion as given; normalisation is the caller's (PyTorch's) job
}

constexpr int kPbThreads = 128;

__global__ void __launch_bounds__(kPbThreads) preprocess_bwd_kernel(const PreprocessBwdParams p) {
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
    const bool rows16 = staged && Mrow == 48 && (reinterpret_cast<size_t>(p.shs) & 15) == 0 &&
                        (!write_dsh || (reinterpret_cast<size_t>(p.dL_dsh) & 15) == 0);
    if (staged && rows_valid > 0) {
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
        store3(p.dL_dcolors, i, acc0.x, acc0.y, acc0.z);
        p.dL_dopacity[idx] = acc0.w;
    }

    float dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float3 dmean = {0.f, 0.f, 0.f};
    float dscale[3] = {0.f, 0.f, 0.f};
    float4 drot = make_float4(0.f, 0.f, 0.f, 0.f);
    float* my_dsh = p.shs ? (staged ? tile + lane * row_stride : (write_dsh ? p.dL_dsh + i * Mrow : nullptr)) : nullptr;
    const float* my_sh = p.shs ? (staged ? tile + lane * row_stride : p.shs + i * Mrow) : nullptr;

    if (visible) {
        const float3 mean = {p.means3D[3 * i], p.means3D[3 * i + 1], p.means3D[3 * i + 2]};
        const float* cov3D = p.cov3D + 6 * i;
        const float* view = p.viewmatrix;
        const float3 dL_dconic = {acc1.z, acc1.w, acc2};

        // ---- conic -> cov2D -> cov3D and view-space mean ----
        float3 t = xform_point_4x3(mean, view);
        const float limx = 1.3f * p.tan_fovx;
        const float limy = 1.3f * p.tan_fovy;
        const float txtz = t.x / t.z;
        const float tytz = t.y / t.z;
        t.x = min(limx, max(-limx, txtz)) * t.z;
        t.y = min(limy, max(-limy, tytz)) * t.z;
        const float x_grad_mul = txtz < -limx || txtz > limx ? 0 : 1;
        const float y_grad_mul = tytz < -limy || tytz > limy ? 0 : 1;
        const float h_x = p.focal_x, h_y = p.focal_y;

        const M3 J(h_x / t.z, 0.0f, -(h_x * t.x) / (t.z * t.z), 0.0f, h_y / t.z, -(h_y * t.y) / (t.z * t.z), 0.f, 0.f, 0.f);
        const M3 Wm(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]);
        const M3 Vrk(cov3D[0], cov3D[1], cov3D[2], cov3D[1], cov3D[3], cov3D[4], cov3D[2], cov3D[4], cov3D[5]);
        const M3 T = Wm * J;
        M3 cov2D = transpose(T) * transpose(Vrk) * T;

        const float a = cov2D[0][0] += 0.3f;
        const float b = cov2D[0][1];
        const float c = cov2D[1][1] += 0.3f;
        const float denom = a * c - b * b;
        float dL_da = 0, dL_db = 0, dL_dc = 0;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        if (denom2inv != 0) {
            dL_da = denom2inv * (-c * c * dL_dconic.x + 2 * b * c * dL_dconic.y + (denom - a * c) * dL_dconic.z);
            dL_dc = denom2inv * (-a * a * dL_dconic.z + 2 * a * b * dL_dconic.y + (denom - a * c) * dL_dconic.x);
            dL_db = denom2inv * 2 * (b * c * dL_dconic.x - (denom + 2 * b * b) * dL_dconic.y + a * b * dL_dconic.z);

            dcov[0] = (T[0][0] * T[0][0] * dL_da + T[0][0] * T[1][0] * dL_db + T[1][0] * T[1][0] * dL_dc);
            dcov[3] = (T[0][1] * T[0][1] * dL_da + T[0][1] * T[1][1] * dL_db + T[1][1] * T[1][1] * dL_dc);
            dcov[5] = (T[0][2] * T[0][2] * dL_da + T[0][2] * T[1][2] * dL_db + T[1][2] * T[1][2] * dL_dc);
            // Off-diagonal entries appear twice in the symmetric matrix.
            dcov[1] = 2 * T[0][0] * T[0][1] * dL_da + (T[0][0] * T[1][1] + T[0][1] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][1] * dL_dc;
            dcov[2] = 2 * T[0][0] * T[0][2] * dL_da + (T[0][0] * T[1][2] + T[0][2] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][2] * dL_dc;
            dcov[4] = 2 * T[0][2] * T[0][1] * dL_da + (T[0][1] * T[1][2] + T[0][2] * T[1][1]) * dL_db + 2 * T[1][1] * T[1][2] * dL_dc;
        }

        const float dL_dT00 = 2 * (T[0][0] * Vrk[0][0] + T[0][1] * Vrk[0][1] + T[0][2] * Vrk[0][2]) * dL_da +
                              (T[1][0] * Vrk[0][0] + T[1][1] * Vrk[0][1] + T[1][2] * Vrk[0][2]) * dL_db;
        const float dL_dT01 = 2 * (T[0][0] * Vrk[1][0] + T[0][1] * Vrk[1][1] + T[0][2] * Vrk[1][2]) * dL_da +
                              (T[1][0] * Vrk[1][0] + T[1][1] * Vrk[1][1] + T[1][2] * Vrk[1][2]) * dL_db;
        const float dL_dT02 = 2 * (T[0][0] * Vrk[2][0] + T[0][1] * Vrk[2][1] + T[0][2] * Vrk[2][2]) * dL_da +
                              (T[1][0] * Vrk[2][0] + T[1][1] * Vrk[2][1] + T[1][2] * Vrk[2][2]) * dL_db;
        const float dL_dT10 = 2 * (T[1][0] * Vrk[0][0] + T[1][1] * Vrk[0][1] + T[1][2] * Vrk[0][2]) * dL_dc +
                              (T[0][0] * Vrk[0][0] + T[0][1] * Vrk[0][1] + T[0][2] * Vrk[0][2]) * dL_db;
        const float dL_dT11 = 2 * (T[1][0] * Vrk[1][0] + T[1][1] * Vrk[1][1] + T[1][2] * Vrk[1][2]) * dL_dc +
                              (T[0][0] * Vrk[1][0] + T[0][1] * Vrk[1][1] + T[0][2] * Vrk[1][2]) * dL_db;
        const float dL_dT12 = 2 * (T[1][0] * Vrk[2][0] + T[1][1] * Vrk[2][1] + T[1][2] * Vrk[2][2]) * dL_dc +
                              (T[0][0] * Vrk[2][0] + T[0][1] * Vrk[2][1] + T[0][2] * Vrk[2][2]) * dL_db;

        const float dL_dJ00 = Wm[0][0] * dL_dT00 + Wm[0][1] * dL_dT01 + Wm[0][2] * dL_dT02;
        const float dL_dJ02 = Wm[2][0] * dL_dT00 + Wm[2][1] * dL_dT01 + Wm[2][2] * dL_dT02;
        const float dL_dJ11 = Wm[1][0] * dL_dT10 + Wm[1][1] * dL_dT11 + Wm[1][2] * dL_dT12;
        const float dL_dJ12 = Wm[2][0] * dL_dT10 + Wm[2][1] * dL_dT11 + Wm[2][2] * dL_dT12;

        const float tz = 1.f / t.z;
        const float tz2 = tz * tz;
        const float tz3 = tz2 * tz;
        const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
        const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
        const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 + (2 * h_y * t.y) * tz3 * dL_dJ12;
        dmean = xform_vec_4x3_T({dL_dtx, dL_dty, dL_dtz}, view);

        // ---- screen-space mean -> 3D mean through the projection ----
        const float* proj = p.projmatrix;
        const float4 m_hom = xform_point_4x4(mean, proj);
        const float m_w = 1.0f / (m_hom.w + 0.0000001f);
        const float mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * m_w * m_w;
        const float mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * m_w * m_w;
        float3 dproj;
        dproj.x = (proj[0] * m_w - proj[3] * mul1) * acc1.x + (proj[1] * m_w - proj[3] * mul2) * acc1.y;
        dproj.y = (proj[4] * m_w - proj[7] * mul1) * acc1.x + (proj[5] * m_w - proj[7] * mul2) * acc1.y;
        dproj.z = (proj[8] * m_w - proj[11] * mul1) * acc1.x + (proj[9] * m_w - proj[11] * mul2) * acc1.y;
        dmean.x += dproj.x;
        dmean.y += dproj.y;
        dmean.z += dproj.z;

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
                rows16 ? sh_backward<true>(p.D, V3(mean.x, mean.y, mean.z), campos, tile + lane * (4 * kShPitch4), p.clamped[idx],
                                           V3(acc0.x, acc0.y, acc0.z), basis, g_rgb)
                       : sh_backward<false>(p.D, V3(mean.x, mean.y, mean.z), campos, my_sh, p.clamped[idx],
                                            V3(acc0.x, acc0.y, acc0.z), basis, g_rgb);
            if (!rows16 && my_dsh) {
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

    if (valid) {
        store3(p.dL_dmeans3D, i, dmean.x, dmean.y, dmean.z);
#pragma unroll
        for (int k = 0; k < 6; ++k) p.dL_dcov3D[6 * i + k] = dcov[k];
        store3(p.dL_dscales, i, dscale[0], dscale[1], dscale[2]);
        store_quat(p.dL_drots, i, p.drot_aligned != 0, drot);
    }
    if (rows16) {
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
    preprocess_bwd_kernel<<<(p.P + kPbThreads - 1) / kPbThreads, kPbThreads, 0, stream>>>(p);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

}  // namespace gs
