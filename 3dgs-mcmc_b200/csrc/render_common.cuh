// Shared pieces of the forward / backward tile renderers.
#pragma once
#include "gs_common.cuh"

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
GS_DEV bool patch_may_contribute(const float mx, const float my, const float A, const float B, const float C,
                                 const float thr, const float X0, const float Y0) {
    if (!(A > 0.0f) || !(C > 0.0f)) return true;
    const float hx = (mx - X0) + 1e-3f, lx = (mx - X0) - (kPatchW + 1e-3f);
    const float hy = (my - Y0) + 1e-3f, ly = (my - Y0) - (kPatchH + 1e-3f);
    const bool inside = (lx <= 0.0f) && (hx >= 0.0f) && (ly <= 0.0f) && (hy >= 0.0f);
    const float nBrC = -__fdividef(B, C), nBrA = -__fdividef(B, A);
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

}  // namespace gs
