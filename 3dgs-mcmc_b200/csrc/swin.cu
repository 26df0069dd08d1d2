// Sliding-window gather + rigid deformation + activations in one pass.
//
// train_swin.py renders frame f from the Gaussians alive at f: SwinGaussianModel.get_basic_para_at
// (scene/gaussian_model.py:598-649) indexes an "immature" (trainable) and a "matured" (frozen) parameter set with
// torch.nonzero index lists, concatenates them tensor by tensor (12 index + 11 cat launches), moves every Gaussian by
// its age-based rigid motion (utils/tempo_utils.py:49-84 rigid_deform with rotvec2quat :4-17, rotvec2mat :19-40,
// quat_mul :42-47: ~40 more launches) and applies the activations — then autograd replays all of it backwards.
// Here: one forward kernel writes the rasterizer's five inputs for the K = Ki + Km active Gaussians directly, one
// backward kernel scatters the gradients back to the rows of the immature set (the matured set is detached,
// gaussian_model.py:481-493).
//
//   age = frame - frame_start[row];  rv = rotvec * age;  a = |rv|;  u = rv / max(a, 1e-12)     (F.normalize)
//   xyz' = R (xyz - cen) + cen + v * age,   R = I + sin(a) [u]x + (1 - cos(a)) [u]x^2          (Rodrigues)
//   rot' = normalize( (cos(a/2), u sin(a/2)) (x) rot )                                          (w, x, y, z)
//   opacity = sigmoid(raw), scaling = exp(raw), feature = cat(f_dc, f_rest)
#include "gs_common.cuh"
#include "kernels.cuh"

namespace gs {

constexpr int kSwinThreads = 128;

GS_DEV float3 cross3(float3 a, float3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
GS_DEV float dotf3(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
GS_DEV float3 ld3(const float* p, size_t row) { return {p[3 * row], p[3 * row + 1], p[3 * row + 2]}; }
GS_DEV void st3(float* p, size_t row, float3 v) { p[3 * row] = v.x; p[3 * row + 1] = v.y; p[3 * row + 2] = v.z; }

// Hamilton product, (w, x, y, z) = (.x, .y, .z, .w) of float4  (tempo_utils.py:42-47)
GS_DEV float4 quat_mul(float4 q1, float4 q2) {
    const float w1 = q1.x, x1 = q1.y, y1 = q1.z, z1 = q1.w, w2 = q2.x, x2 = q2.y, y2 = q2.z, z2 = q2.w;
    return make_float4(w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                       w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2);
}

struct SwinRow {   // which set and row an active Gaussian comes from
    const SwinSet* set;
    size_t row;
};

__global__ void __launch_bounds__(kSwinThreads) swin_gather_fwd_kernel(const __grid_constant__ SwinParams p) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    const int K = p.Ki + p.Km;
    const int lane = threadIdx.x & 31;
    const bool valid = k < K;
    const bool imm = valid && k < p.Ki;
    const SwinSet& S = imm ? p.imm : p.mat;
    size_t row = 0;
    if (valid) row = (size_t)(imm ? p.imm_idx[k] : p.mat_idx[k - p.Ki]);
    if (valid) {
        const float age = p.frame - S.start[row];
        float3 xyz = ld3(S.xyz, row);
        float4 rot = reinterpret_cast<const float4*>(S.rot)[row];
        if (p.deform) {
            const float3 v = ld3(S.v, row), rvec = ld3(S.rotvec, row), cen = ld3(S.cen, row);
            const float3 rv = {rvec.x * age, rvec.y * age, rvec.z * age};
            const float a = sqrtf(dotf3(rv, rv));
            const float inv = 1.f / fmaxf(a, 1e-12f);
            const float3 u = {rv.x * inv, rv.y * inv, rv.z * inv};
            const float s = sinf(a), c = cosf(a), sh = sinf(0.5f * a), ch = cosf(0.5f * a);
            const float3 pt = {xyz.x - cen.x, xyz.y - cen.y, xyz.z - cen.z};
            const float3 up = cross3(u, pt), uup = cross3(u, up);
            xyz = {pt.x + s * up.x + (1.f - c) * uup.x + (cen.x + v.x * age), pt.y + s * up.y + (1.f - c) * uup.y + (cen.y + v.y * age),
                   pt.z + s * up.z + (1.f - c) * uup.z + (cen.z + v.z * age)};
            rot = quat_mul(make_float4(ch, u.x * sh, u.y * sh, u.z * sh), rot);
        }
        const float n = fmaxf(sqrtf(rot.x * rot.x + rot.y * rot.y + rot.z * rot.z + rot.w * rot.w), 1e-12f);
        st3(p.out_xyz, (size_t)k, xyz);
        reinterpret_cast<float4*>(p.out_rot)[k] = make_float4(rot.x / n, rot.y / n, rot.z / n, rot.w / n);
        p.out_opacity[k] = 1.0f / (1.0f + expf(-S.opacity[row]));
        const float3 sc = ld3(S.scaling, row);
        st3(p.out_scaling, (size_t)k, {expf(sc.x), expf(sc.y), expf(sc.z)});
    }
    // ---- features: the warp copies its 32 rows cooperatively (f_dc 3 floats + f_rest 3 (M - 1) floats -> 3 M floats) ----
    const int Mrow = 3 * p.M;
    const int warp_first = k - lane;
    for (int g = 0; g < 32; ++g) {
        if (warp_first + g >= K) break;
        const bool g_imm = warp_first + g < p.Ki;
        const size_t grow = (size_t)__shfl_sync(0xffffffffu, (long long)row, g);
        const SwinSet& G = g_imm ? p.imm : p.mat;
        float* dst = p.out_feat + (size_t)(warp_first + g) * Mrow;
        for (int e = lane; e < Mrow; e += 32) dst[e] = e < 3 ? G.dc[3 * grow + e] : G.rest[(size_t)(Mrow - 3) * grow + (e - 3)];
    }
}

__global__ void __launch_bounds__(kSwinThreads) swin_gather_bwd_kernel(const __grid_constant__ SwinParams p, const __grid_constant__ SwinGrads g) {
    // only the immature rows receive gradients (k < Ki); rows are unique, so plain stores into the zero-filled tensors
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const bool valid = k < p.Ki;
    const SwinSet& S = p.imm;
    size_t row = 0;
    if (valid) row = (size_t)p.imm_idx[k];
    if (valid) {
        const float age = p.frame - S.start[row];
        const float3 gx = g.g_xyz ? ld3(g.g_xyz, (size_t)k) : make_float3(0.f, 0.f, 0.f);
        const float4 gn = g.g_rot ? reinterpret_cast<const float4*>(g.g_rot)[k] : make_float4(0.f, 0.f, 0.f, 0.f);
        const float4 r = reinterpret_cast<const float4*>(S.rot)[row];
        // recompute the forward quantities
        float3 u = {0.f, 0.f, 0.f}, rv = {0.f, 0.f, 0.f};
        float a = 0.f, inv = 0.f, s = 0.f, c = 1.f, sh = 0.f, ch = 1.f;
        float4 q = make_float4(1.f, 0.f, 0.f, 0.f), rr = r;
        float3 pt = {0.f, 0.f, 0.f};
        if (p.deform) {
            const float3 rvec = ld3(S.rotvec, row), cen = ld3(S.cen, row), xyz = ld3(S.xyz, row);
            rv = {rvec.x * age, rvec.y * age, rvec.z * age};
            a = sqrtf(dotf3(rv, rv));
            inv = 1.f / fmaxf(a, 1e-12f);
            u = {rv.x * inv, rv.y * inv, rv.z * inv};
            s = sinf(a); c = cosf(a); sh = sinf(0.5f * a); ch = cosf(0.5f * a);
            pt = {xyz.x - cen.x, xyz.y - cen.y, xyz.z - cen.z};
            q = make_float4(ch, u.x * sh, u.y * sh, u.z * sh);
            rr = quat_mul(q, r);
        }
        // normalize backward: n = rr / max(|rr|, eps)
        const float norm = sqrtf(rr.x * rr.x + rr.y * rr.y + rr.z * rr.z + rr.w * rr.w);
        const float ninv = 1.f / fmaxf(norm, 1e-12f);
        float4 h;
        if (norm > 1e-12f) {
            const float4 nh = make_float4(rr.x * ninv, rr.y * ninv, rr.z * ninv, rr.w * ninv);
            const float d = nh.x * gn.x + nh.y * gn.y + nh.z * gn.z + nh.w * gn.w;
            h = make_float4((gn.x - nh.x * d) * ninv, (gn.y - nh.y * d) * ninv, (gn.z - nh.z * d) * ninv, (gn.w - nh.w * d) * ninv);
        } else {
            h = make_float4(gn.x * ninv, gn.y * ninv, gn.z * ninv, gn.w * ninv);
        }
        float3 d_xyz = gx, d_v = {0.f, 0.f, 0.f}, d_rvec = {0.f, 0.f, 0.f}, d_cen = {0.f, 0.f, 0.f};
        float4 d_rot = h;
        if (p.deform) {
            // position path: xyz' = pt + s (u x pt) + (1 - c) (u x (u x pt)) + cen + v age
            const float3 ug = cross3(u, gx), uug = cross3(u, ug);
            const float3 d_pt = {gx.x - s * ug.x + (1.f - c) * uug.x, gx.y - s * ug.y + (1.f - c) * uug.y, gx.z - s * ug.z + (1.f - c) * uug.z};
            d_xyz = d_pt;
            d_cen = {gx.x - d_pt.x, gx.y - d_pt.y, gx.z - d_pt.z};
            d_v = {gx.x * age, gx.y * age, gx.z * age};
            const float3 up = cross3(u, pt), uup = cross3(u, up);
            float d_s = dotf3(gx, up);
            float d_c = -dotf3(gx, uup);
            const float3 pxg = cross3(pt, gx);
            const float gu = dotf3(gx, u), gp = dotf3(gx, pt), upd = dotf3(u, pt);
            float3 d_u = {s * pxg.x + (1.f - c) * (gx.x * upd + pt.x * gu - 2.f * u.x * gp),
                          s * pxg.y + (1.f - c) * (gx.y * upd + pt.y * gu - 2.f * u.y * gp),
                          s * pxg.z + (1.f - c) * (gx.z * upd + pt.z * gu - 2.f * u.z * gp)};
            // quaternion path: rr = q (x) r
            const float w1 = q.x, x1 = q.y, y1 = q.z, z1 = q.w, w2 = r.x, x2 = r.y, y2 = r.z, z2 = r.w;
            const float dq_w = h.x * w2 + h.y * x2 + h.z * y2 + h.w * z2;
            const float dq_x = -h.x * x2 + h.y * w2 - h.z * z2 + h.w * y2;
            const float dq_y = -h.x * y2 + h.y * z2 + h.z * w2 - h.w * x2;
            const float dq_z = -h.x * z2 - h.y * y2 + h.z * x2 + h.w * w2;
            d_rot = make_float4(h.x * w1 + h.y * x1 + h.z * y1 + h.w * z1, -h.x * x1 + h.y * w1 + h.z * z1 - h.w * y1,
                                -h.x * y1 - h.y * z1 + h.z * w1 + h.w * x1, -h.x * z1 + h.y * y1 - h.z * x1 + h.w * w1);
            const float d_ch = dq_w;
            const float d_sh = u.x * dq_x + u.y * dq_y + u.z * dq_z;
            d_u.x += sh * dq_x; d_u.y += sh * dq_y; d_u.z += sh * dq_z;
            // angle and axis back to the rotation vector
            const float d_a = d_s * c - d_c * s - 0.5f * d_ch * sh + 0.5f * d_sh * ch;
            float3 d_rv;
            if (a > 1e-12f) {
                const float ud = dotf3(u, d_u);
                d_rv = {(d_u.x - u.x * ud) * inv + d_a * u.x, (d_u.y - u.y * ud) * inv + d_a * u.y, (d_u.z - u.z * ud) * inv + d_a * u.z};
            } else {   // F.normalize clamps the norm (a plain scale by 1/eps); torch.norm has zero gradient at 0
                d_rv = {d_u.x * inv, d_u.y * inv, d_u.z * inv};
            }
            d_rvec = {d_rv.x * age, d_rv.y * age, d_rv.z * age};
        }
        st3(g.d_xyz, row, d_xyz);
        reinterpret_cast<float4*>(g.d_rot)[row] = d_rot;
        st3(g.d_v, row, d_v);
        st3(g.d_rotvec, row, d_rvec);
        st3(g.d_cen, row, d_cen);
        const float o = 1.0f / (1.0f + expf(-S.opacity[row]));
        g.d_opacity[row] = (g.g_opacity ? g.g_opacity[k] : 0.f) * o * (1.f - o);
        const float3 sc = ld3(S.scaling, row);
        const float3 gs_ = g.g_scaling ? ld3(g.g_scaling, (size_t)k) : make_float3(0.f, 0.f, 0.f);
        st3(g.d_scaling, row, {gs_.x * expf(sc.x), gs_.y * expf(sc.y), gs_.z * expf(sc.z)});
    }
    if (g.g_feat) {
        const int Mrow = 3 * p.M;
        const int warp_first = k - lane;
        for (int j = 0; j < 32; ++j) {
            if (warp_first + j >= p.Ki) break;
            const size_t grow = (size_t)__shfl_sync(0xffffffffu, (long long)row, j);
            const float* src = g.g_feat + (size_t)(warp_first + j) * Mrow;
            for (int e = lane; e < Mrow; e += 32) {
                if (e < 3) g.d_dc[3 * grow + e] = src[e];
                else g.d_rest[(size_t)(Mrow - 3) * grow + (e - 3)] = src[e];
            }
        }
    }
}

int launch_swin_gather_fwd(const SwinParams& p, cudaStream_t stream) {
    const int K = p.Ki + p.Km;
    if (K <= 0) return GS_OK;
    swin_gather_fwd_kernel<<<(K + kSwinThreads - 1) / kSwinThreads, kSwinThreads, 0, stream>>>(p);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

int launch_swin_gather_bwd(const SwinParams& p, const SwinGrads& g, cudaStream_t stream) {
    if (p.Ki <= 0) return GS_OK;
    swin_gather_bwd_kernel<<<(p.Ki + kSwinThreads - 1) / kSwinThreads, kSwinThreads, 0, stream>>>(p, g);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

}  // namespace gs
