/*
 * oracle_cpu.c — plain-C CPU restatement of the reference's rasterizer + MCMC update path.
 *
 * TEST INFRASTRUCTURE ONLY: this file is the checker, never the product. Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 *
 * Parity status: PINNED on the GPU box against the reference's own CUDA code (oracle/_ref, built from
 * /root/reference by oracle/Makefile) by tests/test_gpu_parity.py::test_against_cpu_oracle, and here against the committed
 * golden vectors in tests/golden/ that were generated from that reference run
 * (tests/golden/make_golden.py). The reference ships no tests or golden vectors of its own
 * (SURVEY.md §4).
 *
 * Every function follows a reference function (paths relative to
 * /root/reference/submodules/diff-gaussian-rasterization/):
 *   orc_preprocess        cuda_rasterizer/forward.cu:155-256 (+ :20-71 SH, :74-113 cov2D, :118-152 cov3D,
 *                         auxiliary.h:41-56 ndc2Pix/getRect, :139-164 in_frustum)
 *   orc_bin               cuda_rasterizer/rasterizer_impl.cu:70-111 duplicateWithKeys, :301-309 stable
 *                         radix sort (cub), :116-138 identifyTileRanges
 *   orc_render_fwd        cuda_rasterizer/forward.cu:261-377
 *   orc_render_bwd        cuda_rasterizer/backward.cu:399-557
 *   orc_preprocess_bwd    cuda_rasterizer/backward.cu:144-274 (cov2D), :346-396 (means, SH :20-139, cov3D :278-341)
 *   orc_mark_visible      cuda_rasterizer/rasterizer_impl.cu:54-66
 *   orc_compute_relocation cuda_rasterizer/utils.cu:6-36
 *   orc_add_noise         /root/reference/train_static.py:132-140, utils/general_utils.py:78-110
 *
 * Arithmetic: FP32 throughout like the kernels, compiled with -ffp-contract=off; where SURVEY.md §8
 * (a-bis) documents the fused operations the reference's SASS performs (ndc2Pix in FP64 with an FMA, the
 * blend exponent, colour accumulation) they are written with explicit fma()/fmaf(). Transcendentals are
 * glibc's, not CUDA's, so float results agree with the GPU to a few ulp, not bit for bit; all integer
 * work (keys, order, ranges, counts) is exact given identical inputs.
 */
#include <float.h>
#include <math.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define TILE 16

static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                               -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                               -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};

/* column-major 3x3, m[j][i] = row i of column j (glm semantics) */
typedef struct { float c[3][3]; } mat3;

static mat3 m3_mul(const mat3* a, const mat3* b) {
    mat3 r;
    for (int j = 0; j < 3; ++j)
        for (int i = 0; i < 3; ++i)
            r.c[j][i] = a->c[0][i] * b->c[j][0] + a->c[1][i] * b->c[j][1] + a->c[2][i] * b->c[j][2];
    return r;
}
static mat3 m3_T(const mat3* a) {
    mat3 r;
    for (int j = 0; j < 3; ++j)
        for (int i = 0; i < 3; ++i) r.c[j][i] = a->c[i][j];
    return r;
}
static mat3 m3_cols(float x0, float y0, float z0, float x1, float y1, float z1, float x2, float y2, float z2) {
    mat3 r = {{{x0, y0, z0}, {x1, y1, z1}, {x2, y2, z2}}};
    return r;
}

static void xform4x3(const float* m, const float* p, float* o) {
    o[0] = m[0] * p[0] + m[4] * p[1] + m[8] * p[2] + m[12];
    o[1] = m[1] * p[0] + m[5] * p[1] + m[9] * p[2] + m[13];
    o[2] = m[2] * p[0] + m[6] * p[1] + m[10] * p[2] + m[14];
}
static void xform4x4(const float* m, const float* p, float* o) {
    xform4x3(m, p, o);
    o[3] = m[3] * p[0] + m[7] * p[1] + m[11] * p[2] + m[15];
}

static float ndc2pix(float v, int S) { return (float)(fma((double)v + 1.0, (double)S, -1.0) * 0.5); }

static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

static void get_rect(const float* p, int r, int gx, int gy, int* mn, int* mx) {
    mn[0] = imin(gx, imax(0, (int)((p[0] - r) / TILE)));
    mn[1] = imin(gy, imax(0, (int)((p[1] - r) / TILE)));
    mx[0] = imin(gx, imax(0, (int)((p[0] + r + TILE - 1) / TILE)));
    mx[1] = imin(gy, imax(0, (int)((p[1] + r + TILE - 1) / TILE)));
}

static mat3 rot_from_quat(const float* q) {
    const float r = q[0], x = q[1], y = q[2], z = q[3];
    return m3_cols(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                   2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                   2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
}

static void cov3d(const float* scale, float mod, const float* q, float* out) {
    mat3 S = m3_cols(mod * scale[0], 0, 0, 0, mod * scale[1], 0, 0, 0, mod * scale[2]);
    mat3 R = rot_from_quat(q);
    mat3 M = m3_mul(&S, &R);
    mat3 Mt = m3_T(&M);
    mat3 Sig = m3_mul(&Mt, &M);
    out[0] = Sig.c[0][0]; out[1] = Sig.c[0][1]; out[2] = Sig.c[0][2];
    out[3] = Sig.c[1][1]; out[4] = Sig.c[1][2]; out[5] = Sig.c[2][2];
}

/* shared by forward and backward: clamped t, J, W, T = W*J, Vrk, cov2D (before +0.3) */
typedef struct { float t[3]; float txtz, tytz, limx, limy; mat3 J, W, T, Vrk, cov; } cov2d_ctx;

static void cov2d_setup(const float* mean, float fx, float fy, float tanx, float tany, const float* c3, const float* view,
                        cov2d_ctx* k) {
    xform4x3(view, mean, k->t);
    k->limx = 1.3f * tanx; k->limy = 1.3f * tany;
    k->txtz = k->t[0] / k->t[2]; k->tytz = k->t[1] / k->t[2];
    k->t[0] = fminf(k->limx, fmaxf(-k->limx, k->txtz)) * k->t[2];
    k->t[1] = fminf(k->limy, fmaxf(-k->limy, k->tytz)) * k->t[2];
    const float* t = k->t;
    k->J = m3_cols(fx / t[2], 0.0f, -(fx * t[0]) / (t[2] * t[2]), 0.0f, fy / t[2], -(fy * t[1]) / (t[2] * t[2]), 0, 0, 0);
    k->W = m3_cols(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]);
    k->T = m3_mul(&k->W, &k->J);
    k->Vrk = m3_cols(c3[0], c3[1], c3[2], c3[1], c3[3], c3[4], c3[2], c3[4], c3[5]);
    mat3 Tt = m3_T(&k->T), Vt = m3_T(&k->Vrk);
    mat3 a = m3_mul(&Tt, &Vt);
    k->cov = m3_mul(&a, &k->T);
}

static void sh_color(int deg, const float* pos, const float* campos, const float* sh, float* rgb, uint8_t* clamped3) {
    float d[3] = {pos[0] - campos[0], pos[1] - campos[1], pos[2] - campos[2]};
    float len = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    const float x = d[0] / len, y = d[1] / len, z = d[2] / len;
    for (int c = 0; c < 3; ++c) {
#define S(k) sh[3 * (k) + c]
        float r = SH_C0 * S(0);
        if (deg > 0) {
            r = r - SH_C1 * y * S(1) + SH_C1 * z * S(2) - SH_C1 * x * S(3);
            if (deg > 1) {
                float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                r = r + SH_C2[0] * xy * S(4) + SH_C2[1] * yz * S(5) + SH_C2[2] * (2.0f * zz - xx - yy) * S(6) +
                    SH_C2[3] * xz * S(7) + SH_C2[4] * (xx - yy) * S(8);
                if (deg > 2) {
                    r = r + SH_C3[0] * y * (3.0f * xx - yy) * S(9) + SH_C3[1] * xy * z * S(10) +
                        SH_C3[2] * y * (4.0f * zz - xx - yy) * S(11) + SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * S(12) +
                        SH_C3[4] * x * (4.0f * zz - xx - yy) * S(13) + SH_C3[5] * z * (xx - yy) * S(14) +
                        SH_C3[6] * x * (xx - 3.0f * yy) * S(15);
                }
            }
        }
#undef S
        r += 0.5f;
        clamped3[c] = r < 0;
        rgb[c] = fmaxf(r, 0.0f);
    }
}

/* ---------------------------------------------------------------------------------------------- */
void orc_preprocess(int P, int D, int M, const float* means3D, const float* scales, float mod, const float* rotations,
                    const float* opacities, const float* shs, const float* cov3D_precomp, const float* colors_precomp,
                    const float* view, const float* proj, const float* campos, int W, int H, float tanx, float tany,
                    int* radii, float* means2D, float* depths, float* cov3Ds, float* rgb, float* conic_opacity,
                    uint32_t* tiles_touched, uint8_t* clamped /* [3P] */) {
    const float fy = H / (2.0f * tany), fx = W / (2.0f * tanx);
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; ++idx) {
        radii[idx] = 0;
        tiles_touched[idx] = 0;
        const float* p = means3D + 3 * (size_t)idx;
        float hom[4], pv[3];
        xform4x4(proj, p, hom);
        const float pw = 1.0f / (hom[3] + 0.0000001f);
        const float pp[2] = {hom[0] * pw, hom[1] * pw};
        xform4x3(view, p, pv);
        if (pv[2] <= 0.2f) continue;
        const float* c3;
        if (cov3D_precomp) c3 = cov3D_precomp + 6 * (size_t)idx;
        else { cov3d(scales + 3 * (size_t)idx, mod, rotations + 4 * (size_t)idx, cov3Ds + 6 * (size_t)idx); c3 = cov3Ds + 6 * (size_t)idx; }
        cov2d_ctx k;
        cov2d_setup(p, fx, fy, tanx, tany, c3, view, &k);
        const float ca = k.cov.c[0][0] + 0.3f, cb = k.cov.c[0][1], cc = k.cov.c[1][1] + 0.3f;
        const float det = ca * cc - cb * cb;
        if (det == 0.0f) continue;
        const float di = 1.f / det;
        const float conic[3] = {cc * di, -cb * di, ca * di};
        const float mid = 0.5f * (ca + cc);
        const float l1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
        const float l2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
        const float my_radius = ceilf(3.f * sqrtf(fmaxf(l1, l2)));
        const float pi[2] = {ndc2pix(pp[0], W), ndc2pix(pp[1], H)};
        int mn[2], mx[2];
        get_rect(pi, (int)my_radius, gx, gy, mn, mx);
        if ((mx[0] - mn[0]) * (mx[1] - mn[1]) == 0) continue;
        if (!colors_precomp) sh_color(D, p, campos, shs + (size_t)idx * M * 3, rgb + 3 * (size_t)idx, clamped + 3 * (size_t)idx);
        depths[idx] = pv[2];
        radii[idx] = (int)my_radius;
        means2D[2 * (size_t)idx] = pi[0]; means2D[2 * (size_t)idx + 1] = pi[1];
        float* co = conic_opacity + 4 * (size_t)idx;
        co[0] = conic[0]; co[1] = conic[1]; co[2] = conic[2]; co[3] = opacities[idx];
        tiles_touched[idx] = (uint32_t)((mx[1] - mn[1]) * (mx[0] - mn[0]));
    }
}

void orc_mark_visible(int P, const float* means3D, const float* view, uint8_t* present) {
    for (int i = 0; i < P; ++i) {
        float pv[3];
        xform4x3(view, means3D + 3 * (size_t)i, pv);
        present[i] = pv[2] > 0.2f;
    }
}

/* Keys, stable sort, ranges. offsets = inclusive scan of tiles_touched (caller supplies; R = offsets[P-1]).
 * keys_out / vals_out have R entries; ranges has 2*tiles entries (zeroed here). */
void orc_bin(int P, int W, int H, const float* means2D, const float* depths, const int* radii, const uint32_t* offsets,
             uint32_t R, uint64_t* keys_out, uint32_t* vals_out, uint32_t* ranges) {
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    uint64_t* k0 = (uint64_t*)malloc(sizeof(uint64_t) * (R ? R : 1));
    uint32_t* v0 = (uint32_t*)malloc(sizeof(uint32_t) * (R ? R : 1));
    for (int idx = 0; idx < P; ++idx) {
        if (radii[idx] <= 0) continue;
        uint32_t off = idx == 0 ? 0 : offsets[idx - 1];
        int mn[2], mx[2];
        get_rect(means2D + 2 * (size_t)idx, radii[idx], gx, gy, mn, mx);
        uint32_t db;
        memcpy(&db, depths + idx, 4);
        for (int y = mn[1]; y < mx[1]; ++y)
            for (int x = mn[0]; x < mx[0]; ++x) {
                k0[off] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | db;
                v0[off] = (uint32_t)idx;
                ++off;
            }
    }
    /* stable LSD radix sort, 16-bit digits over the low 48 bits (tile ids < 2^16 at any supported size) */
    uint64_t* ka = k0; uint32_t* va = v0; uint64_t* kb = keys_out; uint32_t* vb = vals_out;
    size_t* cnt = (size_t*)malloc(sizeof(size_t) * 65537);
    for (int pass = 0; pass < 4; ++pass) {
        const int sh = 16 * pass;
        memset(cnt, 0, sizeof(size_t) * 65537);
        for (uint32_t i = 0; i < R; ++i) cnt[((ka[i] >> sh) & 0xffff) + 1]++;
        for (int d = 0; d < 65536; ++d) cnt[d + 1] += cnt[d];
        for (uint32_t i = 0; i < R; ++i) {
            size_t dst = cnt[(ka[i] >> sh) & 0xffff]++;
            kb[dst] = ka[i]; vb[dst] = va[i];
        }
        uint64_t* tk = ka; ka = kb; kb = tk;
        uint32_t* tv = va; va = vb; vb = tv;
    }
    /* 4 passes: result is back in (k0, v0) */
    memcpy(keys_out, ka, sizeof(uint64_t) * R);
    memcpy(vals_out, va, sizeof(uint32_t) * R);
    free(cnt); free(k0); free(v0);
    memset(ranges, 0, sizeof(uint32_t) * 2 * (size_t)gx * gy);
    for (uint32_t i = 0; i < R; ++i) {
        uint32_t cur = (uint32_t)(keys_out[i] >> 32);
        if (i == 0) ranges[2 * cur] = 0;
        else {
            uint32_t prev = (uint32_t)(keys_out[i - 1] >> 32);
            if (cur != prev) { ranges[2 * prev + 1] = i; ranges[2 * cur] = i; }
        }
        if (i == R - 1) ranges[2 * cur + 1] = R;
    }
}

static inline float blend_power(const float* xy, const float* co, float px, float py, float* dx, float* dy) {
    *dx = xy[0] - px; *dy = xy[1] - py;
    const float a = (co[2] * *dy) * *dy;
    const float b = co[0] * *dx;
    const float s = fmaf(*dx, b, a);
    const float c = (co[1] * *dx) * *dy;
    return fmaf(s, -0.5f, -c);
}

/* tile_step > 1 renders only every tile_step-th tile (bounded CPU sample for the benchmark baseline). */
void orc_render_fwd(int W, int H, const uint32_t* ranges, const uint32_t* point_list, const float* means2D,
                    const float* conic_opacity, const float* colors, const float* bg, float* out_color, float* final_T,
                    uint32_t* n_contrib, int* is_used, int tile_step) {
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    if (tile_step < 1) tile_step = 1;
#pragma omp parallel for schedule(dynamic, 4)
    for (int tile = 0; tile < gx * gy; tile += tile_step) {
        const int tx = tile % gx, ty = tile / gx;
        const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
        for (int ly = 0; ly < TILE; ++ly)
            for (int lx = 0; lx < TILE; ++lx) {
                const int px = tx * TILE + lx, py = ty * TILE + ly;
                if (px >= W || py >= H) continue;
                float T = 1.0f, C[3] = {0, 0, 0};
                uint32_t contributor = 0, last = 0;
                for (uint32_t i = r0; i < r1; ++i) {
                    contributor++;
                    const uint32_t id = point_list[i];
                    const float* co = conic_opacity + 4 * (size_t)id;
                    float dx, dy;
                    const float power = blend_power(means2D + 2 * (size_t)id, co, (float)px, (float)py, &dx, &dy);
                    if (power > 0.0f) continue;
                    const float alpha = fminf(0.99f, co[3] * expf(power));
                    if (alpha < 1.0f / 255.0f) continue;
                    const float test_T = T * (1 - alpha);
                    if (test_T < 0.0001f) break;
                    for (int ch = 0; ch < 3; ++ch) C[ch] = fmaf(T, colors[3 * (size_t)id + ch] * alpha, C[ch]);
                    T = test_T;
                    last = contributor;
                    is_used[id] = 1; /* benign race, same as the reference */
                }
                const size_t pix = (size_t)W * py + px;
                final_T[pix] = T;
                n_contrib[pix] = last;
                for (int ch = 0; ch < 3; ++ch) out_color[(size_t)ch * H * W + pix] = fmaf(T, bg[ch], C[ch]);
            }
    }
}

static inline void atomic_addf(float* p, float v) {
#pragma omp atomic
    *p += v;
}

void orc_render_bwd(int W, int H, const uint32_t* ranges, const uint32_t* point_list, const float* bg, const float* means2D,
                    const float* conic_opacity, const float* colors, const float* final_Ts, const uint32_t* n_contrib,
                    const float* dL_dpix, float* dL_dmean2D /*[P,3]*/, float* dL_dconic /*[P,4]*/, float* dL_dopacity,
                    float* dL_dcolors /*[P,3]*/, int tile_step) {
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
    if (tile_step < 1) tile_step = 1;
#pragma omp parallel for schedule(dynamic, 4)
    for (int tile = 0; tile < gx * gy; tile += tile_step) {
        const int tx = tile % gx, ty = tile / gx;
        const uint32_t r0 = ranges[2 * tile];
        for (int ly = 0; ly < TILE; ++ly)
            for (int lx = 0; lx < TILE; ++lx) {
                const int px = tx * TILE + lx, py = ty * TILE + ly;
                if (px >= W || py >= H) continue;
                const size_t pix = (size_t)W * py + px;
                const float T_final = final_Ts[pix];
                float T = T_final;
                const int last_contributor = (int)n_contrib[pix];
                float accum_rec[3] = {0, 0, 0}, last_color[3] = {0, 0, 0}, last_alpha = 0, dpx[3];
                for (int ch = 0; ch < 3; ++ch) dpx[ch] = dL_dpix[(size_t)ch * H * W + pix];
                float bg_dot = 0;
                for (int ch = 0; ch < 3; ++ch) bg_dot += bg[ch] * dpx[ch];
                for (int pos = last_contributor - 1; pos >= 0; --pos) {
                    const uint32_t id = point_list[r0 + (uint32_t)pos];
                    const float* co = conic_opacity + 4 * (size_t)id;
                    float dx, dy;
                    const float power = blend_power(means2D + 2 * (size_t)id, co, (float)px, (float)py, &dx, &dy);
                    if (power > 0.0f) continue;
                    const float G = expf(power);
                    const float alpha = fminf(0.99f, co[3] * G);
                    if (alpha < 1.0f / 255.0f) continue;
                    T = T / (1.f - alpha);
                    const float dchannel_dcolor = alpha * T;
                    float dL_dalpha = 0.0f;
                    for (int ch = 0; ch < 3; ++ch) {
                        const float c = colors[3 * (size_t)id + ch];
                        accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                        last_color[ch] = c;
                        dL_dalpha += (c - accum_rec[ch]) * dpx[ch];
                        atomic_addf(&dL_dcolors[3 * (size_t)id + ch], dchannel_dcolor * dpx[ch]);
                    }
                    dL_dalpha *= T;
                    last_alpha = alpha;
                    dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;
                    const float dL_dG = co[3] * dL_dalpha;
                    const float gdx = G * dx, gdy = G * dy;
                    const float dG_ddelx = -gdx * co[0] - gdy * co[1];
                    const float dG_ddely = -gdy * co[2] - gdx * co[1];
                    atomic_addf(&dL_dmean2D[3 * (size_t)id], dL_dG * dG_ddelx * ddelx_dx);
                    atomic_addf(&dL_dmean2D[3 * (size_t)id + 1], dL_dG * dG_ddely * ddely_dy);
                    atomic_addf(&dL_dconic[4 * (size_t)id], -0.5f * gdx * dx * dL_dG);
                    atomic_addf(&dL_dconic[4 * (size_t)id + 1], -0.5f * gdx * dy * dL_dG);
                    atomic_addf(&dL_dconic[4 * (size_t)id + 3], -0.5f * gdy * dy * dL_dG);
                    atomic_addf(&dL_dopacity[id], G * dL_dalpha);
                }
            }
    }
}

static void dnormvdv(const float* v, const float* dv, float* o) {
    const float s2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
    const float inv = 1.0f / sqrtf(s2 * s2 * s2);
    o[0] = ((+s2 - v[0] * v[0]) * dv[0] - v[1] * v[0] * dv[1] - v[2] * v[0] * dv[2]) * inv;
    o[1] = (-v[0] * v[1] * dv[0] + (s2 - v[1] * v[1]) * dv[1] - v[2] * v[1] * dv[2]) * inv;
    o[2] = (-v[0] * v[2] * dv[0] - v[1] * v[2] * dv[1] + (s2 - v[2] * v[2]) * dv[2]) * inv;
}

/* Inputs dL_dmean2D[P,3], dL_dconic[P,4], dL_dcolors[P,3] come from orc_render_bwd. Outputs must be zeroed by
 * the caller (like the reference's torch::zeros): dL_dmeans3D[P,3], dL_dcov3D[P,6], dL_dsh[P,M,3],
 * dL_dscales[P,3], dL_drots[P,4]. clamped is [3P] bytes. */
void orc_preprocess_bwd(int P, int D, int M, const float* means3D, const int* radii, const float* shs, const uint8_t* clamped,
                        const float* scales, const float* rotations, float mod, const float* cov3Ds, const float* view,
                        const float* proj, const float* campos, int W, int H, float tanx, float tany, const float* dL_dmean2D,
                        const float* dL_dconic, const float* dL_dcolors, float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh,
                        float* dL_dscales, float* dL_drots) {
    const float h_y = H / (2.0f * tany), h_x = W / (2.0f * tanx);
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; ++idx) {
        if (!(radii[idx] > 0)) continue;
        const size_t i = (size_t)idx;
        const float* mean = means3D + 3 * i;
        /* ---- cov2D backward ---- */
        cov2d_ctx k;
        cov2d_setup(mean, h_x, h_y, tanx, tany, cov3Ds + 6 * i, view, &k);
        const float dcx = dL_dconic[4 * i], dcy = dL_dconic[4 * i + 1], dcz = dL_dconic[4 * i + 3];
        const float xg = (k.txtz < -k.limx || k.txtz > k.limx) ? 0.f : 1.f;
        const float yg = (k.tytz < -k.limy || k.tytz > k.limy) ? 0.f : 1.f;
        const float a = k.cov.c[0][0] + 0.3f, b = k.cov.c[0][1], c = k.cov.c[1][1] + 0.3f;
        const float denom = a * c - b * b;
        float dL_da = 0, dL_db = 0, dL_dc = 0;
        const float d2i = 1.0f / ((denom * denom) + 0.0000001f);
        float* dcov = dL_dcov3D + 6 * i;
#define T_(j, r) k.T.c[j][r]
#define V_(j, r) k.Vrk.c[j][r]
#define W_(j, r) k.W.c[j][r]
        if (d2i != 0) {
            dL_da = d2i * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcz);
            dL_dc = d2i * (-a * a * dcz + 2 * a * b * dcy + (denom - a * c) * dcx);
            dL_db = d2i * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
            dcov[0] = (T_(0, 0) * T_(0, 0) * dL_da + T_(0, 0) * T_(1, 0) * dL_db + T_(1, 0) * T_(1, 0) * dL_dc);
            dcov[3] = (T_(0, 1) * T_(0, 1) * dL_da + T_(0, 1) * T_(1, 1) * dL_db + T_(1, 1) * T_(1, 1) * dL_dc);
            dcov[5] = (T_(0, 2) * T_(0, 2) * dL_da + T_(0, 2) * T_(1, 2) * dL_db + T_(1, 2) * T_(1, 2) * dL_dc);
            dcov[1] = 2 * T_(0, 0) * T_(0, 1) * dL_da + (T_(0, 0) * T_(1, 1) + T_(0, 1) * T_(1, 0)) * dL_db + 2 * T_(1, 0) * T_(1, 1) * dL_dc;
            dcov[2] = 2 * T_(0, 0) * T_(0, 2) * dL_da + (T_(0, 0) * T_(1, 2) + T_(0, 2) * T_(1, 0)) * dL_db + 2 * T_(1, 0) * T_(1, 2) * dL_dc;
            dcov[4] = 2 * T_(0, 2) * T_(0, 1) * dL_da + (T_(0, 1) * T_(1, 2) + T_(0, 2) * T_(1, 1)) * dL_db + 2 * T_(1, 1) * T_(1, 2) * dL_dc;
        } else {
            for (int q = 0; q < 6; ++q) dcov[q] = 0;
        }
        float dT[2][3];
        for (int r = 0; r < 3; ++r) {
            dT[0][r] = 2 * (T_(0, 0) * V_(r, 0) + T_(0, 1) * V_(r, 1) + T_(0, 2) * V_(r, 2)) * dL_da +
                       (T_(1, 0) * V_(r, 0) + T_(1, 1) * V_(r, 1) + T_(1, 2) * V_(r, 2)) * dL_db;
            dT[1][r] = 2 * (T_(1, 0) * V_(r, 0) + T_(1, 1) * V_(r, 1) + T_(1, 2) * V_(r, 2)) * dL_dc +
                       (T_(0, 0) * V_(r, 0) + T_(0, 1) * V_(r, 1) + T_(0, 2) * V_(r, 2)) * dL_db;
        }
        const float dJ00 = W_(0, 0) * dT[0][0] + W_(0, 1) * dT[0][1] + W_(0, 2) * dT[0][2];
        const float dJ02 = W_(2, 0) * dT[0][0] + W_(2, 1) * dT[0][1] + W_(2, 2) * dT[0][2];
        const float dJ11 = W_(1, 0) * dT[1][0] + W_(1, 1) * dT[1][1] + W_(1, 2) * dT[1][2];
        const float dJ12 = W_(2, 0) * dT[1][0] + W_(2, 1) * dT[1][1] + W_(2, 2) * dT[1][2];
#undef T_
#undef V_
#undef W_
        const float tz = 1.f / k.t[2], tz2 = tz * tz, tz3 = tz2 * tz;
        const float dtx = xg * -h_x * tz2 * dJ02;
        const float dty = yg * -h_y * tz2 * dJ12;
        const float dtz = -h_x * tz2 * dJ00 - h_y * tz2 * dJ11 + (2 * h_x * k.t[0]) * tz3 * dJ02 + (2 * h_y * k.t[1]) * tz3 * dJ12;
        float dmean[3] = {view[0] * dtx + view[1] * dty + view[2] * dtz, view[4] * dtx + view[5] * dty + view[6] * dtz,
                          view[8] * dtx + view[9] * dty + view[10] * dtz};
        /* ---- projection path ---- */
        float hom[4];
        xform4x4(proj, mean, hom);
        const float mw = 1.0f / (hom[3] + 0.0000001f);
        const float mul1 = (proj[0] * mean[0] + proj[4] * mean[1] + proj[8] * mean[2] + proj[12]) * mw * mw;
        const float mul2 = (proj[1] * mean[0] + proj[5] * mean[1] + proj[9] * mean[2] + proj[13]) * mw * mw;
        const float g2x = dL_dmean2D[3 * i], g2y = dL_dmean2D[3 * i + 1];
        dmean[0] += (proj[0] * mw - proj[3] * mul1) * g2x + (proj[1] * mw - proj[3] * mul2) * g2y;
        dmean[1] += (proj[4] * mw - proj[7] * mul1) * g2x + (proj[5] * mw - proj[7] * mul2) * g2y;
        dmean[2] += (proj[8] * mw - proj[11] * mul1) * g2x + (proj[9] * mw - proj[11] * mul2) * g2y;
        /* ---- SH backward ---- */
        if (shs) {
            const float* sh = shs + i * M * 3;
            float* dsh = dL_dsh + i * M * 3;
            float diro[3] = {mean[0] - campos[0], mean[1] - campos[1], mean[2] - campos[2]};
            const float len = sqrtf(diro[0] * diro[0] + diro[1] * diro[1] + diro[2] * diro[2]);
            const float x = diro[0] / len, y = diro[1] / len, z = diro[2] / len;
            float dRGB[3], ddir[3] = {0, 0, 0};
            for (int ch = 0; ch < 3; ++ch) dRGB[ch] = dL_dcolors[3 * i + ch] * (clamped[3 * i + ch] ? 0.f : 1.f);
            for (int ch = 0; ch < 3; ++ch) {
#define S(q) sh[3 * (q) + ch]
#define DS(q, basis) dsh[3 * (q) + ch] = (basis) * dRGB[ch]
                float dx_ = 0, dy_ = 0, dz_ = 0;
                DS(0, SH_C0);
                if (D > 0) {
                    DS(1, -SH_C1 * y); DS(2, SH_C1 * z); DS(3, -SH_C1 * x);
                    dx_ = -SH_C1 * S(3); dy_ = -SH_C1 * S(1); dz_ = SH_C1 * S(2);
                    if (D > 1) {
                        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                        DS(4, SH_C2[0] * xy); DS(5, SH_C2[1] * yz); DS(6, SH_C2[2] * (2.f * zz - xx - yy));
                        DS(7, SH_C2[3] * xz); DS(8, SH_C2[4] * (xx - yy));
                        dx_ += SH_C2[0] * y * S(4) + SH_C2[2] * 2.f * -x * S(6) + SH_C2[3] * z * S(7) + SH_C2[4] * 2.f * x * S(8);
                        dy_ += SH_C2[0] * x * S(4) + SH_C2[1] * z * S(5) + SH_C2[2] * 2.f * -y * S(6) + SH_C2[4] * 2.f * -y * S(8);
                        dz_ += SH_C2[1] * y * S(5) + SH_C2[2] * 2.f * 2.f * z * S(6) + SH_C2[3] * x * S(7);
                        if (D > 2) {
                            DS(9, SH_C3[0] * y * (3.f * xx - yy)); DS(10, SH_C3[1] * xy * z);
                            DS(11, SH_C3[2] * y * (4.f * zz - xx - yy)); DS(12, SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
                            DS(13, SH_C3[4] * x * (4.f * zz - xx - yy)); DS(14, SH_C3[5] * z * (xx - yy));
                            DS(15, SH_C3[6] * x * (xx - 3.f * yy));
                            dx_ += (SH_C3[0] * S(9) * 3.f * 2.f * xy + SH_C3[1] * S(10) * yz + SH_C3[2] * S(11) * -2.f * xy +
                                    SH_C3[3] * S(12) * -3.f * 2.f * xz + SH_C3[4] * S(13) * (-3.f * xx + 4.f * zz - yy) +
                                    SH_C3[5] * S(14) * 2.f * xz + SH_C3[6] * S(15) * 3.f * (xx - yy));
                            dy_ += (SH_C3[0] * S(9) * 3.f * (xx - yy) + SH_C3[1] * S(10) * xz + SH_C3[2] * S(11) * (-3.f * yy + 4.f * zz - xx) +
                                    SH_C3[3] * S(12) * -3.f * 2.f * yz + SH_C3[4] * S(13) * -2.f * xy + SH_C3[5] * S(14) * -2.f * yz +
                                    SH_C3[6] * S(15) * -3.f * 2.f * xy);
                            dz_ += (SH_C3[1] * S(10) * xy + SH_C3[2] * S(11) * 4.f * 2.f * yz + SH_C3[3] * S(12) * 3.f * (2.f * zz - xx - yy) +
                                    SH_C3[4] * S(13) * 4.f * 2.f * xz + SH_C3[5] * S(14) * (xx - yy));
                        }
                    }
                }
#undef S
#undef DS
                ddir[0] += dx_ * dRGB[ch]; ddir[1] += dy_ * dRGB[ch]; ddir[2] += dz_ * dRGB[ch];
            }
            float dm[3];
            dnormvdv(diro, ddir, dm);
            dmean[0] += dm[0]; dmean[1] += dm[1]; dmean[2] += dm[2];
        }
        dL_dmeans3D[3 * i] = dmean[0]; dL_dmeans3D[3 * i + 1] = dmean[1]; dL_dmeans3D[3 * i + 2] = dmean[2];
        /* ---- cov3D backward ---- */
        if (scales) {
            const float* q = rotations + 4 * i;
            const float r = q[0], x = q[1], y = q[2], z = q[3];
            mat3 R = rot_from_quat(q);
            const float s[3] = {mod * scales[3 * i], mod * scales[3 * i + 1], mod * scales[3 * i + 2]};
            mat3 S = m3_cols(s[0], 0, 0, 0, s[1], 0, 0, 0, s[2]);
            mat3 Mx = m3_mul(&S, &R);
            mat3 dSig = m3_cols(dcov[0], 0.5f * dcov[1], 0.5f * dcov[2], 0.5f * dcov[1], dcov[3], 0.5f * dcov[4],
                                0.5f * dcov[2], 0.5f * dcov[4], dcov[5]);
            mat3 M2 = Mx;
            for (int j = 0; j < 3; ++j) for (int e = 0; e < 3; ++e) M2.c[j][e] = Mx.c[j][e] * 2.0f;
            mat3 dM = m3_mul(&M2, &dSig);
            mat3 Rt = m3_T(&R), dMt = m3_T(&dM);
            for (int j = 0; j < 3; ++j)
                dL_dscales[3 * i + j] = Rt.c[j][0] * dMt.c[j][0] + Rt.c[j][1] * dMt.c[j][1] + Rt.c[j][2] * dMt.c[j][2];
            for (int j = 0; j < 3; ++j) for (int e = 0; e < 3; ++e) dMt.c[j][e] *= s[j];
#define D_(j, e) dMt.c[j][e]
            float* dq = dL_drots + 4 * i;
            dq[0] = 2 * z * (D_(0, 1) - D_(1, 0)) + 2 * y * (D_(2, 0) - D_(0, 2)) + 2 * x * (D_(1, 2) - D_(2, 1));
            dq[1] = 2 * y * (D_(1, 0) + D_(0, 1)) + 2 * z * (D_(2, 0) + D_(0, 2)) + 2 * r * (D_(1, 2) - D_(2, 1)) - 4 * x * (D_(2, 2) + D_(1, 1));
            dq[2] = 2 * x * (D_(1, 0) + D_(0, 1)) + 2 * r * (D_(2, 0) - D_(0, 2)) + 2 * z * (D_(1, 2) + D_(2, 1)) - 4 * y * (D_(2, 2) + D_(0, 0));
            dq[3] = 2 * r * (D_(0, 1) - D_(1, 0)) + 2 * x * (D_(2, 0) + D_(0, 2)) + 2 * y * (D_(1, 2) + D_(2, 1)) - 4 * z * (D_(1, 1) + D_(0, 0));
#undef D_
        }
    }
}

void orc_compute_relocation(int P, const float* opacity_old, const float* scale_old, const int* N, const float* binoms,
                            int n_max, float* opacity_new, float* scale_new) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int idx = 0; idx < P; ++idx) {
        const int N_idx = N[idx];
        float denom_sum = 0.0f;
        opacity_new[idx] = 1.0f - powf(1.0f - opacity_old[idx], 1.0f / N_idx);
        for (int i = 1; i <= N_idx; ++i)
            for (int k = 0; k <= (i - 1); ++k) {
                const float bin_coeff = binoms[(i - 1) * n_max + k];
                const float term = (float)((pow(-1.0, (double)k) / sqrt((double)(k + 1))) * pow((double)opacity_new[idx], (double)(k + 1)));
                denom_sum = fmaf(bin_coeff, term, denom_sum);
            }
        const float coeff = opacity_old[idx] / denom_sum;
        for (int i = 0; i < 3; ++i) scale_new[idx * 3 + i] = coeff * scale_old[idx * 3 + i];
    }
}

/* xyz += (L L^T) (noise * gate * lr_scale), L = R(q/|q|) diag(exp(s)); gate = sigmoid(100 ((1 - sigmoid(o)) - 0.995)). */
void orc_add_noise(int P, float* xyz, const float* scaling_raw, const float* rotation_raw, const float* opacity_raw,
                   const float* noise, float lr_scale) {
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; ++idx) {
        const size_t i = (size_t)idx;
        const float op = 1.0f / (1.0f + expf(-opacity_raw[i]));
        const float gate = 1.0f / (1.0f + expf(-100.0f * ((1.0f - op) - 0.995f)));
        float n[3];
        for (int c = 0; c < 3; ++c) n[c] = noise[3 * i + c] * gate * lr_scale;
        const float* qr = rotation_raw + 4 * i;
        const float norm = sqrtf(qr[0] * qr[0] + qr[1] * qr[1] + qr[2] * qr[2] + qr[3] * qr[3]);
        const float r = qr[0] / norm, x = qr[1] / norm, y = qr[2] / norm, z = qr[3] / norm;
        const float s[3] = {expf(scaling_raw[3 * i]), expf(scaling_raw[3 * i + 1]), expf(scaling_raw[3 * i + 2])};
        const float Rm[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                                {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                                {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
        float L[3][3], Sg[3][3];
        for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) L[a][b] = Rm[a][b] * s[b];
        for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) Sg[a][b] = L[a][0] * L[b][0] + L[a][1] * L[b][1] + L[a][2] * L[b][2];
        for (int a = 0; a < 3; ++a) xyz[3 * i + a] += Sg[a][0] * n[0] + Sg[a][1] * n[1] + Sg[a][2] * n[2];
    }
}

/* OMP_NUM_THREADS=1 is exported by torchrun to every rank; the timed CPU legs of bench.py override it explicitly. */
void orc_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int orc_num_threads(void) {
    int n = 1;
#ifdef _OPENMP
#pragma omp parallel
    {
#pragma omp master
        n = omp_get_num_threads();
    }
#endif
    return n;
}

/* ------------------------------------------------------------------------------------------------------------------
 * distCUDA2: mean of the squared distances to the three nearest other points. Follows
 * submodules/simple-knn/simple_knn.cu:139-184 (updateKBest<3>, boxMeanDist) without its pruning: plain O(P^2).
 * The reference's device code is compiled with nvcc's default contraction, which turns
 * d.x*d.x + d.y*d.y + d.z*d.z into FMUL dy*dy; FFMA dx*dx + .; FFMA dz*dz + . (read off the SASS of
 * `nvcc 12.9 -O3 -arch=sm_100`); this file is built with -ffp-contract=off, so the fused form is spelled out. A point is excluded by position only (coincident points
 * count with distance 0), as in the reference.
 * ------------------------------------------------------------------------------------------------------------------ */
void orc_knn_mean_dist2(int P, const float* pts, float* out) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; ++i) {
        float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
        const float rx = pts[3 * i], ry = pts[3 * i + 1], rz = pts[3 * i + 2];
        for (int j = 0; j < P; ++j) {
            if (j == i) continue;
            const float dx = pts[3 * j] - rx, dy = pts[3 * j + 1] - ry, dz = pts[3 * j + 2] - rz;
            float dist = fmaf(dz, dz, fmaf(dx, dx, dy * dy));
            for (int k = 0; k < 3; ++k) {
                if (best[k] > dist) { const float t = best[k]; best[k] = dist; dist = t; }
            }
        }
        out[i] = (best[0] + best[1] + best[2]) / 3.0f;
    }
}
