// Internal launcher interface between api.cu and the kernel translation units.
#pragma once
#include "gs_common.cuh"
#include "workspace.cuh"

namespace gs {

struct PreprocessParams {
    int P, D, M;
    const float* __restrict__ means3D;
    const float* __restrict__ scales;
    float scale_modifier;
    const float* __restrict__ rotations;
    const float* __restrict__ opacities;
    const float* __restrict__ shs;
    const float* __restrict__ shs_rest;  // non-null: split layout, shs = features_dc [P][1][3], this = features_rest [P][15][3] (M = 16)
    const float* __restrict__ cov3D_precomp;
    const float* __restrict__ colors_precomp;
    const float* __restrict__ viewmatrix;
    const float* __restrict__ projmatrix;
    const float* __restrict__ campos;
    int W, H;
    float tan_fovx, tan_fovy, focal_x, focal_y;
    uint2 grid;
    int prefiltered;
    int rot_aligned;                     // rotations base is 16-byte aligned (float4 loads)
    // outputs
    int* __restrict__ radii;
    int* __restrict__ is_used;
    float* __restrict__ cov3D;
    uint8_t* __restrict__ clamped;
    SplatRec* __restrict__ rec;          // [P] render records (visible Gaussians only)
    uint64_t* __restrict__ depth_keys;   // [P] (depth_bits << 32 | idx); depth_bits = ~0 when culled (dropped by the sort's first pass)
    uint2* __restrict__ rect;            // [P] (x0 | y0 << 16, w | h << 16) tile rectangle, zero when culled
    uint32_t* __restrict__ num_rendered; // [0] += sum of tiles_touched, [2] += visible Gaussians (zeroed by the caller)
};

struct RenderFwdParams {
    const uint2* __restrict__ ranges;
    const uint32_t* __restrict__ ids;    // Gaussian index of every instance in (tile, depth, index) order, padded
    const SplatRec* __restrict__ rec;    // per-Gaussian render records
    int W, H, tiles_x;
    const float* __restrict__ bg;
    float* __restrict__ final_T;
    uint32_t* __restrict__ n_contrib;
    float* __restrict__ out_color;
    int* __restrict__ is_used;
    uint32_t* __restrict__ tile_max_contrib;
    unsigned long long* counters;        // measurement only (gs_debug_forward_work): null on the product path
};

struct RenderBwdParams {
    const uint2* __restrict__ ranges;
    const uint32_t* __restrict__ ids;
    const SplatRec* __restrict__ rec;
    const uint32_t* __restrict__ tile_max_contrib;
    int W, H, tiles_x;
    const float* __restrict__ bg;
    const float* __restrict__ final_T;
    const uint32_t* __restrict__ n_contrib;
    const float* __restrict__ dL_dpix;
    float4* __restrict__ acc_color_opacity;  // [P] (dL/dr, dL/dg, dL/db, dL/dopacity)
    float4* __restrict__ acc_mean_conic;     // [P] (dL/dmean2D.x, .y, dL/dconic.x, .y)
    float* __restrict__ acc_conic_w;         // [P] dL/dconic.w
};

// Up to 8 device pointers to the same symmetric buffer on the ranks of one node (index = rank), or one NVSwitch
// multicast pointer in ptr[0] with n == 1 (dp_exchange.cu).
struct DpPeers {
    float* ptr[8];
    int n;
};

// Gradient records (dp_exchange.cu, "records" layout): per chunk of kRecChunk consecutive Gaussians one record of
// kRecFloats floats = kRecChunk/32 x (u32 mask, u32 set bits before this word), then 14 floats per set bit in index order:
// masked dL/dRGB 3 | dL/dmean3D 3 | dL/dopacity 1 | dL/dscale 3 | dL/drotation 4.
constexpr int kRecChunk = 128;
constexpr int kRecHeader = 2 * (kRecChunk / 32);
constexpr int kRecWidth = 14;
constexpr int kRecFloats = kRecHeader + kRecWidth * kRecChunk;

struct PreprocessBwdParams {
    int P, D, M;
    const float* __restrict__ means3D;
    const int* __restrict__ radii;
    const float* __restrict__ shs;        // null when colours were precomputed
    const float* __restrict__ shs_rest;   // non-null: split layout (shs = features_dc, this = features_rest; M = 16)
    const uint8_t* __restrict__ clamped;
    const float* __restrict__ scales;     // null when cov3D was precomputed
    const float* __restrict__ rotations;
    float scale_modifier;
    int rot_aligned, drot_aligned;        // rotations / dL_drots bases are 16-byte aligned
    const float* __restrict__ cov3D;      // [6P] (precomputed or from the forward workspace)
    const float* __restrict__ viewmatrix;
    const float* __restrict__ projmatrix;
    const float* __restrict__ campos;
    float focal_x, focal_y, tan_fovx, tan_fovy;
    // accumulators from render backward
    const float4* __restrict__ acc_color_opacity;
    const float4* __restrict__ acc_mean_conic;
    const float* __restrict__ acc_conic_w;
    // outputs (API layouts)
    float* __restrict__ dL_dmeans3D;   // [P,3]
    float* __restrict__ dL_dmeans2D;   // [P,3]
    float* __restrict__ dL_dcolors;    // [P,3]
    float* __restrict__ dL_dopacity;   // [P]
    float* __restrict__ dL_dcov3D;     // [P,6]
    float* __restrict__ dL_dsh;        // [P,M,3]; null with shs non-null = factored exchange (dp_exchange.cu): rows not written
    float* __restrict__ dL_dsh_rest;   // split layout: dL_dsh is [P,1,3] and this [P,15,3]
    float* __restrict__ dL_dscales;    // [P,3]
    float* __restrict__ dL_drots;      // [P,4]
    // View-sharded exchange, "records" layout: push.n > 0 = instead of the dense dL_dmeans3D / dL_dopacity / dL_dscales /
    // dL_drots / dL_dsh arrays (then null, like dL_dcolors / dL_dcov3D may be), every block stores its chunk's record at float
    // offset push_offset + block * kRecFloats of each buffer in push (peer pointers, or one multicast pointer).
    DpPeers push;
    size_t push_offset;
    int push_multicast;
};

// Fused L1 + SSIM loss (loss.cu). win = the normalised 11-tap Gaussian window (sigma 1.5) in FP32.
struct LossParams {
    int C, H, W;
    const float* __restrict__ img;   // [C][H][W] render
    const float* __restrict__ gt;    // [C][H][W] ground truth
    float* __restrict__ d_mu;        // [C][H][W] dS/dmu   (forward out, backward in)
    float* __restrict__ d_e11;       // [C][H][W] dS/de11
    float* __restrict__ d_e12;       // [C][H][W] dS/de12
    float* __restrict__ partial;     // [blocks][2] per-block sums of |x - y| and S
    float* __restrict__ dL_dimg;     // [C][H][W] (backward out)
    float win[11];
};

// Sliding-window gather (swin.cu). One parameter set of the swin model; all pointers are row-major FP32.
struct SwinSet {
    const float* xyz;      // [N][3]
    const float* rot;      // [N][4]   raw quaternion (w, x, y, z)
    const float* dc;       // [N][1][3]
    const float* rest;     // [N][M-1][3]
    const float* opacity;  // [N][1]   raw
    const float* scaling;  // [N][3]   raw
    const float* v;        // [N][3]   rigid velocity
    const float* rotvec;   // [N][3]   rigid rotation vector per frame
    const float* cen;      // [N][3]   rotation centre
    const float* start;    // [N]      frame the age counts from
};
struct SwinParams {
    int Ki, Km, M;
    float frame;
    int deform;
    const long long* imm_idx;   // [Ki] rows of the immature set
    const long long* mat_idx;   // [Km] rows of the matured set
    SwinSet imm, mat;
    float *out_xyz, *out_rot, *out_feat, *out_opacity, *out_scaling;   // [K][3], [K][4], [K][M][3], [K][1], [K][3]
};
struct SwinGrads {
    const float *g_xyz, *g_rot, *g_feat, *g_opacity, *g_scaling;       // gradients of the five outputs (any may be null)
    float *d_xyz, *d_rot, *d_dc, *d_rest, *d_opacity, *d_scaling, *d_v, *d_rotvec, *d_cen;   // immature set, zero-filled
};

// One multi-tensor Adam launch (optim.cu).
struct AdamTensor {
    float* param;
    const float* grad;
    float* m;
    float* v;
    unsigned long long n;
    float neg_step_size;   // -(lr / (1 - beta1^step))
    float bc2_sqrt;        // sqrt(1 - beta2^step)
    int vec_ok;            // all four pointers 16-byte aligned (set by the launcher)
};
struct AdamLaunch {
    AdamTensor t[8];
    unsigned long long first_chunk[8];
    int n_tensors;
    float w1, beta2, w2, eps;   // 1 - beta1, beta2, 1 - beta2, eps
};

int launch_preprocess_fwd(const PreprocessParams& p, cudaStream_t stream);
int launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t stream);
int launch_inclusive_scan(int n, const uint32_t* n_dev, uint32_t* out, uint4* emit, uint32_t* block_sums, cudaStream_t stream);
int launch_fast_binning(int P, const uint32_t* n_dev, uint32_t R, uint2 grid, uint32_t tiles, int bits1, int bits2, const uint4* emit,
                        const uint32_t* offsets, const BinningWS& b, uint2* ranges, cudaStream_t stream);
int launch_depth_sort(uint64_t* keys_a, uint64_t* keys_b, uint32_t n, const uint32_t* n_dev, uint32_t* scratch, const uint2* rect,
                      uint4* emit, cudaStream_t stream);
int launch_small_depth_sort_scan(uint64_t* keys_a, uint64_t* keys_b, uint32_t n, const uint32_t* n_dev, uint32_t* scratch,
                                 const uint2* rect, uint4* emit, uint32_t* offsets, uint32_t* block_sums, cudaStream_t stream,
                                 bool* done);
int launch_strip_ids(uint32_t n, const uint64_t* in, uint32_t* out, cudaStream_t stream);
int launch_emit_instances(int P, const uint32_t* n_dev, const uint4* emit, uint2 grid, uint64_t* inst, cudaStream_t stream);
int launch_radix_sort(uint64_t* keys_a, uint64_t* keys_b, uint32_t n, int first_bit, int passes, char* scratch,
                      uint32_t* error_flag, cudaStream_t stream);
int launch_tile_ranges(uint32_t R, const uint64_t* inst, uint2* ranges, uint32_t tiles, cudaStream_t stream);
int launch_debug_expand(uint32_t tiles, const uint2* ranges, const uint32_t* ids, const SplatRec* rec, uint64_t* keys, uint64_t* raw);
int launch_debug_unpack(int which, uint32_t P, const SplatRec* rec, const uint2* rect, void* dst);
int launch_debug_low32(uint32_t n, const uint64_t* in, uint32_t* out);
int launch_debug_depth_order(uint32_t P, uint32_t n_vis, uint32_t R, const uint64_t* sorted, const uint2* rect, const uint32_t* offsets,
                             uint32_t* order_out, uint32_t* offsets_out);
int launch_render_fwd(const RenderFwdParams& p, int tiles, cudaStream_t stream);
int launch_render_bwd(const RenderBwdParams& p, int tiles, cudaStream_t stream);
int launch_render_work(const RenderBwdParams& p, int tiles, uint32_t* walked, uint32_t* survivors, cudaStream_t stream);
int launch_preprocess_bwd(const PreprocessBwdParams& p, cudaStream_t stream);
int launch_compute_relocation(int P, const float* opacity_old, const float* scale_old, const int* N, const float* binoms,
                              int n_max, float* opacity_new, float* scale_new, cudaStream_t stream);
int launch_add_noise(int P, float* xyz, const float* scaling_raw, const float* rotation_raw, const float* opacity_raw,
                     const float* noise, float lr_scale, unsigned long long seed, unsigned long long offset,
                     cudaStream_t stream);

int launch_adam_step(AdamLaunch& L, cudaStream_t stream);
int launch_loss_forward(const LossParams& p, float* out_means, cudaStream_t stream);
int launch_loss_backward(const LossParams& p, float c_l1, float c_ssim, const float* grad_scale, cudaStream_t stream);
size_t loss_partial_count(int C, int H, int W);
int launch_pair_reduce(const float* partial, int n, double inv_a, double inv_b, float* out, cudaStream_t stream);
int l1_block_count(size_t n);
int launch_l1_forward(size_t n, const float* x, const float* y, float* partial, float* out_mean, cudaStream_t stream);
int launch_l1_backward(size_t n, const float* x, const float* y, const float* grad_scale, float* dL_dx, cudaStream_t stream);
int launch_activate_fwd(int P, const float* raw_o, const float* raw_s, const float* raw_q, float* o, float* s, float* q,
                        float* partial, float* out_means, cudaStream_t stream);
int launch_activate_bwd(int P, const float* raw_o, const float* raw_s, const float* raw_q, const float* g_o, const float* g_s,
                        const float* g_q, float c_o, float c_s, const float* reg_grad, float* d_o, float* d_s, float* d_q,
                        cudaStream_t stream);
size_t activate_partial_count(int P);
int launch_swin_gather_fwd(const SwinParams& p, cudaStream_t stream);
int launch_swin_gather_bwd(const SwinParams& p, const SwinGrads& g, cudaStream_t stream);
size_t knn_workspace_bytes(int P);
int launch_knn_mean_dist2(int P, const float* points, float* out, char* ws, cudaStream_t stream);
int launch_color_grad_push(int P, const int* radii, const uint8_t* clamped, const float4* acc, const DpPeers& dst,
                           size_t slot_offset, bool multicast, cudaStream_t stream);
int launch_color_grad_push_sparse(int P, const int* radii, const uint8_t* clamped, const float4* acc, const DpPeers& dst,
                                  size_t slot_offset, bool multicast, int sm_count, cudaStream_t stream);
size_t dp_sparse_slot_floats(int P);
int launch_grad_record_combine(int P, int D, int M, int V, const float* means3D, const float* campos, const float* gathered,
                               size_t slot_stride, float* dL_dmeans3D, float* dL_dopacity, float* dL_dscales, float* dL_drots,
                               float* dL_dsh, float* dL_dsh_rest, cudaStream_t stream);
int launch_sh_grad_combine(int P, int D, int M, int V, const float* means3D, const float* campos, const float* gathered,
                           size_t slot_stride, bool sparse, float* dL_dsh, cudaStream_t stream);
int launch_slab_all_reduce(const DpPeers& peers, bool multicast, size_t n_floats, int rank, int world, int sm_count,
                           cudaStream_t stream);

}  // namespace gs
