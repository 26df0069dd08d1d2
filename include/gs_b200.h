/*
 * gs_b200.h — C ABI of lib3dgs_mcmc_b200.so: the B200 (sm_100a) Gaussian-splatting rasterizer and
 * MCMC update kernels.
 *
 * This is the drop-in boundary for the hot path of NeutrinoLiu/3dgs-mcmc. Every entry point takes
 * plain device pointers, sizes, scalars and a CUDA stream (as void*); no torch types. Each one
 * replaces one native entry of the reference's pybind module `_C`
 * (submodules/diff-gaussian-rasterization/ext.cpp:15-20):
 *
 *   gs_forward_preprocess + gs_forward_render  <-  rasterize_gaussians
 *        = RasterizeGaussiansCUDA (rasterize_points.cu:36-118) -> CudaRasterizer::Rasterizer::forward
 *          (cuda_rasterizer/rasterizer.h:33-58, rasterizer_impl.cu:198-338). The reference sizes its
 *          binning buffer from a std::function callback after a blocking D2H copy of the instance count
 *          (rasterizer_impl.cu:282-286); a C ABI has no callbacks, so the forward is two calls: the
 *          first ends by copying R to a pinned host word, the caller allocates gs_binning_bytes(R).
 *   gs_backward            <-  rasterize_gaussians_backward
 *        = RasterizeGaussiansBackwardCUDA (rasterize_points.cu:120-199) -> Rasterizer::backward
 *          (rasterizer.h:60-86, rasterizer_impl.cu:342-436)
 *   gs_mark_visible        <-  mark_visible (rasterize_points.cu:201-220, rasterizer_impl.cu:141-153)
 *   gs_compute_relocation  <-  compute_relocation (rasterize_points.cu:222-248, utils.h:8-19, utils.cu:6-51)
 *   gs_add_noise           additive: fuses the inline torch code of train_static.py:132-140
 *
 * Conventions (same as the reference): all float data FP32, row-contiguous; an absent
 * optional input is a NULL pointer; viewmatrix / projmatrix are 16 floats, element [4*i+j] of the
 * transposed matrix; quaternions are (w,x,y,z); SH is [P][M][3]; the image is CHW; radii==0 means
 * culled. Workspaces are opaque byte buffers owned by the caller; the library keeps no state.
 * Every function returns 0 on success, a positive cudaError_t value if a CUDA call failed, or a
 * negative GS_ERR_* code for bad arguments. Work is enqueued on `stream`; nothing synchronises
 * except where stated.
 */
#ifndef GS_B200_H_INCLUDED
#define GS_B200_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GS_B200_OK 0
#define GS_B200_ERR_ARG (-1)
#define GS_B200_ERR_WORKSPACE (-2)
#define GS_B200_ERR_LIMIT (-3) /* input larger than an entry point's fixed capacity */

/* Library / ABI version (major*100 + minor). */
int gs_abi_version(void);

/* Human-readable text for a return code of this library. */
const char* gs_error_string(int code);

/* ---- workspace sizes (bytes). Replace required<GeometryState/ImageState/BinningState>()
 *      (cuda_rasterizer/rasterizer_impl.h:65-72). ---- */
size_t gs_geom_bytes(int P);
size_t gs_image_bytes(int W, int H);
size_t gs_binning_bytes(unsigned int R, int W, int H, int P);
size_t gs_backward_scratch_bytes(int P);

/* Forward, stage 1: preprocess (cull, project, covariance, conic, radius, SH colour), then the depth order of
 * the Gaussians (radix sort of depth_bits|index) and the inclusive scan of tiles_touched in that order.
 * Writes radii[P] and zero-fills is_used[P]. If num_rendered_host != NULL (must be pinned host memory) the
 * instance count R is copied there asynchronously on `stream` right after the preprocess kernel; if
 * count_ready_event != NULL (a handle from gs_host_event_create) it is recorded after that copy, so the host can
 * gs_host_event_wait() for R and size the binning workspace while the depth sort is still running. Without
 * an event, synchronise the stream before reading R. */
int gs_forward_preprocess(void* stream, int P, int D, int M,
                          const float* means3D, const float* shs, const float* colors_precomp,
                          const float* opacities, const float* scales, float scale_modifier,
                          const float* rotations, const float* cov3D_precomp,
                          const float* viewmatrix, const float* projmatrix, const float* campos,
                          int W, int H, float tan_fovx, float tan_fovy, int prefiltered,
                          int* radii, int* is_used, void* geom_ws, unsigned int* num_rendered_host,
                          void* count_ready_event);

/* Host-side wait handle for the instance count (a CUDA event without timing). The reference blocks the whole
 * stream with a synchronous cudaMemcpy at this point (rasterizer_impl.cu:282). */
void* gs_host_event_create(void);
int gs_host_event_wait(void* event);
int gs_host_event_destroy(void* event);

/* Forward, stage 2: tile-instance emission in depth order, stable radix sort by tile, tile ranges, packed splat
 * stream, tile render. The resulting per-tile lists equal the reference's (tile|depth)-sorted lists exactly.
 * Writes out_color[3][H][W] and sets is_used[id]=1 for every blended Gaussian. */
int gs_forward_render(void* stream, int P, unsigned int R, int W, int H,
                      const float* background, const float* colors_precomp, const int* radii,
                      void* geom_ws, void* binning_ws, void* image_ws,
                      float* out_color, int* is_used);

/* Backward of the whole rasterizer. Every output is fully written (no pre-zeroing needed):
 * dL_dmeans2D[P][3], dL_dcolors[P][3], dL_dopacity[P], dL_dmeans3D[P][3], dL_dcov3D[P][6],
 * dL_dsh[P][M][3] (may be NULL when shs is NULL), dL_dscales[P][3], dL_drotations[P][4]. */
int gs_backward(void* stream, int P, int D, int M, unsigned int R, int W, int H,
                const float* background, const float* means3D, const float* shs, const float* colors_precomp,
                const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
                const float* viewmatrix, const float* projmatrix, const float* campos,
                float tan_fovx, float tan_fovy, const int* radii,
                const void* geom_ws, const void* binning_ws, const void* image_ws,
                const float* dL_dpix, void* scratch,
                float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D,
                float* dL_dcov3D, float* dL_dsh, float* dL_dscales, float* dL_drotations);

/* The two halves of gs_backward, for callers that overlap an inter-GPU exchange with the per-Gaussian half
 * (Rasterizer::backward runs the same two groups back to back: BACKWARD::render, then BACKWARD::preprocess,
 * rasterizer_impl.cu:391-435). gs_backward(...) == gs_backward_render(...) then gs_backward_geometry(...) on the
 * same scratch. In gs_backward_geometry dL_dsh may be NULL while shs is not: the SH gradient rows are then not
 * written (the caller rebuilds them with gs_dp_sh_grad_combine). */
int gs_backward_render(void* stream, int P, unsigned int R, int W, int H, const float* background,
                       const void* geom_ws, const void* binning_ws, const void* image_ws,
                       const float* dL_dpix, void* scratch);
int gs_backward_geometry(void* stream, int P, int D, int M, int W, int H,
                         const float* means3D, const float* shs, const float* scales, float scale_modifier,
                         const float* rotations, const float* cov3D_precomp,
                         const float* viewmatrix, const float* projmatrix, const float* campos,
                         float tan_fovx, float tan_fovy, const int* radii,
                         const void* geom_ws, const void* scratch,
                         float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D,
                         float* dL_dcov3D, float* dL_dsh, float* dL_dscales, float* dL_drotations);

/* ---- view-sharded data parallelism (additive; the reference is single-GPU). One process per GPU,
 *      Gaussians replicated, one camera view per rank. The SH gradient of one view is rank one,
 *      dL_dsh[k][c] = basis_k(dir) * dL_dRGB[c] (cuda_rasterizer/backward.cu:20-139), so instead of all-reducing
 *      3M floats per Gaussian the ranks all-gather the clamp-masked dL_dRGB (3 floats) and rebuild the sum. ----
 *
 * gs_dp_color_grad_push: after gs_backward_render, write this rank's masked dL_dRGB [P][3] at float offset
 *   slot_offset (a multiple of 4) of every buffer in dst_ptrs[n_dst] (peer-mapped device pointers, one per rank,
 *   16-byte aligned), or, with multicast != 0, through the single NVSwitch multicast pointer dst_ptrs[0]
 *   (multimem.st). The kernel does not wait for anything; order it against the consumers with a barrier.
 * gs_dp_sh_grad_combine: dL_dsh[P][M][3] = sum over views v (in order) of basis(dir(means3D, campos_views[v]))
 *   (x) color_grads[v * slot_stride + 3 i ...]; products rounded separately, so the result equals adding the
 *   per-view dL_dsh tensors in view order, on every rank bit for bit.
 * gs_dp_slab_all_reduce: in-place sum of a float slab that every rank holds at the same symmetric address.
 *   slab_ptrs[world] are the peer-mapped pointers in rank order (or one multicast pointer with multicast != 0);
 *   this rank reduces slice `rank` of `world` and stores it to every rank. n_floats must be a multiple of 4.
 *   The caller brackets the call with cross-rank barriers (all slabs written before; all slices landed after).
 *
 * layout (both calls must use the same one; gs_dp_slot_floats(P, layout) = floats one view's slot needs):
 *   GS_DP_DENSE   [P][3] floats, zeros included.
 *   GS_DP_SPARSE  per chunk of 256 Gaussians a 784-float record: 8 x (u32 mask, u32 set bits before this word) followed by
 *                 the chunk's non-zero triples packed in index order. Only the header and the non-zero triples are
 *                 stored (a Gaussian that was culled or never blended has an exactly zero gradient), from a bounded
 *                 number of CTAs. Results are bit-identical to GS_DP_DENSE. slot_offset / slot_stride multiples of 4. */
#define GS_DP_DENSE 0
#define GS_DP_SPARSE 1
#define GS_DP_RECORDS 2 /* gs_backward_geometry_push / gs_dp_grad_record_combine below */
size_t gs_dp_slot_floats(int P, int layout);
int gs_dp_color_grad_push(void* stream, int P, const int* radii, const void* geom_ws, const void* scratch,
                          float* const* dst_ptrs, int n_dst, int multicast, size_t slot_offset, int layout);
int gs_dp_sh_grad_combine(void* stream, int P, int D, int M, int V, const float* means3D,
                          const float* campos_views, const float* color_grads, size_t slot_stride, int layout,
                          float* dL_dsh);
int gs_dp_slab_all_reduce(void* stream, float* const* slab_ptrs, int n_ptrs, int multicast, size_t n_floats,
                          int rank, int world);

/* "Records" exchange (the default over symmetric memory): no all-reduce at all. Where a view's colour gradient is zero
 * (culled, or never blended) so are its gradients w.r.t. the mean, opacity, scale and rotation, so only the Gaussians with
 * a gradient travel, whole:
 * gs_backward_geometry_push: gs_backward_geometry for the shs + scales/rotations path, but instead of the dense parameter
 *   gradients every 128-Gaussian chunk leaves as a record of 1800 floats at float offset slot_offset + chunk * 1800 of every
 *   buffer in dst_ptrs[n_dst] (or through one multicast pointer): 4 x (u32 mask, u32 set bits before this word), then per
 *   set bit 14 floats {masked dL_dRGB 3, dL_dmean3D 3, dL_dopacity 1, dL_dscale 3, dL_drotation 4} in index order. Only the
 *   header and the packed records are stored. dL_dmeans2D [P][3] (this view's own) is still written densely.
 * gs_dp_grad_record_combine: from V such slots (records + v * slot_stride, slot_stride >= gs_dp_slot_floats(P,
 *   GS_DP_RECORDS)) the gradients summed over the views in view order - bit-identical on every rank, and equal to adding the
 *   per-view outputs of gs_backward in that order: dL_dmeans3D [P][3], dL_dopacity [P], dL_dscales [P][3],
 *   dL_drotations [P][4], dL_dsh [P][M][3] (basis(dir_v) (x) dL_dRGB_v; with dL_dsh_rest non-NULL: dL_dsh [P][1][3] and
 *   dL_dsh_rest [P][M-1][3], see the split SH entry points below). Only rows some view touched are written:
 *   outputs_are_zero != 0 promises the outputs hold zeros already (zero them off the critical path), 0 makes the call clear
 *   them first. V <= 32. */
/* gs_backward_render that also records cleared_event (from gs_host_event_create; may be NULL) on the stream right after the
 * accumulator clear, before the tile kernel: gs_stream_wait_event(side_stream, cleared_event) lets memory-bound side work
 * (the zero fill of the records exchange's outputs) start under the compute-bound tile kernel instead of beside the clear. */
int gs_backward_render_ev(void* stream, int P, unsigned int R, int W, int H, const float* background,
                          const void* geom_ws, const void* binning_ws, const void* image_ws, const float* dL_dpix,
                          void* scratch, void* cleared_event);
int gs_stream_wait_event(void* stream, void* event);
int gs_backward_geometry_push(void* stream, int P, int D, int M, int W, int H, const float* means3D, const float* shs,
                              const float* shs_rest, const float* scales, float scale_modifier, const float* rotations,
                              const float* viewmatrix, const float* projmatrix, const float* campos, float tan_fovx,
                              float tan_fovy, const int* radii, const void* geom_ws, const void* scratch,
                              float* dL_dmeans2D, float* const* dst_ptrs, int n_dst, int multicast, size_t slot_offset);
int gs_dp_grad_record_combine(void* stream, int P, int D, int M, int V, const float* means3D,
                              const float* campos_views, const float* records, size_t slot_stride,
                              int outputs_are_zero, float* dL_dmeans3D, float* dL_dopacity, float* dL_dscales,
                              float* dL_drotations, float* dL_dsh, float* dL_dsh_rest);

/* Split SH tensors (additive). The model stores features_dc [P][1][3] and features_rest [P][15][3] apart and the
 * reference concatenates them on every call of get_features (scene/gaussian_model_static.py:139-144: a 576 MB copy at
 * P = 3M, and autograd splits the gradient again). These entry points read - and differentiate into - the two tensors as
 * they lie: the same kernels, the same arithmetic instructions, bit-identical results to the concatenated layout.
 * M is 16 (degree-3 storage); all four SH pointers 16-byte aligned. In gs_backward_geometry_push / gs_dp_grad_record_combine
 * above, shs_rest / dL_dsh_rest non-NULL select the same layout (shs = features_dc, dL_dsh = its [P][1][3] gradient). */
int gs_forward_preprocess_split_sh(void* stream, int P, int D, const float* means3D, const float* features_dc,
                                   const float* features_rest, const float* opacities, const float* scales,
                                   float scale_modifier, const float* rotations, const float* viewmatrix,
                                   const float* projmatrix, const float* campos, int W, int H, float tan_fovx,
                                   float tan_fovy, int prefiltered, int* radii, int* is_used, void* geom_ws,
                                   unsigned int* num_rendered_host, void* count_ready_event);
int gs_backward_geometry_split_sh(void* stream, int P, int D, int W, int H, const float* means3D,
                                  const float* features_dc, const float* features_rest, const float* scales,
                                  float scale_modifier, const float* rotations, const float* viewmatrix,
                                  const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy,
                                  const int* radii, const void* geom_ws, const void* scratch, float* dL_dmeans2D,
                                  float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D,
                                  float* dL_dfeatures_dc, float* dL_dfeatures_rest, float* dL_dscales,
                                  float* dL_drotations);

/* present[i] = (view * means3D[i]).z > 0.2 ; one byte per Gaussian (bool). */
int gs_mark_visible(void* stream, int P, const float* means3D, const float* viewmatrix,
                    const float* projmatrix, unsigned char* present);

/* MCMC relocation (Eq. 9): opacity_new[P], scale_new[3P]. N holds int32 counts (caller clamps to [1, n_max-1]). */
int gs_compute_relocation(void* stream, int P, const float* opacity_old, const float* scale_old,
                          const int* N, const float* binoms, int n_max,
                          float* opacity_new, float* scale_new);

/* xyz[P][3] += Sigma(scaling_raw, rotation_raw) * (n * gate(opacity_raw) * lr_scale), lr_scale = noise_lr * xyz_lr.
 * noise[P][3] supplies the standard-normal deviates n; if NULL they are generated in-kernel from (seed, offset). */
int gs_add_noise(void* stream, int P, float* xyz, const float* scaling_raw, const float* rotation_raw,
                 const float* opacity_raw, const float* noise, float lr_scale,
                 unsigned long long seed, unsigned long long offset);

/* Multi-tensor Adam step (additive): one launch applies torch.optim.Adam's update (no weight decay,
 * no amsgrad) to up to 8 parameter tensors - the reference's six parameter groups, scene/gaussian_model_static.py:191-200,
 * stepped at train_static.py:129-130. Arrays are indexed by tensor: device pointers params / grads / exp_avg /
 * exp_avg_sq, element counts, per-tensor learning rate and step count (>= 1, already incremented, as torch keeps one
 * per parameter). Hyper-parameters are doubles, as torch keeps them: 1 - beta and the bias corrections are formed in
 * double and rounded to FP32 once. */
int gs_adam_step(void* stream, int n_tensors, float* const* params, const float* const* grads,
                 float* const* exp_avg, float* const* exp_avg_sq, const unsigned long long* numel,
                 const double* lr, const int* step, double beta1, double beta2, double eps);

/* Fused photometric loss (additive): the loss of train_static.py:94-97,
 *   loss = (1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image, gt)),
 * with l1_loss / ssim of utils/loss_utils.py:21-22, :40-67 (11x11 Gaussian window sigma 1.5, zero padding, per channel).
 * Forward: writes out_means[2] (device) = { mean |image - gt|, mean ssim_map } and keeps the three derivative maps in
 * ws (gs_loss_workspace_bytes). Backward: dL_dimage[C][H][W] = up * (c_l1 * sign(image - gt) + c_ssim * d(sum ssim_map)/dimage)
 * with up = *grad_scale (device scalar, the upstream gradient of the loss) or 1 when NULL; for the loss above
 * c_l1 = (1 - lambda) / (C H W), c_ssim = -lambda / (C H W). The ground truth gets no gradient. */
size_t gs_loss_workspace_bytes(int C, int H, int W);
int gs_photometric_loss_forward(void* stream, int C, int H, int W, const float* image, const float* gt,
                                void* ws, float* out_means);
int gs_photometric_loss_backward(void* stream, int C, int H, int W, const float* image, const float* gt,
                                 const void* ws, float c_l1, float c_ssim, const float* grad_scale,
                                 float* dL_dimage);

/* Plain L1 loss (additive): out_mean[0] = mean |x - y| over n floats (utils/loss_utils.py:21-22,
 * torch.abs(network_output - gt).mean()); out_mean is device float[2] ([1] is written as 0). ws: gs_l1_workspace_bytes(n)
 * of device scratch for the per-block partial sums, which are added in a fixed order in double. Backward:
 * dL_dx = *grad_scale (device scalar, NULL = 1) * sign(x - y) / n, sign(0) = 0 as torch's abs backward. */
size_t gs_l1_workspace_bytes(size_t n);
int gs_l1_loss_forward(void* stream, size_t n, const float* x, const float* y, void* ws, float* out_mean);
int gs_l1_loss_backward(void* stream, size_t n, const float* x, const float* y, const float* grad_scale,
                        float* dL_dx);

/* gs_l1_loss_forward without a workspace argument, for callers that cannot give it one. This entry point has no
 * counterpart in the reference (which gets all of its temporaries from torch's allocator); it exists for hosts that are
 * not allowed to allocate device memory at the point of the call: a plugin inside an engine that keeps the device
 * allocator to itself, a frame loop that must not hit cudaMalloc's device-wide synchronisation, a process with a memory
 * budget fixed at start-up. The per-block partial sums go into a 16 KB array that is part of the library image (every GPU
 * that loads the library holds it until the process exits); the call itself allocates and frees nothing.
 *   - n <= GS_L1_STATIC_MAX_N, else GS_B200_ERR_LIMIT. Within the limit the result equals gs_l1_loss_forward's bit for bit.
 *   - `stream` must belong to the current device. The first call on each device looks up the array's device address
 *     (cudaGetSymbolAddress); issue that first call before starting any stream capture.
 *   - All calls on one device share the array. Calls that may overlap on the GPU (different streams without an
 *     ordering between them) race on it; serialise them. Nothing in the array has to be cleared between calls: every
 *     slot a call reads it has written first. */
#define GS_L1_STATIC_MAX_N ((size_t)8388608) /* 2048 blocks x 4096 floats; 3 x 1080 x 1920 = 6 220 800 fits */
int gs_l1_loss_forward_static(void* stream, size_t n, const float* x, const float* y, float* out_mean);

/* Fused parameter activations + regularisers (additive): opacity = sigmoid(raw), scaling = exp(raw),
 * rotation = raw / max(|raw|, 1e-12) (scene/gaussian_model_static.py:43-56, :126-146) in one pass, which also writes
 * out_means[2] (device) = { mean |opacity|, mean |scaling| }, the two regulariser terms of train_static.py:99-100.
 * Backward: d_raw_* from the gradients w.r.t. the activated tensors (any may be NULL = zero) plus the regularisers'
 * own gradient, c_opacity * up * sign(opacity) and c_scaling * up * sign(scaling) with up = *reg_grad (device scalar;
 * NULL = 0); for reg = w_o * mean|opacity| + w_s * mean|scaling|: c_opacity = w_o / P, c_scaling = w_s / (3 P).
 * rotation pointers must be 16-byte aligned. */
size_t gs_activation_workspace_bytes(int P);
int gs_activate_forward(void* stream, int P, const float* raw_opacity, const float* raw_scaling,
                        const float* raw_rotation, float* opacity, float* scaling, float* rotation,
                        void* ws, float* out_means);
int gs_activate_backward(void* stream, int P, const float* raw_opacity, const float* raw_scaling,
                         const float* raw_rotation, const float* g_opacity, const float* g_scaling,
                         const float* g_rotation, float c_opacity, float c_scaling, const float* reg_grad,
                         float* d_raw_opacity, float* d_raw_scaling, float* d_raw_rotation);

/* distCUDA2: mean_dist2[i] = mean of the squared distances from points[i] to its three nearest other
 * points; replaces simple_knn._C.distCUDA2 = SimpleKNN::knn (submodules/simple-knn/spatial.cu:15-26,
 * simple_knn.cu:186-220), called at model initialisation (scene/gaussian_model_static.py:168). points[P][3],
 * mean_dist2[P], ws of gs_knn_workspace_bytes(P) bytes. Bit-identical to the reference (exact neighbours, same
 * distance expression, same averaging). */
size_t gs_knn_workspace_bytes(int P);
int gs_knn_mean_dist2(void* stream, int P, const float* points, float* mean_dist2, void* ws);

/* Sliding-window gather + rigid deformation + activations (additive): the rasterizer inputs of the
 * Gaussians alive at `frame`, as SwinGaussianModel.get_basic_para_at builds them (scene/gaussian_model.py:598-649 with
 * utils/tempo_utils.py:4-84 rigid_deform and the activations of gaussian_model.py:43-56).
 * A parameter set is a host array of 10 device pointers in this order:
 *   xyz[N][3], rotation[N][4] (raw, w x y z), features_dc[N][1][3], features_rest[N][M-1][3], opacity[N][1] (raw),
 *   scaling[N][3] (raw), rigid_v[N][3], rigid_rotvec[N][3], rigid_rotcen[N][3], frame_start[N] (float).
 * Forward: active Gaussian k < Ki is row imm_idx[k] of the immature set, k >= Ki row mat_idx[k - Ki] of the matured
 * set (int64 indices, as torch.nonzero returns them); age = frame - frame_start[row]; deform == 0 skips the motion
 * (rigid_deform(skip=True)). Outputs for K = Ki + Km: xyz[K][3], rotation[K][4] (normalised), feature[K][M][3]
 * (cat(dc, rest)), opacity[K][1] (sigmoid), scaling[K][3] (exp).
 * Backward: out_grads = 5 device pointers (gradients of the outputs in that order, any may be NULL), imm_grads = 9
 * zero-filled device tensors shaped like the first nine members of the immature set; rows imm_idx[k] are written
 * (the matured set is detached in the reference and gets no gradient). */
int gs_swin_gather_forward(void* stream, int Ki, int Km, int M, float frame, int deform,
                           const long long* imm_idx, const long long* mat_idx,
                           const float* const* imm_set, const float* const* mat_set,
                           float* out_xyz, float* out_rotation, float* out_feature, float* out_opacity,
                           float* out_scaling);
int gs_swin_gather_backward(void* stream, int Ki, int M, float frame, int deform, const long long* imm_idx,
                            const float* const* imm_set, const float* const* out_grads,
                            float* const* imm_grads);

/* ---- measurement hooks (bench.py): per-stage timing with CUDA events recorded on the caller's stream between
 *      the stages of each call, and a counter of kernel launches made by this library. Process-global; not used
 *      by the compute path. gs_profile_enable(n) arms n event marks (0 disables); gs_profile_collect() waits for
 *      the last mark, writes per-stage summed milliseconds and call counts ([gs_stage_count()] each) and rearms. ---- */
int gs_profile_enable(int max_marks);
int gs_profile_collect(float* ms_sum, int* calls);
int gs_stage_count(void);
const char* gs_stage_name(int stage);
unsigned long long gs_launch_count(void);

/* Test / debug hook: copy one internal array of the workspaces into dst (device memory). `which` uses the
 * numbering of the reference's state structs so parity tests can compare array by array:
 *   0 depths f32[P] | 1 clamped u8[P] (bit c = channel c) | 3 means2D f32[2P] | 4 cov3D f32[6P]
 *   5 conic_opacity f32[4P] | 6 rgb f32[3P] | 7 tiles_touched u32[P]
 *   8 inclusive scan of tiles_touched in depth order u32[P] | 9 gaussian ids in (depth, index) order u32[P]
 *   10 sorted gaussian ids u32[R] | 12 sorted keys (tile << 32 | depth_bits) u64[R], both in the reference's format
 *   13 sorted instances as (tile << 32 | id) u64[R] | 2 per-Gaussian render records 48 B x P
 *   20 final_T f32[WH] | 21 n_contrib u32[WH] | 22 ranges u32[2*tiles] | 23 tile_max_contrib u32[tiles]
 *   30 tile-sort error flag u32[1] | 31 depth-sort error flag u32[1]
 * Returns the array's size in bytes (copied only when it fits in max_bytes), or -1. Synchronous. */
long long gs_debug_fetch(int which, int P, unsigned int R, int W, int H,
                         const void* geom_ws, const void* binning_ws, const void* image_ws,
                         void* dst, size_t max_bytes);

/* Measurement hook (bench.py's roofline note, tools/tile_work.py): per tile of the last forward, the number of
 * (8x4 pixel patch, splat) list positions the tile renderers walk in the backward (walked_per_tile[tiles]; x 32 lanes
 * = the "evaluated pairs" figure of the roofline, counted BEFORE the conservative patch cull) and how many of them
 * survive the cull and are evaluated per pixel (evaluated_per_tile[tiles]). Device arrays; asynchronous on `stream`. */
int gs_debug_render_work(void* stream, int P, unsigned int R, int W, int H, const void* geom_ws, const void* binning_ws,
                         const void* image_ws, unsigned int* walked_per_tile, unsigned int* evaluated_per_tile);

/* Measurement hook: while device_counters (5 x u64 in device memory, zeroed by the caller) is non-null, gs_forward_render
 * launches the tallying build of the forward tile kernel, which adds to it: [0] warp-chunks walked, [1] records patch-tested,
 * [2] (warp, record) pairs evaluated per pixel, [3] (pixel, record) blends, [4] tile-level survivors summed over warp-chunks.
 * Pass NULL to return to the product kernel. Process-global; not used by the compute path. */
int gs_debug_forward_work(unsigned long long* device_counters);

#ifdef __cplusplus
}
#endif
#endif /* GS_B200_H_INCLUDED */
