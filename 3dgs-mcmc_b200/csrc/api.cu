// C ABI of lib3dgs_mcmc_b200.so (declared in include/gs_b200.h): argument checking, workspace carving
// and stage orchestration. No torch types, no global state; everything runs on the caller's stream.
#include "../../include/gs_b200.h"

#include <cmath>
#include <cstdio>
#include <vector>

#include "kernels.cuh"

using namespace gs;

#include <cstdlib>

#include <nvtx3/nvToolsExt.h>
namespace gs {
// On for large scenes only: measured 3.073 -> 3.034 ms/step at 3M Gaussians / 1080p, but 0.528 -> 0.536 ms at 300K / 800x800,
// where the chain is bound by the host's launch rate and cudaLaunchKernelEx with an attribute costs the host more per launch.
// GS_PDL=0 / 1 forces it off / on.
static thread_local bool t_pdl_large = false;
void pdl_hint(long long gaussians) { t_pdl_large = gaussians >= 1000000; }
bool pdl_enabled() {
    static const int forced = [] { const char* e = getenv("GS_PDL"); return e ? (e[0] == '0' ? 0 : 1) : -1; }();
    return forced < 0 ? t_pdl_large : forced == 1;
}
}  // namespace gs

namespace {

inline uint32_t tiles_of(int W, int H, uint2* grid) {
    const uint32_t gx = (uint32_t)((W + GS_TILE_X - 1) / GS_TILE_X), gy = (uint32_t)((H + GS_TILE_Y - 1) / GS_TILE_Y);
    if (grid) *grid = make_uint2(gx, gy);
    return gx * gy;
}

struct BwdScratch {
    float4* acc_color_opacity;
    float4* acc_mean_conic;
    float* acc_conic_w;
    size_t bytes;
    static BwdScratch carve(char* base, size_t P) {
        Carver c(base);
        BwdScratch s;
        s.acc_color_opacity = c.take<float4>(P);
        s.acc_mean_conic = c.take<float4>(P);
        s.acc_conic_w = c.take<float>(P);
        s.bytes = align_up(c.off, kAlign);
        return s;
    }
};


// ---- optional per-stage timing (CUDA events on the caller's stream) and a kernel-launch counter ----
enum Stage {
    ST_PREPROCESS = 0, ST_DEPTH_SORT, ST_SCAN, ST_DUPLICATE, ST_SORT, ST_PACK, ST_RENDER_FWD, ST_BWD_ZERO, ST_RENDER_BWD,
    ST_PREPROCESS_BWD, ST_RELOCATION, ST_ADD_NOISE, ST_DP_PUSH, ST_DP_COMBINE, ST_DP_ALL_REDUCE, ST_ADAM, ST_LOSS_FWD, ST_LOSS_BWD, ST_ACT_FWD, ST_ACT_BWD, ST_KNN, ST_SWIN_FWD, ST_SWIN_BWD, ST_COUNT
};
const char* const kStageNames[ST_COUNT] = {"preprocess_fwd", "depth_sort", "scan", "emit_instances", "tile_sort", "tile_ranges",
                                           "render_fwd", "bwd_zero", "render_bwd", "preprocess_bwd",
                                           "compute_relocation", "add_noise", "dp_color_grad_push", "dp_sh_grad_combine",
                                           "dp_slab_all_reduce", "adam_step", "loss_fwd", "loss_bwd", "activate_fwd",
                                           "activate_bwd", "knn_mean_dist2", "swin_gather_fwd", "swin_gather_bwd"};
struct Profiler {
    bool on = false;
    std::vector<cudaEvent_t> ev;
    std::vector<int> stage;   // stage[i] = what ran between ev[i-1] and ev[i]; -1 = region start
    size_t n = 0;
    unsigned long long launches = 0;
} g_prof;

inline void prof_mark(cudaStream_t stream, int stage) {
    if (!g_prof.on || g_prof.n >= g_prof.ev.size()) return;
    cudaEventRecord(g_prof.ev[g_prof.n], stream);
    g_prof.stage[g_prof.n] = stage;
    ++g_prof.n;
}
inline void count_launches(int k) { g_prof.launches += (unsigned long long)k; }

// NVTX ranges per stage (tracing): with GS_NVTX=1 in the environment every stage's launches are issued
// inside a host-side range named like the stage (gs_stage_name), so a timeline tool (nsys) labels the kernels by stage. The
// reference has no tracing of its own (its only analogue is CHECK_CUDA, cuda_rasterizer/auxiliary.h:166-173). Off by
// default; the NVTX calls are no-ops anyway unless a tool is attached.
const bool g_nvtx = [] { const char* e = getenv("GS_NVTX"); return e && atoi(e) != 0; }();
struct NvtxScope {
    bool open = false;
    NvtxScope() = default;
    explicit NvtxScope(int s) { stage(s); }
    NvtxScope(const NvtxScope&) = delete;
    NvtxScope& operator=(const NvtxScope&) = delete;
    void stage(int s) {
        if (!g_nvtx) return;
        if (open) nvtxRangePop();
        nvtxRangePushA(kStageNames[s]);
        open = true;
    }
    ~NvtxScope() { if (open) nvtxRangePop(); }
};
unsigned long long* g_fwd_counters = nullptr;   // set by gs_debug_forward_work for the next forward(s); measurement only

}  // namespace

extern "C" {

int gs_abi_version(void) { return 107; }

const char* gs_error_string(int code) {
    if (code == GS_OK) return "ok";
    if (code == GS_ERR_ARG) return "gs_b200: invalid argument";
    if (code == GS_ERR_WORKSPACE) return "gs_b200: workspace too small or misaligned";
    if (code == GS_ERR_LIMIT) return "gs_b200: input larger than the fixed capacity of this entry point";
    if (code > 0) return cudaGetErrorString((cudaError_t)code);
    return "gs_b200: unknown error";
}


int gs_profile_enable(int max_marks) {
    for (cudaEvent_t e : g_prof.ev) cudaEventDestroy(e);
    g_prof.ev.clear(); g_prof.stage.clear(); g_prof.n = 0; g_prof.on = false;
    if (max_marks <= 0) return GS_OK;
    g_prof.ev.resize((size_t)max_marks);
    g_prof.stage.assign((size_t)max_marks, -1);
    for (auto& e : g_prof.ev) GS_CUDA_TRY(cudaEventCreate(&e));
    g_prof.on = true;
    return GS_OK;
}

int gs_profile_collect(float* ms_sum, int* calls) {
    for (int i = 0; i < ST_COUNT; ++i) { ms_sum[i] = 0.f; calls[i] = 0; }
    if (g_prof.n == 0) return GS_OK;
    GS_CUDA_TRY(cudaEventSynchronize(g_prof.ev[g_prof.n - 1]));
    for (size_t i = 1; i < g_prof.n; ++i) {
        const int st = g_prof.stage[i];
        if (st < 0) continue;
        float ms = 0.f;
        GS_CUDA_TRY(cudaEventElapsedTime(&ms, g_prof.ev[i - 1], g_prof.ev[i]));
        ms_sum[st] += ms;
        calls[st] += 1;
    }
    g_prof.n = 0;
    return GS_OK;
}

int gs_stage_count(void) { return ST_COUNT; }
const char* gs_stage_name(int stage) { return (stage >= 0 && stage < ST_COUNT) ? kStageNames[stage] : "?"; }
unsigned long long gs_launch_count(void) { return g_prof.launches; }

size_t gs_geom_bytes(int P) { return GeomWS::carve(nullptr, (size_t)(P > 0 ? P : 0)).bytes + kAlign; }
size_t gs_image_bytes(int W, int H) { return ImageWS::carve(nullptr, (size_t)W, (size_t)H).bytes + kAlign; }
size_t gs_binning_bytes(unsigned int R, int W, int H, int P) {
    return BinningWS::carve(nullptr, R, tiles_of(W, H, nullptr), (size_t)(P > 0 ? P : 0)).bytes + kAlign;
}
size_t gs_backward_scratch_bytes(int P) { return BwdScratch::carve(nullptr, (size_t)(P > 0 ? P : 0)).bytes + kAlign; }

static inline char* aligned_base(const void* p) {
    return reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(p), kAlign));
}

// split SH layout (features_dc [P][1][3] + features_rest [P][15][3]): both 16-byte aligned, M = 16
static bool split_sh_ok(const float* dc, const float* rest, int M) {
    return dc && rest && M == 16 && is_aligned16(dc) && is_aligned16(rest);
}

static int forward_preprocess_impl(void* stream_, int P, int D, int M, const float* means3D, const float* shs, const float* shs_rest,
                                   const float* colors_precomp, const float* opacities, const float* scales,
                                   float scale_modifier, const float* rotations, const float* cov3D_precomp,
                                   const float* viewmatrix, const float* projmatrix, const float* campos, int W, int H,
                                   float tan_fovx, float tan_fovy, int prefiltered, int* radii, int* is_used, void* geom_ws,
                                   unsigned int* num_rendered_host, void* count_ready_event) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (P < 0 || W <= 0 || H <= 0 || D < 0 || D > 3) return GS_ERR_ARG;
    if (P > 0 && shs_rest && (colors_precomp || !split_sh_ok(shs, shs_rest, M))) return GS_ERR_ARG;
    if (P > 0 && (!means3D || !opacities || !viewmatrix || !projmatrix || !radii || !is_used || !geom_ws)) return GS_ERR_ARG;
    if (P > 0 && !colors_precomp && (!shs || !campos || M < (D + 1) * (D + 1))) return GS_ERR_ARG;
    if (P > 0 && !cov3D_precomp && (!scales || !rotations)) return GS_ERR_ARG;
    if (P == 0) {
        if (num_rendered_host) *num_rendered_host = 0;
        if (count_ready_event) GS_CUDA_TRY(cudaEventRecord((cudaEvent_t)count_ready_event, stream));
        return GS_OK;
    }
    GeomWS g = GeomWS::carve(aligned_base(geom_ws), (size_t)P);
    pdl_hint(P);

    PreprocessParams pp;
    pp.P = P; pp.D = D; pp.M = M;
    pp.means3D = means3D; pp.scales = scales; pp.scale_modifier = scale_modifier; pp.rotations = rotations;
    pp.opacities = opacities; pp.shs = shs; pp.shs_rest = shs_rest; pp.cov3D_precomp = cov3D_precomp; pp.colors_precomp = colors_precomp;
    pp.viewmatrix = viewmatrix; pp.projmatrix = projmatrix; pp.campos = campos;
    pp.W = W; pp.H = H; pp.tan_fovx = tan_fovx; pp.tan_fovy = tan_fovy;
    pp.focal_y = H / (2.0f * tan_fovy);
    pp.focal_x = W / (2.0f * tan_fovx);
    tiles_of(W, H, &pp.grid);
    pp.prefiltered = prefiltered;
    pp.rot_aligned = is_aligned16(rotations) ? 1 : 0;
    pp.radii = radii; pp.is_used = is_used; pp.cov3D = g.cov3D; pp.clamped = g.clamped;
    pp.rec = g.rec; pp.depth_keys = g.dkeys_a; pp.num_rendered = g.num_rendered;
    pp.rect = g.rect;
    prof_mark(stream, -1);
    NvtxScope nv(ST_PREPROCESS);
    GS_CUDA_TRY(cudaMemsetAsync(g.num_rendered, 0, 16, stream));
    int rc = launch_preprocess_fwd(pp, stream);
    if (rc) return rc;
    prof_mark(stream, ST_PREPROCESS);
    // The instance count is known here; the host can size the binning workspace while the depth sort runs.
    if (num_rendered_host)
        GS_CUDA_TRY(cudaMemcpyAsync(num_rendered_host, g.num_rendered, sizeof(unsigned int), cudaMemcpyDeviceToHost, stream));
    if (count_ready_event) GS_CUDA_TRY(cudaEventRecord((cudaEvent_t)count_ready_event, stream));
    // Depth order of the Gaussians: stable sort of (depth_bits << 32 | idx) on the depth field, then the scan of the tile
    // counts in that order. Small scenes: both in one cooperative kernel; else the kernel chain.
    bool fused = false;
    nv.stage(ST_DEPTH_SORT);
    rc = launch_small_depth_sort_scan(g.dkeys_a, g.dkeys_b, (uint32_t)P, g.num_rendered + 2, reinterpret_cast<uint32_t*>(g.dsort_scratch),
                                      g.rect, g.emit, g.point_offsets, g.scan_block_sums, stream, &fused);
    if (rc) return rc;
    if (fused) {
        prof_mark(stream, ST_DEPTH_SORT);
        count_launches(2);
        return GS_OK;
    }
    rc = launch_depth_sort(g.dkeys_a, g.dkeys_b, (uint32_t)P, g.num_rendered + 2, reinterpret_cast<uint32_t*>(g.dsort_scratch), g.rect, g.emit, stream);
    if (rc) return rc;
    prof_mark(stream, ST_DEPTH_SORT);
    nv.stage(ST_SCAN);
    rc = launch_inclusive_scan(P, g.num_rendered + 2, g.point_offsets, g.emit, g.scan_block_sums, stream);
    if (rc) return rc;
    prof_mark(stream, ST_SCAN);
    count_launches(1 + 3 * kDepthSortPasses + 2);
    return GS_OK;
}

int gs_forward_preprocess(void* stream_, int P, int D, int M, const float* means3D, const float* shs,
                          const float* colors_precomp, const float* opacities, const float* scales,
                          float scale_modifier, const float* rotations, const float* cov3D_precomp,
                          const float* viewmatrix, const float* projmatrix, const float* campos, int W, int H,
                          float tan_fovx, float tan_fovy, int prefiltered, int* radii, int* is_used, void* geom_ws,
                          unsigned int* num_rendered_host, void* count_ready_event) {
    return forward_preprocess_impl(stream_, P, D, M, means3D, shs, nullptr, colors_precomp, opacities, scales, scale_modifier,
                                   rotations, cov3D_precomp, viewmatrix, projmatrix, campos, W, H, tan_fovx, tan_fovy, prefiltered,
                                   radii, is_used, geom_ws, num_rendered_host, count_ready_event);
}

int gs_forward_preprocess_split_sh(void* stream_, int P, int D, const float* means3D, const float* features_dc,
                                   const float* features_rest, const float* opacities, const float* scales,
                                   float scale_modifier, const float* rotations, const float* viewmatrix,
                                   const float* projmatrix, const float* campos, int W, int H, float tan_fovx, float tan_fovy,
                                   int prefiltered, int* radii, int* is_used, void* geom_ws, unsigned int* num_rendered_host,
                                   void* count_ready_event) {
    if (P > 0 && (!features_dc || !features_rest)) return GS_ERR_ARG;
    return forward_preprocess_impl(stream_, P, D, 16, means3D, features_dc, features_rest, nullptr, opacities, scales,
                                   scale_modifier, rotations, nullptr, viewmatrix, projmatrix, campos, W, H, tan_fovx, tan_fovy,
                                   prefiltered, radii, is_used, geom_ws, num_rendered_host, count_ready_event);
}

int gs_forward_render(void* stream_, int P, unsigned int R, int W, int H, const float* background,
                      const float* colors_precomp, const int* radii, void* geom_ws, void* binning_ws, void* image_ws,
                      float* out_color, int* is_used) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (P < 0 || W <= 0 || H <= 0 || !background || !out_color || !image_ws) return GS_ERR_ARG;
    if (P > 0 && (!geom_ws || !binning_ws || !radii || !is_used)) return GS_ERR_ARG;
    if (R >= (1u << 30)) return GS_ERR_ARG;
    uint2 grid;
    const uint32_t tiles = tiles_of(W, H, &grid);
    ImageWS img = ImageWS::carve(aligned_base(image_ws), (size_t)W, (size_t)H);
    GeomWS g = GeomWS::carve(aligned_base(geom_ws), (size_t)P);
    BinningWS b = BinningWS::carve(aligned_base(binning_ws), R, tiles, (size_t)(P > 0 ? P : 0));
    if (P == 0) R = 0;
    pdl_hint(P);

    int rc = GS_OK;
    prof_mark(stream, -1);
    NvtxScope nv(ST_SORT);
    if (P > 0) GS_CUDA_TRY(cudaMemsetAsync(b.error_flag, 0, 16, stream));
    if (P > 0 && bin_fast_path(tiles, (size_t)P)) {
        // Two fused binning passes (count -> scan -> scatter each); the tile ranges fall out of the second count.
        int bits1, bits2;
        bin_digit_bits(tiles, &bits1, &bits2);
        rc = launch_fast_binning(P, g.num_rendered + 2, R, grid, tiles, bits1, bits2, g.emit, g.point_offsets, b, img.ranges, stream);
        if (rc) return rc;
        prof_mark(stream, ST_SORT);
        count_launches(R > 0 ? 7 : 0);
    } else {
        // Generic path (tile ids wider than two binning digits): emission, multi-pass radix sort, range detection.
        const int passes = sort_passes(tiles);
        if (R > 0) {
            nv.stage(ST_DUPLICATE);
            rc = launch_emit_instances(P, g.num_rendered + 2, g.emit, grid, b.inst_a, stream);
            if (rc) return rc;
            prof_mark(stream, ST_DUPLICATE);
            nv.stage(ST_SORT);
            rc = launch_radix_sort(b.inst_a, b.inst_b, R, 32, passes, b.sort_scratch, b.error_flag, stream);
            if (rc) return rc;
            rc = launch_strip_ids(R, b.sorted(passes), b.ids, stream);
            if (rc) return rc;
            prof_mark(stream, ST_SORT);
            count_launches(4 + passes);
        }
        nv.stage(ST_PACK);
        rc = launch_tile_ranges(R, b.sorted(passes), img.ranges, tiles, stream);
        if (rc) return rc;
        prof_mark(stream, ST_PACK);
        count_launches(1);
    }
    (void)colors_precomp;  // already folded into the render records by the preprocess stage

    RenderFwdParams rp;
    rp.ranges = img.ranges; rp.ids = b.ids; rp.rec = g.rec; rp.W = W; rp.H = H; rp.tiles_x = (int)grid.x; rp.bg = background;
    rp.final_T = img.final_T; rp.n_contrib = img.n_contrib; rp.out_color = out_color; rp.is_used = is_used;
    rp.tile_max_contrib = img.tile_max_contrib;
    rp.counters = g_fwd_counters;
    nv.stage(ST_RENDER_FWD);
    rc = launch_render_fwd(rp, (int)tiles, stream);
    prof_mark(stream, ST_RENDER_FWD);
    count_launches(1);
    return rc;
}

int gs_backward_render(void* stream_, int P, unsigned int R, int W, int H, const float* background, const void* geom_ws,
                       const void* binning_ws, const void* image_ws, const float* dL_dpix, void* scratch) {
    return gs_backward_render_ev(stream_, P, R, W, H, background, geom_ws, binning_ws, image_ws, dL_dpix, scratch, nullptr);
}

int gs_stream_wait_event(void* stream_, void* event) {
    if (!event) return GS_ERR_ARG;
    GS_CUDA_TRY(cudaStreamWaitEvent((cudaStream_t)stream_, (cudaEvent_t)event, 0));
    return GS_OK;
}

int gs_backward_render_ev(void* stream_, int P, unsigned int R, int W, int H, const float* background, const void* geom_ws,
                          const void* binning_ws, const void* image_ws, const float* dL_dpix, void* scratch,
                          void* cleared_event) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (P < 0 || W <= 0 || H <= 0) return GS_ERR_ARG;
    if (P == 0) return GS_OK;
    if (!geom_ws || !binning_ws || !image_ws || !dL_dpix || !scratch || !background) return GS_ERR_ARG;
    uint2 grid;
    const uint32_t tiles = tiles_of(W, H, &grid);
    ImageWS img = ImageWS::carve(aligned_base(image_ws), (size_t)W, (size_t)H);
    GeomWS g = GeomWS::carve(aligned_base(geom_ws), (size_t)P);
    BinningWS b = BinningWS::carve(aligned_base(binning_ws), R, tiles, (size_t)(P > 0 ? P : 0));
    BwdScratch sc = BwdScratch::carve(aligned_base(scratch), (size_t)P);
    prof_mark(stream, -1);
    NvtxScope nv(ST_BWD_ZERO);
    GS_CUDA_TRY(cudaMemsetAsync(sc.acc_color_opacity, 0, sc.bytes, stream));
    prof_mark(stream, ST_BWD_ZERO);
    if (cleared_event) GS_CUDA_TRY(cudaEventRecord((cudaEvent_t)cleared_event, stream));
    if (R > 0) {
        RenderBwdParams rb;
        rb.ranges = img.ranges; rb.ids = b.ids; rb.rec = g.rec; rb.tile_max_contrib = img.tile_max_contrib;
        rb.W = W; rb.H = H; rb.tiles_x = (int)grid.x; rb.bg = background; rb.final_T = img.final_T;
        rb.n_contrib = img.n_contrib; rb.dL_dpix = dL_dpix;
        rb.acc_color_opacity = sc.acc_color_opacity; rb.acc_mean_conic = sc.acc_mean_conic; rb.acc_conic_w = sc.acc_conic_w;
        nv.stage(ST_RENDER_BWD);
        const int rc = launch_render_bwd(rb, (int)tiles, stream);
        if (rc) return rc;
        prof_mark(stream, ST_RENDER_BWD);
        count_launches(1);
    }
    return GS_OK;
}

int gs_debug_forward_work(unsigned long long* device_counters) {
    g_fwd_counters = device_counters;
    return GS_OK;
}

int gs_debug_render_work(void* stream_, int P, unsigned int R, int W, int H, const void* geom_ws, const void* binning_ws,
                         const void* image_ws, unsigned int* walked_per_tile, unsigned int* evaluated_per_tile) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (P <= 0 || W <= 0 || H <= 0 || !geom_ws || !binning_ws || !image_ws || !walked_per_tile || !evaluated_per_tile) return GS_ERR_ARG;
    uint2 grid;
    const uint32_t tiles = tiles_of(W, H, &grid);
    ImageWS img = ImageWS::carve(aligned_base(image_ws), (size_t)W, (size_t)H);
    GeomWS g = GeomWS::carve(aligned_base(geom_ws), (size_t)P);
    BinningWS b = BinningWS::carve(aligned_base(binning_ws), R, tiles, (size_t)(P > 0 ? P : 0));
    RenderBwdParams rb = {};
    rb.ranges = img.ranges; rb.ids = b.ids; rb.rec = g.rec; rb.W = W; rb.H = H; rb.tiles_x = (int)grid.x; rb.n_contrib = img.n_contrib;
    return launch_render_work(rb, (int)tiles, walked_per_tile, evaluated_per_tile, stream);
}

static int fill_peers(DpPeers* peers, float* const* ptrs, int n);

// push: null = the dense outputs; else the "records" exchange (the five parameter gradients leave as chunk records)
static int backward_geometry_impl(void* stream_, int P, int D, int M, int W, int H, const float* means3D, const float* shs,
                                  const float* shs_rest, float* dL_dsh_rest, const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
                                  const float* viewmatrix, const float* projmatrix, const float* campos, float tan_fovx,
                                  float tan_fovy, const int* radii, const void* geom_ws, const void* scratch, float* dL_dmeans2D,
                                  float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh,
                                  float* dL_dscales, float* dL_drotations, const DpPeers* push, size_t push_offset,
                                  int push_multicast) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (P < 0 || W <= 0 || H <= 0 || D < 0 || D > 3) return GS_ERR_ARG;
    if (P == 0) return GS_OK;
    if (!means3D || !radii || !geom_ws || !scratch || !viewmatrix || !projmatrix || !dL_dmeans2D) return GS_ERR_ARG;
    if (!push && (!dL_dcolors || !dL_dopacity || !dL_dmeans3D || !dL_dcov3D || !dL_dscales || !dL_drotations)) return GS_ERR_ARG;
    pdl_hint(P);
    if (push && (!shs || !scales || !rotations || cov3D_precomp || (push_offset & 3))) return GS_ERR_ARG;
    if (shs && !campos) return GS_ERR_ARG;
    if (shs && !dL_dsh && 3 * M > 48) return GS_ERR_ARG;   // factored mode needs the staged SH rows (sh_stage.cuh: kShMaxRow)
    if (shs_rest && (!split_sh_ok(shs, shs_rest, M) || (dL_dsh != nullptr) != (dL_dsh_rest != nullptr))) return GS_ERR_ARG;
    if (shs_rest && dL_dsh && !split_sh_ok(dL_dsh, dL_dsh_rest, M)) return GS_ERR_ARG;
    GeomWS g = GeomWS::carve(aligned_base(geom_ws), (size_t)P);
    BwdScratch sc = BwdScratch::carve(aligned_base(scratch), (size_t)P);
    prof_mark(stream, -1);
    NvtxScope nv(ST_PREPROCESS_BWD);
    PreprocessBwdParams pb;
    pb.P = P; pb.D = D; pb.M = M; pb.means3D = means3D; pb.radii = radii; pb.shs = shs; pb.clamped = g.clamped;
    pb.shs_rest = shs_rest; pb.dL_dsh_rest = dL_dsh_rest;
    pb.scales = cov3D_precomp ? nullptr : scales; pb.rotations = rotations; pb.scale_modifier = scale_modifier;
    pb.rot_aligned = is_aligned16(rotations) ? 1 : 0;
    pb.drot_aligned = is_aligned16(dL_drotations) ? 1 : 0;
    pb.cov3D = cov3D_precomp ? cov3D_precomp : g.cov3D;
    pb.viewmatrix = viewmatrix; pb.projmatrix = projmatrix; pb.campos = campos;
    pb.focal_y = H / (2.0f * tan_fovy);
    pb.focal_x = W / (2.0f * tan_fovx);
    pb.tan_fovx = tan_fovx; pb.tan_fovy = tan_fovy;
    pb.acc_color_opacity = sc.acc_color_opacity; pb.acc_mean_conic = sc.acc_mean_conic; pb.acc_conic_w = sc.acc_conic_w;
    pb.dL_dmeans3D = dL_dmeans3D; pb.dL_dmeans2D = dL_dmeans2D; pb.dL_dcolors = dL_dcolors; pb.dL_dopacity = dL_dopacity;
    pb.dL_dcov3D = dL_dcov3D; pb.dL_dsh = dL_dsh; pb.dL_dscales = dL_dscales; pb.dL_drots = dL_drotations;
    pb.push.n = 0; pb.push_offset = 0; pb.push_multicast = 0;
    if (push) { pb.push = *push; pb.push_offset = push_offset; pb.push_multicast = push_multicast; }
    const int rc = launch_preprocess_bwd(pb, stream);
    prof_mark(stream, ST_PREPROCESS_BWD);
    count_launches(1);
    return rc;
}

int gs_backward_geometry(void* stream_, int P, int D, int M, int W, int H, const float* means3D, const float* shs,
                         const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
                         const float* viewmatrix, const float* projmatrix, const float* campos, float tan_fovx,
                         float tan_fovy, const int* radii, const void* geom_ws, const void* scratch, float* dL_dmeans2D,
                         float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh,
                         float* dL_dscales, float* dL_drotations) {
    return backward_geometry_impl(stream_, P, D, M, W, H, means3D, shs, nullptr, nullptr, scales, scale_modifier, rotations,
                                  cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, radii, geom_ws, scratch,
                                  dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations,
                                  nullptr, 0, 0);
}

int gs_backward_geometry_split_sh(void* stream_, int P, int D, int W, int H, const float* means3D, const float* features_dc,
                                  const float* features_rest, const float* scales, float scale_modifier, const float* rotations,
                                  const float* viewmatrix, const float* projmatrix, const float* campos, float tan_fovx,
                                  float tan_fovy, const int* radii, const void* geom_ws, const void* scratch, float* dL_dmeans2D,
                                  float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D,
                                  float* dL_dfeatures_dc, float* dL_dfeatures_rest, float* dL_dscales, float* dL_drotations) {
    if (P > 0 && (!features_dc || !features_rest || !dL_dfeatures_dc || !dL_dfeatures_rest)) return GS_ERR_ARG;
    return backward_geometry_impl(stream_, P, D, 16, W, H, means3D, features_dc, features_rest, dL_dfeatures_rest, scales,
                                  scale_modifier, rotations, nullptr, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, radii,
                                  geom_ws, scratch, dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dfeatures_dc,
                                  dL_dscales, dL_drotations, nullptr, 0, 0);
}

int gs_backward_geometry_push(void* stream_, int P, int D, int M, int W, int H, const float* means3D, const float* shs,
                              const float* shs_rest, const float* scales, float scale_modifier, const float* rotations, const float* viewmatrix,
                              const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy, const int* radii,
                              const void* geom_ws, const void* scratch, float* dL_dmeans2D, float* const* dst_ptrs, int n_dst,
                              int multicast, size_t slot_offset) {
    DpPeers peers;
    const int rc = fill_peers(&peers, dst_ptrs, n_dst);
    if (rc) return rc;
    if (multicast && n_dst != 1) return GS_ERR_ARG;
    return backward_geometry_impl(stream_, P, D, M, W, H, means3D, shs, shs_rest, nullptr, scales, scale_modifier, rotations, nullptr,
                                  viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, radii, geom_ws, scratch, dL_dmeans2D, nullptr,
                                  nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, &peers, slot_offset, multicast);
}

int gs_backward(void* stream_, int P, int D, int M, unsigned int R, int W, int H, const float* background,
                const float* means3D, const float* shs, const float* colors_precomp, const float* scales,
                float scale_modifier, const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy, const int* radii,
                const void* geom_ws, const void* binning_ws, const void* image_ws, const float* dL_dpix, void* scratch,
                float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D,
                float* dL_dsh, float* dL_dscales, float* dL_drotations) {
    (void)colors_precomp;
    if (P < 0 || W <= 0 || H <= 0 || D < 0 || D > 3) return GS_ERR_ARG;
    if (P == 0) return GS_OK;
    if (!means3D || !radii || !geom_ws || !binning_ws || !image_ws || !dL_dpix || !scratch || !background ||
        !viewmatrix || !projmatrix)
        return GS_ERR_ARG;
    if (!dL_dmeans2D || !dL_dcolors || !dL_dopacity || !dL_dmeans3D || !dL_dcov3D || !dL_dscales || !dL_drotations)
        return GS_ERR_ARG;
    if (shs && (!dL_dsh || !campos)) return GS_ERR_ARG;
    int rc = gs_backward_render(stream_, P, R, W, H, background, geom_ws, binning_ws, image_ws, dL_dpix, scratch);
    if (rc) return rc;
    return gs_backward_geometry(stream_, P, D, M, W, H, means3D, shs, scales, scale_modifier, rotations, cov3D_precomp,
                                viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, radii, geom_ws, scratch, dL_dmeans2D,
                                dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations);
}

static int fill_peers(DpPeers* peers, float* const* ptrs, int n) {
    if (!ptrs || n < 1 || n > 8) return GS_ERR_ARG;
    for (int r = 0; r < 8; ++r) peers->ptr[r] = r < n ? ptrs[r] : nullptr;
    for (int r = 0; r < n; ++r)
        if (!ptrs[r] || (reinterpret_cast<size_t>(ptrs[r]) & 15)) return GS_ERR_ARG;
    peers->n = n;
    return GS_OK;
}

size_t gs_dp_slot_floats(int P, int layout) {
    if (P <= 0) return 0;
    if (layout == GS_DP_RECORDS) return (size_t)kRecFloats * (size_t)((P + kRecChunk - 1) / kRecChunk);
    return layout == GS_DP_SPARSE ? dp_sparse_slot_floats(P) : ((size_t)3 * P + 3) / 4 * 4;
}

int gs_dp_grad_record_combine(void* stream_, int P, int D, int M, int V, const float* means3D, const float* campos_views,
                              const float* records, size_t slot_stride, int outputs_are_zero, float* dL_dmeans3D,
                              float* dL_dopacity, float* dL_dscales, float* dL_drotations, float* dL_dsh, float* dL_dsh_rest) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (P < 0 || D < 0 || D > 3 || V < 1) return GS_ERR_ARG;
    if (P == 0) return GS_OK;
    if (!means3D || !campos_views || !records || !dL_dmeans3D || !dL_dopacity || !dL_dscales || !dL_drotations || !dL_dsh)
        return GS_ERR_ARG;
    if (slot_stride < gs_dp_slot_floats(P, GS_DP_RECORDS) || (slot_stride & 3) || (reinterpret_cast<size_t>(records) & 15))
        return GS_ERR_ARG;
    prof_mark(stream, -1);
    NvtxScope nv(ST_DP_COMBINE);
    if (!outputs_are_zero) {
        GS_CUDA_TRY(cudaMemsetAsync(dL_dmeans3D, 0, (size_t)P * 3 * sizeof(float), stream));
        GS_CUDA_TRY(cudaMemsetAsync(dL_dopacity, 0, (size_t)P * sizeof(float), stream));
        GS_CUDA_TRY(cudaMemsetAsync(dL_dscales, 0, (size_t)P * 3 * sizeof(float), stream));
        GS_CUDA_TRY(cudaMemsetAsync(dL_drotations, 0, (size_t)P * 4 * sizeof(float), stream));
        GS_CUDA_TRY(cudaMemsetAsync(dL_dsh, 0, (size_t)P * (dL_dsh_rest ? 1 : M) * 3 * sizeof(float), stream));
        if (dL_dsh_rest) GS_CUDA_TRY(cudaMemsetAsync(dL_dsh_rest, 0, (size_t)P * (M - 1) * 3 * sizeof(float), stream));
    }
    const int rc = launch_grad_record_combine(P, D, M, V, means3D, campos_views, records, slot_stride, dL_dmeans3D, dL_dopacity,
                                              dL_dscales, dL_drotations, dL_dsh, dL_dsh_rest, stream);
    prof_mark(stream, ST_DP_COMBINE);
    count_launches(outputs_are_zero ? 1 : 6);
    return rc;
}

int gs_dp_color_grad_push(void* stream_, int P, const int* radii, const void* geom_ws, const void* scratch,
                          float* const* dst_ptrs, int n_dst, int multicast, size_t slot_offset, int layout) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (P < 0 || (slot_offset & 3) || (layout != GS_DP_DENSE && layout != GS_DP_SPARSE)) return GS_ERR_ARG;
    if (P == 0) return GS_OK;
    if (!radii || !geom_ws || !scratch) return GS_ERR_ARG;
    DpPeers peers;
    int rc = fill_peers(&peers, dst_ptrs, n_dst);
    if (rc) return rc;
    if (multicast && n_dst != 1) return GS_ERR_ARG;
    GeomWS g = GeomWS::carve(aligned_base(geom_ws), (size_t)P);
    BwdScratch sc = BwdScratch::carve(aligned_base(scratch), (size_t)P);
    prof_mark(stream, -1);
    NvtxScope nv(ST_DP_PUSH);
    if (layout == GS_DP_SPARSE) {
        int dev = 0, sms = 0;
        GS_CUDA_TRY(cudaGetDevice(&dev));
        GS_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        rc = launch_color_grad_push_sparse(P, radii, g.clamped, sc.acc_color_opacity, peers, slot_offset, multicast != 0, sms, stream);
    } else {
        rc = launch_color_grad_push(P, radii, g.clamped, sc.acc_color_opacity, peers, slot_offset, multicast != 0, stream);
    }
    prof_mark(stream, ST_DP_PUSH);
    count_launches(1);
    return rc;
}

int gs_dp_sh_grad_combine(void* stream_, int P, int D, int M, int V, const float* means3D, const float* campos_views,
                          const float* color_grads, size_t slot_stride, int layout, float* dL_dsh) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (P < 0 || D < 0 || D > 3 || V < 1 || (layout != GS_DP_DENSE && layout != GS_DP_SPARSE)) return GS_ERR_ARG;
    if (P == 0) return GS_OK;
    if (!means3D || !campos_views || !color_grads || !dL_dsh) return GS_ERR_ARG;
    if (slot_stride < (layout == GS_DP_SPARSE ? dp_sparse_slot_floats(P) : (size_t)3 * P)) return GS_ERR_ARG;
    if (layout == GS_DP_SPARSE && ((reinterpret_cast<size_t>(color_grads) & 15) || (slot_stride & 3))) return GS_ERR_ARG;
    prof_mark(stream, -1);
    NvtxScope nv(ST_DP_COMBINE);
    const int rc = launch_sh_grad_combine(P, D, M, V, means3D, campos_views, color_grads, slot_stride, layout == GS_DP_SPARSE,
                                          dL_dsh, stream);
    prof_mark(stream, ST_DP_COMBINE);
    count_launches(1);
    return rc;
}

int gs_dp_slab_all_reduce(void* stream_, float* const* slab_ptrs, int n_ptrs, int multicast, size_t n_floats, int rank,
                          int world) {
    cudaStream_t stream = (cudaStream_t)stream_;
    DpPeers peers;
    int rc = fill_peers(&peers, slab_ptrs, n_ptrs);
    if (rc) return rc;
    if (multicast ? n_ptrs != 1 : n_ptrs != world) return GS_ERR_ARG;
    int dev = 0, sms = 0;
    GS_CUDA_TRY(cudaGetDevice(&dev));
    GS_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    prof_mark(stream, -1);
    NvtxScope nv(ST_DP_ALL_REDUCE);
    rc = launch_slab_all_reduce(peers, multicast != 0, n_floats, rank, world, sms, stream);
    prof_mark(stream, ST_DP_ALL_REDUCE);
    count_launches(1);
    return rc;
}

int gs_adam_step(void* stream_, int n_tensors, float* const* params, const float* const* grads, float* const* exp_avg,
                 float* const* exp_avg_sq, const unsigned long long* numel, const double* lr, const int* step, double beta1,
                 double beta2, double eps) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (n_tensors < 0 || n_tensors > 8) return GS_ERR_ARG;
    if (n_tensors == 0) return GS_OK;
    if (!params || !grads || !exp_avg || !exp_avg_sq || !numel || !lr || !step) return GS_ERR_ARG;
    AdamLaunch L;
    L.n_tensors = 0;
    // torch computes these scalars in double and hands them to float kernels (torch/optim/adam.py, non-capturable path)
    L.w1 = (float)(1.0 - beta1);
    L.beta2 = (float)beta2;
    L.w2 = (float)(1.0 - beta2);
    L.eps = (float)eps;
    for (int i = 0; i < n_tensors; ++i) {
        if (numel[i] == 0) continue;
        if (!params[i] || !grads[i] || !exp_avg[i] || !exp_avg_sq[i] || step[i] < 1) return GS_ERR_ARG;
        AdamTensor& T = L.t[L.n_tensors++];
        T.param = params[i]; T.grad = grads[i]; T.m = exp_avg[i]; T.v = exp_avg_sq[i]; T.n = numel[i];
        const double bc1 = 1.0 - pow(beta1, (double)step[i]);
        const double bc2 = 1.0 - pow(beta2, (double)step[i]);
        T.neg_step_size = (float)(-(lr[i] / bc1));
        T.bc2_sqrt = (float)sqrt(bc2);
        T.vec_ok = 0;
    }
    prof_mark(stream, -1);
    NvtxScope nv(ST_ADAM);
    const int rc = launch_adam_step(L, stream);
    prof_mark(stream, ST_ADAM);
    count_launches(1);
    return rc;
}

namespace {
struct LossWS {
    float *d_mu, *d_e11, *d_e12, *partial;
    size_t bytes;
    static LossWS carve(char* base, int C, int H, int W) {
        Carver c(base);
        LossWS w;
        const size_t n = (size_t)C * H * W;
        w.d_mu = c.take<float>(n);
        w.d_e11 = c.take<float>(n);
        w.d_e12 = c.take<float>(n);
        w.partial = c.take<float>(2 * loss_partial_count(C, H, W));
        w.bytes = align_up(c.off, kAlign);
        return w;
    }
};
int fill_loss_params(LossParams* p, int C, int H, int W, const float* image, const float* gt, const void* ws) {
    if (C <= 0 || H <= 0 || W <= 0 || C > 65535 || !image || !gt || !ws) return GS_ERR_ARG;
    LossWS w = LossWS::carve(aligned_base(ws), C, H, W);
    p->C = C; p->H = H; p->W = W; p->img = image; p->gt = gt;
    p->d_mu = w.d_mu; p->d_e11 = w.d_e11; p->d_e12 = w.d_e12; p->partial = w.partial; p->dL_dimg = nullptr;
    // utils/loss_utils.py:27-29: torch.Tensor([exp(-(x - 5)^2 / (2 sigma^2))]) (python doubles -> FP32), divided by its FP32 sum
    // (torch's sum of the 11 floats is the correctly rounded one: accumulate in double, round once)
    float g[11];
    double sum = 0.0;
    for (int i = 0; i < 11; ++i) { g[i] = (float)exp(-(double)((i - 5) * (i - 5)) / (2.0 * 1.5 * 1.5)); sum += (double)g[i]; }
    for (int i = 0; i < 11; ++i) p->win[i] = g[i] / (float)sum;
    return GS_OK;
}
}  // namespace

size_t gs_loss_workspace_bytes(int C, int H, int W) {
    if (C <= 0 || H <= 0 || W <= 0) return kAlign;
    return LossWS::carve(nullptr, C, H, W).bytes + kAlign;
}

int gs_photometric_loss_forward(void* stream_, int C, int H, int W, const float* image, const float* gt, void* ws,
                                float* out_means) {
    cudaStream_t stream = (cudaStream_t)stream_;
    LossParams p;
    int rc = fill_loss_params(&p, C, H, W, image, gt, ws);
    if (rc) return rc;
    if (!out_means) return GS_ERR_ARG;
    prof_mark(stream, -1);
    NvtxScope nv(ST_LOSS_FWD);
    rc = launch_loss_forward(p, out_means, stream);
    prof_mark(stream, ST_LOSS_FWD);
    count_launches(2);
    return rc;
}

int gs_photometric_loss_backward(void* stream_, int C, int H, int W, const float* image, const float* gt, const void* ws,
                                 float c_l1, float c_ssim, const float* grad_scale, float* dL_dimage) {
    cudaStream_t stream = (cudaStream_t)stream_;
    LossParams p;
    int rc = fill_loss_params(&p, C, H, W, image, gt, ws);
    if (rc) return rc;
    if (!dL_dimage) return GS_ERR_ARG;
    p.dL_dimg = dL_dimage;
    prof_mark(stream, -1);
    NvtxScope nv(ST_LOSS_BWD);
    rc = launch_loss_backward(p, c_l1, c_ssim, grad_scale, stream);
    prof_mark(stream, ST_LOSS_BWD);
    count_launches(1);
    return rc;
}

size_t gs_l1_workspace_bytes(size_t n) { return align_up(2 * sizeof(float) * (size_t)l1_block_count(n), kAlign) + kAlign; }

int gs_l1_loss_forward(void* stream_, size_t n, const float* x, const float* y, void* ws, float* out_mean) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (n == 0 || !x || !y || !ws || !out_mean) return GS_ERR_ARG;
    float* partial = reinterpret_cast<float*>(align_up(reinterpret_cast<size_t>(ws), kAlign));
    prof_mark(stream, -1);
    NvtxScope nv(ST_LOSS_FWD);
    const int rc = launch_l1_forward(n, x, y, partial, out_mean, stream);
    prof_mark(stream, ST_LOSS_FWD);
    count_launches(2);
    return rc;
}

int gs_l1_loss_backward(void* stream_, size_t n, const float* x, const float* y, const float* grad_scale, float* dL_dx) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (n == 0 || !x || !y || !dL_dx) return GS_ERR_ARG;
    prof_mark(stream, -1);
    NvtxScope nv(ST_LOSS_BWD);
    const int rc = launch_l1_backward(n, x, y, grad_scale, dL_dx, stream);
    prof_mark(stream, ST_LOSS_BWD);
    count_launches(1);
    return rc;
}

size_t gs_activation_workspace_bytes(int P) {
    return align_up(2 * sizeof(float) * activate_partial_count(P > 0 ? P : 0), kAlign) + kAlign;
}

int gs_activate_forward(void* stream_, int P, const float* raw_opacity, const float* raw_scaling, const float* raw_rotation,
                        float* opacity, float* scaling, float* rotation, void* ws, float* out_means) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (P < 0) return GS_ERR_ARG;
    if (P == 0) return GS_OK;
    if (!raw_opacity || !raw_scaling || !raw_rotation || !opacity || !scaling || !rotation || !ws || !out_means) return GS_ERR_ARG;
    if ((reinterpret_cast<size_t>(raw_rotation) | reinterpret_cast<size_t>(rotation)) & 15) return GS_ERR_ARG;
    prof_mark(stream, -1);
    NvtxScope nv(ST_ACT_FWD);
    const int rc = launch_activate_fwd(P, raw_opacity, raw_scaling, raw_rotation, opacity, scaling, rotation,
                                       reinterpret_cast<float*>(aligned_base(ws)), out_means, stream);
    prof_mark(stream, ST_ACT_FWD);
    count_launches(2);
    return rc;
}

int gs_activate_backward(void* stream_, int P, const float* raw_opacity, const float* raw_scaling, const float* raw_rotation,
                         const float* g_opacity, const float* g_scaling, const float* g_rotation, float c_opacity,
                         float c_scaling, const float* reg_grad, float* d_raw_opacity, float* d_raw_scaling,
                         float* d_raw_rotation) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (P < 0) return GS_ERR_ARG;
    if (P == 0) return GS_OK;
    if (!raw_opacity || !raw_scaling || !raw_rotation || !d_raw_opacity || !d_raw_scaling || !d_raw_rotation) return GS_ERR_ARG;
    if ((reinterpret_cast<size_t>(raw_rotation) | reinterpret_cast<size_t>(d_raw_rotation) | reinterpret_cast<size_t>(g_rotation)) & 15)
        return GS_ERR_ARG;
    prof_mark(stream, -1);
    NvtxScope nv(ST_ACT_BWD);
    const int rc = launch_activate_bwd(P, raw_opacity, raw_scaling, raw_rotation, g_opacity, g_scaling, g_rotation, c_opacity,
                                       c_scaling, reg_grad, d_raw_opacity, d_raw_scaling, d_raw_rotation, stream);
    prof_mark(stream, ST_ACT_BWD);
    count_launches(1);
    return rc;
}

namespace {
int fill_swin_set(SwinSet* s, const float* const* ptrs, bool required) {
    if (!ptrs) return required ? GS_ERR_ARG : GS_OK;
    const float** dst[10] = {&s->xyz, &s->rot, &s->dc, &s->rest, &s->opacity, &s->scaling, &s->v, &s->rotvec, &s->cen, &s->start};
    for (int i = 0; i < 10; ++i) {
        if (required && !ptrs[i]) return GS_ERR_ARG;
        *dst[i] = ptrs[i];
    }
    if (reinterpret_cast<size_t>(s->rot) & 15) return GS_ERR_ARG;
    return GS_OK;
}
int fill_swin_params(SwinParams* p, int Ki, int Km, int M, float frame, int deform, const long long* imm_idx,
                     const long long* mat_idx, const float* const* imm_set, const float* const* mat_set) {
    if (Ki < 0 || Km < 0 || M < 1) return GS_ERR_ARG;
    SwinSet zero = {};
    p->Ki = Ki; p->Km = Km; p->M = M; p->frame = frame; p->deform = deform; p->imm_idx = imm_idx; p->mat_idx = mat_idx;
    p->imm = zero; p->mat = zero;
    if ((Ki > 0 && !imm_idx) || (Km > 0 && !mat_idx)) return GS_ERR_ARG;
    int rc = fill_swin_set(&p->imm, imm_set, Ki > 0);
    if (rc) return rc;
    return fill_swin_set(&p->mat, mat_set, Km > 0);
}
}  // namespace

int gs_swin_gather_forward(void* stream_, int Ki, int Km, int M, float frame, int deform, const long long* imm_idx,
                           const long long* mat_idx, const float* const* imm_set, const float* const* mat_set,
                           float* out_xyz, float* out_rotation, float* out_feature, float* out_opacity, float* out_scaling) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SwinParams p;
    int rc = fill_swin_params(&p, Ki, Km, M, frame, deform, imm_idx, mat_idx, imm_set, mat_set);
    if (rc) return rc;
    if (Ki + Km == 0) return GS_OK;
    if (!out_xyz || !out_rotation || !out_feature || !out_opacity || !out_scaling || (reinterpret_cast<size_t>(out_rotation) & 15))
        return GS_ERR_ARG;
    p.out_xyz = out_xyz; p.out_rot = out_rotation; p.out_feat = out_feature; p.out_opacity = out_opacity; p.out_scaling = out_scaling;
    prof_mark(stream, -1);
    NvtxScope nv(ST_SWIN_FWD);
    rc = launch_swin_gather_fwd(p, stream);
    prof_mark(stream, ST_SWIN_FWD);
    count_launches(1);
    return rc;
}

int gs_swin_gather_backward(void* stream_, int Ki, int M, float frame, int deform, const long long* imm_idx,
                            const float* const* imm_set, const float* const* out_grads, float* const* imm_grads) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SwinParams p;
    int rc = fill_swin_params(&p, Ki, 0, M, frame, deform, imm_idx, nullptr, imm_set, nullptr);
    if (rc) return rc;
    if (Ki == 0) return GS_OK;
    if (!out_grads || !imm_grads) return GS_ERR_ARG;
    p.out_xyz = p.out_rot = p.out_feat = p.out_opacity = p.out_scaling = nullptr;
    SwinGrads g;
    g.g_xyz = out_grads[0]; g.g_rot = out_grads[1]; g.g_feat = out_grads[2]; g.g_opacity = out_grads[3]; g.g_scaling = out_grads[4];
    for (int i = 0; i < 9; ++i)
        if (!imm_grads[i]) return GS_ERR_ARG;
    g.d_xyz = imm_grads[0]; g.d_rot = imm_grads[1]; g.d_dc = imm_grads[2]; g.d_rest = imm_grads[3]; g.d_opacity = imm_grads[4];
    g.d_scaling = imm_grads[5]; g.d_v = imm_grads[6]; g.d_rotvec = imm_grads[7]; g.d_cen = imm_grads[8];
    if ((reinterpret_cast<size_t>(g.d_rot) | reinterpret_cast<size_t>(g.g_rot)) & 15) return GS_ERR_ARG;
    prof_mark(stream, -1);
    NvtxScope nv(ST_SWIN_BWD);
    rc = launch_swin_gather_bwd(p, g, stream);
    prof_mark(stream, ST_SWIN_BWD);
    count_launches(1);
    return rc;
}

size_t gs_knn_workspace_bytes(int P) { return knn_workspace_bytes(P) + kAlign; }

int gs_knn_mean_dist2(void* stream_, int P, const float* points, float* mean_dist2, void* ws) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (P < 0) return GS_ERR_ARG;
    if (P == 0) return GS_OK;
    if (!points || !mean_dist2 || !ws) return GS_ERR_ARG;
    pdl_hint(P);
    prof_mark(stream, -1);
    NvtxScope nv(ST_KNN);
    const int rc = launch_knn_mean_dist2(P, points, mean_dist2, aligned_base(ws), stream);
    prof_mark(stream, ST_KNN);
    count_launches(6 + 3 * 4);
    return rc;
}

void* gs_host_event_create(void) {
    cudaEvent_t ev = nullptr;
    if (cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    return (void*)ev;
}
int gs_host_event_wait(void* event) {
    if (!event) return GS_ERR_ARG;
    GS_CUDA_TRY(cudaEventSynchronize((cudaEvent_t)event));
    return GS_OK;
}
int gs_host_event_destroy(void* event) {
    if (!event) return GS_OK;
    GS_CUDA_TRY(cudaEventDestroy((cudaEvent_t)event));
    return GS_OK;
}

int gs_mark_visible(void* stream_, int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                    unsigned char* present) {
    (void)projmatrix;  // the reference's frustum test only uses the view-space depth
    if (P < 0) return GS_ERR_ARG;
    if (P > 0 && (!means3D || !viewmatrix || !present)) return GS_ERR_ARG;
    return launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream_);
}

int gs_compute_relocation(void* stream_, int P, const float* opacity_old, const float* scale_old, const int* N,
                          const float* binoms, int n_max, float* opacity_new, float* scale_new) {
    if (P < 0 || n_max <= 0) return GS_ERR_ARG;
    if (P > 0 && (!opacity_old || !scale_old || !N || !binoms || !opacity_new || !scale_new)) return GS_ERR_ARG;
    prof_mark((cudaStream_t)stream_, -1);
    NvtxScope nv(ST_RELOCATION);
    int rc = launch_compute_relocation(P, opacity_old, scale_old, N, binoms, n_max, opacity_new, scale_new, (cudaStream_t)stream_);
    prof_mark((cudaStream_t)stream_, ST_RELOCATION);
    count_launches(P > 0 ? 1 : 0);
    return rc;
}

int gs_add_noise(void* stream_, int P, float* xyz, const float* scaling_raw, const float* rotation_raw,
                 const float* opacity_raw, const float* noise, float lr_scale, unsigned long long seed,
                 unsigned long long offset) {
    if (P < 0) return GS_ERR_ARG;
    if (P > 0 && (!xyz || !scaling_raw || !rotation_raw || !opacity_raw)) return GS_ERR_ARG;
    prof_mark((cudaStream_t)stream_, -1);
    NvtxScope nv(ST_ADD_NOISE);
    int rc = launch_add_noise(P, xyz, scaling_raw, rotation_raw, opacity_raw, noise, lr_scale, seed, offset, (cudaStream_t)stream_);
    prof_mark((cudaStream_t)stream_, ST_ADD_NOISE);
    count_launches(P > 0 ? 1 : 0);
    return rc;
}

long long gs_debug_fetch(int which, int P, unsigned int R, int W, int H, const void* geom_ws, const void* binning_ws,
                         const void* image_ws, void* dst, size_t max_bytes) {
    const void* src = nullptr;
    size_t n = 0;
    uint2 grid;
    const uint32_t tiles = tiles_of(W, H, &grid);
    const size_t Ps = (size_t)P, Rs = (size_t)R, Npix = (size_t)W * (size_t)H;
    if (which < 10 || which == 31) {
        if (!geom_ws) return -1;
        GeomWS g = GeomWS::carve(aligned_base(geom_ws), Ps);
        switch (which) {
            case 31: src = g.num_rendered + 1; n = 4; break;
            case 0: case 3: case 5: case 6: case 7: {
                // the reference's per-Gaussian arrays, unpacked from the render records / tile rectangles (zeros for
                // Gaussians that are not visible: the forward writes no record for them)
                static const size_t width[8] = {4, 0, 0, 8, 0, 16, 12, 4};
                n = Ps * width[which];
                if (dst && n > 0 && n <= max_bytes) {
                    if (cudaDeviceSynchronize() != cudaSuccess) return -1;
                    if (launch_debug_unpack(which, (uint32_t)P, g.rec, g.rect, dst)) return -1;
                }
                return (long long)n;
            }
            case 1: src = g.clamped; n = Ps; break;
            case 4: src = g.cov3D; n = Ps * 24; break;
            case 8:   // inclusive scan of the tile counts in depth order; entries of the culled Gaussians (last) repeat R
            case 9: { // Gaussian indices in (depth, index) order; culled ones last
                n = Ps * 4;
                if (dst && n > 0 && n <= max_bytes) {
                    if (cudaDeviceSynchronize() != cudaSuccess) return -1;
                    uint32_t cnt[4] = {0, 0, 0, 0};
                    if (cudaMemcpy(cnt, g.num_rendered, sizeof(cnt), cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
                    if (launch_debug_depth_order((uint32_t)P, cnt[2], cnt[0], g.dkeys_a, g.rect, g.point_offsets,
                                                 which == 9 ? (uint32_t*)dst : nullptr, which == 8 ? (uint32_t*)dst : nullptr)) return -1;
                }
                return (long long)n;
            }
            case 2: src = g.rec; n = Ps * sizeof(SplatRec); break;
            // (case 9: above)
            default: return -1;
        }
    } else if (which < 20 || which == 30) {
        if (!binning_ws) return -1;
        BinningWS b = BinningWS::carve(aligned_base(binning_ws), Rs, tiles, Ps);
        if (which == 12 || which == 13) {
            // Views of the sorted list in the reference's key format (tile << 32 | depth_bits) / as raw instances
            // (tile << 32 | idx), rebuilt from the tile ranges and the sorted ids.
            n = Rs * 8;
            if (dst && n > 0 && n <= max_bytes) {
                if (!geom_ws || !image_ws) return -1;
                GeomWS g = GeomWS::carve(aligned_base(geom_ws), Ps);
                ImageWS im = ImageWS::carve(aligned_base(image_ws), (size_t)W, (size_t)H);
                if (cudaDeviceSynchronize() != cudaSuccess) return -1;
                if (launch_debug_expand(tiles, im.ranges, b.ids, g.rec, which == 12 ? (uint64_t*)dst : nullptr,
                                        which == 13 ? (uint64_t*)dst : nullptr))
                    return -1;
            }
            return (long long)n;
        }
        switch (which) {
            case 10: src = b.ids; n = Rs * 4; break;   // sorted Gaussian ids (the reference's point_list)
            case 30: src = b.error_flag; n = 4; break;
            default: return -1;
        }
    } else {
        if (!image_ws) return -1;
        ImageWS s = ImageWS::carve(aligned_base(image_ws), (size_t)W, (size_t)H);
        switch (which) {
            case 20: src = s.final_T; n = Npix * 4; break;
            case 21: src = s.n_contrib; n = Npix * 4; break;
            case 22: src = s.ranges; n = (size_t)tiles * 8; break;
            case 23: src = s.tile_max_contrib; n = (size_t)tiles * 4; break;
            default: return -1;
        }
    }
    if (dst && n > 0 && n <= max_bytes) {
        if (cudaMemcpy(dst, src, n, cudaMemcpyDeviceToDevice) != cudaSuccess) return -1;
    }
    return (long long)n;
}

}  // extern "C"
