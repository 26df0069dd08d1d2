// distCUDA2 for sm_100a: for every point, the mean of the squared distances to its three nearest
// OTHER points. Replaces submodules/simple-knn (simple_knn.cu:45-220, spatial.cu:15-26), which the model calls once at
// initialisation (scene/gaussian_model_static.py:168) and which must exist for the model class to import at all.
//
// The reference sorts the points along a 30-bit Morton curve, cuts the order into boxes of 1024 and lets every point
// test EVERY box (P x P/1024 box tests, then a scan of each surviving box). Same result, different search: the
// Morton-sorted order is a linear octree (a cell of level L is a contiguous run of codes sharing a 3L-bit prefix, found
// by binary search), so each point
//   1. bounds its third-nearest distance r with members of the finest cell around it that holds three other points,
//   2. walks the octree depth first from the root, dropping every cell whose box is farther than the current
//      third-best distance (never an overestimate, so nothing nearer is lost), scanning small cells, splitting large ones.
// Work is O(P log P) instead of O(P^2 / 1024). The three distances are exact, are evaluated with the reference's
// expression (d = other - ref; d.x d.x + d.y d.y + d.z d.z under nvcc's default contraction) and are averaged the same
// way ((b0 + b1 + b2) / 3.0f in ascending order), so the output is bit-identical to the reference's.
// Like the reference, a point is excluded only by its own position in the order: coincident points count with distance 0.
#include <cfloat>

#include "gs_common.cuh"
#include "kernels.cuh"
#include "workspace.cuh"

namespace gs {

constexpr int kKnnThreads = 256;
constexpr int kKnnLevels = 10;      // 10 bits per axis

struct KnnBounds {
    float3 minn, maxx;
};

__global__ void __launch_bounds__(kKnnThreads) knn_bounds_partial_kernel(int P, const float* __restrict__ pts, KnnBounds* __restrict__ partial) {
    __shared__ KnnBounds s[kKnnThreads / 32];
    float3 lo = {FLT_MAX, FLT_MAX, FLT_MAX}, hi = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
        const float x = pts[3 * (size_t)i], y = pts[3 * (size_t)i + 1], z = pts[3 * (size_t)i + 2];
        lo.x = fminf(lo.x, x); lo.y = fminf(lo.y, y); lo.z = fminf(lo.z, z);
        hi.x = fmaxf(hi.x, x); hi.y = fmaxf(hi.y, y); hi.z = fmaxf(hi.z, z);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo.x = fminf(lo.x, __shfl_xor_sync(0xffffffffu, lo.x, o)); lo.y = fminf(lo.y, __shfl_xor_sync(0xffffffffu, lo.y, o));
        lo.z = fminf(lo.z, __shfl_xor_sync(0xffffffffu, lo.z, o)); hi.x = fmaxf(hi.x, __shfl_xor_sync(0xffffffffu, hi.x, o));
        hi.y = fmaxf(hi.y, __shfl_xor_sync(0xffffffffu, hi.y, o)); hi.z = fmaxf(hi.z, __shfl_xor_sync(0xffffffffu, hi.z, o));
    }
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = {lo, hi};
    __syncthreads();
    if (threadIdx.x == 0) {
        KnnBounds b = s[0];
        for (int w = 1; w < kKnnThreads / 32; ++w) {
            b.minn.x = fminf(b.minn.x, s[w].minn.x); b.minn.y = fminf(b.minn.y, s[w].minn.y); b.minn.z = fminf(b.minn.z, s[w].minn.z);
            b.maxx.x = fmaxf(b.maxx.x, s[w].maxx.x); b.maxx.y = fmaxf(b.maxx.y, s[w].maxx.y); b.maxx.z = fmaxf(b.maxx.z, s[w].maxx.z);
        }
        partial[blockIdx.x] = b;
    }
}

__global__ void knn_bounds_final_kernel(int n, const KnnBounds* __restrict__ partial, KnnBounds* __restrict__ out) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    KnnBounds b = partial[0];
    for (int i = 1; i < n; ++i) {
        b.minn.x = fminf(b.minn.x, partial[i].minn.x); b.minn.y = fminf(b.minn.y, partial[i].minn.y); b.minn.z = fminf(b.minn.z, partial[i].minn.z);
        b.maxx.x = fmaxf(b.maxx.x, partial[i].maxx.x); b.maxx.y = fmaxf(b.maxx.y, partial[i].maxx.y); b.maxx.z = fmaxf(b.maxx.z, partial[i].maxx.z);
    }
    *out = b;
}

GS_DEV uint32_t spread3(uint32_t x) {   // 10 bits -> every third bit
    x = (x | (x << 16)) & 0x030000FFu;
    x = (x | (x << 8)) & 0x0300F00Fu;
    x = (x | (x << 4)) & 0x030C30C3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}
GS_DEV uint32_t quantise(float v, float lo, float hi) {
    const float ext = hi - lo;
    if (!(ext > 0.f)) return 0u;                       // degenerate axis: one cell
    const float t = (v - lo) / ext * 1023.f;
    return (uint32_t)min(1023, max(0, (int)t));
}

__global__ void __launch_bounds__(kKnnThreads) knn_keys_kernel(int P, const float* __restrict__ pts, const KnnBounds* __restrict__ bounds,
                                                               uint64_t* __restrict__ keys) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const KnnBounds b = *bounds;
    const uint32_t qx = quantise(pts[3 * (size_t)i], b.minn.x, b.maxx.x);
    const uint32_t qy = quantise(pts[3 * (size_t)i + 1], b.minn.y, b.maxx.y);
    const uint32_t qz = quantise(pts[3 * (size_t)i + 2], b.minn.z, b.maxx.z);
    const uint32_t code = spread3(qx) | (spread3(qy) << 1) | (spread3(qz) << 2);
    keys[i] = ((uint64_t)code << 32) | (uint32_t)i;
}

// Sorted keys -> code[] (u32), and the points gathered into curve order as float4 (x, y, z, original index bits).
__global__ void __launch_bounds__(kKnnThreads) knn_gather_kernel(int P, const uint64_t* __restrict__ keys, const float* __restrict__ pts,
                                                                 uint32_t* __restrict__ codes, float4* __restrict__ sorted) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const uint64_t k = keys[i];
    const uint32_t id = (uint32_t)k;
    codes[i] = (uint32_t)(k >> 32);
    sorted[i] = make_float4(pts[3 * (size_t)id], pts[3 * (size_t)id + 1], pts[3 * (size_t)id + 2], __uint_as_float(id));
}

GS_DEV uint32_t lower_bound_u32(const uint32_t* __restrict__ a, uint32_t n, uint32_t v) {   // first i with a[i] >= v
    uint32_t lo = 0, hi = n;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (__ldg(a + mid) < v) lo = mid + 1; else hi = mid;
    }
    return lo;
}

GS_DEV void k_best3(float dist, float* best) {   // simple_knn.cu updateKBest<3>
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        if (best[j] > dist) {
            const float t = best[j];
            best[j] = dist;
            dist = t;
        }
    }
}

GS_DEV float quantised_coord(float v, float lo, float hi) {   // the real-valued cell coordinate quantise() truncates
    const float ext = hi - lo;
    if (!(ext > 0.f)) return 0.f;
    return fminf(fmaxf((v - lo) / ext * 1023.f, 0.f), 1023.f);
}

__global__ void __launch_bounds__(kKnnThreads) knn_search_kernel(int P, const uint32_t* __restrict__ codes, const float4* __restrict__ sorted,
                                                                 const KnnBounds* __restrict__ bounds, float* __restrict__ out) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const KnnBounds bd = *bounds;
    const float4 me = sorted[idx];
    const uint32_t code = codes[idx];
    const float ext[3] = {bd.maxx.x - bd.minn.x, bd.maxx.y - bd.minn.y, bd.maxx.z - bd.minn.z};
    const float tp[3] = {quantised_coord(me.x, bd.minn.x, bd.maxx.x), quantised_coord(me.y, bd.minn.y, bd.maxx.y),
                         quantised_coord(me.z, bd.minn.z, bd.maxx.z)};
    float unit[3];   // real length of one finest-level cell per axis (0 on a degenerate axis)
#pragma unroll
    for (int a = 0; a < 3; ++a) unit[a] = ext[a] > 0.f ? ext[a] * (1.0f / 1023.f) : 0.f;
    // 1. an upper bound on the third-nearest distance: climb to the finest level at which the point's own cell holds
    //    at least three other points and take the three best of (up to 96 of) its members nearest along the curve.
    //    Any three other points bound the answer; these are spatially close by construction. (The reference bounds with
    //    the six neighbours along the curve only, simple_knn.cu:147-156, which can sit across a jump of the curve.)
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    {
        uint32_t clo = 0, chi = (uint32_t)P;
        for (int s = 0; s <= kKnnLevels; ++s) {
            const uint32_t prefix = code >> (3 * s);
            clo = lower_bound_u32(codes, (uint32_t)P, prefix << (3 * s));
            const uint64_t end = ((uint64_t)prefix + 1ull) << (3 * s);
            chi = end > 0x3FFFFFFFull ? (uint32_t)P : lower_bound_u32(codes, (uint32_t)P, (uint32_t)end);
            if (chi - clo >= 4u) break;
        }
        const uint32_t w0i = (uint32_t)max((int)clo, idx - 48), w1i = min(chi, (uint32_t)idx + 49u);
        for (uint32_t i = w0i; i < w1i; ++i) {
            if ((int)i == idx) continue;
            const float4 o = sorted[i];
            const float3 d = {o.x - me.x, o.y - me.y, o.z - me.z};
            k_best3(d.x * d.x + d.y * d.y + d.z * d.z, best);
        }
    }
    const float reject = best[2];
    best[0] = best[1] = best[2] = FLT_MAX;
    // 2. depth-first walk of the implicit octree from the root with that bound: a cell is dropped when its box is
    //    farther than the current third-best distance (box distance evaluated in cell units with a 1e-3-cell allowance
    //    for the rounding of the quantisation, so it never overestimates), scanned when it is small, split into its
    //    eight children (sub-ranges found by binary search inside the parent's range) otherwise. The child holding the
    //    query is visited first.
    constexpr int kStack = 80;            // <= 7 siblings per level x 10 levels + 8
    constexpr uint32_t kLeaf = 48;        // scan cells of at most this many points
    uint32_t st_first[kStack], st_lo[kStack], st_hi[kStack];
    int st_sh[kStack];
    int sp = 0;
    st_first[0] = 0u; st_lo[0] = 0u; st_hi[0] = (uint32_t)P; st_sh[0] = kKnnLevels; sp = 1;
    while (sp > 0) {
        --sp;
        const uint32_t first = st_first[sp], lo = st_lo[sp], hi = st_hi[sp];
        const int sh = st_sh[sp];
        float box2 = 0.f;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            uint32_t x = (first >> a) & 0x09249249u;
            x = (x | (x >> 2)) & 0x030C30C3u;
            x = (x | (x >> 4)) & 0x0300F00Fu;
            x = (x | (x >> 8)) & 0x030000FFu;
            x = (x | (x >> 16)) & 0x000003FFu;
            const float c0 = (float)x, c1 = (float)(x + (1u << sh));
            float dt = tp[a] < c0 ? c0 - tp[a] : (tp[a] >= c1 ? tp[a] - c1 : 0.f);
            dt = fmaxf(dt - 1e-3f, 0.f) * unit[a] * 0.99999f;
            box2 = fmaf(dt, dt, box2);
        }
        if (box2 > reject || box2 > best[2]) continue;
        if (hi - lo <= kLeaf || sh == 0) {
            for (uint32_t i = lo; i < hi; ++i) {
                if ((int)i == idx) continue;
                const float4 o = sorted[i];
                const float3 d = {o.x - me.x, o.y - me.y, o.z - me.z};
                const float dist = d.x * d.x + d.y * d.y + d.z * d.z;
                k_best3(dist, best);
            }
            continue;
        }
        const int csh = sh - 1;
        const uint32_t step = 1u << (3 * csh);
        const bool mine = (code >> (3 * sh)) == (first >> (3 * sh));
        const int kq = mine ? (int)((code >> (3 * csh)) & 7u) : -1;
        uint32_t edge[9];
        edge[0] = lo;
        edge[8] = hi;
#pragma unroll
        for (int k = 1; k < 8; ++k) edge[k] = edge[k - 1] + lower_bound_u32(codes + edge[k - 1], hi - edge[k - 1], first + (uint32_t)k * step);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            if (k == kq || edge[k + 1] == edge[k]) continue;
            st_first[sp] = first + (uint32_t)k * step; st_lo[sp] = edge[k]; st_hi[sp] = edge[k + 1]; st_sh[sp] = csh; ++sp;
        }
        if (kq >= 0 && edge[kq + 1] > edge[kq]) {
            st_first[sp] = first + (uint32_t)kq * step; st_lo[sp] = edge[kq]; st_hi[sp] = edge[kq + 1]; st_sh[sp] = csh; ++sp;
        }
    }
    out[__float_as_uint(me.w)] = (best[0] + best[1] + best[2]) / 3.0f;
}

size_t knn_workspace_bytes(int P) {
    Carver c(nullptr);
    const size_t n = (size_t)(P > 0 ? P : 0);
    c.take<uint64_t>(n); c.take<uint64_t>(n); c.take<uint32_t>(n); c.take<float4>(n);
    c.take<KnnBounds>(1024 + 1);
    c.take<uint32_t>(256 * ((n + kBinTile - 1) / kBinTile) + 256 + 64);   // + the depth sort's done counter
    return align_up(c.off, kAlign);
}

int launch_knn_mean_dist2(int P, const float* points, float* out, char* ws, cudaStream_t stream) {
    if (P <= 0) return GS_OK;
    Carver c(ws);
    const size_t n = (size_t)P;
    uint64_t* keys_a = c.take<uint64_t>(n);
    uint64_t* keys_b = c.take<uint64_t>(n);
    uint32_t* codes = c.take<uint32_t>(n);
    float4* sorted = c.take<float4>(n);
    KnnBounds* bounds = c.take<KnnBounds>(1024 + 1);
    uint32_t* scratch = c.take<uint32_t>(256 * ((n + kBinTile - 1) / kBinTile) + 256 + 64);   // + the depth sort's done counter
    const int blocks = (P + kKnnThreads - 1) / kKnnThreads;
    const int rblocks = blocks < 1024 ? blocks : 1024;
    knn_bounds_partial_kernel<<<rblocks, kKnnThreads, 0, stream>>>(P, points, bounds + 1);
    knn_bounds_final_kernel<<<1, 32, 0, stream>>>(rblocks, bounds + 1, bounds);
    knn_keys_kernel<<<blocks, kKnnThreads, 0, stream>>>(P, points, bounds, keys_a);
    GS_LAUNCH_CHECK();
    const int rc = launch_depth_sort(keys_a, keys_b, (uint32_t)P, nullptr, scratch, nullptr, nullptr, stream);
    if (rc) return rc;
    knn_gather_kernel<<<blocks, kKnnThreads, 0, stream>>>(P, keys_a, points, codes, sorted);
    knn_search_kernel<<<blocks, kKnnThreads, 0, stream>>>(P, codes, sorted, bounds, out);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

}  // namespace gs
