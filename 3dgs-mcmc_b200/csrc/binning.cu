// Binning for sm_100a: inclusive scan of tiles_touched, tile-key duplication, a stable LSD radix
// sort of (tile << 32 | depth_bits, gaussian_idx) pairs in the onesweep style (one upfront
// histogram of every digit, then one kernel per 8-bit pass with decoupled look-back), tile ranges
// and the tile-ordered packed splat stream the render kernels consume.
//
// Results are bit-identical to the reference's duplicateWithKeys + cub::DeviceRadixSort::SortPairs +
// identifyTileRanges (cuda_rasterizer/rasterizer_impl.cu:70-138, :278-318): the emission order is
// Gaussian-index-major then y-major / x, and the sort is stable, so the sorted order is unique.
#include "gs_common.cuh"
#include "kernels.cuh"
#include "workspace.cuh"

namespace gs {

// ------------------------------------------------------------------------------------------------
// Inclusive scan (u32), three small kernels: block sums -> scan of block sums -> apply.
// ------------------------------------------------------------------------------------------------
constexpr int kScanThreads = 256;
constexpr int kScanItems = kScanBlock / kScanThreads;  // 8

GS_DEV uint32_t warp_incl_scan(uint32_t v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t n = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += n;
    }
    return v;
}

// Block-wide exclusive scan of one value per thread (blockDim.x <= 1024). Returns exclusive prefix,
// total in *total.
GS_DEV uint32_t block_excl_scan(uint32_t v, uint32_t* smem_warp /*[32]*/, uint32_t* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nwarps = (blockDim.x + 31) >> 5;
    const uint32_t incl = warp_incl_scan(v, lane);
    if (lane == 31) smem_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = lane < nwarps ? smem_warp[lane] : 0;
        uint32_t wi = warp_incl_scan(w, lane);
        smem_warp[lane] = wi - w;  // exclusive per warp
        if (lane == 31) smem_warp[32] = wi;
    }
    __syncthreads();
    const uint32_t r = smem_warp[warp] + incl - v;
    *total = smem_warp[32];
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(kScanThreads) scan_block_sums_kernel(const uint32_t* __restrict__ in, int n,
                                                                       uint32_t* __restrict__ block_sums) {
    __shared__ uint32_t sw[33];
    const size_t base = (size_t)blockIdx.x * kScanBlock;
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
        const size_t k = base + (size_t)i * kScanThreads + threadIdx.x;
        if (k < (size_t)n) s += in[k];
    }
    uint32_t total;
    block_excl_scan(s, sw, &total);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

// Single block: exclusive scan of block_sums in place; writes the grand total to *total_out.
__global__ void __launch_bounds__(1024) scan_spine_kernel(uint32_t* __restrict__ block_sums, int nb,
                                                          uint32_t* __restrict__ total_out) {
    __shared__ uint32_t sw[33];
    uint32_t carry = 0;
    for (int base = 0; base < nb; base += 1024) {
        const int k = base + threadIdx.x;
        const uint32_t v = k < nb ? block_sums[k] : 0;
        uint32_t total;
        const uint32_t ex = block_excl_scan(v, sw, &total);
        if (k < nb) block_sums[k] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) *total_out = carry;
}

__global__ void __launch_bounds__(kScanThreads) scan_apply_kernel(const uint32_t* __restrict__ in, int n,
                                                                  const uint32_t* __restrict__ block_offsets,
                                                                  uint32_t* __restrict__ out) {
    __shared__ uint32_t sw[33];
    // Blocked arrangement: thread t owns items [t*8, t*8+8) of the block so the scan order is the index order.
    const size_t base = (size_t)blockIdx.x * kScanBlock + (size_t)threadIdx.x * kScanItems;
    uint32_t v[kScanItems];
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
        v[i] = (base + i < (size_t)n) ? in[base + i] : 0;
        s += v[i];
    }
    uint32_t total;
    uint32_t run = block_excl_scan(s, sw, &total) + block_offsets[blockIdx.x];
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
        run += v[i];
        if (base + i < (size_t)n) out[base + i] = run;
    }
}

int launch_inclusive_scan(const uint32_t* in, int n, uint32_t* out, uint32_t* block_sums, uint32_t* total_out,
                          cudaStream_t stream) {
    if (n <= 0) {
        GS_CUDA_TRY(cudaMemsetAsync(total_out, 0, sizeof(uint32_t), stream));
        return GS_OK;
    }
    const int nb = (n + kScanBlock - 1) / kScanBlock;
    scan_block_sums_kernel<<<nb, kScanThreads, 0, stream>>>(in, n, block_sums);
    scan_spine_kernel<<<1, 1024, 0, stream>>>(block_sums, nb, total_out);
    scan_apply_kernel<<<nb, kScanThreads, 0, stream>>>(in, n, block_sums, out);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

// ------------------------------------------------------------------------------------------------
// Tile-key duplication. One lane per Gaussian; splats covering many tiles are emitted by the whole
// warp (coalesced 8-byte key stores) instead of one lane looping serially.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t kCoopTiles = 16;

__global__ void __launch_bounds__(256) duplicate_keys_kernel(int P, const float2* __restrict__ means2D,
                                                             const float* __restrict__ depths,
                                                             const uint32_t* __restrict__ offsets,
                                                             const int* __restrict__ radii, uint2 grid,
                                                             uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    uint32_t off = 0, w = 0, h = 0, depth_bits = 0;
    uint2 rmin = {0, 0}, rmax = {0, 0};
    if (idx < P) {
        const int r = radii[idx];
        if (r > 0) {
            off = (idx == 0) ? 0 : offsets[idx - 1];
            tile_rect(means2D[idx], r, rmin, rmax, grid);
            w = rmax.x - rmin.x;
            h = rmax.y - rmin.y;
            depth_bits = __float_as_uint(depths[idx]);
        }
    }
    const uint32_t n = w * h;
    // Small footprints: serial emission by the owning lane.
    if (n > 0 && n <= kCoopTiles) {
        uint32_t o = off;
        for (uint32_t y = rmin.y; y < rmax.y; ++y)
            for (uint32_t x = rmin.x; x < rmax.x; ++x) {
                keys[o] = ((uint64_t)(y * grid.x + x) << 32) | depth_bits;
                vals[o] = (uint32_t)idx;
                ++o;
            }
    }
    // Large footprints: the warp emits them together.
    uint32_t big = __ballot_sync(0xffffffffu, n > kCoopTiles);
    while (big) {
        const int src = __ffs(big) - 1;
        big &= big - 1;
        const uint32_t s_off = __shfl_sync(0xffffffffu, off, src);
        const uint32_t s_w = __shfl_sync(0xffffffffu, w, src);
        const uint32_t s_n = __shfl_sync(0xffffffffu, n, src);
        const uint32_t s_x0 = __shfl_sync(0xffffffffu, rmin.x, src);
        const uint32_t s_y0 = __shfl_sync(0xffffffffu, rmin.y, src);
        const uint32_t s_d = __shfl_sync(0xffffffffu, depth_bits, src);
        const uint32_t s_idx = (uint32_t)(idx - lane + src);
        for (uint32_t t = lane; t < s_n; t += 32) {
            const uint32_t ty = t / s_w, tx = t - ty * s_w;
            keys[s_off + t] = ((uint64_t)((s_y0 + ty) * grid.x + (s_x0 + tx)) << 32) | s_d;
            vals[s_off + t] = s_idx;
        }
    }
}

int launch_duplicate_keys(int P, const float2* means2D, const float* depths, const uint32_t* offsets,
                          const int* radii, uint2 grid, uint64_t* keys, uint32_t* vals, cudaStream_t stream) {
    if (P <= 0) return GS_OK;
    duplicate_keys_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, means2D, depths, offsets, radii, grid, keys, vals);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

// ------------------------------------------------------------------------------------------------
// Onesweep-style stable LSD radix sort, 8-bit digits.
//   radix_histogram_kernel : digit histograms of every pass in one read of the keys.
//   radix_base_kernel      : exclusive scan of each pass's 256 bins.
//   radix_pass_kernel      : per pass; ranks 4096 keys per block with warp match, resolves the
//                            cross-block prefix by decoupled look-back, scatters through shared memory.
// Look-back state: one u32 per (tile, digit): bits 31:30 flag (0 empty, 1 local count, 2 inclusive
// prefix), bits 29:0 value. Tiles are handed out by an atomic ticket so every predecessor of a tile
// is already resident (forward progress); spins are bounded and raise an error flag instead of hanging.
// ------------------------------------------------------------------------------------------------
constexpr int kSortThreads = 256;
constexpr int kSortItems = kSortTile / kSortThreads;  // 16
constexpr int kMaxPasses = 8;
constexpr uint32_t kFlagLocal = 1u << 30;
constexpr uint32_t kFlagGlobal = 2u << 30;
constexpr uint32_t kValueMask = (1u << 30) - 1;

__global__ void __launch_bounds__(256) radix_histogram_kernel(const uint64_t* __restrict__ keys, uint32_t R, int passes,
                                                              uint32_t* __restrict__ hist /*[passes][256]*/) {
    __shared__ uint32_t sh[kMaxPasses * kRadixBins];
    for (int i = threadIdx.x; i < passes * kRadixBins; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    // Warp-aggregated counting: the high digits (tile bits, depth exponent) are heavily repeated inside
    // a warp, so one lane per distinct digit adds the whole group's count.
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31;
    for (size_t base = (size_t)blockIdx.x * blockDim.x; base < R; base += stride) {
        const size_t i = base + threadIdx.x;
        const bool valid = i < R;
        const uint32_t active = __ballot_sync(0xffffffffu, valid);
        if (valid) {
            const uint64_t k = keys[i];
            for (int p = 0; p < passes; ++p) {
                const uint32_t d = (uint32_t)((k >> (p * kRadixBits)) & (kRadixBins - 1));
                const uint32_t peers = __match_any_sync(active, d);
                if (lane == __ffs(peers) - 1) atomicAdd(&sh[p * kRadixBins + d], (uint32_t)__popc(peers));
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * kRadixBins; i += blockDim.x) {
        const uint32_t v = sh[i];
        if (v) atomicAdd(&hist[i], v);
    }
}

__global__ void __launch_bounds__(kRadixBins) radix_base_kernel(uint32_t* __restrict__ hist, int passes) {
    __shared__ uint32_t sw[33];
    for (int p = 0; p < passes; ++p) {
        const uint32_t v = hist[p * kRadixBins + threadIdx.x];
        uint32_t total;
        const uint32_t ex = block_excl_scan(v, sw, &total);
        hist[p * kRadixBins + threadIdx.x] = ex;
    }
}

struct RadixPassParams {
    const uint64_t* __restrict__ keys_in;
    const uint32_t* __restrict__ vals_in;
    uint64_t* __restrict__ keys_out;
    uint32_t* __restrict__ vals_out;
    uint32_t R;
    int shift;
    const uint32_t* __restrict__ digit_base;  // [256] exclusive global digit offsets for this pass
    uint32_t* status;                          // [ntiles][256] look-back state (zeroed)
    uint32_t* ticket;                          // [1] tile ticket counter (zeroed)
    uint32_t* error_flag;                      // [1]
};

GS_DEV uint32_t ld_relaxed(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
GS_DEV void st_relaxed(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__global__ void __launch_bounds__(kSortThreads) radix_pass_kernel(const RadixPassParams p) {
    __shared__ uint64_t s_keys[kSortTile];           // 32 KB
    __shared__ uint32_t s_warp_hist[kSortThreads / 32][kRadixBins];  // 8 KB; becomes per-warp exclusive offsets
    __shared__ uint32_t s_local_base[kRadixBins];    // start of each digit's run inside the tile
    __shared__ uint32_t s_global_off[kRadixBins];    // global position of that run
    __shared__ uint32_t s_scan[33];
    __shared__ uint32_t s_tile;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(p.ticket, 1u);
    for (int i = tid; i < (kSortThreads / 32) * kRadixBins; i += kSortThreads) (&s_warp_hist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const size_t tile_base = (size_t)tile * kSortTile;
    const uint32_t tile_n = (uint32_t)min((size_t)kSortTile, (size_t)p.R - tile_base);

    // 1. Load keys warp-striped: item i of this thread is element warp*512 + i*32 + lane of the tile.
    uint64_t key[kSortItems];
    uint32_t rank[kSortItems];
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const uint32_t e = warp * (32 * kSortItems) + i * 32 + lane;
        key[i] = e < tile_n ? p.keys_in[tile_base + e] : ~0ull;
    }
    // 2. Rank inside the warp, in element order (stable).
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const uint32_t e = warp * (32 * kSortItems) + i * 32 + lane;
        const bool valid = e < tile_n;
        const uint32_t d = valid ? (uint32_t)((key[i] >> p.shift) & (kRadixBins - 1)) : 0xffffffffu;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (valid && lane == leader) {
            old = s_warp_hist[warp][d];
            s_warp_hist[warp][d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        rank[i] = old + __popc(peers & lt_mask);
        __syncwarp();
    }
    __syncthreads();

    // 3. Thread d owns digit d: exclusive prefix over warps, block count, look-back.
    {
        const int d = tid;
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < kSortThreads / 32; ++w) {
            const uint32_t t = s_warp_hist[w][d];
            s_warp_hist[w][d] = run;
            run += t;
        }
        const uint32_t count = run;
        uint32_t* my_status = p.status + (size_t)tile * kRadixBins + d;
        if (tile > 0) st_relaxed(my_status, kFlagLocal | count);

        uint32_t total;
        const uint32_t lbase = block_excl_scan(count, s_scan, &total);
        s_local_base[d] = lbase;

        uint32_t excl = 0;
        if (tile > 0) {
            int t = (int)tile - 1;
            uint32_t spins = 0;
            while (true) {
                const uint32_t s = ld_relaxed(p.status + (size_t)t * kRadixBins + d);
                const uint32_t flag = s & ~kValueMask;
                if (flag == 0) {
                    if (++spins > (1u << 22)) { atomicExch(p.error_flag, 1u); break; }
                    continue;
                }
                excl += s & kValueMask;
                if (flag == kFlagGlobal || t == 0) break;
                --t;
            }
        }
        st_relaxed(my_status, kFlagGlobal | ((excl + count) & kValueMask));
        s_global_off[d] = p.digit_base[d] + excl;
    }
    __syncthreads();

    // 4. Scatter keys into shared memory in sorted order, remember each item's slot.
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const uint32_t e = warp * (32 * kSortItems) + i * 32 + lane;
        if (e < tile_n) {
            const uint32_t d = (uint32_t)((key[i] >> p.shift) & (kRadixBins - 1));
            rank[i] = s_local_base[d] + s_warp_hist[warp][d] + rank[i];
            s_keys[rank[i]] = key[i];
        }
    }
    __syncthreads();
    // 5. Write keys out: consecutive threads write consecutive slots of a digit run.
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const uint32_t s = i * kSortThreads + tid;
        if (s < tile_n) {
            const uint64_t k = s_keys[s];
            const uint32_t d = (uint32_t)((k >> p.shift) & (kRadixBins - 1));
            p.keys_out[s_global_off[d] + (s - s_local_base[d])] = k;
        }
    }
    __syncthreads();
    // 6. Same for values, reusing the key staging area (digits re-derived from the staged keys' slots).
    uint32_t* s_vals = reinterpret_cast<uint32_t*>(s_keys);          // first 16 KB
    uint8_t* s_digit = reinterpret_cast<uint8_t*>(s_vals + kSortTile);  // next 4 KB
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const uint32_t e = warp * (32 * kSortItems) + i * 32 + lane;
        if (e < tile_n) {
            s_vals[rank[i]] = p.vals_in[tile_base + e];
            s_digit[rank[i]] = (uint8_t)((key[i] >> p.shift) & (kRadixBins - 1));
        }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const uint32_t s = i * kSortThreads + tid;
        if (s < tile_n) {
            const uint32_t d = s_digit[s];
            p.vals_out[s_global_off[d] + (s - s_local_base[d])] = s_vals[s];
        }
    }
}

// Sorts R pairs. Result lands in (keys_a, vals_a) when `passes` is even, else in (keys_b, vals_b).
int launch_radix_sort(uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, uint32_t R, int passes,
                      char* scratch, uint32_t* error_flag, cudaStream_t stream) {
    if (R == 0) return GS_OK;
    if (R >= (1u << 30) || passes > kMaxPasses) return GS_ERR_ARG;
    const size_t ntiles = ((size_t)R + kSortTile - 1) / kSortTile;
    const size_t hist_bytes = align_up((size_t)passes * kRadixBins * 4, kAlign);
    const size_t ticket_bytes = align_up((size_t)passes * 4 + 16, kAlign);
    const size_t status_bytes = (size_t)passes * ntiles * kRadixBins * 4;
    uint32_t* hist = reinterpret_cast<uint32_t*>(scratch);
    uint32_t* tickets = reinterpret_cast<uint32_t*>(scratch + hist_bytes);
    uint32_t* status = reinterpret_cast<uint32_t*>(scratch + hist_bytes + ticket_bytes);
    GS_CUDA_TRY(cudaMemsetAsync(scratch, 0, hist_bytes + ticket_bytes + status_bytes, stream));

    const int hist_blocks = (int)min((size_t)148 * 8, ((size_t)R + 1023) / 1024);
    radix_histogram_kernel<<<hist_blocks, 256, 0, stream>>>(keys_a, R, passes, hist);
    radix_base_kernel<<<1, kRadixBins, 0, stream>>>(hist, passes);
    for (int pass = 0; pass < passes; ++pass) {
        RadixPassParams rp;
        const bool even = (pass & 1) == 0;
        rp.keys_in = even ? keys_a : keys_b;
        rp.vals_in = even ? vals_a : vals_b;
        rp.keys_out = even ? keys_b : keys_a;
        rp.vals_out = even ? vals_b : vals_a;
        rp.R = R;
        rp.shift = pass * kRadixBits;
        rp.digit_base = hist + pass * kRadixBins;
        rp.status = status + (size_t)pass * ntiles * kRadixBins;
        rp.ticket = tickets + pass;
        rp.error_flag = error_flag;
        radix_pass_kernel<<<(unsigned)ntiles, kSortThreads, 0, stream>>>(rp);
    }
    GS_LAUNCH_CHECK();
    return GS_OK;
}

// ------------------------------------------------------------------------------------------------
// Tile ranges + packed splat stream, one thread per sorted instance.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ranges_and_pack_kernel(uint32_t R, const uint64_t* __restrict__ keys,
                                                              const uint32_t* __restrict__ vals,
                                                              const float2* __restrict__ means2D,
                                                              const float4* __restrict__ conic_opacity,
                                                              const float* __restrict__ colors /*[3P]*/,
                                                              uint2* __restrict__ ranges, SplatRec* __restrict__ packed) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    const uint32_t tile = (uint32_t)(keys[i] >> 32);
    if (i == 0) {
        ranges[tile].x = 0;
    } else {
        const uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
        if (tile != prev) {
            ranges[prev].y = i;
            ranges[tile].x = i;
        }
    }
    if (i == R - 1) ranges[tile].y = R;

    const uint32_t id = vals[i];
    const float2 xy = means2D[id];
    const float4 co = conic_opacity[id];
    // Conservative reject threshold: power < thr  ==>  opacity * expf(power) < 1/255 (with 1e-3 slack
    // in the exponent, far above the expf / logf / product rounding errors), so the render kernels can
    // skip expf for such pairs without changing any decision.
    float thr;
    if (co.w > 0.0f) {
        const float lt = logf(1.0f / (255.0f * co.w));
        thr = lt - 0.001f - 1e-6f * fabsf(lt);
    } else {
        thr = co.w == co.w ? 1.0f : -INFINITY;  // non-positive opacity never passes; NaN takes the exact path
    }
    SplatRec rec;
    rec.x = xy.x; rec.y = xy.y; rec.A = co.x; rec.B = co.y;
    rec.C = co.z; rec.opacity = co.w; rec.thr = thr; rec.id = (int)id;
    rec.r = colors[3 * (size_t)id]; rec.g = colors[3 * (size_t)id + 1]; rec.b = colors[3 * (size_t)id + 2]; rec.pad = 0.f;
    float4* dst = reinterpret_cast<float4*>(packed + i);
    const float4* src = reinterpret_cast<const float4*>(&rec);
    dst[0] = src[0];
    dst[1] = src[1];
    dst[2] = src[2];
}

int launch_ranges_and_pack(uint32_t R, const uint64_t* keys, const uint32_t* vals, const float2* means2D,
                           const float4* conic_opacity, const float* colors, uint2* ranges, uint32_t tiles,
                           SplatRec* packed, cudaStream_t stream) {
    GS_CUDA_TRY(cudaMemsetAsync(ranges, 0, (size_t)tiles * sizeof(uint2), stream));
    if (R == 0) return GS_OK;
    ranges_and_pack_kernel<<<(R + 255) / 256, 256, 0, stream>>>(R, keys, vals, means2D, conic_opacity, colors, ranges, packed);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

}  // namespace gs
