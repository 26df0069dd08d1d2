// Binning for sm_100a: depth ordering of the Gaussians, tile-instance emission, a stable onesweep
// radix sort of the instances by tile, tile ranges and the per-tile ranges of the sorted list
// the render kernels walk.
//
// The reference sorts R instance pairs (tile << 32 | depth_bits, gaussian_idx) with a 6-pass CUB radix
// sort over 44-45 significant bits (cuda_rasterizer/rasterizer_impl.cu:70-138, :278-318; ~152 B of
// HBM traffic per instance). The sorted order is unique — (tile, depth, gaussian index), because the
// emission is index-ordered and the sort is stable — so any algorithm that produces that order is
// bit-identical. This file gets there with far less traffic:
//   1. sort the P Gaussians once by (depth_bits, index)      — 4 passes over P keys (P << R, L2-resident);
//   2. scan tiles_touched in that order and emit every Gaussian's tile instances in that order, as
//      u64 (tile << 32 | gaussian_idx);
//   3. stable-sort the R instances by the tile field only     — ceil(log2(tiles)/8) = 2 passes at 1080p;
//      stability preserves the (depth, index) order inside each tile.
// That is 8 B (histogram) + 2 x 16 B per instance instead of 8 + 6 x 24 B. Both sorts use the same
// kernels: one histogram of every digit up front, then one kernel per 8-bit digit with decoupled
// look-back (onesweep), keys only, u64.
#include "gs_common.cuh"
#include "kernels.cuh"
#include "workspace.cuh"

namespace gs {

// ------------------------------------------------------------------------------------------------
// Block / warp scan helpers
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

// Block-wide exclusive scan of one value per thread (blockDim.x <= 1024). Returns the exclusive prefix,
// total in *total. smem_warp must hold 33 words.
GS_DEV uint32_t block_excl_scan(uint32_t v, uint32_t* smem_warp, uint32_t* total) {
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

// ------------------------------------------------------------------------------------------------
// Inclusive scan of tiles_touched taken in depth order: out[r] = sum_{r' <= r} in[order[r'] & 0xffffffff].
// Three small kernels: block sums -> scan of block sums -> apply. `order` may be null (identity).
// ------------------------------------------------------------------------------------------------
GS_DEV uint32_t scan_load(const uint32_t* __restrict__ in, const uint64_t* __restrict__ order, size_t k) {
    return order ? in[(uint32_t)order[k]] : in[k];
}

__global__ void __launch_bounds__(kScanThreads) scan_block_sums_kernel(const uint32_t* __restrict__ in,
                                                                       const uint64_t* __restrict__ order, int n,
                                                                       uint32_t* __restrict__ block_sums) {
    __shared__ uint32_t sw[33];
    const size_t base = (size_t)blockIdx.x * kScanBlock;
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
        const size_t k = base + (size_t)i * kScanThreads + threadIdx.x;
        if (k < (size_t)n) s += scan_load(in, order, k);
    }
    uint32_t total;
    block_excl_scan(s, sw, &total);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

// Single block: exclusive scan of block_sums in place.
__global__ void __launch_bounds__(1024) scan_spine_kernel(uint32_t* __restrict__ block_sums, int nb) {
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
}

__global__ void __launch_bounds__(kScanThreads) scan_apply_kernel(const uint32_t* __restrict__ in,
                                                                  const uint64_t* __restrict__ order, int n,
                                                                  const uint32_t* __restrict__ block_offsets,
                                                                  uint32_t* __restrict__ out) {
    __shared__ uint32_t sw[33];
    // Blocked arrangement: thread t owns items [t*8, t*8+8) of the block so the scan order is the list order.
    const size_t base = (size_t)blockIdx.x * kScanBlock + (size_t)threadIdx.x * kScanItems;
    uint32_t v[kScanItems];
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
        v[i] = (base + i < (size_t)n) ? scan_load(in, order, base + i) : 0;
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

int launch_inclusive_scan(const uint32_t* in, const uint64_t* order, int n, uint32_t* out, uint32_t* block_sums,
                          cudaStream_t stream) {
    if (n <= 0) return GS_OK;
    const int nb = (n + kScanBlock - 1) / kScanBlock;
    scan_block_sums_kernel<<<nb, kScanThreads, 0, stream>>>(in, order, n, block_sums);
    scan_spine_kernel<<<1, 1024, 0, stream>>>(block_sums, nb);
    scan_apply_kernel<<<nb, kScanThreads, 0, stream>>>(in, order, n, block_sums, out);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

// ------------------------------------------------------------------------------------------------
// Tile-instance emission in depth order. Thread r owns the Gaussian of depth rank r; splats covering
// many tiles are emitted by the whole warp (coalesced 8-byte stores) instead of one lane looping.
// Instance = (tile << 32) | gaussian_idx.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t kCoopTiles = 16;

__global__ void __launch_bounds__(256) emit_instances_kernel(int P, const uint64_t* __restrict__ depth_order,
                                                             const float2* __restrict__ means2D,
                                                             const uint32_t* __restrict__ offsets,
                                                             const int* __restrict__ radii, uint2 grid,
                                                             uint64_t* __restrict__ inst) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    uint32_t off = 0, w = 0, h = 0, idx = 0;
    uint2 rmin = {0, 0}, rmax = {0, 0};
    if (r < P) {
        const uint64_t dk = depth_order[r];
        idx = (uint32_t)dk;
        const int rad = (uint32_t)(dk >> 32) != 0xffffffffu ? radii[idx] : 0;
        if (rad > 0) {
            off = (r == 0) ? 0 : offsets[r - 1];
            tile_rect(means2D[idx], rad, rmin, rmax, grid);
            w = rmax.x - rmin.x;
            h = rmax.y - rmin.y;
        }
    }
    const uint32_t n = w * h;
    // Small footprints: serial emission by the owning lane.
    if (n > 0 && n <= kCoopTiles) {
        uint32_t o = off;
        for (uint32_t y = rmin.y; y < rmax.y; ++y)
            for (uint32_t x = rmin.x; x < rmax.x; ++x) inst[o++] = ((uint64_t)(y * grid.x + x) << 32) | idx;
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
        const uint32_t s_idx = __shfl_sync(0xffffffffu, idx, src);
        for (uint32_t t = lane; t < s_n; t += 32) {
            const uint32_t ty = t / s_w, tx = t - ty * s_w;
            inst[s_off + t] = ((uint64_t)((s_y0 + ty) * grid.x + (s_x0 + tx)) << 32) | s_idx;
        }
    }
}

int launch_emit_instances(int P, const uint64_t* depth_order, const float2* means2D, const uint32_t* offsets,
                          const int* radii, uint2 grid, uint64_t* inst, cudaStream_t stream) {
    if (P <= 0) return GS_OK;
    emit_instances_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, depth_order, means2D, offsets, radii, grid, inst);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

// ------------------------------------------------------------------------------------------------
// Onesweep-style stable LSD radix sort of u64 keys on the digits [first_bit + 8 d, first_bit + 8 d + 8),
// d = 0 .. passes-1 (bits outside that window are carried along untouched: they are the payload).
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

__global__ void __launch_bounds__(256) radix_histogram_kernel(const uint64_t* __restrict__ keys, uint32_t n, int first_bit,
                                                              int passes, uint32_t* __restrict__ hist /*[passes][256]*/) {
    __shared__ uint32_t sh[kMaxPasses * kRadixBins];
    for (int i = threadIdx.x; i < passes * kRadixBins; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    // Warp-aggregated counting: high digits (tile rows, depth exponent) repeat heavily inside a warp, so
    // one lane per distinct digit adds the whole group's count.
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31;
    for (size_t base = (size_t)blockIdx.x * blockDim.x; base < n; base += stride) {
        const size_t i = base + threadIdx.x;
        const bool valid = i < n;
        const uint32_t active = __ballot_sync(0xffffffffu, valid);
        if (valid) {
            const uint64_t k = keys[i] >> first_bit;
            for (int p = 0; p < passes; ++p) {
                const uint32_t d = (uint32_t)(k >> (p * kRadixBits)) & (kRadixBins - 1);
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
    uint64_t* __restrict__ keys_out;
    uint32_t n;
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

__global__ void __launch_bounds__(kSortThreads, 3) radix_pass_kernel(const RadixPassParams p) {
    __shared__ uint64_t s_keys[kSortTile];                           // 32 KB
    __shared__ uint32_t s_warp_hist[kSortThreads / 32][kRadixBins];  // 8 KB; becomes per-warp exclusive offsets
    __shared__ uint32_t s_local_base[kRadixBins];                    // start of each digit's run inside the tile
    __shared__ uint32_t s_global_off[kRadixBins];                    // global position of that run
    __shared__ uint32_t s_scan[33];
    __shared__ uint32_t s_tile;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(p.ticket, 1u);
    for (int i = tid; i < (kSortThreads / 32) * kRadixBins; i += kSortThreads) (&s_warp_hist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const size_t tile_base = (size_t)tile * kSortTile;
    const uint32_t tile_n = (uint32_t)min((size_t)kSortTile, (size_t)p.n - tile_base);

    // 1. Load keys warp-striped: item i of this thread is element warp*512 + i*32 + lane of the tile.
    uint64_t key[kSortItems];
    uint16_t rank[kSortItems];
    const uint32_t warp_base = warp * (32 * kSortItems) + lane;
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const uint32_t e = warp_base + i * 32;
        key[i] = e < tile_n ? p.keys_in[tile_base + e] : ~0ull;
    }
    // 2. Rank inside the warp, in element order (stable).
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const uint32_t e = warp_base + i * 32;
        const bool valid = e < tile_n;
        const uint32_t d = valid ? (uint32_t)(key[i] >> p.shift) & (kRadixBins - 1) : 0xffffffffu;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (valid && lane == leader) {
            old = s_warp_hist[warp][d];
            s_warp_hist[warp][d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        rank[i] = (uint16_t)(old + __popc(peers & lt_mask));
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

    // 4. Scatter keys into shared memory in sorted order.
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const uint32_t e = warp_base + i * 32;
        if (e < tile_n) {
            const uint32_t d = (uint32_t)(key[i] >> p.shift) & (kRadixBins - 1);
            s_keys[s_local_base[d] + s_warp_hist[warp][d] + rank[i]] = key[i];
        }
    }
    __syncthreads();
    // 5. Write out: consecutive threads write consecutive slots of a digit run.
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const uint32_t s = i * kSortThreads + tid;
        if (s < tile_n) {
            const uint64_t k = s_keys[s];
            const uint32_t d = (uint32_t)(k >> p.shift) & (kRadixBins - 1);
            p.keys_out[s_global_off[d] + (s - s_local_base[d])] = k;
        }
    }
}

// Sorts n keys. The result lands in keys_a when `passes` is even, else in keys_b.
int launch_radix_sort(uint64_t* keys_a, uint64_t* keys_b, uint32_t n, int first_bit, int passes, char* scratch,
                      uint32_t* error_flag, cudaStream_t stream) {
    if (n == 0 || passes == 0) return GS_OK;
    if (n >= (1u << 30) || passes > kMaxPasses || first_bit + passes * kRadixBits > 64) return GS_ERR_ARG;
    const size_t ntiles = ((size_t)n + kSortTile - 1) / kSortTile;
    const size_t hist_bytes = align_up((size_t)passes * kRadixBins * 4, kAlign);
    const size_t ticket_bytes = align_up((size_t)passes * 4 + 16, kAlign);
    const size_t status_bytes = (size_t)passes * ntiles * kRadixBins * 4;
    uint32_t* hist = reinterpret_cast<uint32_t*>(scratch);
    uint32_t* tickets = reinterpret_cast<uint32_t*>(scratch + hist_bytes);
    uint32_t* status = reinterpret_cast<uint32_t*>(scratch + hist_bytes + ticket_bytes);
    GS_CUDA_TRY(cudaMemsetAsync(scratch, 0, hist_bytes + ticket_bytes + status_bytes, stream));

    const int hist_blocks = (int)min((size_t)148 * 8, ((size_t)n + 1023) / 1024);
    radix_histogram_kernel<<<hist_blocks, 256, 0, stream>>>(keys_a, n, first_bit, passes, hist);
    radix_base_kernel<<<1, kRadixBins, 0, stream>>>(hist, passes);
    for (int pass = 0; pass < passes; ++pass) {
        RadixPassParams rp;
        const bool even = (pass & 1) == 0;
        rp.keys_in = even ? keys_a : keys_b;
        rp.keys_out = even ? keys_b : keys_a;
        rp.n = n;
        rp.shift = first_bit + pass * kRadixBits;
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
// Tile ranges, one thread per sorted instance (identifyTileRanges, rasterizer_impl.cu:116-138).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) tile_ranges_kernel(uint32_t R, const uint64_t* __restrict__ inst,
                                                          uint2* __restrict__ ranges) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    const uint32_t tile = (uint32_t)(inst[i] >> 32);
    if (i == 0) {
        ranges[tile].x = 0;
    } else {
        const uint32_t prev = (uint32_t)(inst[i - 1] >> 32);
        if (tile != prev) {
            ranges[prev].y = i;
            ranges[tile].x = i;
        }
    }
    if (i == R - 1) ranges[tile].y = R;
}

int launch_tile_ranges(uint32_t R, const uint64_t* inst, uint2* ranges, uint32_t tiles, cudaStream_t stream) {
    GS_CUDA_TRY(cudaMemsetAsync(ranges, 0, (size_t)tiles * sizeof(uint2), stream));
    if (R == 0) return GS_OK;
    tile_ranges_kernel<<<(R + 255) / 256, 256, 0, stream>>>(R, inst, ranges);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

// ------------------------------------------------------------------------------------------------
// Debug / parity views of the sorted instance list in the reference's formats (gs_debug_fetch):
// keys (tile << 32 | depth_bits) and the Gaussian index list.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) debug_expand_kernel(uint32_t R, const uint64_t* __restrict__ inst,
                                                           const float* __restrict__ depths, uint64_t* __restrict__ keys,
                                                           uint32_t* __restrict__ ids) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    const uint64_t k = inst[i];
    const uint32_t id = (uint32_t)k;
    if (keys) keys[i] = (k & 0xffffffff00000000ull) | __float_as_uint(depths[id]);
    if (ids) ids[i] = id;
}

int launch_debug_expand(uint32_t R, const uint64_t* inst, const float* depths, uint64_t* keys, uint32_t* ids) {
    if (R == 0) return GS_OK;
    debug_expand_kernel<<<(R + 255) / 256, 256>>>(R, inst, depths, keys, ids);
    GS_LAUNCH_CHECK();
    GS_CUDA_TRY(cudaDeviceSynchronize());
    return GS_OK;
}

__global__ void __launch_bounds__(256) debug_low32_kernel(uint32_t n, const uint64_t* __restrict__ in, uint32_t* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (uint32_t)in[i];
}

int launch_debug_low32(uint32_t n, const uint64_t* in, uint32_t* out) {
    if (n == 0) return GS_OK;
    debug_low32_kernel<<<(n + 255) / 256, 256>>>(n, in, out);
    GS_LAUNCH_CHECK();
    GS_CUDA_TRY(cudaDeviceSynchronize());
    return GS_OK;
}

}  // namespace gs
