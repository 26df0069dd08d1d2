// Binning for sm_100a: depth ordering of the visible Gaussians, the depth-ordered emission table, a stable
// two-digit partition of the tile instances by tile id, and the per-tile ranges the render kernels walk.
//
// The reference sorts R instance pairs (tile << 32 | depth_bits, gaussian_idx) with a 6-pass CUB radix
// sort over 44-45 significant bits (cuda_rasterizer/rasterizer_impl.cu:70-138, :278-318; ~152 B of
// HBM traffic per instance). The sorted order is unique - (tile, depth, gaussian index), because the
// emission is index-ordered and the sort is stable - so any algorithm that produces that order is
// bit-identical. This file gets there with far less work:
//   1. sort the Gaussians once by (depth_bits, index): four count -> scan -> scatter passes over P keys
//      (launch_depth_sort; culled Gaussians carry a sentinel depth and are dropped by the first pass; the last
//      pass writes the depth-ordered emission table (idx, rectangle, tile count));
//   2. scan the tile counts in that order (launch_inclusive_scan): every Gaussian's instances own a known range
//      of emission slots;
//   3. stable-partition the instances by tile id in two digits (launch_fast_binning): pass 1 expands the slots
//      in registers from the emission table (no instance list in emission order is ever written) and its block
//      histograms come analytically from the rectangles; pass 2's count also yields the per-tile counts, from
//      which the tile ranges follow by one scan. No inter-block dependency inside any kernel.
// Images of more than 65536 tiles take the generic path further down (emission + onesweep radix sort with
// decoupled look-back + range detection), which is also what GS_FORCE_GENERIC_BINNING=1 exercises.
#include "gs_common.cuh"
#include "kernels.cuh"
#include "workspace.cuh"

#include <cooperative_groups.h>
#include <atomic>
#include <cstdlib>
#include <mutex>

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
// Inclusive scan of the per-Gaussian tile counts taken in depth order:
//     out[r] = sum_{r' <= r} tiles(order[r'] & 0xffffffff),  tiles = w * h of the Gaussian's tile rectangle.
// Two small kernels: block sums -> apply (each block first adds up the sums of the blocks before it), both streaming over the depth-ordered emission table
// (idx, x0 | y0 << 16, w | h << 16, tile count) that the LAST pass of the depth sort wrote while it had every key's final
// position in hand (one random gather of the rectangle per Gaussian, instead of one in each of two scan kernels); the apply
// kernel replaces the count by the first emission slot: the table the first binning pass expands.
// ------------------------------------------------------------------------------------------------
GS_DEV uint32_t rect_tiles(const uint2 rc) { return (rc.y & 0xffffu) * (rc.y >> 16); }

__global__ void __launch_bounds__(kScanThreads) scan_block_sums_kernel(const uint4* __restrict__ emit, int n_host,
                                                                       const uint32_t* __restrict__ n_dev,
                                                                       uint32_t* __restrict__ block_sums) {
    pdl_trigger();
    pdl_wait();
    const int n = n_dev ? (int)*n_dev : n_host;   // the grid is sized for n_host >= n; blocks past the end sum nothing
    __shared__ uint32_t sw[33];
    const size_t base = (size_t)blockIdx.x * kScanBlock;
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
        const size_t k = base + (size_t)i * kScanThreads + threadIdx.x;
        if (k < (size_t)n) s += emit[k].w;
    }
    uint32_t total;
    block_excl_scan(s, sw, &total);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

// emit[r].w: tile count in, first emission slot out; out[r] = first slot of rank r + 1.
// Warp w of a block owns 256 consecutive entries and walks them 32 at a time (lane-consecutive 16-byte loads and stores:
// the table moves as whole, coalesced uint4s), carrying the running sum from round to round.
__global__ void __launch_bounds__(kScanThreads) scan_apply_kernel(int n_host, const uint32_t* __restrict__ n_dev,
                                                                  const uint32_t* __restrict__ block_sums,
                                                                  uint32_t* __restrict__ out, uint4* __restrict__ emit) {
    pdl_trigger();
    pdl_wait();
    const int n = n_dev ? (int)*n_dev : n_host;
    if ((size_t)blockIdx.x * kScanBlock >= (size_t)n) return;
    __shared__ uint32_t sw[kScanThreads / 32], sb[kScanThreads / 32];
    static_assert(kScanItems * 32 * (kScanThreads / 32) == kScanBlock, "one warp per kScanItems * 32 entries");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t base = (size_t)blockIdx.x * kScanBlock + (size_t)warp * (kScanItems * 32) + lane;
    uint4 E[kScanItems];
    uint32_t excl[kScanItems];
    uint32_t carry = 0;
#pragma unroll
    for (int r = 0; r < kScanItems; ++r) E[r] = base + r * 32 < (size_t)n ? emit[base + r * 32] : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
    for (int r = 0; r < kScanItems; ++r) {
        const uint32_t incl = warp_incl_scan(E[r].w, lane);
        excl[r] = carry + incl - E[r].w;
        carry += __shfl_sync(0xffffffffu, incl, 31);
    }
    // offset of this block = sum of the sums of the blocks before it (at most a few thousand words, L2 resident): every block
    // adds them up itself instead of a one-block "spine" kernel between the two passes (one launch fewer per step)
    uint32_t before = 0;
    for (uint32_t j = threadIdx.x; j < blockIdx.x; j += kScanThreads) before += __ldg(block_sums + j);
    before = __reduce_add_sync(0xffffffffu, before);
    if (lane == 0) { sw[warp] = carry; sb[warp] = before; }
    __syncthreads();
    uint32_t woff = 0;
#pragma unroll
    for (int w = 0; w < kScanThreads / 32; ++w) woff += sb[w] + (w < warp ? sw[w] : 0u);
#pragma unroll
    for (int r = 0; r < kScanItems; ++r) {
        if (base + r * 32 < (size_t)n) {
            const uint32_t run = woff + excl[r], v = E[r].w;
            E[r].w = run;
            emit[base + r * 32] = E[r];
            out[base + r * 32] = run + v;
        }
    }
}

// emit: the depth-ordered table (idx, x0 | y0 << 16, w | h << 16, tile count) written by the last pass of launch_depth_sort.
// n_dev (optional): the entry count on the device (<= n, which then only sizes the grids).
int launch_inclusive_scan(int n, const uint32_t* n_dev, uint32_t* out, uint4* emit, uint32_t* block_sums, cudaStream_t stream) {
    if (n <= 0) return GS_OK;
    const int nb = (n + kScanBlock - 1) / kScanBlock;
    launch_pdl(scan_block_sums_kernel, dim3(nb), dim3(kScanThreads), 0, stream, emit, n, n_dev, block_sums);
    launch_pdl(scan_apply_kernel, dim3(nb), dim3(kScanThreads), 0, stream, n, n_dev, block_sums, out, emit);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

// ------------------------------------------------------------------------------------------------
// Tile-instance emission in depth order. Thread r owns the Gaussian of depth rank r; splats covering
// many tiles are emitted by the whole warp (coalesced 8-byte stores) instead of one lane looping.
// Instance = (tile << 32) | gaussian_idx.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t kCoopTiles = 16;

__global__ void __launch_bounds__(256) emit_instances_kernel(int P_host, const uint32_t* __restrict__ n_dev,
                                                             const uint4* __restrict__ emit, uint2 grid,
                                                             uint64_t* __restrict__ inst) {
    const int P = n_dev ? (int)*n_dev : P_host;   // depth ranks that exist (the visible Gaussians)
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    uint32_t off = 0, w = 0, h = 0, idx = 0, x0 = 0, y0 = 0;
    if (r < P) {
        const uint4 E = emit[r];   // (idx, x0 | y0 << 16, w | h << 16, first slot) of the Gaussian of depth rank r
        idx = E.x; x0 = E.y & 0xffffu; y0 = E.y >> 16; w = E.z & 0xffffu; h = E.z >> 16; off = E.w;
    }
    const uint32_t n = w * h;
    // Small footprints: serial emission by the owning lane.
    if (n > 0 && n <= kCoopTiles) {
        uint32_t o = off;
        for (uint32_t y = y0; y < y0 + h; ++y)
            for (uint32_t x = x0; x < x0 + w; ++x) inst[o++] = ((uint64_t)(y * grid.x + x) << 32) | idx;
    }
    // Large footprints: the warp emits them together.
    uint32_t big = __ballot_sync(0xffffffffu, n > kCoopTiles);
    while (big) {
        const int src = __ffs(big) - 1;
        big &= big - 1;
        const uint32_t s_off = __shfl_sync(0xffffffffu, off, src);
        const uint32_t s_w = __shfl_sync(0xffffffffu, w, src);
        const uint32_t s_n = __shfl_sync(0xffffffffu, n, src);
        const uint32_t s_x0 = __shfl_sync(0xffffffffu, x0, src);
        const uint32_t s_y0 = __shfl_sync(0xffffffffu, y0, src);
        const uint32_t s_idx = __shfl_sync(0xffffffffu, idx, src);
        for (uint32_t t = lane; t < s_n; t += 32) {
            const uint32_t ty = t / s_w, tx = t - ty * s_w;
            inst[s_off + t] = ((uint64_t)((s_y0 + ty) * grid.x + (s_x0 + tx)) << 32) | s_idx;
        }
    }
}

int launch_emit_instances(int P, const uint32_t* n_dev, const uint4* emit, uint2 grid, uint64_t* inst, cudaStream_t stream) {
    if (P <= 0) return GS_OK;
    emit_instances_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, n_dev, emit, grid, inst);
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

// ================================================================================================
// Fast path: two fused, lookback-free binning passes (count -> scan -> scatter each).
//
//   pass 1: block b owns emission slots [b*kBinTile, (b+1)*kBinTile) of the depth-ordered emission (no instance
//           list is ever written in emission order: the slots are expanded in registers from the emission table).
//           bin_count1_kernel  - digit histogram of every block, ANALYTICALLY from the tile rectangles (a row of a
//                                rectangle is a run of consecutive tile ids = a circular run of low digits: two
//                                shared-memory atomics per row on a difference array);
//           bin_scan_kernel    - exclusive scan over blocks for every digit (one CTA per digit) + digit totals;
//           bin_scatter_kernel<1> - expands, ranks by the low tile digit, scatters u64 keys (tile << 32 | idx).
//   pass 2: bin_count2_kernel  - high-digit histogram of every block of those keys, and the per-tile instance
//                                counts (shared-memory aggregated) from which the tile ranges follow by one scan
//                                (identifyTileRanges, rasterizer_impl.cu:116-138, without a pass over the list);
//           bin_scan_kernel    - as above (+ one extra CTA: tile counts -> ranges);
//           bin_scatter_kernel<2> - ranks by the high digit, scatters the Gaussian ids (u32).
// Both passes are stable, so every tile's list ends up in (depth, index) order: the reference's order. There is no
// inter-block dependency inside any kernel (no tickets, flags, spins or fences).
// ================================================================================================
constexpr int kBinThreads = 256;
// Depth sorts of at most this many blocks let the count kernel's last block do the digit scan (no bin_scan_kernel launch).
// 0 = never: measured at 300K Gaussians (74 blocks) the serial tail (fence, ticket, one block scanning 256 rows) costs more
// than the 256-CTA scan kernel and its launch: depth sort 0.096 -> 0.157 ms. Kept for the record and for GS_FUSE_SCAN.
constexpr uint32_t kFuseScanMaxBlocks = 0;
constexpr int kBinItems = kBinTile / kBinThreads;  // 16
constexpr int kTileSpan = 8;                        // low digits per block that bin_count2_kernel aggregates in shared memory
static_assert(kBinItems == 16 && kBinThreads == 256, "the binning kernels are written for 256 threads x 16 slots");


// Lanes of the warp whose digit equals this lane's, from BITS ballots. (__match_any_sync is not used: on sm_100a MATCH.ANY
// runs on one per-SM unit at ~3.5 cycles per DISTINCT value in the warp - measured with tools/microbench/match_vs_ballot.cu -
// which bounded the binning passes at >100 cycles per warp-item; BITS ballots cost 4 issue slots each, spread over the
// four schedulers.)
template <int BITS>
GS_DEV uint32_t peer_mask(const uint32_t d) {
    uint32_t peers = 0xffffffffu;
#pragma unroll
    for (int k = 0; k < BITS; ++k) {
        const bool bit = (d & (1u << k)) != 0u;
        const uint32_t b = __ballot_sync(0xffffffffu, bit);
        peers &= ~(b ^ (bit ? 0xffffffffu : 0u));
    }
    return peers;
}

// Depth rank of the Gaussian owning the first / last emission slot of every pass-1 block.
__global__ void __launch_bounds__(256) block_bounds_kernel(int P_host, const uint32_t* __restrict__ n_dev,
                                                            const uint32_t* __restrict__ offsets, uint32_t R,
                                                            uint32_t* __restrict__ first, uint32_t* __restrict__ last) {
    pdl_trigger();
    pdl_wait();
    const int P = n_dev ? (int)*n_dev : P_host;
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= P) return;
    const uint32_t s1 = offsets[r], s0 = r ? offsets[r - 1] : 0u;
    if (s1 <= s0) return;
    for (uint32_t b = s0 / kBinTile; b <= (s1 - 1) / kBinTile; ++b) {
        if (b * (uint32_t)kBinTile >= s0) first[b] = (uint32_t)r;
        const uint32_t e = min((b + 1) * (uint32_t)kBinTile, R) - 1u;
        if (e < s1) last[b] = (uint32_t)r;
    }
}

// t / w and t % w for t < 65536 * w... in practice t < number of tiles <= 65536: the float product is at least
// 0.5 / w away from an integer, far above its rounding error.
GS_DEV uint32_t div_small(uint32_t t, uint32_t w, float rw) { (void)w; return (uint32_t)(((float)t + 0.5f) * rw); }
// 1 / w for div_small by one MUFU.RCP. (t + 0.5) / w is at least 0.5 / w away from an integer, a relative margin of
// 0.5 / (t + 0.5) > 7.6e-6 for t < 65536 (the fast path's tile-count limit); rcp.approx (1 ulp) and the two roundings of
// div_small add up to < 3e-7, so the truncation is exact. __frcp_rn costs ~20 instructions and a branch per slot.
GS_DEV float rcp_small(uint32_t w) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"((float)w));
    return r;
}

// hist[d][b] = number of slots of block b whose tile id has low digit d, from the rectangles alone.
__global__ void __launch_bounds__(kBinThreads) bin_count1_kernel(const uint4* __restrict__ emit, const uint32_t* __restrict__ block_first,
                                                                  const uint32_t* __restrict__ block_last, uint32_t R, uint32_t grid_x,
                                                                  int bits, uint32_t* __restrict__ hist, uint32_t nb) {
    pdl_trigger();
    pdl_wait();
    __shared__ int s_diff[256];
    __shared__ int s_all;
    __shared__ uint32_t s_scan[33];
    const int tid = threadIdx.x;
    const uint32_t b = blockIdx.x;
    const int bins = 1 << bits;
    const uint32_t mask = (uint32_t)bins - 1u;
    s_diff[tid] = 0;
    if (tid == 0) s_all = 0;
    __syncthreads();
    const uint32_t slot_begin = b * (uint32_t)kBinTile, slot_end = min(R, slot_begin + (uint32_t)kBinTile);
    const uint32_t r0 = block_first[b], ng = block_last[b] - r0 + 1u;
    // Four lanes per Gaussian, each taking every fourth row of its clipped rectangle: a warp's trip count is the longest
    // quarter-column of its 8 Gaussians instead of the tallest of 32 (mean height ~4 rows, maximum over 32 lanes ~12).
    constexpr int kSub = 4;
    const uint32_t sub = tid & (kSub - 1);
    for (uint32_t i = tid / kSub; i < ng; i += kBinThreads / kSub) {
        const uint4 E = __ldg(emit + r0 + i);
        const uint32_t w = E.z & 0xffffu, h = E.z >> 16, x0 = E.y & 0xffffu, y0 = E.y >> 16;
        // clip the Gaussian's slots [E.w, E.w + w h) to the block: local positions [ta, tb)
        const uint32_t ta = max(slot_begin, E.w) - E.w, tb = min(slot_end, E.w + w * h) - E.w;
        const float rw = rcp_small(w);
        const uint32_t ya = div_small(ta, w, rw), yb = div_small(tb - 1, w, rw);
        const uint32_t xa = ta - ya * w, xb = tb - 1 - yb * w + 1;
        for (uint32_t y = ya + sub; y <= yb; y += kSub) {
            const uint32_t xs = y == ya ? xa : 0u, xe = y == yb ? xb : w;
            const uint32_t len = xe - xs;
            const uint32_t s0 = ((y0 + y) * grid_x + x0 + xs) & mask;   // digit of the first tile of this row segment
            const uint32_t full = len >> bits, rem = len & mask;       // a run of `len` consecutive digits (mod bins)
            if (full) atomicAdd(&s_all, (int)full);
            if (rem) {
                atomicAdd(&s_diff[s0], 1);
                const uint32_t e = s0 + rem;
                if (e < (uint32_t)bins) {
                    atomicAdd(&s_diff[e], -1);
                } else if (e > (uint32_t)bins) {
                    atomicAdd(&s_diff[0], 1);
                    atomicAdd(&s_diff[e - bins], -1);
                }
            }
        }
    }
    __syncthreads();
    const uint32_t v = (uint32_t)s_diff[tid];
    uint32_t total;
    const uint32_t ex = block_excl_scan(v, s_scan, &total);   // wrap-around arithmetic: the partial sums are >= 0
    if (tid < bins) hist[(size_t)tid * nb + b] = ex + v + (uint32_t)s_all;
}

// CTA d < bins: in-place exclusive scan of hist[d][0..nb) and totals[d]. CTA `bins` (launched only for pass 2):
// ranges[t] = [start, start + count) from the per-tile counts ((0, 0) for an empty tile, as the reference leaves them).
constexpr int kScanPer = 8;
__global__ void __launch_bounds__(1024) bin_scan_kernel(uint32_t* __restrict__ hist, uint32_t nb, int bins, uint32_t* __restrict__ totals,
                                                         const uint32_t* __restrict__ tile_counts, uint32_t tiles, uint2* __restrict__ ranges) {
    pdl_trigger();
    pdl_wait();
    __shared__ uint32_t sw[33];
    const int tid = threadIdx.x;
    const bool do_ranges = (int)blockIdx.x == bins;
    const uint32_t n = do_ranges ? tiles : nb;
    uint32_t* row = do_ranges ? nullptr : hist + (size_t)blockIdx.x * nb;
    uint32_t carry = 0;
    for (uint32_t base = 0; base < n; base += 1024 * kScanPer) {
        uint32_t v[kScanPer], s = 0;
#pragma unroll
        for (int i = 0; i < kScanPer; ++i) {
            const uint32_t k = base + tid * kScanPer + i;
            v[i] = k < n ? (do_ranges ? tile_counts[k] : row[k]) : 0u;
            s += v[i];
        }
        uint32_t total;
        uint32_t run = carry + block_excl_scan(s, sw, &total);
#pragma unroll
        for (int i = 0; i < kScanPer; ++i) {
            const uint32_t k = base + tid * kScanPer + i;
            if (k < n) {
                if (do_ranges) ranges[k] = v[i] ? make_uint2(run, run + v[i]) : make_uint2(0u, 0u);
                else row[k] = run;
            }
            run += v[i];
        }
        carry += total;
    }
    if (!do_ranges && tid == 0) totals[blockIdx.x] = carry;
}

// Pass-1 output word: (high tile digit << kshift) | idx with kshift = 32 - bits2. The low digit of element e is implied by
// its position: pass 1 wrote digit d's run at [start[d], start[d + 1]), start = exclusive scan of the pass-1 digit totals.
//
// hist[d][b] = number of keys of block b with high digit d; tile_counts[tile] += per-tile counts of the block.
__global__ void __launch_bounds__(kBinThreads) bin_count2_kernel(const uint32_t* __restrict__ keys, uint32_t R, int bits1, int bits2,
                                                                  const uint32_t* __restrict__ totals1, uint32_t* __restrict__ hist,
                                                                  uint32_t nb, uint32_t* __restrict__ tile_counts) {
    pdl_trigger();
    pdl_wait();
    __shared__ uint32_t s_t[kTileSpan][256];
    __shared__ uint32_t s_start[257];     // exclusive scan of the pass-1 digit totals (+ R at [bins1])
    __shared__ uint32_t s_scan[33];
    __shared__ uint32_t s_lo[2];
    const int tid = threadIdx.x;
    const uint32_t b = blockIdx.x;
    const uint32_t slot_begin = b * (uint32_t)kBinTile, n = min((uint32_t)kBinTile, R - slot_begin);
    const int bins1 = 1 << bits1;
    const int kshift = 32 - bits2;
    for (int i = tid; i < kTileSpan * 256; i += kBinThreads) (&s_t[0][0])[i] = 0;
    {
        const uint32_t v = tid < bins1 ? totals1[tid] : 0u;
        uint32_t total;
        const uint32_t ex = block_excl_scan(v, s_scan, &total);
        s_start[tid] = ex;                       // entries >= bins1 hold R: the searches below stop there
        if (tid == 0) s_start[256] = total;
    }
    __syncthreads();
    // low digits of the block's first and last slot: the largest d with start[d] <= slot (thread d checks its own run)
    if (tid < bins1) {
        const uint32_t lo = s_start[tid], hi = s_start[tid + 1];
        if (lo <= slot_begin && slot_begin < hi) s_lo[0] = (uint32_t)tid;
        const uint32_t last = slot_begin + n - 1u;
        if (lo <= last && last < hi) s_lo[1] = (uint32_t)tid;
    }
    __syncthreads();
    // The keys are sorted by the low digit: the block holds low digits lmin..lmax only.
    const uint32_t lmin = s_lo[0], span = s_lo[1] - lmin + 1u;
    const bool compact = span <= (uint32_t)kTileSpan;
    const bool single = span == 1u;   // the usual case: a low digit's run is ~R / 2^bits1 keys, far longer than a block
    // all 16 loads in flight before the first shared-memory atomic
    uint32_t hv[kBinItems];
#pragma unroll
    for (int i = 0; i < kBinItems; ++i) {
        const uint32_t e = i * kBinThreads + tid;
        hv[i] = e < n ? keys[(size_t)slot_begin + e] >> kshift : 0u;
    }
    if (single) {
        // one low digit in the whole block: the rows of s_t serve as per-warp copies of the high-digit histogram
        static_assert(kTileSpan == kBinThreads / 32, "per-warp copies use one row of s_t each");
#pragma unroll
        for (int i = 0; i < kBinItems; ++i)
            if (i * kBinThreads + tid < (int)n) atomicAdd(&s_t[tid >> 5][hv[i]], 1u);
    } else {
#pragma unroll 4
        for (int i = 0; i < kBinItems; ++i) {
            const uint32_t e = i * kBinThreads + tid;
            if (e < n) {
                const uint32_t pos = slot_begin + e;
                const uint32_t hi = hv[i];
                if (compact) {
                    uint32_t l = 0;
#pragma unroll
                    for (int k = 1; k < kTileSpan; ++k) l += (k < (int)span && s_start[lmin + k] <= pos) ? 1u : 0u;
                    atomicAdd(&s_t[l][hi], 1u);
                } else {
                    // many short runs in one block (tiny scenes): binary search for the run that holds pos
                    uint32_t lo_d = lmin, hi_d = lmin + span - 1u;
                    while (lo_d < hi_d) {
                        const uint32_t mid = (lo_d + hi_d + 1u) >> 1;
                        if (s_start[mid] <= pos) lo_d = mid; else hi_d = mid - 1u;
                    }
                    atomicAdd(&s_t[0][hi], 1u);
                    atomicAdd(tile_counts + ((hi << bits1) | lo_d), 1u);
                }
            }
        }
    }
    __syncthreads();
    const int bins2 = 1 << bits2;
    if (tid < bins2) {
        uint32_t c = 0;
        if (single) {
#pragma unroll
            for (int w = 0; w < kTileSpan; ++w) c += s_t[w][tid];
            if (c) atomicAdd(tile_counts + (((uint32_t)tid << bits1) | lmin), c);
        } else {
            const int rows = compact ? (int)span : 1;
            for (int l = 0; l < rows; ++l) {
                const uint32_t t = s_t[l][tid];
                c += t;
                if (compact && t) atomicAdd(tile_counts + (((uint32_t)tid << bits1) | (lmin + l)), t);
            }
        }
        hist[(size_t)tid * nb + b] = c;
    }
}

struct BinScatterParams {
    // pass 1 input
    const uint4* __restrict__ emit;          // [P] depth-ordered (idx, x0 | y0 << 16, w | h << 16, first slot)
    const uint32_t* __restrict__ offsets;    // [P] inclusive scan (first slot of rank r+1)
    const uint32_t* __restrict__ block_first;
    const uint32_t* __restrict__ block_last;
    uint32_t grid_x;
    // pass 0 (depth ordering of the Gaussians): u64 keys in and out, digit in the upper word; the last pass also gathers
    // every Gaussian's tile rectangle and writes the depth-ordered emission table (rect / emit_out non-null)
    const uint64_t* __restrict__ keys_in;
    uint64_t* __restrict__ keys_out;
    const uint2* __restrict__ rect;
    uint4* __restrict__ emit_out;
    // pass 1 output / pass 2 input: (high tile digit << kshift) | idx
    const uint32_t* __restrict__ keys32_in;
    uint32_t* __restrict__ keys32_out;
    int kshift;                              // 32 - bits2
    uint32_t* __restrict__ ids_out;          // pass 2 output
    uint32_t n;                              // R (pass 0: number of keys, or its upper bound when n_dev is set)
    const uint32_t* __restrict__ n_dev;      // pass 0, optional: the key count on the device (the visible Gaussians)
    int drop_sentinel;                       // pass 0, first pass: keys whose depth word is ~0 (culled) are not carried on
    uint32_t nb;
    int shift, bits;                         // digit = (tile >> shift) & (2^bits - 1)
    const uint32_t* __restrict__ scanned;    // [2^bits][nb] exclusive scan over blocks (bin_scan_kernel)
    const uint32_t* __restrict__ totals;     // [2^bits] digit totals
};

template <int PASS>
struct BinKey { using type = uint32_t; };
template <>
struct BinKey<0> { using type = uint64_t; };

// PASS 0: u64 keys (depth_bits << 32 | idx), digit = BITS bits of the upper word at p.shift      (depth ordering)
// PASS 1: slots expanded from the emission table, digit = low BITS bits of the tile id; writes (high digit << kshift) | idx
// PASS 2: those 32-bit words, digit = word >> kshift (the high tile digit); writes the Gaussian index
//
// Stable in-warp ranks (kAtomicPeers, the default): every lane ORs its lane bit into a per-warp, per-digit mask word in shared
// memory (ATOMS.OR: one instruction whatever the number of digit bits, and independent of the order in which the hardware
// resolves lanes that hit the same word), reads the word back - that IS the set of lanes holding the same digit - and the
// group's lowest lane advances the digit's running count and clears the word. ~15 instructions per slot against ~55 for
// the BITS ballots + leader election of peer_mask<BITS> (kept as the other template branch; GS_BIN_RANK=ballot selects it).
template <int PASS, int BITS, bool kAtomicPeers, bool kFull>
GS_DEV void bin_scatter_body(const BinScatterParams& p, const uint32_t n_keys, typename BinKey<PASS>::type* __restrict__ s_keys,
                             uint8_t* __restrict__ s_dig, uint32_t (*s_hist)[256], uint32_t* __restrict__ s_gbase,
                             uint32_t* __restrict__ s_scan) {
    using Key = typename BinKey<PASS>::type;
    // per-warp peer masks [warps][256]: s_keys is idle while the ranks are formed (pass 1: its first 8 KB are the owner
    // table, the masks take the second 8 KB)
    uint32_t* const s_peer = reinterpret_cast<uint32_t*>(s_keys) + (PASS == 1 ? kBinTile / 2 : 0);
    static_assert(sizeof(Key) * kBinTile >= (PASS == 1 ? 8192 : 0) + (kBinThreads / 32) * 256 * 4, "peer masks must fit in s_keys");

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int bins = 1 << BITS;
    constexpr uint32_t dmask = (uint32_t)bins - 1u;
    for (int i = tid; i < (kBinThreads / 32) * 256; i += kBinThreads) (&s_hist[0][0])[i] = 0;
    if (kAtomicPeers)
        for (int i = tid; i < (kBinThreads / 32) * 256; i += kBinThreads) s_peer[i] = 0;
    if (PASS == 1)
        for (int i = tid; i < kBinTile / 8; i += kBinThreads) reinterpret_cast<uint4*>(s_keys)[i] = make_uint4(0u, 0u, 0u, 0u);
    __syncthreads();
    const uint32_t tile = blockIdx.x;
    const uint32_t slot_begin = tile * (uint32_t)kBinTile;
    const uint32_t tile_n = kFull ? (uint32_t)kBinTile : min((uint32_t)kBinTile, n_keys - slot_begin);
    const uint32_t warp_base = warp * (32 * kBinItems) + lane;   // item i of this thread is slot warp_base + 32 i

    // ---- 1. keys into registers, warp-striped; dig[k >> 2] byte (k & 3) = digit of item k ----
    Key key[kBinItems];
    uint32_t dig[kBinItems / 4];
#pragma unroll
    for (int q = 0; q < kBinItems / 4; ++q) dig[q] = 0u;
    if (PASS == 1) {
        uint16_t* s_owner = reinterpret_cast<uint16_t*>(s_keys);
        const uint32_t r0 = p.block_first[tile];
        const uint32_t ng = p.block_last[tile] - r0 + 1u;   // Gaussians whose slots intersect this block (<= kBinTile + 1)
        // head flags: local Gaussian i > 0 starts at slot offsets[r0 + i - 1] (>= slot_begin); Gaussian 0 owns slot 0
        for (uint32_t i = 1 + tid; i < ng; i += kBinThreads) s_owner[p.offsets[r0 + i - 1] - slot_begin] = (uint16_t)i;
        __syncthreads();
        // inclusive max-scan in a blocked arrangement: thread t owns slots [16 t, 16 t + 16)
        uint4 q[2];
        q[0] = reinterpret_cast<uint4*>(s_owner)[2 * tid];
        q[1] = reinterpret_cast<uint4*>(s_owner)[2 * tid + 1];
        uint32_t* qw = reinterpret_cast<uint32_t*>(q);
        uint32_t v[16];
        uint32_t m = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            m = max(m, qw[j] & 0xffffu); v[2 * j] = m;
            m = max(m, qw[j] >> 16);     v[2 * j + 1] = m;
        }
        uint32_t inc = m;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc = max(inc, up);
        }
        if (lane == 31) s_scan[warp] = inc;
        __syncthreads();
        uint32_t carry = __shfl_up_sync(0xffffffffu, inc, 1);
        if (lane == 0) carry = 0;
        for (int w = 0; w < warp; ++w) carry = max(carry, s_scan[w]);
#pragma unroll
        for (int j = 0; j < 8; ++j) qw[j] = max(v[2 * j], carry) | (max(v[2 * j + 1], carry) << 16);
        reinterpret_cast<uint4*>(s_owner)[2 * tid] = q[0];
        reinterpret_cast<uint4*>(s_owner)[2 * tid + 1] = q[1];
        __syncthreads();
#pragma unroll
        for (int i = 0; i < kBinItems; ++i) {
            const uint32_t e = warp_base + i * 32;
            key[i] = 0u;
            if (kFull || e < tile_n) {
                const uint4 E = __ldg(p.emit + r0 + s_owner[e]);
                const uint32_t t = slot_begin + e - E.w, w = E.z & 0xffffu;
                const uint32_t ty = div_small(t, w, rcp_small(w));
                const uint32_t tx = t - ty * w;
                const uint32_t tl = ((E.y >> 16) + ty) * p.grid_x + (E.y & 0xffffu) + tx;
                key[i] = (Key)(((tl >> BITS) << p.kshift) | E.x);
                dig[i >> 2] |= (tl & dmask) << (8 * (i & 3));
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < kBinItems; ++i) {
            const uint32_t e = warp_base + i * 32;
            uint32_t d = 0u;
            if (PASS == 0) {
                key[i] = (kFull || e < tile_n) ? (Key)p.keys_in[(size_t)slot_begin + e] : (Key)~0ull;
                d = (uint32_t)((uint64_t)key[i] >> (32 + p.shift)) & dmask;
            } else {
                key[i] = (kFull || e < tile_n) ? (Key)p.keys32_in[(size_t)slot_begin + e] : (Key)0u;
                d = (uint32_t)key[i] >> p.kshift;
            }
            dig[i >> 2] |= (d & 0xffu) << (8 * (i & 3));
        }
    }

    // ---- 2. stable rank inside the warp: peer mask per item from ballots, running per-warp digit counters kept by
    //         one leader lane per distinct digit (plain LDS / STS; __syncwarp orders them between items) ----
    const uint32_t lt_mask = (1u << lane) - 1u;
    uint16_t rank[kBinItems];
    uint32_t* const wh = s_hist[warp];
    // does slot k of this thread hold a key? (first depth-sort pass: sentinel keys - culled Gaussians, and the ~0 padding
    // past the end - are dropped here, so the later passes, the scans and the emission table see the visible ones only)
    auto item_valid = [&](const int k) -> bool {
        if (kFull) return true;
        if (PASS == 0 && p.drop_sentinel) return (uint32_t)((uint64_t)key[k] >> 32) != 0xffffffffu;
        return warp_base + k * 32 < tile_n;
    };
    if (kAtomicPeers) {
        // same 256-word stride per warp as s_hist: a digit's mask and its running count sit at a constant distance
        const ptrdiff_t peer_off = (s_peer + warp * 256) - wh;
        const uint32_t lane_bit = 1u << lane;
#pragma unroll
        for (int k = 0; k < kBinItems; ++k) {
            const bool valid = item_valid(k);
            uint32_t* const hp = wh + ((dig[k >> 2] >> (8 * (k & 3))) & 0xffu);
            if (valid) atomicOr(hp + peer_off, lane_bit);
            __syncwarp();
            uint32_t peers = 0u, old = 0u;
            if (valid) {
                peers = hp[peer_off];
                old = hp[0];
            }
            __syncwarp();
            const uint32_t before = peers & lt_mask;
            if (valid && before == 0u) {   // lowest lane of the group
                hp[0] = old + (uint32_t)__popc(peers);
                hp[peer_off] = 0u;
            }
            rank[k] = (uint16_t)(old + __popc(before));
            __syncwarp();
        }
    } else {
#pragma unroll
        for (int k = 0; k < kBinItems; ++k) {
            const bool valid = item_valid(k);
            const uint32_t d = (dig[k >> 2] >> (8 * (k & 3))) & 0xffu;   // invalid slots are masked out below
            const uint32_t peers = peer_mask<BITS>(d) & __ballot_sync(0xffffffffu, valid);
            const int leader = __ffs(peers) - 1;
            uint32_t old = 0;
            if (valid && lane == leader) {
                old = wh[d];
                wh[d] = old + (uint32_t)__popc(peers);
            }
            old = __shfl_sync(0xffffffffu, old, leader & 31);
            rank[k] = (uint16_t)(old + __popc(peers & lt_mask));
            __syncwarp();
        }
    }
    __syncthreads();   // (also: every read of s_owner is done, s_keys may be overwritten)

    // ---- 3. thread d owns digit d: position of each warp's group inside the block, global base of the digit's run ----
    uint32_t count = 0, tot_d = 0;
    if (tid < bins) {
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < kBinThreads / 32; ++w) {
            const uint32_t t = s_hist[w][tid];
            s_hist[w][tid] = run;
            run += t;
        }
        count = run;
        tot_d = p.totals[tid];
    }
    uint32_t total;
    const uint32_t lbase = block_excl_scan(count, s_scan, &total);   // start of digit d's run inside the block
    const uint32_t block_keys = total;                               // keys of this block that are carried on
    const uint32_t dbase = block_excl_scan(tot_d, s_scan, &total);   // start of digit d in the output
    if (tid < bins) {
#pragma unroll
        for (int w = 0; w < kBinThreads / 32; ++w) s_hist[w][tid] += lbase;
        s_gbase[tid] = dbase + p.scanned[(size_t)tid * p.nb + tile] - lbase;
    }
    __syncthreads();

    // ---- 4. scatter into shared memory in sorted order, then write out: consecutive threads write consecutive slots ----
#pragma unroll
    for (int i = 0; i < kBinItems; ++i) {
        const uint32_t e = warp_base + i * 32;
        if (item_valid(i)) {
            const uint32_t d = (dig[i >> 2] >> (8 * (i & 3))) & 0xffu;
            const uint32_t at = s_hist[warp][d] + rank[i];
            // pass 2 reads its word again (L1 / L2 hit) instead of carrying 16 registers through the ranking: the freed
            // registers buy a fifth resident CTA, which the latency-bound ranking (dependent shared-memory round trips) needs
            if (PASS == 2) s_keys[at] = (Key)__ldg(p.keys32_in + (size_t)slot_begin + e);
            else s_keys[at] = key[i];
            if (PASS == 1) s_dig[at] = (uint8_t)d;
        }
    }
    __syncthreads();
    const uint32_t idmask = PASS == 2 ? ((p.kshift < 32 ? (1u << p.kshift) : 0u) - 1u) : 0u;
#pragma unroll
    for (int i = 0; i < kBinItems; ++i) {
        const uint32_t s = i * kBinThreads + tid;
        if (kFull || s < block_keys) {
            const Key k = s_keys[s];
            uint32_t d;
            if (PASS == 0) d = (uint32_t)((uint64_t)k >> (32 + p.shift)) & dmask;
            else if (PASS == 1) d = s_dig[s];
            else d = (uint32_t)k >> p.kshift;
            const uint32_t pos = s_gbase[d] + s;
            if (PASS == 0) {
                p.keys_out[pos] = (uint64_t)k;
                if (p.emit_out) {
                    const uint32_t idx = (uint32_t)k;
                    const uint2 rc = __ldg(p.rect + idx);
                    p.emit_out[pos] = make_uint4(idx, rc.x, rc.y, rect_tiles(rc));
                }
            } else if (PASS == 1) p.keys32_out[pos] = (uint32_t)k;
            else p.ids_out[pos] = (uint32_t)k & idmask;
        }
    }
}

// kFull: every slot of the block exists (all blocks but the last) - no validity predicates anywhere.
template <int PASS, int BITS, int MINB = 4, bool kAtomicPeers = true>
__global__ void __launch_bounds__(kBinThreads, MINB) bin_scatter_kernel(const BinScatterParams p) {
    pdl_trigger();
    pdl_wait();
    using Key = typename BinKey<PASS>::type;
    __shared__ __align__(16) Key s_keys[kBinTile];                   // 32 / 16 KB (pass 1 first uses it as u16 owner[kBinTile])
    __shared__ uint8_t s_dig[PASS == 1 ? kBinTile : 16];             // pass 1: the low digit travels beside the word, not in it
    __shared__ uint32_t s_hist[kBinThreads / 32][256];               // per-warp digit counts -> block positions
    __shared__ uint32_t s_gbase[256];                                // global position of slot 0 of each digit's run
    __shared__ uint32_t s_scan[33];
    const uint32_t n_keys = (PASS == 0 && p.n_dev) ? *p.n_dev : p.n;
    if (blockIdx.x * (uint32_t)kBinTile >= n_keys) return;   // the grid was sized for an upper bound
    if ((blockIdx.x + 1u) * (uint32_t)kBinTile <= n_keys && !(PASS == 0 && p.drop_sentinel)) bin_scatter_body<PASS, BITS, kAtomicPeers, true>(p, n_keys, s_keys, s_dig, s_hist, s_gbase, s_scan);
    else bin_scatter_body<PASS, BITS, kAtomicPeers, false>(p, n_keys, s_keys, s_dig, s_hist, s_gbase, s_scan);
}

// hist[d][b] = number of keys of block b whose digit (key >> shift) & (2^bits - 1) is d. Generic counterpart of
// bin_count1/2 for plain u64 keys (the depth ordering of the Gaussians).
__global__ void __launch_bounds__(kBinThreads) bin_count_kernel(const uint64_t* __restrict__ keys, uint32_t n_host,
                                                                 const uint32_t* __restrict__ n_dev, int drop_sentinel, int shift, int bits,
                                                                 uint32_t* __restrict__ hist, uint32_t nb, uint32_t* __restrict__ totals, uint32_t* __restrict__ done_counter) {
    pdl_trigger();
    pdl_wait();
    __shared__ uint32_t s_h[kBinThreads / 32][256];
    const int tid = threadIdx.x, warp = tid >> 5;
    const uint32_t b = blockIdx.x;
    const uint32_t n = n_dev ? *n_dev : n_host;   // blocks past the end count nothing (their histogram columns are zero)
    const uint32_t slot_begin = b * (uint32_t)kBinTile, cnt = slot_begin < n ? min((uint32_t)kBinTile, n - slot_begin) : 0u;
    const uint32_t dmask = (1u << bits) - 1u;
    for (int i = tid; i < (kBinThreads / 32) * 256; i += kBinThreads) (&s_h[0][0])[i] = 0;
    __syncthreads();
    // all 16 loads in flight before the first shared-memory atomic (the kernel is latency bound: 733 blocks, one wave)
    uint64_t kv[kBinItems];
#pragma unroll
    for (int i = 0; i < kBinItems; ++i) {
        const uint32_t e = i * kBinThreads + tid;
        kv[i] = e < cnt ? keys[(size_t)slot_begin + e] : 0ull;
    }
#pragma unroll
    for (int i = 0; i < kBinItems; ++i) {
        const uint32_t e = i * kBinThreads + tid;
        if (e < cnt && !(drop_sentinel && (uint32_t)(kv[i] >> 32) == 0xffffffffu)) atomicAdd(&s_h[warp][(uint32_t)(kv[i] >> shift) & dmask], 1u);
    }
    __syncthreads();
    if (tid < (1 << bits)) {
        uint32_t c = 0;
#pragma unroll
        for (int w = 0; w < kBinThreads / 32; ++w) c += s_h[w][tid];
        hist[(size_t)tid * nb + b] = c;
    }
    // Small inputs (done_counter non-null, nb <= kFuseScanMaxBlocks): the LAST block to finish scans all digit rows in place
    // (thread d: exclusive scan of hist[d][0..nb), totals[d]) - what bin_scan_kernel does with 256 CTAs, without its launch.
    // At 300K Gaussians the chain is bound by the host's launch rate, not by the work.
    if (done_counter == nullptr) return;
    __shared__ bool s_last;
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = atomicAdd(done_counter, 1u) == nb - 1u;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (tid < (1 << bits)) {
        uint32_t* row = hist + (size_t)tid * nb;
        uint32_t run = 0;
        for (uint32_t j = 0; j < nb; j += 8) {
            uint32_t v[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) v[q] = j + q < nb ? __ldcg(row + j + q) : 0u;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                if (j + q < nb) row[j + q] = run;
                run += v[q];
            }
        }
        totals[tid] = run;
    }
    if (tid == 0) *done_counter = 0u;   // ready for the next pass
}

template <int PASS, int BITS>
static void launch_scatter(const BinScatterParams& sp, uint32_t nb, cudaStream_t stream) {
    // GS_BIN_MINB=5 (measurement): 5 resident CTAs per SM at 48 registers (more spills) instead of 4 at 64
    static const int minb = [] { const char* e = getenv(PASS == 2 ? "GS_BIN_MINB2" : "GS_BIN_MINB"); return e ? atoi(e) : 4; }();
    // GS_BIN_RANK=ballot (measurement): in-warp ranks from BITS ballots instead of the shared-memory peer masks. The depth
    // ordering (pass 0) always uses ballots: its upper digits repeat heavily inside a warp (a handful of exponents), and lanes
    // that hit the same mask word serialise in ATOMS (measured 24 -> 32 us per pass with the masks).
    static const bool ballot = [] { const char* e = getenv("GS_BIN_RANK"); return e && e[0] == 'b'; }();
    if (ballot || PASS == 0) launch_pdl(bin_scatter_kernel<PASS, BITS, 4, false>, dim3(nb), dim3(kBinThreads), 0, stream, sp);
    else if (minb == 6) launch_pdl(bin_scatter_kernel<PASS, BITS, 6, true>, dim3(nb), dim3(kBinThreads), 0, stream, sp);
    else if (minb == 5) launch_pdl(bin_scatter_kernel<PASS, BITS, 5, true>, dim3(nb), dim3(kBinThreads), 0, stream, sp);
    else if (minb == 3) launch_pdl(bin_scatter_kernel<PASS, BITS, 3, true>, dim3(nb), dim3(kBinThreads), 0, stream, sp);
    else launch_pdl(bin_scatter_kernel<PASS, BITS, 4, true>, dim3(nb), dim3(kBinThreads), 0, stream, sp);
}
template <int PASS>
static int dispatch_scatter(int bits, const BinScatterParams& sp, uint32_t nb, cudaStream_t stream) {
    switch (bits) {
        case 1: launch_scatter<PASS, 1>(sp, nb, stream); break;
        case 2: launch_scatter<PASS, 2>(sp, nb, stream); break;
        case 3: launch_scatter<PASS, 3>(sp, nb, stream); break;
        case 4: launch_scatter<PASS, 4>(sp, nb, stream); break;
        case 5: launch_scatter<PASS, 5>(sp, nb, stream); break;
        case 6: launch_scatter<PASS, 6>(sp, nb, stream); break;
        case 7: launch_scatter<PASS, 7>(sp, nb, stream); break;
        case 8: launch_scatter<PASS, 8>(sp, nb, stream); break;
        default: return GS_ERR_ARG;
    }
    return GS_OK;
}
int launch_fast_binning(int P, const uint32_t* n_dev, uint32_t R, uint2 grid, uint32_t tiles, int bits1, int bits2, const uint4* emit,
                        const uint32_t* offsets, const BinningWS& b, uint2* ranges, cudaStream_t stream) {
    GS_CUDA_TRY(cudaMemsetAsync(b.tile_counts, 0, (size_t)tiles * sizeof(uint32_t), stream));
    if (R == 0 || P <= 0) {
        GS_CUDA_TRY(cudaMemsetAsync(ranges, 0, (size_t)tiles * sizeof(uint2), stream));
        return GS_OK;
    }
    const uint32_t nb = (R + kBinTile - 1) / kBinTile;
    const int bins1 = 1 << bits1, bins2 = 1 << bits2;
    launch_pdl(block_bounds_kernel, dim3((P + 255) / 256), dim3(256), 0, stream, P, n_dev, offsets, R, b.block_first, b.block_last);
    launch_pdl(bin_count1_kernel, dim3(nb), dim3(kBinThreads), 0, stream, emit, b.block_first, b.block_last, R, grid.x, bits1, b.hist1, nb);
    launch_pdl(bin_scan_kernel, dim3(bins1), dim3(1024), 0, stream, b.hist1, nb, bins1, b.totals, nullptr, 0u, nullptr);
    BinScatterParams sp;
    sp.emit = emit; sp.offsets = offsets; sp.block_first = b.block_first; sp.block_last = b.block_last; sp.grid_x = grid.x;
    sp.keys_in = nullptr; sp.keys_out = nullptr; sp.rect = nullptr; sp.emit_out = nullptr; sp.n_dev = nullptr; sp.drop_sentinel = 0; sp.keys32_in = nullptr; sp.keys32_out = b.keys1; sp.kshift = 32 - bits2;
    sp.ids_out = nullptr; sp.n = R; sp.nb = nb;
    sp.shift = 0; sp.bits = bits1; sp.scanned = b.hist1; sp.totals = b.totals;
    int rc = dispatch_scatter<1>(bits1, sp, nb, stream);
    if (rc) return rc;
    launch_pdl(bin_count2_kernel, dim3(nb), dim3(kBinThreads), 0, stream, b.keys1, R, bits1, bits2, b.totals, b.hist2, nb, b.tile_counts);
    launch_pdl(bin_scan_kernel, dim3(bins2 + 1), dim3(1024), 0, stream, b.hist2, nb, bins2, b.totals + 256, b.tile_counts, tiles, ranges);
    sp.keys32_in = b.keys1; sp.keys32_out = nullptr; sp.ids_out = b.ids;
    sp.shift = bits1; sp.bits = bits2; sp.scanned = b.hist2; sp.totals = b.totals + 256;
    rc = dispatch_scatter<2>(bits2, sp, nb, stream);
    if (rc) return rc;
    GS_LAUNCH_CHECK();
    return GS_OK;
}

// ------------------------------------------------------------------------------------------------
// Small scenes: the whole depth ordering (4 x count -> scan -> scatter) AND the tile-count scan in ONE cooperative kernel.
// At 300K Gaussians the fourteen launches of the chain cost ~0.11 ms for ~0.03 ms of work: every kernel is a few microseconds
// behind a launch and a drain. With at most one wave of blocks (cooperative launch: co-residency is guaranteed by the
// driver, or the launch fails and the caller takes the multi-kernel path) the phases are separated by grid-wide barriers
// instead. Same arithmetic, same per-block code (bin_scatter_body), same results bit for bit.
// ------------------------------------------------------------------------------------------------
struct SmallSortParams {
    uint64_t* keys_a;
    uint64_t* keys_b;
    uint32_t n;                 // keys in keys_a (sentinels included when n_dev is set)
    const uint32_t* n_dev;      // optional: the number of real keys (device)
    uint32_t* hist;             // [256][nb]
    uint32_t* totals;           // [256]
    uint32_t nb;
    const uint2* rect;
    uint4* emit;
    uint32_t* offsets;          // [n] inclusive scan of the tile counts in depth order
    uint32_t* block_sums;       // [ceil(n / kScanBlock)]
};

__global__ void __launch_bounds__(kBinThreads, 2) small_depth_sort_scan_kernel(const SmallSortParams q) {
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    __shared__ __align__(16) uint64_t s_keys[kBinTile];
    __shared__ uint8_t s_dig[16];
    __shared__ uint32_t s_hist[kBinThreads / 32][256];
    __shared__ uint32_t s_gbase[256];
    __shared__ uint32_t s_scan[33];
    static_assert(kScanThreads == kBinThreads && kBinTile == 2 * kScanBlock, "one sort block = two scan blocks, same thread count");
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t b = blockIdx.x, nb = q.nb;
    const uint32_t n_real = q.n_dev ? *q.n_dev : q.n;

    for (int pass = 0; pass < 4; ++pass) {
        uint64_t* in = (pass & 1) ? q.keys_b : q.keys_a;
        uint64_t* out = (pass & 1) ? q.keys_a : q.keys_b;
        const bool drop = q.n_dev != nullptr && pass == 0;
        const uint32_t n_in = pass == 0 ? q.n : n_real;
        const int shift = 32 + 8 * pass;
        // ---- count: this block's digit histogram -> hist[d][b] ----
        for (int i = tid; i < (kBinThreads / 32) * 256; i += kBinThreads) (&s_hist[0][0])[i] = 0;
        __syncthreads();
        {
            const uint32_t slot_begin = b * (uint32_t)kBinTile, cnt = slot_begin < n_in ? min((uint32_t)kBinTile, n_in - slot_begin) : 0u;
            uint64_t kv[kBinItems];
#pragma unroll
            for (int i = 0; i < kBinItems; ++i) {
                const uint32_t e = i * kBinThreads + tid;
                kv[i] = e < cnt ? in[(size_t)slot_begin + e] : ~0ull;
            }
#pragma unroll
            for (int i = 0; i < kBinItems; ++i) {
                const uint32_t e = i * kBinThreads + tid;
                if (e < cnt && !(drop && (uint32_t)(kv[i] >> 32) == 0xffffffffu)) atomicAdd(&s_hist[warp][(uint32_t)(kv[i] >> shift) & 0xffu], 1u);
            }
        }
        __syncthreads();
        {
            uint32_t c = 0;
#pragma unroll
            for (int w = 0; w < kBinThreads / 32; ++w) c += s_hist[w][tid];
            q.hist[(size_t)tid * nb + b] = c;
        }
        grid.sync();
        // ---- scan: every digit row in place (exclusive over the blocks), totals[d]. A row has nb <= 160 words, so one WARP
        // scans a row with shuffles (no block barrier) and the grid's 8 * nb warps share the 256 rows; with a single block
        // a row is one word and thread d settles row d directly. (Round 2 gave each row to a whole block: at nb = 1 that was
        // 256 block-wide scans of one element per pass, 0.46 ms in the driver's 4K-Gaussian capture.) ----
        static_assert(kBinThreads == 256, "thread d owns digit row d in the single-block case");
        if (nb == 1u) {
            const uint32_t v = q.hist[tid];
            q.totals[tid] = v;
            q.hist[tid] = 0u;
        } else {
            constexpr uint32_t kWarps = kBinThreads / 32;
            for (uint32_t d = b * kWarps + (uint32_t)warp; d < 256u; d += nb * kWarps) {
                uint32_t* row = q.hist + (size_t)d * nb;
                uint32_t carry = 0;
                for (uint32_t base = 0; base < nb; base += 32u) {
                    const bool live = base + (uint32_t)lane < nb;
                    const uint32_t v = live ? row[base + lane] : 0u;
                    const uint32_t incl = warp_incl_scan(v, lane);
                    if (live) row[base + lane] = carry + incl - v;
                    carry += __shfl_sync(0xffffffffu, incl, 31);
                }
                if (lane == 0) q.totals[d] = carry;
            }
        }
        grid.sync();
        // ---- scatter (the multi-kernel path's per-block code) ----
        {
            BinScatterParams sp = {};
            sp.keys_in = in; sp.keys_out = out; sp.n = q.n; sp.nb = nb; sp.shift = 8 * pass; sp.bits = 8;
            sp.drop_sentinel = drop ? 1 : 0;
            sp.scanned = q.hist; sp.totals = q.totals;
            if (pass == 3) { sp.rect = q.rect; sp.emit_out = q.emit; }
            const uint32_t n_keys = n_in;
            if (b * (uint32_t)kBinTile < n_keys) {
                if ((b + 1u) * (uint32_t)kBinTile <= n_keys && !drop) bin_scatter_body<0, 8, false, true>(sp, n_keys, s_keys, s_dig, s_hist, s_gbase, s_scan);
                else bin_scatter_body<0, 8, false, false>(sp, n_keys, s_keys, s_dig, s_hist, s_gbase, s_scan);
            }
        }
        grid.sync();
    }
    if (q.emit == nullptr) return;   // sort only (every block takes this branch together)

    // ---- inclusive scan of the tile counts in depth order: emit[r].w count in, first slot out; offsets[r] ----
    uint32_t* const sw = s_scan;                       // 8 warp sums
    uint32_t* const sb = &s_gbase[0];                  // 8 partial "blocks before" sums
    const int n = (int)n_real;
    uint32_t excl[2][kScanItems];
    uint4 E[2][kScanItems];
    uint32_t wcarry[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const size_t base = ((size_t)2 * b + h) * kScanBlock + (size_t)warp * (kScanItems * 32) + lane;
        uint32_t carry = 0;
#pragma unroll
        for (int r = 0; r < kScanItems; ++r) E[h][r] = base + r * 32 < (size_t)n ? q.emit[base + r * 32] : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
        for (int r = 0; r < kScanItems; ++r) {
            const uint32_t incl = warp_incl_scan(E[h][r].w, lane);
            excl[h][r] = carry + incl - E[h][r].w;
            carry += __shfl_sync(0xffffffffu, incl, 31);
        }
        wcarry[h] = carry;
        if (lane == 0) sw[warp] = carry;
        __syncthreads();
        if (tid == 0) {
            uint32_t t = 0;
            for (int w = 0; w < kBinThreads / 32; ++w) t += sw[w];
            q.block_sums[2 * b + h] = t;
        }
        __syncthreads();
    }
    grid.sync();
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const uint32_t sblock = 2 * b + h;
        const size_t base = (size_t)sblock * kScanBlock + (size_t)warp * (kScanItems * 32) + lane;
        uint32_t before = 0;
        for (uint32_t j = tid; j < sblock; j += kBinThreads) before += q.block_sums[j];
        before = __reduce_add_sync(0xffffffffu, before);
        __syncthreads();
        if (lane == 0) { sw[warp] = wcarry[h]; sb[warp] = before; }
        __syncthreads();
        uint32_t woff = 0;
#pragma unroll
        for (int w = 0; w < kBinThreads / 32; ++w) woff += sb[w] + (w < warp ? sw[w] : 0u);
#pragma unroll
        for (int r = 0; r < kScanItems; ++r) {
            if (base + r * 32 < (size_t)n) {
                const uint32_t run = woff + excl[h][r], v = E[h][r].w;
                E[h][r].w = run;
                q.emit[base + r * 32] = E[h][r];
                q.offsets[base + r * 32] = run + v;
            }
        }
    }
}

constexpr int kMaxCoopDevices = 64;
static std::mutex g_coop_mutex;
static std::atomic<int> g_coop_capacity[kMaxCoopDevices] = {
#define GS_M1x8 {-1}, {-1}, {-1}, {-1}, {-1}, {-1}, {-1}, {-1}
    GS_M1x8, GS_M1x8, GS_M1x8, GS_M1x8, GS_M1x8, GS_M1x8, GS_M1x8, GS_M1x8
#undef GS_M1x8
};

// Returns GS_OK with *done = true when the cooperative kernel was launched; *done = false when the input is too large for one
// co-resident wave (or the launch was refused): the caller then runs launch_depth_sort + launch_inclusive_scan.
int launch_small_depth_sort_scan(uint64_t* keys_a, uint64_t* keys_b, uint32_t n, const uint32_t* n_dev, uint32_t* scratch,
                                 const uint2* rect, uint4* emit, uint32_t* offsets, uint32_t* block_sums, cudaStream_t stream,
                                 bool* done) {
    *done = false;
    if (n == 0) return GS_OK;
    static const int enabled = [] { const char* e = getenv("GS_SMALL_SORT"); return e ? atoi(e) : 1; }();
    const uint32_t nb = (n + kBinTile - 1) / kBinTile;
    // Co-resident block capacity of the cooperative kernel, per device (a process may drive several), looked up once under a
    // lock from the return codes of the queries themselves; 0 = not available on that device (also remembered after the
    // first refused launch, so a refusal is not retried on every step).
    int dev = 0;
    GS_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxCoopDevices) return GS_OK;
    int capacity = g_coop_capacity[dev].load(std::memory_order_acquire);
    if (capacity < 0) {
        std::lock_guard<std::mutex> lock(g_coop_mutex);
        capacity = g_coop_capacity[dev].load(std::memory_order_relaxed);
        if (capacity < 0) {
            int sms = 0, per_sm = 0, coop = 0;
            const bool ok = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess &&
                            cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev) == cudaSuccess &&
                            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, small_depth_sort_scan_kernel, kBinThreads, 0) == cudaSuccess;
            capacity = (ok && coop) ? sms * per_sm : 0;
            g_coop_capacity[dev].store(capacity, std::memory_order_release);
        }
    }
    // one block per SM at most: beyond that the separate kernels (more parallel scans) win
    if (!enabled || (int)nb > capacity || nb > 160u) return GS_OK;
    SmallSortParams q;
    q.keys_a = keys_a; q.keys_b = keys_b; q.n = n; q.n_dev = n_dev; q.hist = scratch; q.totals = scratch + (size_t)256 * nb; q.nb = nb;
    q.rect = rect; q.emit = emit; q.offsets = offsets; q.block_sums = block_sums;
    void* args[] = {(void*)&q};
    const cudaError_t e = cudaLaunchCooperativeKernel((const void*)small_depth_sort_scan_kernel, dim3(nb), dim3(kBinThreads), args, 0, stream);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();   // refused (e.g. MPS / partitioned device): not an error of ours, take the other path
        g_coop_capacity[dev].store(0, std::memory_order_release);   // ... from now on, without asking again
        return GS_OK;
    }
    *done = true;
    return GS_OK;
}

// Stable LSD sort of n u64 keys on the 32 bits [32, 64) in four 8-bit passes (count -> scan -> scatter each, no
// look-back): the depth ordering of the Gaussians. The sorted keys end up in keys_a. scratch: (256 * nb + 256 + 1) u32.
// rect / emit (optional, both or neither): the last pass also writes emit[rank] = (idx, rect[idx].x, rect[idx].y, tiles).
// n_dev (optional): the input's n keys include sentinels (depth word ~0); *n_dev of them are real. The sentinels are dropped
// by the first pass and the sorted real keys end up in keys_a[0, *n_dev).
int launch_depth_sort(uint64_t* keys_a, uint64_t* keys_b, uint32_t n, const uint32_t* n_dev, uint32_t* scratch, const uint2* rect,
                      uint4* emit, cudaStream_t stream) {
    if (n == 0) return GS_OK;
    const uint32_t nb = (n + kBinTile - 1) / kBinTile;
    uint32_t* hist = scratch;
    uint32_t* totals = scratch + (size_t)256 * nb;
    uint32_t* done_counter = totals + 256;                  // one word, self-resetting; zeroed here once per call
    static const uint32_t fuse_max = [] { const char* e = getenv("GS_FUSE_SCAN"); return e ? (uint32_t)atoi(e) : kFuseScanMaxBlocks; }();
    const bool fuse_scan = nb <= fuse_max;
    if (fuse_scan) GS_CUDA_TRY(cudaMemsetAsync(done_counter, 0, sizeof(uint32_t), stream));
    for (int pass = 0; pass < 4; ++pass) {
        const uint64_t* in = (pass & 1) ? keys_b : keys_a;
        uint64_t* out = (pass & 1) ? keys_a : keys_b;
        // n_dev set: the input holds n keys of which *n_dev are real and the rest carry the sentinel depth ~0. The first pass
        // reads all n and drops the sentinels; from then on only *n_dev keys exist.
        const int drop = (n_dev != nullptr && pass == 0) ? 1 : 0;
        const uint32_t* pass_n_dev = pass == 0 ? nullptr : n_dev;
        if (fuse_scan) {
            launch_pdl(bin_count_kernel, dim3(nb), dim3(kBinThreads), 0, stream, in, n, pass_n_dev, drop, 32 + 8 * pass, 8, hist, nb, totals, done_counter);
        } else {
            launch_pdl(bin_count_kernel, dim3(nb), dim3(kBinThreads), 0, stream, in, n, pass_n_dev, drop, 32 + 8 * pass, 8, hist, nb, (uint32_t*)nullptr, (uint32_t*)nullptr);
            launch_pdl(bin_scan_kernel, dim3(256), dim3(1024), 0, stream, hist, nb, 256, totals, nullptr, 0u, nullptr);
        }
        BinScatterParams sp = {};
        sp.keys_in = in; sp.keys_out = out; sp.n = n; sp.n_dev = pass_n_dev; sp.drop_sentinel = drop; sp.nb = nb; sp.shift = 8 * pass; sp.bits = 8;
        sp.scanned = hist; sp.totals = totals;
        if (pass == 3) { sp.rect = rect; sp.emit_out = emit; }
        launch_scatter<0, 8>(sp, nb, stream);
    }
    GS_LAUNCH_CHECK();
    return GS_OK;
}

// Generic path tail: keep only the Gaussian index of every sorted instance.
__global__ void __launch_bounds__(256) strip_ids_kernel(uint32_t n, const uint64_t* __restrict__ in, uint32_t* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (uint32_t)in[i];
}

int launch_strip_ids(uint32_t n, const uint64_t* in, uint32_t* out, cudaStream_t stream) {
    if (n == 0) return GS_OK;
    strip_ids_kernel<<<(n + 255) / 256, 256, 0, stream>>>(n, in, out);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

// ------------------------------------------------------------------------------------------------
// Debug / parity views of the sorted instance list in the reference's formats (gs_debug_fetch):
// keys (tile << 32 | depth_bits) and raw instances (tile << 32 | idx), rebuilt from the ranges and the ids.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) debug_expand_kernel(uint32_t tiles, const uint2* __restrict__ ranges,
                                                           const uint32_t* __restrict__ ids, const SplatRec* __restrict__ rec,
                                                           uint64_t* __restrict__ keys, uint64_t* __restrict__ raw) {
    const uint32_t tile = blockIdx.x;
    if (tile >= tiles) return;
    const uint2 rg = ranges[tile];
    for (uint32_t i = rg.x + threadIdx.x; i < rg.y; i += blockDim.x) {
        const uint32_t id = ids[i];
        if (keys) keys[i] = ((uint64_t)tile << 32) | __float_as_uint(rec[id].depth);
        if (raw) raw[i] = ((uint64_t)tile << 32) | id;
    }
}

int launch_debug_expand(uint32_t tiles, const uint2* ranges, const uint32_t* ids, const SplatRec* rec, uint64_t* keys, uint64_t* raw) {
    if (tiles == 0) return GS_OK;
    debug_expand_kernel<<<tiles, 256>>>(tiles, ranges, ids, rec, keys, raw);
    GS_LAUNCH_CHECK();
    GS_CUDA_TRY(cudaDeviceSynchronize());
    return GS_OK;
}

// The reference's per-Gaussian state arrays unpacked from the render records (which: 0 depths, 3 means2D, 5 conic_opacity,
// 6 rgb) and the tile rectangles (7 tiles_touched). Gaussians without a record (culled) read as zeros.
__global__ void __launch_bounds__(256) debug_unpack_kernel(int which, uint32_t P, const SplatRec* __restrict__ rec,
                                                            const uint2* __restrict__ rect, void* __restrict__ dst) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const uint2 rc = rect[i];
    const bool vis = rc.y != 0u;
    SplatRec r = {};
    if (vis) r = rec[i];
    float* f = reinterpret_cast<float*>(dst);
    switch (which) {
        case 0: f[i] = r.depth; break;
        case 3: f[2 * i] = r.x; f[2 * i + 1] = r.y; break;
        case 5: f[4 * i] = r.A; f[4 * i + 1] = r.B; f[4 * i + 2] = r.C; f[4 * i + 3] = r.opacity; break;
        case 6: f[3 * i] = r.r; f[3 * i + 1] = r.g; f[3 * i + 2] = r.b; break;
        case 7: reinterpret_cast<uint32_t*>(dst)[i] = rect_tiles(rc); break;
        default: break;
    }
}

int launch_debug_unpack(int which, uint32_t P, const SplatRec* rec, const uint2* rect, void* dst) {
    if (P == 0) return GS_OK;
    debug_unpack_kernel<<<(P + 255) / 256, 256>>>(which, P, rec, rect, dst);
    GS_LAUNCH_CHECK();
    GS_CUDA_TRY(cudaDeviceSynchronize());
    return GS_OK;
}

__global__ void __launch_bounds__(256) debug_low32_kernel(uint32_t n, const uint64_t* __restrict__ in, uint32_t* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (uint32_t)in[i];
}

// Tests read the depth order and its scan as P-sized arrays with the culled Gaussians last (in index order, zero tiles each).
// One block: the visible prefix is the sorted keys' low words, the culled ones are compacted behind it.
__global__ void __launch_bounds__(1024) debug_depth_order_kernel(uint32_t P, uint32_t n_vis, uint32_t R, const uint64_t* __restrict__ sorted,
                                                                 const uint2* __restrict__ rect, const uint32_t* __restrict__ offsets,
                                                                 uint32_t* __restrict__ order_out, uint32_t* __restrict__ offsets_out) {
    __shared__ uint32_t sw[33];
    for (uint32_t i = threadIdx.x; i < P; i += 1024) {
        if (order_out && i < n_vis) order_out[i] = (uint32_t)sorted[i];
        if (offsets_out) offsets_out[i] = i < n_vis ? offsets[i] : R;
    }
    if (!order_out) return;
    uint32_t carry = n_vis;
    for (uint32_t base = 0; base < P; base += 1024) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t c = (i < P && rect_tiles(rect[i]) == 0u) ? 1u : 0u;
        uint32_t total;
        const uint32_t ex = block_excl_scan(c, sw, &total);
        if (c) order_out[carry + ex] = i;
        carry += total;
    }
}

int launch_debug_depth_order(uint32_t P, uint32_t n_vis, uint32_t R, const uint64_t* sorted, const uint2* rect, const uint32_t* offsets,
                             uint32_t* order_out, uint32_t* offsets_out) {
    if (P == 0) return GS_OK;
    debug_depth_order_kernel<<<1, 1024>>>(P, n_vis, R, sorted, rect, offsets, order_out, offsets_out);
    GS_LAUNCH_CHECK();
    GS_CUDA_TRY(cudaDeviceSynchronize());
    return GS_OK;
}

int launch_debug_low32(uint32_t n, const uint64_t* in, uint32_t* out) {
    if (n == 0) return GS_OK;
    debug_low32_kernel<<<(n + 255) / 256, 256>>>(n, in, out);
    GS_LAUNCH_CHECK();
    GS_CUDA_TRY(cudaDeviceSynchronize());
    return GS_OK;
}

}  // namespace gs
