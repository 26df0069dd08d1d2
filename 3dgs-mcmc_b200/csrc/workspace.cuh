// Opaque workspace layouts (the three byte tensors Python keeps between forward and backward).
//
// The reference carves its geomBuffer / binningBuffer / imgBuffer with a bump allocator
// (cuda_rasterizer/rasterizer_impl.h:21-73, rasterizer_impl.cu:155-194). The layouts below are
// ours: SoA arrays aligned to 256 B, plus a tile-ordered packed splat stream (48 B records) that the
// render kernels consume with 1-D bulk copies (cp.async.bulk) instead of indirect gathers.
#pragma once
#include "gs_common.cuh"

namespace gs {

constexpr size_t kAlign = 256;
constexpr int kScanBlock = 2048;  // items per scan block (256 threads x 8)
constexpr int kRadixBits = 8;
constexpr int kRadixBins = 1 << kRadixBits;
constexpr int kSortTile = 256 * 16;   // keys per sort block (256 threads x 16 keys)
constexpr int kDepthSortPasses = 4;   // 32 depth bits, 8 per pass

// Scratch of one radix sort of n keys: per-pass digit histograms, tile tickets, look-back status words.
__host__ inline size_t radix_scratch_bytes(size_t n, int passes) {
    const size_t ntiles = (n + kSortTile - 1) / kSortTile;
    return align_up((size_t)passes * kRadixBins * 4, kAlign) + align_up((size_t)passes * 4 + 16, kAlign) +
           (size_t)passes * ntiles * kRadixBins * 4 + kAlign;
}

struct Carver {
    char* base;
    size_t off;
    __host__ explicit Carver(char* b) : base(b), off(0) {}
    template <typename T>
    __host__ T* take(size_t count) {
        off = align_up(off, kAlign);
        T* p = reinterpret_cast<T*>(base + off);
        off += count * sizeof(T);
        return p;
    }
};

// One 48-byte record per visible Gaussian: everything the tile renderers need about a splat.
struct __align__(16) SplatRec {
    float x, y, A, B;        // q0: pixel mean, conic.x, conic.y
    float C, opacity, thr;   // q1: conic.z, opacity, conservative log-threshold for alpha >= 1/255
    int id;                  //     Gaussian index
    float r, g, b, depth;    // q2: colour, view-space depth
};
static_assert(sizeof(SplatRec) == 48, "SplatRec must be 48 bytes");

// Per-Gaussian state written by preprocess, read by binning / render / backward.
struct GeomWS {
    float* depths;            // [P]   view-space z
    float2* means2D;          // [P]   pixel coordinates
    float4* conic_opacity;    // [P]   (A, B, C, opacity)
    float* rgb;               // [3P]  clamped SH colour
    float* cov3D;             // [6P]  upper triangle of Sigma
    SplatRec* rec;            // [P]   48-byte render record of every visible Gaussian (what the tile renderers gather)
    uint8_t* clamped;         // [P]   bit c set when channel c was clamped to 0
    uint32_t* tiles_touched;  // [P]
    uint32_t* point_offsets;  // [P]   inclusive scan of tiles_touched taken in depth order
    uint32_t* scan_block_sums;  // [ceil(P/kScanBlock)+1]
    uint32_t* num_rendered;   // [0] device copy of R, [1] depth-sort error flag
    uint64_t* dkeys_a;        // [P]   (depth_bits << 32 | idx), culled Gaussians carry depth_bits = ~0; sorted in place
    uint64_t* dkeys_b;        // [P]   ping-pong partner (4 passes: the sorted keys end up in dkeys_a)
    char* dsort_scratch;      // radix-sort histograms / tickets / look-back status for the depth sort
    size_t dsort_scratch_bytes;
    size_t bytes;

    __host__ static GeomWS carve(char* base, size_t P) {
        Carver c(base);
        GeomWS g;
        g.depths = c.take<float>(P);
        g.means2D = c.take<float2>(P);
        g.conic_opacity = c.take<float4>(P);
        g.rgb = c.take<float>(3 * P);
        g.cov3D = c.take<float>(6 * P);
        g.rec = c.take<SplatRec>(P);
        g.clamped = c.take<uint8_t>(P);
        g.tiles_touched = c.take<uint32_t>(P);
        g.point_offsets = c.take<uint32_t>(P);
        g.scan_block_sums = c.take<uint32_t>((P + kScanBlock - 1) / kScanBlock + 1);
        g.num_rendered = c.take<uint32_t>(4);
        g.dkeys_a = c.take<uint64_t>(P);
        g.dkeys_b = c.take<uint64_t>(P);
        g.dsort_scratch_bytes = radix_scratch_bytes(P, kDepthSortPasses);
        g.dsort_scratch = c.take<char>(g.dsort_scratch_bytes);
        g.bytes = align_up(c.off, kAlign);
        return g;
    }
};

// Per-pixel / per-tile state.
struct ImageWS {
    float* final_T;       // [W*H]
    uint32_t* n_contrib;  // [W*H]  1-based list position of the last blended splat
    uint2* ranges;        // [tiles] [start, end) into the sorted instance list
    uint32_t* tile_max_contrib;  // [tiles] max n_contrib over the tile (backward starts there)
    size_t bytes;

    __host__ static ImageWS carve(char* base, size_t W, size_t H) {
        const size_t tiles = ((W + GS_TILE_X - 1) / GS_TILE_X) * ((H + GS_TILE_Y - 1) / GS_TILE_Y);
        Carver c(base);
        ImageWS s;
        s.final_T = c.take<float>(W * H);
        s.n_contrib = c.take<uint32_t>(W * H);
        s.ranges = c.take<uint2>(tiles);
        s.tile_max_contrib = c.take<uint32_t>(tiles);
        s.bytes = align_up(c.off, kAlign);
        return s;
    }
};


struct BinningWS {
    uint64_t* inst_a;   // [R] instances (tile << 32 | gaussian_idx) in emission (depth) order; ping-pong partner of inst_b
    uint64_t* inst_b;   // [R]
    uint32_t* error_flag;  // [1] set by the sort if a look-back spin ran out
    char* sort_scratch; // histograms, tile tickets, look-back status words
    size_t sort_scratch_bytes;
    size_t bytes;
    // After the sort the instances are in inst_a when `passes` is even, else in inst_b.
    __host__ static BinningWS carve(char* base, size_t R, int passes) {
        const size_t Rp = R > 0 ? R : 1;
        Carver c(base);
        BinningWS b;
        b.inst_a = c.take<uint64_t>(Rp + 2);   // +2: the renderers' 16-byte-aligned bulk copies may read one entry past the end
        b.inst_b = c.take<uint64_t>(Rp + 2);
        b.error_flag = c.take<uint32_t>(4);
        b.sort_scratch_bytes = radix_scratch_bytes(Rp, passes);
        b.sort_scratch = c.take<char>(b.sort_scratch_bytes);
        b.bytes = align_up(c.off, kAlign);
        return b;
    }
    __host__ const uint64_t* sorted(int passes) const { return (passes & 1) ? inst_b : inst_a; }
};

// Number of 8-bit radix passes that cover the tile field of the instances for `tiles` tiles.
__host__ inline int sort_passes(uint32_t tiles) {
    int bits = 0;
    while ((1u << bits) < tiles && bits < 31) ++bits;   // tile ids fit in `bits` bits
    if (bits == 0) bits = 1;
    return (bits + kRadixBits - 1) / kRadixBits;
}

}  // namespace gs
