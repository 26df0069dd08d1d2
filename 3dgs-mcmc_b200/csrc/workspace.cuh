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

// Per-Gaussian state written by preprocess, read by binning / render / backward.
struct GeomWS {
    float* depths;            // [P]   view-space z
    float2* means2D;          // [P]   pixel coordinates
    float4* conic_opacity;    // [P]   (A, B, C, opacity)
    float* rgb;               // [3P]  clamped SH colour
    float* cov3D;             // [6P]  upper triangle of Sigma
    uint8_t* clamped;         // [P]   bit c set when channel c was clamped to 0
    uint32_t* tiles_touched;  // [P]
    uint32_t* point_offsets;  // [P]   inclusive scan of tiles_touched
    uint32_t* scan_block_sums;  // [ceil(P/kScanBlock)+1]
    uint32_t* num_rendered;   // [1]   device copy of R
    size_t bytes;

    __host__ static GeomWS carve(char* base, size_t P) {
        Carver c(base);
        GeomWS g;
        g.depths = c.take<float>(P);
        g.means2D = c.take<float2>(P);
        g.conic_opacity = c.take<float4>(P);
        g.rgb = c.take<float>(3 * P);
        g.cov3D = c.take<float>(6 * P);
        g.clamped = c.take<uint8_t>(P);
        g.tiles_touched = c.take<uint32_t>(P);
        g.point_offsets = c.take<uint32_t>(P);
        g.scan_block_sums = c.take<uint32_t>((P + kScanBlock - 1) / kScanBlock + 1);
        g.num_rendered = c.take<uint32_t>(4);
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

// One 48-byte record per (tile, splat) instance, in sorted order.
struct __align__(16) SplatRec {
    float x, y, A, B;        // q0: pixel mean, conic.x, conic.y
    float C, opacity, thr;   // q1: conic.z, opacity, conservative log-threshold for alpha >= 1/255
    int id;                  //     Gaussian index
    float r, g, b, pad;      // q2: colour
};
static_assert(sizeof(SplatRec) == 48, "SplatRec must be 48 bytes");

constexpr int kRadixBits = 8;
constexpr int kRadixBins = 1 << kRadixBits;
constexpr int kSortTile = 256 * 16;  // keys per sort block (256 threads x 16 keys)

struct BinningWS {
    uint64_t* keys_a;   // [R] unsorted keys; ping-pong partner of keys_b
    uint64_t* keys_b;   // [R]
    uint32_t* vals_a;   // [R]
    uint32_t* vals_b;   // [R]
    SplatRec* packed;   // [R] tile-ordered records (what render fwd/bwd stream)
    uint32_t* error_flag;  // [1] set by the sort if a look-back spin ran out
    char* sort_scratch; // histograms, tile tickets, look-back status words
    size_t sort_scratch_bytes;
    size_t bytes;
    // After the sort the sorted pairs are in (keys_a, vals_a) when `passes` is even, else (keys_b, vals_b).
    __host__ static BinningWS carve(char* base, size_t R, int passes) {
        const size_t Rp = R > 0 ? R : 1;
        const size_t sort_tiles = (Rp + kSortTile - 1) / kSortTile;
        Carver c(base);
        BinningWS b;
        b.packed = c.take<SplatRec>(Rp);
        b.keys_a = c.take<uint64_t>(Rp);
        b.keys_b = c.take<uint64_t>(Rp);
        b.vals_a = c.take<uint32_t>(Rp);
        b.vals_b = c.take<uint32_t>(Rp);
        b.error_flag = c.take<uint32_t>(4);
        b.sort_scratch_bytes = align_up((size_t)passes * kRadixBins * 4, kAlign) + align_up((size_t)passes * 4 + 16, kAlign) +
                               (size_t)passes * sort_tiles * kRadixBins * 4 + kAlign;
        b.sort_scratch = c.take<char>(b.sort_scratch_bytes);
        b.bytes = align_up(c.off, kAlign);
        return b;
    }
};

// Number of 8-bit radix passes needed for keys (tile << 32 | depth_bits) with `tiles` tiles.
__host__ inline int sort_passes(uint32_t tiles) {
    int bits = 0;
    while ((1u << bits) < tiles && bits < 31) ++bits;   // tile ids fit in `bits` bits
    if (bits == 0) bits = 1;
    return (32 + bits + kRadixBits - 1) / kRadixBits;
}

}  // namespace gs
