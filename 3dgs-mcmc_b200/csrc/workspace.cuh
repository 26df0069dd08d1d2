// Opaque workspace layouts (the three byte tensors Python keeps between forward and backward).
//
// The reference carves its geomBuffer / binningBuffer / imgBuffer with a bump allocator
// (cuda_rasterizer/rasterizer_impl.h:21-73, rasterizer_impl.cu:155-194). The layouts below are
// ours: SoA arrays aligned to 256 B, plus a tile-ordered packed splat stream (48 B records) that the
// render kernels consume with 1-D bulk copies (cp.async.bulk) instead of indirect gathers.
#pragma once
#include <cstdlib>

#include "gs_common.cuh"

namespace gs {

constexpr size_t kAlign = 256;
constexpr int kScanBlock = 2048;  // items per scan block (256 threads x 8)
constexpr int kRadixBits = 8;
constexpr int kRadixBins = 1 << kRadixBits;
constexpr int kSortTile = 256 * 16;   // keys per sort block (256 threads x 16 keys)
constexpr int kDepthSortPasses = 4;   // 32 depth bits, 8 per pass
constexpr int kBinTile = 256 * 16;    // instance slots per block of the two tile-binning passes
constexpr int kBinMaxBits = 8;        // digit width limit of one binning pass (one thread per bin, 256 threads)

// Tile-grid helpers shared by the workspace layouts and the launch code.
__host__ inline uint32_t grid_tiles_x(size_t W) { return (uint32_t)((W + GS_TILE_X - 1) / GS_TILE_X); }
__host__ inline uint32_t grid_tiles_y(size_t H) { return (uint32_t)((H + GS_TILE_Y - 1) / GS_TILE_Y); }
// Bits needed for a tile id, split over the two binning passes (low digit first). 0 bits2 never happens: both
// passes always run (the second also strips the keys down to Gaussian ids).
__host__ inline void bin_digit_bits(uint32_t tiles, int* bits1, int* bits2) {
    int bits = 2;
    while ((1u << bits) < tiles && bits < 31) ++bits;
    *bits1 = (bits + 1) / 2;
    *bits2 = bits - *bits1;
}
// GS_FORCE_GENERIC_BINNING=1 selects the generic path everywhere (tests exercise it on small images that way).
// The fast path carries an instance between its two passes as ONE 32-bit word, (high tile digit << (32 - bits2)) | idx
// (the low digit is implied by the position after pass 1), so the Gaussian index must fit in 32 - bits2 bits: 67M
// Gaussians at 1080p, 16.7M at the 65536-tile limit. Anything larger takes the generic path.
__host__ inline bool bin_fast_path(uint32_t tiles, size_t P) {
    static const bool force_generic = [] { const char* e = getenv("GS_FORCE_GENERIC_BINNING"); return e && e[0] == '1'; }();
    int b1, b2;
    bin_digit_bits(tiles, &b1, &b2);
    return b1 <= kBinMaxBits && P <= ((size_t)1 << (32 - b2)) && !force_generic;
}

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
    float* cov3D;             // [6P]  upper triangle of Sigma (read again by the per-Gaussian backward)
    SplatRec* rec;            // [P]   48-byte render record of every visible Gaussian: pixel mean, conic, opacity, colour,
                              //       depth - what the tile renderers gather; the reference's separate depths / means2D /
                              //       conic_opacity / rgb arrays are NOT kept (gs_debug_fetch unpacks them from here)
    uint8_t* clamped;         // [P]   bit c set when channel c was clamped to 0
    uint32_t* point_offsets;  // [P]   inclusive scan of tiles_touched taken in depth order
    uint32_t* scan_block_sums;  // [ceil(P/kScanBlock)+1]
    uint32_t* num_rendered;   // [0] device copy of R, [1] depth-sort error flag, [2] number of visible Gaussians (= depth keys)
    uint2* rect;              // [P]   (x0 | y0 << 16, w | h << 16) tile rectangle; w = h = 0 when culled
    uint4* emit;              // [P]   depth-ordered emission table: (idx, x0 | y0 << 16, w | h << 16, first slot)
    uint64_t* dkeys_a;        // [P]   (depth_bits << 32 | idx), culled Gaussians carry depth_bits = ~0; sorted in place
    uint64_t* dkeys_b;        // [P]   ping-pong partner (4 passes: the sorted keys end up in dkeys_a)
    char* dsort_scratch;      // radix-sort histograms / tickets / look-back status for the depth sort
    size_t dsort_scratch_bytes;
    size_t bytes;

    __host__ static GeomWS carve(char* base, size_t P) {
        Carver c(base);
        GeomWS g;
        g.cov3D = c.take<float>(6 * P);
        g.rec = c.take<SplatRec>(P);
        g.clamped = c.take<uint8_t>(P);
        g.point_offsets = c.take<uint32_t>(P);
        g.scan_block_sums = c.take<uint32_t>((P + kScanBlock - 1) / kScanBlock + 1);
        g.num_rendered = c.take<uint32_t>(4);
        g.rect = c.take<uint2>(P);
        g.emit = c.take<uint4>(P);
        g.dkeys_a = c.take<uint64_t>(P);
        g.dkeys_b = c.take<uint64_t>(P);
        // depth sort (launch_depth_sort): [256][nb] block histograms + 256 digit totals
        g.dsort_scratch_bytes = (256 * ((P + kBinTile - 1) / kBinTile) + 256) * sizeof(uint32_t) + kAlign;
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


// Number of 8-bit radix passes that cover the tile field of the instances for `tiles` tiles (generic path).
__host__ inline int sort_passes(uint32_t tiles) {
    int bits = 0;
    while ((1u << bits) < tiles && bits < 31) ++bits;   // tile ids fit in `bits` bits
    if (bits == 0) bits = 1;
    return (bits + kRadixBits - 1) / kRadixBits;
}

// Per-instance state. The render kernels only ever read `ids` (sorted Gaussian indices) and ImageWS::ranges.
//   fast path (tile ids of up to 2 * kBinMaxBits bits, i.e. images up to 65536 tiles): two fused binning passes;
//   generic path (larger images): emission + multi-pass radix sort of u64 instances, then stripped to ids.
struct BinningWS {
    uint32_t* ids;          // [R + 8] Gaussian index of every instance in (tile, depth, index) order; padded because the
                            //         renderers' 16-byte-aligned bulk copies may start 3 entries early / end 3 late
    uint32_t* error_flag;   // [4] set if a look-back spin ran out
    // fast path
    uint32_t* keys1;        // [R] pass-1 output (high tile digit << (32 - bits2)) | idx, stably sorted by the low tile digit
    uint32_t* block_first;  // [nb] depth rank of the Gaussian owning the first slot of pass-1 block b
    uint32_t* block_last;   // [nb] ... owning its last slot
    uint32_t* totals;       // [2][256] digit totals of the two passes
    uint32_t* tile_counts;  // [tiles] instances per tile (zeroed per call)
    uint32_t* hist1;        // [bins1][nb] per-block digit counts, scanned in place over blocks (pass 1)
    uint32_t* hist2;        // [bins2][nb] (pass 2)
    size_t nb;
    // generic path
    uint64_t* inst_a;       // [R + 2] instances (tile << 32 | idx) in emission order; ping-pong partner of inst_b
    uint64_t* inst_b;       // [R + 2]
    char* sort_scratch;
    size_t sort_scratch_bytes;
    size_t bytes;

    __host__ static BinningWS carve(char* base, size_t R, uint32_t tiles, size_t P) {
        const size_t Rp = R > 0 ? R : 1;
        Carver c(base);
        BinningWS b = {};
        b.ids = c.take<uint32_t>(Rp + 8);
        b.error_flag = c.take<uint32_t>(4);
        if (bin_fast_path(tiles, P)) {
            int bits1, bits2;
            bin_digit_bits(tiles, &bits1, &bits2);
            b.nb = (Rp + kBinTile - 1) / kBinTile;
            b.keys1 = c.take<uint32_t>(Rp);
            b.block_first = c.take<uint32_t>(b.nb);
            b.block_last = c.take<uint32_t>(b.nb);
            b.totals = c.take<uint32_t>(2 * 256);
            b.tile_counts = c.take<uint32_t>(tiles);
            b.hist1 = c.take<uint32_t>(b.nb << bits1);
            b.hist2 = c.take<uint32_t>(b.nb << bits2);
        } else {
            const int passes = sort_passes(tiles);
            b.inst_a = c.take<uint64_t>(Rp + 2);
            b.inst_b = c.take<uint64_t>(Rp + 2);
            b.sort_scratch_bytes = radix_scratch_bytes(Rp, passes);
            b.sort_scratch = c.take<char>(b.sort_scratch_bytes);
        }
        b.bytes = align_up(c.off, kAlign);
        return b;
    }
    __host__ const uint64_t* sorted(int passes) const { return (passes & 1) ? inst_b : inst_a; }
};

}  // namespace gs
