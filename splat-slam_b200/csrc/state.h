// Private layout of the three opaque state buffers that travel between forward and backward.
// The reference bump-allocates GeometryState / ImageState / BinningState out of byte tensors
// (DGR/cuda_rasterizer/rasterizer_impl.h:21-73, rasterizer_impl.cu:155-194); the layout is private to the
// library, so this build uses its own: SoA arrays sized for coalesced access, 256-B aligned sub-arrays,
// and all scratch (scan partials, radix-sort histograms and look-back states) carved from the same caller-owned
// buffers so the library never allocates device memory.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace splat {

constexpr int kPreBlock = 256;       // Gaussians per preprocess / duplicate block
constexpr int kSortItems = 16;       // pairs per thread in the radix-sort kernels
constexpr int kSortThreads = 256;
constexpr int kSortTile = kSortItems * kSortThreads;  // 4096 pairs per block
constexpr int kRadixBits = 8;
constexpr int kRadix = 1 << kRadixBits;

// 32-bit words of scratch launch_sort_pairs needs for n pairs and up to 4 passes:
// [4][256] digit histograms + 64 (tickets, error flag) + [4][ntiles][256] look-back tile states
// where launch_sort_pairs leaves its "look-back timed out" flag inside the scratch (checked by the consuming kernel)
__host__ __device__ inline const uint32_t* sort_error_flag(const uint32_t* scratch) { return scratch + 4 * kRadix + 8; }

inline size_t sort_scratch_words(size_t n) {
    const size_t ntiles = (n + kSortTile - 1) / kSortTile;
    return 4 * kRadix + 64 + 4 * (ntiles ? ntiles : 1) * kRadix;
}

struct Carver {
    char* base;
    size_t off;
    explicit Carver(void* p) : base(reinterpret_cast<char*>(p)), off(0) {}
    template <typename T>
    T* take(size_t count) {
        off = (off + 255) & ~size_t(255);
        T* p = reinterpret_cast<T*>(base + off);
        off += count * sizeof(T);
        return p;
    }
    size_t bytes() const { return (off + 255) & ~size_t(255); }
};

struct GeomState {
    float* depths;           // [P]   view-space z (its bit pattern is the low half of the reference's sort key)
    float2* means2D;         // [P]   pixel-space centre
    float4* conic_opacity;   // [P]   (conic.x, conic.y, conic.z, opacity)
    float* rgb;              // [3P]  SH-evaluated colour (unused when colors_precomp is given)
    float* cov3D;            // [6P]  world covariance (unused when cov3D_precomp is given)
    float4* rec;             // [3P]  48-byte render record per Gaussian (copied per instance after the sort)
    uint8_t* clamped;        // [P]   bit k set <=> colour channel k was clamped at 0
    uint32_t* tiles_touched; // [P]
    uint32_t* dkey;          // [P]   depth bits (0xFFFFFFFF for culled Gaussians) -> sorted
    uint32_t* order;         // [P]   Gaussian ids -> ids in (depth, id) order
    uint32_t* dkey_alt;      // [P]   ping-pong buffers of the depth sort
    uint32_t* order_alt;     // [P]
    uint32_t* sort_scratch;  // [sort_scratch_words(P)]
    uint32_t* block_sums;    // [nblk+1] instance counts per 256 depth-sorted Gaussians, then their exclusive scan
    uint32_t* num_rendered;  // [1]   R = total instances (device copy)
    static GeomState carve(void* buf, size_t P, size_t* bytes) {
        Carver c(buf);
        GeomState g;
        g.depths = c.take<float>(P);
        g.means2D = c.take<float2>(P);
        g.conic_opacity = c.take<float4>(P);
        g.rgb = c.take<float>(3 * P);
        g.cov3D = c.take<float>(6 * P);
        g.rec = c.take<float4>(3 * P);
        g.clamped = c.take<uint8_t>(P);
        g.tiles_touched = c.take<uint32_t>(P);
        g.dkey = c.take<uint32_t>(P);
        g.order = c.take<uint32_t>(P);
        g.dkey_alt = c.take<uint32_t>(P);
        g.order_alt = c.take<uint32_t>(P);
        g.sort_scratch = c.take<uint32_t>(sort_scratch_words(P));
        g.block_sums = c.take<uint32_t>((P + kPreBlock - 1) / kPreBlock + 1);
        g.num_rendered = c.take<uint32_t>(4);
        if (bytes) *bytes = c.bytes();
        return g;
    }
};

struct ImageState {
    float* final_T;      // [N] transmittance left after the last blended Gaussian
    uint32_t* n_contrib; // [N] 1-based index (in the tile's list) of the last blended Gaussian
    uint2* ranges;       // [tiles] [start,end) of each tile in the sorted instance list
    static ImageState carve(void* buf, size_t N, size_t tiles, size_t* bytes) {
        Carver c(buf);
        ImageState s;
        s.final_T = c.take<float>(N);
        s.n_contrib = c.take<uint32_t>(N);
        s.ranges = c.take<uint2>(tiles);
        if (bytes) *bytes = c.bytes();
        return s;
    }
};

struct BinningState {
    float* records;          // [R * SPLAT_REC_FLOATS] per-instance render records in sorted order: all the backward reads
    uint32_t* point_list;    // [R] Gaussian id per instance (sort values)
    uint32_t* tkey;          // [R] tile id per instance (sort keys)
    uint32_t* tkey_alt;      // [R] ping-pong buffers of the tile sort
    uint32_t* vals_alt;      // [R]
    uint32_t* sort_scratch;  // [sort_scratch_words(R)]
    uint64_t* keys;          // [R] the reference's sorted (tile << 32 | depth bits), only carved (LAST, so that nothing
                             //     else moves) when the caller asked for the sort output to be kept for inspection
    static BinningState carve(void* buf, size_t R, size_t* bytes, bool with_keys) {
        Carver c(buf);
        BinningState b;
        const size_t Rp = R ? R : 1;
        b.records = c.take<float>(Rp * 12);
        b.point_list = c.take<uint32_t>(Rp);
        b.tkey = c.take<uint32_t>(Rp);
        b.tkey_alt = c.take<uint32_t>(Rp);
        b.vals_alt = c.take<uint32_t>(Rp);
        b.sort_scratch = c.take<uint32_t>(sort_scratch_words(Rp));
        b.keys = with_keys ? c.take<uint64_t>(Rp) : nullptr;
        if (bytes) *bytes = c.bytes();
        return b;
    }
};

}  // namespace splat
