// Private layout of the three opaque state buffers that travel between forward and backward.
// The reference bump-allocates GeometryState / ImageState / BinningState out of byte tensors
// (DGR/cuda_rasterizer/rasterizer_impl.h:21-73, rasterizer_impl.cu:155-194); the layout is private to the
// library, so this build uses its own: SoA arrays sized for coalesced access, 256-B aligned sub-arrays,
// and all scratch (scan partials, radix-sort histograms) carved from the same caller-owned buffers so the
// library never allocates device memory.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace splat {

constexpr int kPreBlock = 256;       // Gaussians per preprocess / duplicate block
constexpr int kSortItems = 16;       // keys per thread in the radix-sort kernels
constexpr int kSortThreads = 256;
constexpr int kSortTile = kSortItems * kSortThreads;  // 4096 keys per block
constexpr int kRadixBits = 8;
constexpr int kRadix = 1 << kRadixBits;

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
    float* depths;           // [P]   view-space z (its bit pattern is the low half of the sort key)
    float2* means2D;         // [P]   pixel-space centre
    float4* conic_opacity;   // [P]   (conic.x, conic.y, conic.z, opacity)
    float* rgb;              // [3P]  SH-evaluated colour (unused when colors_precomp is given)
    float* cov3D;            // [6P]  world covariance (unused when cov3D_precomp is given)
    float4* rec;             // [3P]  48-byte render record per Gaussian (copied per instance after the sort)
    uint8_t* clamped;        // [P]   bit k set <=> colour channel k was clamped at 0
    uint32_t* tiles_touched; // [P]
    uint32_t* block_sums;    // [nblk+1] per-256-Gaussian-block tile counts, then their exclusive scan
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
    uint64_t* keys;        // [R] sorted (tile << 32 | depth bits)
    uint32_t* point_list;  // [R] sorted Gaussian ids
    uint64_t* keys_alt;    // [R] ping-pong buffer
    uint32_t* vals_alt;    // [R]
    float* records;        // [R * SPLAT_REC_FLOATS] per-instance render records in sorted order
    uint32_t* totals;      // [8 * kRadix] per-pass global digit histograms, then 8 tile tickets + error flag (64 slots)
    uint32_t* hist;        // [8][ntiles][kRadix] look-back tile states of the radix-sort passes (directly after totals)
    static constexpr size_t kTotalsWords = 8 * kRadix + 64;   // 8448 B: keeps `hist` contiguous with `totals`
    static size_t sort_blocks(size_t R) { return (R + kSortTile - 1) / kSortTile; }
    // bytes of sort scratch that must be zero before a sort with `passes` passes (totals + the used tile states)
    static size_t sort_scratch_bytes(size_t R, int passes) {
        return sizeof(uint32_t) * (kTotalsWords + (size_t)passes * sort_blocks(R ? R : 1) * kRadix);
    }
    static BinningState carve(void* buf, size_t R, size_t* bytes) {
        Carver c(buf);
        BinningState b;
        const size_t Rp = R ? R : 1;
        b.keys = c.take<uint64_t>(Rp);
        b.point_list = c.take<uint32_t>(Rp);
        b.keys_alt = c.take<uint64_t>(Rp);
        b.vals_alt = c.take<uint32_t>(Rp);
        b.records = c.take<float>(Rp * 12);
        b.totals = c.take<uint32_t>(kTotalsWords);
        b.hist = c.take<uint32_t>(8 * kRadix * sort_blocks(Rp));
        if (bytes) *bytes = c.bytes();
        return b;
    }
};

}  // namespace splat
