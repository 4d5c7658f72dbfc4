// Binning: instance offsets, (tile|depth) key generation, stable 64-bit LSD radix sort, tile ranges and the
// gather of per-instance render records.
// Replaces cub::DeviceScan::InclusiveSum (DGR/cuda_rasterizer/rasterizer_impl.cu:280), duplicateWithKeys (:70-111),
// getHigherMsb + cub::DeviceRadixSort::SortPairs (:35-50,303-311), cudaMemset + identifyTileRanges (:313-321,116-138).
#include <cuda_fp16.h>
#include "common.cuh"
#include "kernels.h"

namespace splat {

// ---- offsets -----------------------------------------------------------------------------------------
// Exclusive scan of the per-block instance counts written by preprocess_forward_kernel; the per-Gaussian
// offsets themselves never reach HBM (duplicate_with_keys_kernel rebuilds them inside the block).
__global__ void __launch_bounds__(1024) scan_block_sums_kernel(int nblk, uint32_t* __restrict__ sums,
                                                               uint32_t* __restrict__ total_out) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < nblk; base += 1024) {
        const int i = base + threadIdx.x;
        const uint32_t v = (i < nblk) ? sums[i] : 0u;
        uint32_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += t;
            }
            s_warp[lane] = w;  // inclusive over warps
        }
        __syncthreads();
        const uint32_t carry = s_carry;
        const uint32_t warp_off = warp ? s_warp[warp - 1] : 0u;
        if (i < nblk) sums[i] = carry + warp_off + inc - v;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = carry + s_warp[31];
        __syncthreads();
    }
    if (threadIdx.x == 0) { sums[nblk] = s_carry; total_out[0] = s_carry; }
}

void launch_scan_block_sums(int P, const GeomState& g, cudaStream_t s) {
    const int nblk = (P + kPreBlock - 1) / kPreBlock;
    scan_block_sums_kernel<<<1, 1024, 0, s>>>(nblk, g.block_sums, g.num_rendered);
}

// ---- key generation ----------------------------------------------------------------------------------
// One block per 256 Gaussians. The block rebuilds its local offsets by a shared-memory scan, then ALL threads
// emit the block's instances cooperatively (instance k -> owner by binary search) so that the 12-byte
// key/value writes are fully coalesced and a screen-filling Gaussian costs its neighbours nothing.
// Emission order per Gaussian is row-major over its tile rectangle, as in the reference, so ties under the
// stable sort resolve identically.
__global__ void __launch_bounds__(kPreBlock)
duplicate_with_keys_kernel(int P, int grid_x, int grid_y, const int* __restrict__ radii, GeomState g,
                           uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out) {
    __shared__ uint32_t s_incl[kPreBlock];
    __shared__ uint32_t s_rect[kPreBlock];   // x0 | y0 << 12 | width << 24 ... widths can reach 4095 -> use two arrays
    __shared__ uint32_t s_w[kPreBlock];
    __shared__ uint32_t s_depth[kPreBlock];
    __shared__ uint32_t s_warp[kPreBlock / 32];
    const int idx = blockIdx.x * kPreBlock + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t cnt = 0;
    if (idx < P) {
        const int radius = radii[idx];
        if (radius > 0) {
            const float2 p = g.means2D[idx];
            uint32_t x0, y0, x1, y1;
            get_rect_ref(p.x, p.y, radius, grid_x, grid_y, x0, y0, x1, y1);
            cnt = (x1 - x0) * (y1 - y0);
            s_rect[threadIdx.x] = x0 | (y0 << 16);
            s_w[threadIdx.x] = x1 - x0;
            s_depth[threadIdx.x] = __float_as_uint(g.depths[idx]);
        }
    }
    uint32_t inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t woff = 0;
#pragma unroll
    for (int w = 0; w < kPreBlock / 32; ++w) woff += (w < warp) ? s_warp[w] : 0u;
    s_incl[threadIdx.x] = woff + inc;
    __syncthreads();
    const uint32_t total = s_incl[kPreBlock - 1];
    const uint32_t base = g.block_sums[blockIdx.x];
    for (uint32_t k = threadIdx.x; k < total; k += kPreBlock) {
        // first owner o with s_incl[o] > k
        int lo = 0, hi = kPreBlock - 1;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int mid = (lo + hi) >> 1;
            if (s_incl[mid] > k) hi = mid; else lo = mid + 1;
        }
        const int o = lo;
        const uint32_t w = s_w[o];
        const uint32_t cnt_o = s_incl[o] - (o ? s_incl[o - 1] : 0u);
        const uint32_t local = k - (s_incl[o] - cnt_o);
        const uint32_t ry = local / w, rx = local - ry * w;
        const uint32_t rect = s_rect[o];
        const uint32_t tile = ((rect >> 16) + ry) * (uint32_t)grid_x + (rect & 0xffffu) + rx;
        keys_out[base + k] = ((uint64_t)tile << 32) | (uint64_t)s_depth[o];
        vals_out[base + k] = (uint32_t)(blockIdx.x * kPreBlock + o);
    }
}

void launch_duplicate_with_keys(int P, int grid_x, int grid_y, const int* radii, const GeomState& g,
                                uint64_t* keys_out, uint32_t* vals_out, cudaStream_t s) {
    const int nblk = (P + kPreBlock - 1) / kPreBlock;
    duplicate_with_keys_kernel<<<nblk, kPreBlock, 0, s>>>(P, grid_x, grid_y, radii, g, keys_out, vals_out);
}

// ---- 64-bit (tile|depth) radix sort ------------------------------------------------------------------
// Stable LSD sort, 8-bit digits, over the low 32 + ceil(log2(tiles)) key bits only (the reference passes the
// same end_bit to CUB, rasterizer_impl.cu:303-311). Per pass: block histograms -> digit-major offset scan ->
// stable scatter with warp-synchronous match_any ranking.

// Number of significant tile-id bits, same value as the reference's getHigherMsb (rasterizer_impl.cu:35-50).
static int higher_msb(uint32_t n) {
    int msb = 0;
    while (n >> msb) ++msb;   // smallest msb with n >> msb == 0
    return msb;
}

int sort_passes(int grid_x, int grid_y) {
    const int bits = 32 + higher_msb((uint32_t)(grid_x * grid_y));
    return (bits + kRadixBits - 1) / kRadixBits;
}

__global__ void __launch_bounds__(kSortThreads)
radix_hist_kernel(const uint64_t* __restrict__ keys, int R, int shift, int nblk, uint32_t* __restrict__ hist,
                  uint32_t* __restrict__ totals) {
    __shared__ uint32_t s_h[kRadix];
    s_h[threadIdx.x] = 0;
    __syncthreads();
    const int base = blockIdx.x * kSortTile;
    const int lane = threadIdx.x & 31;
#pragma unroll 4
    for (int i = 0; i < kSortItems; ++i) {
        const int idx = base + i * kSortThreads + threadIdx.x;
        const bool valid = idx < R;
        const uint32_t d = valid ? (uint32_t)(keys[idx] >> shift) & (kRadix - 1) : kRadix;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        if (valid && lane == __ffs(peers) - 1) atomicAdd(&s_h[d], __popc(peers));
    }
    __syncthreads();
    const uint32_t c = s_h[threadIdx.x];
    hist[(size_t)threadIdx.x * nblk + blockIdx.x] = c;
    if (c) atomicAdd(&totals[threadIdx.x], c);
}

// block d: global base of digit d (sum of totals below d) + exclusive scan of its row of block counts
__global__ void __launch_bounds__(kSortThreads)
radix_scan_kernel(int nblk, uint32_t* __restrict__ hist, const uint32_t* __restrict__ totals) {
    __shared__ uint32_t s_warp[kSortThreads / 32];
    __shared__ uint32_t s_carry;
    const int d = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t v = (threadIdx.x < d) ? totals[threadIdx.x] : 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) s_warp[warp] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t s = 0;
        for (int w = 0; w < kSortThreads / 32; ++w) s += s_warp[w];
        s_carry = s;
    }
    __syncthreads();
    uint32_t* row = hist + (size_t)d * nblk;
    for (int b0 = 0; b0 < nblk; b0 += kSortThreads) {
        const int i = b0 + threadIdx.x;
        const uint32_t x = (i < nblk) ? row[i] : 0u;
        uint32_t inc = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        uint32_t woff = 0, tot = 0;
#pragma unroll
        for (int w = 0; w < kSortThreads / 32; ++w) { const uint32_t sw = s_warp[w]; woff += (w < warp) ? sw : 0u; tot += sw; }
        const uint32_t carry = s_carry;
        if (i < nblk) row[i] = carry + woff + inc - x;
        __syncthreads();
        if (threadIdx.x == 0) s_carry = carry + tot;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(kSortThreads)
radix_scatter_kernel(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                     uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int R, int shift, int nblk,
                     const uint32_t* __restrict__ hist) {
    constexpr int kWarps = kSortThreads / 32;
    constexpr int kPerWarp = kSortTile / kWarps;   // contiguous keys per warp
    constexpr int kSteps = kPerWarp / 32;
    __shared__ uint32_t s_cnt[kWarps][kRadix];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < kWarps * kRadix; i += kSortThreads) (&s_cnt[0][0])[i] = 0;
    __syncthreads();
    const int wbase = blockIdx.x * kSortTile + warp * kPerWarp;
    uint64_t key[kSteps];
    uint32_t rank[kSteps];
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int st = 0; st < kSteps; ++st) {
        const int idx = wbase + st * 32 + lane;
        const bool valid = idx < R;
        key[st] = valid ? keys_in[idx] : 0ull;
        const uint32_t d = valid ? (uint32_t)(key[st] >> shift) & (kRadix - 1) : kRadix;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(peers) - 1;
        uint32_t prev = 0;
        if (valid && lane == leader) { prev = s_cnt[warp][d]; s_cnt[warp][d] = prev + __popc(peers); }
        prev = __shfl_sync(0xffffffffu, prev, leader);
        rank[st] = prev + __popc(peers & lt);
        __syncwarp();
    }
    __syncthreads();
    {   // per digit: global offset of this block, then exclusive scan over the block's warps
        const int d = threadIdx.x;
        uint32_t off = hist[(size_t)d * nblk + blockIdx.x];
#pragma unroll
        for (int w = 0; w < kWarps; ++w) { const uint32_t c = s_cnt[w][d]; s_cnt[w][d] = off; off += c; }
    }
    __syncthreads();
#pragma unroll
    for (int st = 0; st < kSteps; ++st) {
        const int idx = wbase + st * 32 + lane;
        if (idx < R) {
            const uint32_t d = (uint32_t)(key[st] >> shift) & (kRadix - 1);
            const uint32_t pos = s_cnt[warp][d] + rank[st];
            keys_out[pos] = key[st];
            vals_out[pos] = vals_in[idx];
        }
    }
}

// Sorts (b.keys_alt/b.vals_alt or b.keys/b.point_list, whichever launch_duplicate_with_keys filled for this
// pass count) so that the result always lands in b.keys / b.point_list.
void launch_radix_sort(int R, int passes, const BinningState& b, cudaStream_t s) {
    if (R <= 0) return;
    const int nblk = (int)BinningState::sort_blocks((size_t)R);
    cudaMemsetAsync(b.totals, 0, sizeof(uint32_t) * 8 * kRadix, s);
    // with an odd pass count the unsorted input starts in the alt buffers, with an even count in the main ones
    uint64_t* kin = (passes & 1) ? b.keys_alt : b.keys;
    uint32_t* vin = (passes & 1) ? b.vals_alt : b.point_list;
    uint64_t* kout = (passes & 1) ? b.keys : b.keys_alt;
    uint32_t* vout = (passes & 1) ? b.point_list : b.vals_alt;
    for (int p = 0; p < passes; ++p) {
        const int shift = p * kRadixBits;
        radix_hist_kernel<<<nblk, kSortThreads, 0, s>>>(kin, R, shift, nblk, b.hist, b.totals + p * kRadix);
        radix_scan_kernel<<<kRadix, kSortThreads, 0, s>>>(nblk, b.hist, b.totals + p * kRadix);
        radix_scatter_kernel<<<nblk, kSortThreads, 0, s>>>(kin, vin, kout, vout, R, shift, nblk, b.hist);
        uint64_t* tk = kin; kin = kout; kout = tk;
        uint32_t* tv = vin; vin = vout; vout = tv;
    }
}

// ---- tile ranges + render records --------------------------------------------------------------------
// One thread per sorted instance: detects tile boundaries (identifyTileRanges) and gathers the instance's
// Gaussian into a 48-byte record {x, y, conic.x, conic.y | conic.z, opacity, depth, id | r, g, b, -} so the
// render kernels stream contiguous records with bulk async copies instead of four dependent gathers.
// Conservative axis-aligned half extents (hx, hy), packed as two round-up fp16, of the region where this instance
// can pass the render kernels' alpha >= 1/255 test: alpha = o*exp(-q(d)) >= 1/255 needs q(d) <= tau = ln(255*o), and
// {q <= tau} with q(d) = 0.5*(A dx^2 + C dy^2) + B dx dy is an ellipse whose bounding box has half widths
// sqrt(2 tau C/det), sqrt(2 tau A/det). Margins (+0.01 on tau, x1.001, +0.05 px, fp16 round-up) cover float rounding
// in power/expf, so a pixel the reference would blend is never outside the box. Degenerate conics -> "everywhere".
__device__ __forceinline__ uint32_t footprint_half_extents(float4 co) {
    const float A = co.x, B = co.y, C = co.z, o = co.w;
    float hx = 65504.0f, hy = 65504.0f;
    const float tau0 = logf(255.0f * o) + 0.02f;
    const float det = A * C - B * B;
    if (tau0 < 0.0f) {
        hx = hy = -65504.0f;                      // opacity below 1/255: can never contribute
    } else if (det > 0.0f && A > 0.0f && C > 0.0f) {
        // the terms of `power` reach 2*tau*A*C/det near the ellipse, so its float rounding error scales with the
        // anisotropy A*C/det: widen tau accordingly (negligible for ordinary splats, safe for needle-like ones)
        const float tau = tau0 * (1.0f + 4e-6f * (A * C / det));
        const float ex = sqrtf(2.0f * tau * C / det) * 1.001f + 0.05f;
        const float ey = sqrtf(2.0f * tau * A / det) * 1.001f + 0.05f;
        if (ex < 65504.0f) hx = ex;
        if (ey < 65504.0f) hy = ey;
    }
    const __half2 h = __halves2half2(__float2half_ru(hx), __float2half_ru(hy));
    return *reinterpret_cast<const uint32_t*>(&h);
}

__global__ void __launch_bounds__(256)
ranges_and_records_kernel(int R, const uint64_t* __restrict__ keys, const uint32_t* __restrict__ point_list,
                          const float* __restrict__ colors, GeomState g, uint2* __restrict__ ranges,
                          float4* __restrict__ records) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    const uint32_t tile = (uint32_t)(keys[i] >> 32);
    if (i == 0) {
        ranges[tile].x = 0;
    } else {
        const uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
        if (tile != prev) { ranges[prev].y = i; ranges[tile].x = i; }
    }
    if (i == R - 1) ranges[tile].y = R;
    const uint32_t id = point_list[i];
    const float2 xy = g.means2D[id];
    const float4 co = g.conic_opacity[id];
    const float d = g.depths[id];
    const float r = colors[3 * id], gg = colors[3 * id + 1], b = colors[3 * id + 2];
    records[3 * (size_t)i + 0] = make_float4(xy.x, xy.y, co.x, co.y);
    records[3 * (size_t)i + 1] = make_float4(co.z, co.w, d, __uint_as_float(id));
    records[3 * (size_t)i + 2] = make_float4(r, gg, b, __uint_as_float(footprint_half_extents(co)));
}

void launch_ranges_and_records(int R, int tiles, const float* colors_precomp, const GeomState& g,
                               const BinningState& b, uint2* ranges, cudaStream_t s) {
    cudaMemsetAsync(ranges, 0, sizeof(uint2) * (size_t)tiles, s);
    if (R <= 0) return;
    const float* colors = colors_precomp ? colors_precomp : g.rgb;
    ranges_and_records_kernel<<<(R + 255) / 256, 256, 0, s>>>(R, b.keys, b.point_list, colors, g, ranges,
                                                              reinterpret_cast<float4*>(b.records));
}

}  // namespace splat
