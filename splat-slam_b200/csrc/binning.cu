// Binning: instance offsets and tile-key generation in depth order (tile ranges and the per-instance render records
// come out of the last sort pass). Together with radix_sort.cu this replaces cub::DeviceScan::InclusiveSum (DGR/cuda_rasterizer/rasterizer_impl.cu:280),
// duplicateWithKeys (:70-111), cub::DeviceRadixSort::SortPairs (:303-311), cudaMemset + identifyTileRanges
// (:313-321,116-138).
#include <cstdio>
#include "common.cuh"
#include "kernels.h"

namespace splat {

// ---- offsets -----------------------------------------------------------------------------------------
// Instance count of every block of 256 Gaussians taken in depth-sorted order; the block that finishes last turns the
// counts into their exclusive scan (the per-Gaussian offsets themselves never reach HBM: duplicate_with_keys_kernel
// rebuilds them inside the block). `state`: word 0 receives the total, word 2 is the ticket (zero on entry).
__global__ void __launch_bounds__(kPreBlock)
sorted_block_sums_kernel(int P, int nblk, const uint32_t* __restrict__ order, const uint32_t* __restrict__ tiles_touched,
                         uint32_t* __restrict__ sums, uint32_t* __restrict__ state) {
    __shared__ uint32_t s_warp[kPreBlock / 32];
    __shared__ bool s_last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int i = blockIdx.x * kPreBlock + threadIdx.x;
    uint32_t v = (i < P) ? tiles_touched[order[i]] : 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) s_warp[warp] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t s = 0;
#pragma unroll
        for (int w = 0; w < kPreBlock / 32; ++w) s += s_warp[w];
        sums[blockIdx.x] = s;
        __threadfence();
        s_last = (atomicAdd(&state[2], 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // exclusive scan of sums[0..nblk): every thread takes a contiguous chunk
    const int per = (nblk + kPreBlock - 1) / kPreBlock;
    const int lo = min(nblk, (int)threadIdx.x * per), hi = min(nblk, lo + per);
    uint32_t chunk = 0;
    for (int k = lo; k < hi; ++k) chunk += __ldcg(sums + k);
    uint32_t inc = chunk;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    __syncthreads();
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t woff = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kPreBlock / 32; ++w) { woff += (w < warp) ? s_warp[w] : 0u; total += s_warp[w]; }
    uint32_t running = woff + inc - chunk;
    for (int k = lo; k < hi; ++k) {
        const uint32_t c = __ldcg(sums + k);
        sums[k] = running;
        running += c;
    }
    if (threadIdx.x == 0) { sums[nblk] = total; state[0] = total; state[2] = 0u; }
}

// Depth sort of the Gaussians (4 digit passes over P items, see radix_sort.cu) and the instance offsets in that order.
void launch_depth_sort_and_offsets(int P, const GeomState& g, cudaStream_t s) {
    launch_sort_pairs(g.dkey, g.order, g.dkey_alt, g.order_alt, P, 4, g.sort_scratch, s);   // even: result in place
    const int nblk = (P + kPreBlock - 1) / kPreBlock;
    sorted_block_sums_kernel<<<nblk, kPreBlock, 0, s>>>(P, nblk, g.order, g.tiles_touched, g.block_sums, g.num_rendered);
}

// ---- key generation ----------------------------------------------------------------------------------
// One block per 256 Gaussians of the DEPTH-SORTED order. The block rebuilds its local offsets by a shared-memory
// scan, then ALL threads emit the block's instances cooperatively (instance k -> owner by binary search) so that the
// 8-byte (tile, id) writes are fully coalesced and a screen-filling Gaussian costs its neighbours nothing.
// Emission order per Gaussian is row-major over its tile rectangle, as in the reference.
__global__ void __launch_bounds__(kPreBlock)
duplicate_with_keys_kernel(int P, int grid_x, int grid_y, const int* __restrict__ radii, GeomState g,
                           uint32_t* __restrict__ tkeys_out, uint32_t* __restrict__ vals_out,
                           uint2* __restrict__ ranges, int tiles) {
    // preset of the tile ranges for the last sort pass (see radix_onesweep_kernel<true>): done here, by the first
    // blocks, instead of by a memset node of its own
    for (int t = blockIdx.x * kPreBlock + threadIdx.x; t < tiles; t += gridDim.x * kPreBlock)
        ranges[t] = make_uint2(0xFFFFFFFFu, 0xFFFFFFFFu);
    // a depth sort whose look-back gave up must not go unnoticed: fail the stream loudly
    if (blockIdx.x == 0 && threadIdx.x == 0 && *sort_error_flag(g.sort_scratch) != 0u) {
        printf("splat_raster: depth sort look-back timed out\n");
        __trap();
    }
    __shared__ uint32_t s_incl[kPreBlock];
    __shared__ uint32_t s_rect[kPreBlock];   // x0 | y0 << 16
    __shared__ uint32_t s_w[kPreBlock];
    __shared__ uint32_t s_gid[kPreBlock];
    __shared__ uint32_t s_warp[kPreBlock / 32];
    const int i = blockIdx.x * kPreBlock + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t cnt = 0;
    if (i < P) {
        const uint32_t gid = g.order[i];
        const int radius = radii[gid];
        if (radius > 0) {
            const float2 p = g.means2D[gid];
            uint32_t x0, y0, x1, y1;
            get_rect_ref(p.x, p.y, radius, grid_x, grid_y, x0, y0, x1, y1);
            cnt = (x1 - x0) * (y1 - y0);
            s_rect[threadIdx.x] = x0 | (y0 << 16);
            s_w[threadIdx.x] = x1 - x0;
            s_gid[threadIdx.x] = gid;
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
        tkeys_out[base + k] = ((rect >> 16) + ry) * (uint32_t)grid_x + (rect & 0xffffu) + rx;
        vals_out[base + k] = s_gid[o];
    }
}

void launch_duplicate_with_keys(int P, int grid_x, int grid_y, const int* radii, const GeomState& g,
                                uint32_t* tkeys_out, uint32_t* vals_out, uint2* ranges, cudaStream_t s) {
    const int nblk = (P + kPreBlock - 1) / kPreBlock;
    duplicate_with_keys_kernel<<<nblk, kPreBlock, 0, s>>>(P, grid_x, grid_y, radii, g, tkeys_out, vals_out, ranges,
                                                          grid_x * grid_y);
}

// ---- tile ranges ---------------------------------------------------------------------------------------
// Tile ranges and the per-instance render records are produced by the last pass of the instance sort
// (radix_sort.cu: radix_onesweep_kernel<true>); here only the preset every range starts from.
void launch_reset_ranges(int tiles, uint2* ranges, cudaStream_t s) {
    cudaMemsetAsync(ranges, 0xFF, sizeof(uint2) * (size_t)tiles, s);
}

}  // namespace splat
