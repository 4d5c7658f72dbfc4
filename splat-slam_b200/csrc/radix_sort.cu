// Radix sort of (32-bit key, 32-bit value) pairs, hand-written for sm_100a, and the way it is used to realise the
// reference's 64-bit (tile | depth) key sort.
// Replaces getHigherMsb + cub::DeviceRadixSort::SortPairs (DGR/cuda_rasterizer/rasterizer_impl.cu:35-50,303-311).
//
// The reference sorts R instance keys (tile << 32 | depth bits) with a stable LSD sort over 32 + ceil(log2(tiles))
// bits. An LSD sort is a sequence of stable passes, low digits first, so it factors exactly:
//   1. the four depth-digit passes only depend on the Gaussian, not on the instance -> run them ONCE PER GAUSSIAN
//      (P items instead of R ~ 4P): sort (depth bits, Gaussian id), culled Gaussians keyed 0xFFFFFFFF to the end;
//   2. emit the instances in that order (duplicate_with_keys_kernel walks the sorted Gaussians);
//   3. run the remaining tile-digit passes (two for <= 65536 tiles) stably over the R instances.
// Ties (same tile, same depth bits) resolve by Gaussian index in both formulations, so the resulting point list,
// keys and tile ranges are bit-identical to the reference's (tests/test_gpu_parity.py).
//
// Each pass is ONE kernel ("one sweep"): it reads its 4096-pair tile once, ranks it with warp-synchronous match_any
// (no atomics; keys stay in registers), gets the tile's global digit offsets by decoupled look-back over the
// preceding tiles' published counts (windows of 16 independent loads), reorders the pairs through shared memory and
// writes each digit run with coalesced stores. The digit totals of all passes come from one up-front histogram
// kernel; every tile scans the 256 totals of its pass itself. Tiles take their index from an atomic ticket, so a tile only waits for tiles that are already
// resident and the look-back cannot deadlock; a bounded spin turns a would-be hang into an error flag.
#include "common.cuh"
#include "kernels.h"

namespace splat {

namespace {
constexpr uint32_t kFlagAgg = 1u << 30;   // tile's own digit count is published
constexpr uint32_t kFlagInc = 2u << 30;   // inclusive prefix (all tiles up to this one) is published
constexpr uint32_t kFlagMask = 3u << 30;
constexpr uint32_t kValMask = ~kFlagMask;
constexpr int kWarps = kSortThreads / 32;
constexpr int kPerWarp = kSortTile / kWarps;   // contiguous pairs per warp (512)
constexpr int kSteps = kPerWarp / 32;          // 16
}  // namespace

// Number of significant tile-id bits, same value as the reference's getHigherMsb (rasterizer_impl.cu:35-50).
static int higher_msb(uint32_t n) {
    int msb = 0;
    while (n >> msb) ++msb;   // smallest msb with n >> msb == 0
    return msb;
}

int tile_sort_passes(int grid_x, int grid_y) {
    return (higher_msb((uint32_t)(grid_x * grid_y)) + kRadixBits - 1) / kRadixBits;
}

// ---- all digit histograms from one read of the keys ------------------------------------------------------
__global__ void __launch_bounds__(kSortThreads)
radix_global_hist_kernel(const uint32_t* __restrict__ keys, int n, int passes, uint32_t* __restrict__ ghist) {
    __shared__ uint32_t s_h[4][kRadix];
    for (int i = threadIdx.x; i < 4 * kRadix; i += kSortThreads) (&s_h[0][0])[i] = 0;
    __syncthreads();
    const int stride = (int)gridDim.x * kSortThreads;
    for (int base = (int)blockIdx.x * kSortThreads + (int)threadIdx.x; base < n; base += 4 * stride) {
        uint32_t k[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) k[u] = (base + u * stride < n) ? keys[base + u * stride] : 0u;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (base + u * stride < n) {
                for (int p = 0; p < passes; ++p) atomicAdd(&s_h[p][(k[u] >> (p * kRadixBits)) & (kRadix - 1)], 1u);
            }
        }
    }
    __syncthreads();
    for (int p = 0; p < passes; ++p) {
        const uint32_t c = s_h[p][threadIdx.x];
        if (c) atomicAdd(&ghist[p * kRadix + threadIdx.x], c);
    }
}

// ---- one pass ---------------------------------------------------------------------------------------------
// kEmitRecords: the LAST pass of the instance sort. Its scatter already knows every instance's final position, so
// instead of writing the sorted (tile, Gaussian id) pairs and having another kernel read them back, it gathers each
// Gaussian's 48-byte render record straight into sorted order (what ranges_and_records_kernel did in round 1) and
// marks the tile ranges: equal tiles are adjacent inside a sort tile, the first and last instance of every such run
// lower ranges[tile].x to their position / ranges[tile].y to ~(position + 1) with atomicMin (both start at
// 0xFFFFFFFF; render_forward_kernel rewrites its tile's entry as [start, end), or (0, 0) for an empty tile, which is
// identifyTileRanges' result, DGR/cuda_rasterizer/rasterizer_impl.cu:116-138).
template <bool kEmitRecords>
__global__ void __launch_bounds__(kSortThreads, 3)
radix_onesweep_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                      uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int n, int shift,
                      const uint32_t* __restrict__ gbase, uint32_t* tile_state, uint32_t* ticket,
                      uint32_t* error_flag, int ntiles, RecordSink sink) {
    // per-warp digit counts -> per-warp local positions; with kEmitRecords the same words later stage the records
    // of 32 instances per warp (96 float4 = 384 words each)
    constexpr int kCntWords = kEmitRecords ? kWarps * 384 : kWarps * kRadix;
    __shared__ __align__(16) uint32_t s_cnt_raw[kCntWords];
    uint32_t (*s_cnt)[kRadix] = reinterpret_cast<uint32_t (*)[kRadix]>(s_cnt_raw);
    __shared__ __align__(16) uint2 s_kv[kSortTile];  // reorder buffer (key, value)
    __shared__ uint32_t s_goff[kRadix];              // global offset of the digit run minus its local base
    __shared__ uint32_t s_scan[kWarps];
    __shared__ uint32_t s_tile;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    for (int i = threadIdx.x; i < kWarps * kRadix; i += kSortThreads) (&s_cnt[0][0])[i] = 0;
    __syncthreads();
    const int tile = (int)s_tile;
    if (tile >= ntiles) return;
    const int wbase = tile * kSortTile + warp * kPerWarp;

    // all loads are issued before the ranking loop: __syncwarp() in that loop is a memory barrier and would
    // otherwise serialise one global round trip per step
    uint32_t key[kSteps], val[kSteps], rank[kSteps];
#pragma unroll
    for (int st = 0; st < kSteps; ++st) {
        const int idx = wbase + st * 32 + lane;
        key[st] = (idx < n) ? keys_in[idx] : 0u;
    }
#pragma unroll
    for (int st = 0; st < kSteps; ++st) {
        const int idx = wbase + st * 32 + lane;
        val[st] = (idx < n) ? vals_in[idx] : 0u;
    }
    // rank[st] = number of earlier keys of this warp with the same digit (stable within the warp's chunk)
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int st = 0; st < kSteps; ++st) {
        const int idx = wbase + st * 32 + lane;
        const bool valid = idx < n;
        const uint32_t d = valid ? (key[st] >> shift) & (kRadix - 1) : (uint32_t)kRadix;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(peers) - 1;
        uint32_t prev = 0;
        if (valid && lane == leader) { prev = s_cnt[warp][d]; s_cnt[warp][d] = prev + __popc(peers); }
        prev = __shfl_sync(0xffffffffu, prev, leader);
        rank[st] = prev + __popc(peers & lt);
        __syncwarp();
    }
    __syncthreads();

    // thread d owns digit d: tile count, exclusive prefix over this tile's warps, look-back over earlier tiles
    const int d = threadIdx.x;
    uint32_t c = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) { const uint32_t t = s_cnt[w][d]; s_cnt[w][d] = c; c += t; }
    volatile uint32_t* state = tile_state + (size_t)tile * kRadix;
    uint32_t prefix = 0;
    if (tile == 0) {
        state[d] = c | kFlagInc;
    } else {
        state[d] = c | kFlagAgg;
        // Look back kLook predecessors at a time: the loads are independent, so a window costs one L2 round trip
        // instead of kLook (when a whole wave of tiles starts together none has an inclusive prefix yet and the walk
        // is hundreds of tiles long).
        constexpr int kLook = 16;
        bool found = false;
        for (int t = tile - 1; t >= 0 && !found; t -= kLook) {
            uint32_t v[kLook];
#pragma unroll
            for (int i = 0; i < kLook; ++i)
                v[i] = (t - i >= 0) ? (uint32_t)*(volatile uint32_t*)(tile_state + (size_t)(t - i) * kRadix + d)
                                    : (uint32_t)(2u << 30);   // before tile 0: inclusive prefix 0
#pragma unroll
            for (int i = 0; i < kLook; ++i) {
                if (found) break;
                if ((v[i] & kFlagMask) == 0u) {
                    volatile uint32_t* ps = tile_state + (size_t)(t - i) * kRadix + d;
                    int spins = 0;
                    do {
                        v[i] = *ps;
                        if (++spins > (1 << 22)) { atomicExch(error_flag, 1u); break; }
                    } while ((v[i] & kFlagMask) == 0u);
                }
                prefix += v[i] & kValMask;
                found = (v[i] & kFlagInc) != 0u;
            }
        }
        state[d] = (prefix + c) | kFlagInc;
    }
    // local base of digit d inside the tile: exclusive scan of c over the 256 digits
    uint32_t inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_scan[warp] = inc;
    __syncthreads();
    uint32_t woff = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) woff += (w < warp) ? s_scan[w] : 0u;
    const uint32_t local_base = woff + inc - c;
    // global base of digit d = exclusive scan of the pass's digit totals (raw histogram in, scanned here by every
    // tile: 256 values, cheaper than a launch of its own)
    const uint32_t gtot = gbase[d];
    uint32_t ginc = gtot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, ginc, o);
        if (lane >= o) ginc += t;
    }
    __syncthreads();                      // s_scan is reused
    if (lane == 31) s_scan[warp] = ginc;
    __syncthreads();
    uint32_t gwoff = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) gwoff += (w < warp) ? s_scan[w] : 0u;
    s_goff[d] = (gwoff + ginc - gtot) + prefix - local_base;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) s_cnt[w][d] += local_base;
    __syncthreads();

    // reorder the pairs through shared memory, then write the digit runs with coalesced stores
#pragma unroll
    for (int st = 0; st < kSteps; ++st) {
        const int idx = wbase + st * 32 + lane;
        if (idx < n) {
            const uint32_t dg = (key[st] >> shift) & (kRadix - 1);
            s_kv[s_cnt[warp][dg] + rank[st]] = make_uint2(key[st], val[st]);
        }
    }
    __syncthreads();
    const int n_valid = min(kSortTile, n - tile * kSortTile);
    if constexpr (!kEmitRecords) {
        for (int i = threadIdx.x; i < n_valid; i += kSortThreads) {
            const uint2 kv = s_kv[i];
            const uint32_t pos = s_goff[(kv.x >> shift) & (kRadix - 1)] + i;
            keys_out[pos] = kv.x;
            vals_out[pos] = kv.y;
        }
    } else {
        // (Gathering four instances per thread at a time and storing straight from registers, without the transposition,
        // was measured and is slower: 0.182 vs 0.173 ms for the pass.)
        float4* s_t = reinterpret_cast<float4*>(s_cnt_raw) + warp * 96;   // the digit counts are dead by now
        for (int base = 0; base < n_valid; base += kSortThreads) {
            const int i = base + (int)threadIdx.x;
            uint32_t pos = 0;
            if (i < n_valid) {
                const uint2 kv = s_kv[i];
                pos = s_goff[(kv.x >> shift) & (kRadix - 1)] + (uint32_t)i;
                // first / last instance of a run of equal tiles inside this sort tile (~0 is never a tile id)
                const uint32_t prev = (i > 0) ? s_kv[i - 1].x : 0xFFFFFFFFu;
                const uint32_t next = (i + 1 < n_valid) ? s_kv[i + 1].x : 0xFFFFFFFFu;
                if (prev != kv.x) atomicMin(&sink.ranges[kv.x].x, pos);
                if (next != kv.x) atomicMin(&sink.ranges[kv.x].y, ~(pos + 1u));
                const float4* src = sink.gauss_rec + 3 * (size_t)kv.y;
                const float4 r0 = __ldg(src), r1 = __ldg(src + 1), r2 = __ldg(src + 2);
                s_t[lane * 3 + 0] = r0;
                s_t[lane * 3 + 1] = r1;
                s_t[lane * 3 + 2] = r2;
                if (sink.keys64 != nullptr) {      // inspection only (splat_raster_keep_sort_output)
                    sink.keys64[pos] = ((uint64_t)kv.x << 32) | (uint64_t)__float_as_uint(r2.w);
                    sink.point_list[pos] = kv.y;
                }
            }
            __syncwarp();
            // 32 records = 96 float4, mostly contiguous in the output (a digit run is): lane l writes parts l, l+32, l+64
            const int wbase = base + warp * 32;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const int e = k * 32 + lane;
                const int el = e / 3, part = e - 3 * el;
                const uint32_t p = __shfl_sync(0xffffffffu, pos, el);
                if (wbase + el < n_valid) sink.records[3 * (size_t)p + part] = s_t[e];
            }
            __syncwarp();
        }
    }
}

// Stable LSD sort of n (key, value) pairs on key bits [0, 8*passes). Ping-pongs between (k0, v0) and (k1, v1): the
// input is read from (k0, v0); the result lands in (k0, v0) for an even pass count and in (k1, v1) for an odd one.
// With `sink` the last pass emits render records and tile ranges instead of the sorted pairs (see the kernel).
// scratch: SortScratch::words(n) zero-initialised-by-us 32-bit words.
void launch_sort_pairs(uint32_t* k0, uint32_t* v0, uint32_t* k1, uint32_t* v1, int n, int passes, uint32_t* scratch,
                       cudaStream_t s, const RecordSink* sink) {
    if (n <= 0 || passes <= 0) return;
    const int ntiles = (n + kSortTile - 1) / kSortTile;
    // scratch: [4][256] digit histograms, [4] tickets, error flag (64 words), then [passes][ntiles][256] tile states
    cudaMemsetAsync(scratch, 0, sizeof(uint32_t) * (4 * kRadix + 64 + (size_t)passes * ntiles * kRadix), s);
    uint32_t* ghist = scratch;
    uint32_t* tickets = scratch + 4 * kRadix;
    uint32_t* error_flag = tickets + 8;
    uint32_t* states = scratch + 4 * kRadix + 64;
    const int hist_blocks = ntiles < 592 ? ntiles : 592;   // 148 SMs x 4
    radix_global_hist_kernel<<<hist_blocks, kSortThreads, 0, s>>>(k0, n, passes, ghist);
    uint32_t *kin = k0, *vin = v0, *kout = k1, *vout = v1;
    for (int p = 0; p < passes; ++p) {
        if (sink != nullptr && p == passes - 1)
            radix_onesweep_kernel<true><<<ntiles, kSortThreads, 0, s>>>(
                kin, vin, kout, vout, n, p * kRadixBits, ghist + p * kRadix, states + (size_t)p * ntiles * kRadix,
                tickets + p, error_flag, ntiles, *sink);
        else
            radix_onesweep_kernel<false><<<ntiles, kSortThreads, 0, s>>>(
                kin, vin, kout, vout, n, p * kRadixBits, ghist + p * kRadix, states + (size_t)p * ntiles * kRadix,
                tickets + p, error_flag, ntiles, RecordSink());
        uint32_t* tk = kin; kin = kout; kout = tk;
        uint32_t* tv = vin; vin = vout; vout = tv;
    }
}

}  // namespace splat
