// 64-bit (tile | depth) key / 32-bit value radix sort, hand-written for sm_100a.
// Replaces getHigherMsb + cub::DeviceRadixSort::SortPairs (DGR/cuda_rasterizer/rasterizer_impl.cu:35-50,303-311).
//
// Stable LSD sort with 8-bit digits over the low 32 + ceil(log2(tiles)) key bits only (the reference passes the same
// end_bit to CUB). "One sweep" structure: one up-front kernel builds the global digit histograms of ALL passes from a
// single read of the keys; each pass is then ONE kernel that reads its 4096-key tile once, ranks it with
// warp-synchronous match_any (no atomics), obtains the tile's global digit offsets by decoupled look-back over the
// preceding tiles' published counts, reorders keys and values through shared memory and writes each digit run with
// fully coalesced stores. Traffic per pass: 12 B read + 12 B written per instance.
//
// Tiles take their index from an atomic ticket, so a tile only ever waits for tiles that are already resident:
// the look-back cannot deadlock. A bounded spin turns a would-be hang (e.g. scratch not cleared) into an error flag.
#include "common.cuh"
#include "kernels.h"

namespace splat {

namespace {
constexpr uint32_t kFlagAgg = 1u << 30;   // tile's own digit count is published
constexpr uint32_t kFlagInc = 2u << 30;   // inclusive prefix (all tiles up to this one) is published
constexpr uint32_t kFlagMask = 3u << 30;
constexpr uint32_t kValMask = ~kFlagMask;
constexpr int kWarps = kSortThreads / 32;
constexpr int kPerWarp = kSortTile / kWarps;   // contiguous keys per warp (512)
constexpr int kSteps = kPerWarp / 32;          // 16
}  // namespace

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

// ---- all digit histograms from one read of the keys ------------------------------------------------------
__global__ void __launch_bounds__(kSortThreads)
radix_global_hist_kernel(const uint64_t* __restrict__ keys, int R, int passes, uint32_t* __restrict__ ghist) {
    __shared__ uint32_t s_h[8][kRadix];
    for (int i = threadIdx.x; i < 8 * kRadix; i += kSortThreads) (&s_h[0][0])[i] = 0;
    __syncthreads();
    // keys arrive in emission order (by Gaussian), so neighbouring lanes rarely share a digit: plain shared-memory
    // atomics beat warp aggregation here; four keys per thread in flight hide the load latency
    const int stride = (int)gridDim.x * kSortThreads;
    for (int base = (int)blockIdx.x * kSortThreads + (int)threadIdx.x; base < R; base += 4 * stride) {
        uint64_t k[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) k[u] = (base + u * stride < R) ? keys[base + u * stride] : 0ull;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (base + u * stride < R) {
                for (int p = 0; p < passes; ++p)
                    atomicAdd(&s_h[p][(uint32_t)(k[u] >> (p * kRadixBits)) & (kRadix - 1)], 1u);
            }
        }
    }
    __syncthreads();
    for (int p = 0; p < passes; ++p) {
        const uint32_t c = s_h[p][threadIdx.x];
        if (c) atomicAdd(&ghist[p * kRadix + threadIdx.x], c);
    }
}

// block p: exclusive scan of pass p's 256 digit totals (in place)
__global__ void __launch_bounds__(kRadix) radix_global_scan_kernel(uint32_t* __restrict__ ghist) {
    __shared__ uint32_t s_warp[kRadix / 32];
    uint32_t* h = ghist + blockIdx.x * kRadix;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t v = h[threadIdx.x];
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t woff = 0;
#pragma unroll
    for (int w = 0; w < kRadix / 32; ++w) woff += (w < warp) ? s_warp[w] : 0u;
    h[threadIdx.x] = woff + inc - v;
}

// ---- one pass ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kSortThreads)
radix_onesweep_kernel(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                      uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int R, int shift,
                      const uint32_t* __restrict__ gbase, uint32_t* tile_state, uint32_t* ticket,
                      uint32_t* error_flag, int ntiles) {
    __shared__ uint32_t s_cnt[kWarps][kRadix];     // per-warp digit counts -> per-warp local positions
    __shared__ __align__(16) uint64_t s_keys[kSortTile];   // reorder buffer (keys, then reused for values)
    __shared__ uint32_t s_goff[kRadix];            // global offset of the digit run minus its local base
    __shared__ uint8_t s_dig[kSortTile];           // digit of each reordered slot (for the value pass)
    __shared__ uint32_t s_scan[kWarps];
    __shared__ uint32_t s_tile;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    for (int i = threadIdx.x; i < kWarps * kRadix; i += kSortThreads) (&s_cnt[0][0])[i] = 0;
    __syncthreads();
    const int tile = (int)s_tile;
    if (tile >= ntiles) return;
    const int wbase = tile * kSortTile + warp * kPerWarp;

    // rank: keys stay in registers; rank[st] = number of earlier keys of this warp with the same digit
    uint64_t key[kSteps];
    uint32_t rank[kSteps];
    const uint32_t lt = (1u << lane) - 1u;
    // all 16 loads are issued before the ranking loop: __syncwarp() in that loop is a memory barrier and would
    // otherwise serialise one global round trip per step
#pragma unroll
    for (int st = 0; st < kSteps; ++st) {
        const int idx = wbase + st * 32 + lane;
        key[st] = (idx < R) ? keys_in[idx] : 0ull;
    }
#pragma unroll
    for (int st = 0; st < kSteps; ++st) {
        const int idx = wbase + st * 32 + lane;
        const bool valid = idx < R;
        const uint32_t d = valid ? (uint32_t)(key[st] >> shift) & (kRadix - 1) : (uint32_t)kRadix;
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
        // instead of kLook. (When a whole wave of tiles starts together none of them has an inclusive prefix yet and
        // the walk is hundreds of tiles long; one dependent load per tile made that walk the cost of the pass.)
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
    s_goff[d] = gbase[d] + prefix - local_base;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) s_cnt[w][d] += local_base;
    __syncthreads();

    // reorder keys through shared memory; rank[] becomes the local slot
#pragma unroll
    for (int st = 0; st < kSteps; ++st) {
        const int idx = wbase + st * 32 + lane;
        if (idx < R) {
            const uint32_t dg = (uint32_t)(key[st] >> shift) & (kRadix - 1);
            const uint32_t pos = s_cnt[warp][dg] + rank[st];
            rank[st] = pos;
            s_keys[pos] = key[st];
        }
    }
    __syncthreads();
    const int n_valid = min(kSortTile, R - tile * kSortTile);
    for (int i = threadIdx.x; i < n_valid; i += kSortThreads) {
        const uint64_t k = s_keys[i];
        const uint32_t dg = (uint32_t)(k >> shift) & (kRadix - 1);
        s_dig[i] = (uint8_t)dg;
        keys_out[s_goff[dg] + i] = k;
    }
    __syncthreads();
    uint32_t* s_vals = reinterpret_cast<uint32_t*>(s_keys);
#pragma unroll
    for (int st = 0; st < kSteps; ++st) {
        const int idx = wbase + st * 32 + lane;
        if (idx < R) s_vals[rank[st]] = vals_in[idx];
    }
    __syncthreads();
    for (int i = threadIdx.x; i < n_valid; i += kSortThreads) vals_out[s_goff[s_dig[i]] + i] = s_vals[i];
}

// Sorts (b.keys_alt/b.vals_alt or b.keys/b.point_list, whichever the key generation filled for this pass count) so
// that the result always lands in b.keys / b.point_list.
void launch_radix_sort(int R, int passes, const BinningState& b, cudaStream_t s) {
    if (R <= 0) return;
    const int ntiles = (int)BinningState::sort_blocks((size_t)R);
    // scratch: [8][256] digit histograms, [8] tickets + error flag, then [passes][ntiles][256] tile states
    cudaMemsetAsync(b.totals, 0, BinningState::sort_scratch_bytes((size_t)R, passes), s);
    uint32_t* ghist = b.totals;
    uint32_t* tickets = b.totals + 8 * kRadix;
    uint32_t* error_flag = tickets + 8;
    uint32_t* states = b.hist;
    const int hist_blocks = ntiles < 592 ? ntiles : 592;   // 148 SMs x 4
    const int sweep_blocks = ntiles;                       // tiles take their index from the ticket, not blockIdx
    radix_global_hist_kernel<<<hist_blocks, kSortThreads, 0, s>>>(
        (passes & 1) ? b.keys_alt : b.keys, R, passes, ghist);
    radix_global_scan_kernel<<<passes, kRadix, 0, s>>>(ghist);
    // with an odd pass count the unsorted input starts in the alt buffers, with an even count in the main ones
    uint64_t* kin = (passes & 1) ? b.keys_alt : b.keys;
    uint32_t* vin = (passes & 1) ? b.vals_alt : b.point_list;
    uint64_t* kout = (passes & 1) ? b.keys : b.keys_alt;
    uint32_t* vout = (passes & 1) ? b.point_list : b.vals_alt;
    for (int p = 0; p < passes; ++p) {
        radix_onesweep_kernel<<<sweep_blocks, kSortThreads, 0, s>>>(kin, vin, kout, vout, R, p * kRadixBits,
                                                                   ghist + p * kRadix,
                                                                   states + (size_t)p * ntiles * kRadix, tickets + p,
                                                                   error_flag, ntiles);
        uint64_t* tk = kin; kin = kout; kout = tk;
        uint32_t* tv = vin; vin = vout; vout = tv;
    }
}

int radix_sort_kernel_count(int passes) { return passes + 2; }

}  // namespace splat
