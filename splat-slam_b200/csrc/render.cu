// Tile renderers: forward front-to-back alpha compositing of colour/depth/opacity and the backward pass with
// back-to-front recomposition. Replaces DGR/cuda_rasterizer/forward.cu:263-392 and backward.cu:541-787.
//
// B200 design: one CTA per 16x16 tile (tile ids/ranges stay bit-compatible with the reference). Each warp owns
// an 8x4 pixel patch so whole-warp rejection of a Gaussian is likely. Per-instance 48-byte records (built in
// sorted order by ranges_and_records_kernel) are streamed into a two-stage shared-memory ring with
// cp.async.bulk (TMA 1-D bulk copy, SASS UBLKCP) completing on an mbarrier; the consumer loop reads them with
// broadcast LDS.128. The backward reduces the 10 per-Gaussian gradient terms across the warp with a
// recursive-halving shuffle tree (12 SHFL instead of 50) and issues one RED per term per warp.
#include <cuda_fp16.h>
#include "common.cuh"
#include "kernels.h"

namespace splat {

constexpr int kBatch = 128;                              // records per pipeline stage

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// one thread: arm the barrier with the byte count and launch the bulk copy global -> shared
__device__ __forceinline__ void bulk_load(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    } while (!ok);
}

// pixel owned by (warp, lane): warps tile the 16x16 block as 2 x 4 patches of 8 x 4 pixels
__device__ __forceinline__ void pixel_of_thread(int tile_x, int tile_y, uint32_t& px, uint32_t& py) {
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    px = tile_x * SPLAT_TILE + (warp & 1) * 8 + (lane & 7);
    py = tile_y * SPLAT_TILE + (warp >> 1) * 4 + (lane >> 3);
}

// pixel rectangle [x0, x1] x [y0, y1] (inclusive, as floats) of this warp's 8x4 patch
__device__ __forceinline__ float4 patch_of_warp(int tile_x, int tile_y) {
    const uint32_t warp = threadIdx.x >> 5;
    const float x0 = (float)(tile_x * SPLAT_TILE + (warp & 1) * 8), y0 = (float)(tile_y * SPLAT_TILE + (warp >> 1) * 4);
    return make_float4(x0, y0, x0 + 7.0f, y0 + 3.0f);
}

// Per-warp binning of one staged batch: each lane tests cnt/32 records against the warp's pixel patch using the
// conservative footprint box stored in the record (ranges_and_records_kernel) and the survivors' batch indices are
// compacted, in order, into `list`. A (warp, instance) pair that cannot reach alpha >= 1/255 anywhere in the patch
// then costs ~0.5 instructions instead of a full power/exp evaluation by all 32 lanes. Comparisons are written so
// that NaN coordinates count as "inside" (the reference would still evaluate them). Returns the survivor count.
__device__ __forceinline__ int bin_batch_for_warp(const float4* rec, int cnt, float4 patch, uint8_t* list,
                                                  uint32_t lane) {
    int n = 0;
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < kBatch / 32; ++r) {
        const int j = r * 32 + (int)lane;
        bool live = false;
        if (j < cnt) {
            const float4 r0 = rec[3 * j];
            const uint32_t hw = __float_as_uint(rec[3 * j + 2].w);
            const float2 h = __half22float2(*reinterpret_cast<const __half2*>(&hw));
            live = !(r0.x < patch.x - h.x) && !(r0.x > patch.z + h.x) && !(r0.y < patch.y - h.y) &&
                   !(r0.y > patch.w + h.y);
        }
        const uint32_t m = __ballot_sync(0xffffffffu, live);
        if (live) list[n + __popc(m & lt)] = (uint8_t)j;
        n += __popc(m);
    }
    __syncwarp();
    return n;
}

// =====================================================================================================
// forward
// =====================================================================================================
__global__ void __launch_bounds__(SPLAT_TILE_PIXELS)
render_forward_kernel(int W, int H, int grid_x, const uint2* __restrict__ ranges, const float* __restrict__ records,
                      const float* __restrict__ bg, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                      float* __restrict__ out_color, float* __restrict__ out_depth, float* __restrict__ out_opacity,
                      int* __restrict__ n_touched) {
    __shared__ __align__(128) float4 s_rec[2][kBatch * 3];
    __shared__ __align__(8) uint64_t s_bar[2];
    __shared__ int s_cnt[2][kBatch];
    __shared__ uint8_t s_list[SPLAT_TILE_PIXELS / 32][kBatch];
    __shared__ uint32_t s_tv[SPLAT_TILE_PIXELS / 32][kBatch];

    const int tile = blockIdx.x;
    const int tile_x = tile % grid_x, tile_y = tile / grid_x;
    uint32_t px, py;
    pixel_of_thread(tile_x, tile_y, px, py);
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const uint32_t pix_id = (uint32_t)W * py + px;
    const float pixfx = (float)px, pixfy = (float)py;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float4 patch = patch_of_warp(tile_x, tile_y);

    const uint2 range = ranges[tile];
    const int n = (int)(range.y - range.x);
    const int nb = (n + kBatch - 1) / kBatch;
    const float4* src = reinterpret_cast<const float4*>(records) + 3 * (size_t)range.x;

    if (threadIdx.x == 0) { mbar_init(&s_bar[0], 1); mbar_init(&s_bar[1], 1); fence_mbar_init(); }
    if (threadIdx.x < kBatch) { s_cnt[0][threadIdx.x] = 0; s_cnt[1][threadIdx.x] = 0; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int b = 0; b < 2 && b < nb; ++b) {
            const int cnt = min(kBatch, n - b * kBatch);
            bulk_load(&s_rec[b][0], src + 3 * (size_t)b * kBatch, (uint32_t)cnt * 48u, &s_bar[b]);
        }
    }

    bool done = !inside;
    float T = 1.0f, C0 = 0.0f, C1 = 0.0f, C2 = 0.0f, Dp = 0.0f;
    uint32_t last_contributor = 0;

    int b = 0;
    for (; b < nb; ++b) {
        // whole tile saturated -> stop (forward.cu:318-320)
        if (__syncthreads_count(done) == SPLAT_TILE_PIXELS) break;
        const int stage = b & 1;
        const int cnt = min(kBatch, n - b * kBatch);
        mbar_wait(&s_bar[stage], (uint32_t)((b >> 1) & 1));
        const float4* rec = &s_rec[stage][0];
        const int n_live = __all_sync(0xffffffffu, done) ? 0 : bin_batch_for_warp(rec, cnt, patch, s_list[warp], lane);
        int k = 0;
        for (; k < n_live; ++k) {
            // survivors of the binning almost always touch some lane, so everything up to the blend is evaluated
            // unconditionally and a single ballot per instance serves both the blend and the n_touched count
            const int j = s_list[warp][k];
            const float4 r0 = rec[3 * j];          // x, y, conic.x, conic.y
            const float4 r1 = rec[3 * j + 1];      // conic.z, opacity, depth, id
            const float dx = fsub(r0.x, pixfx), dy = fsub(r0.y, pixfy);
            const float power = gauss_power_ref(dx, dy, r0.z, r0.w, r1.x);
            // alpha = min(0.99, opacity * exp(power)); 1/255 cut; transmittance test (forward.cu:350-363)
            const float alpha = fminf(0.99f, fmul(r1.y, expf(power)));
            const float test_T = fmul(T, fsub(1.0f, alpha));
            bool act = !done && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
            if (act && test_T < 0.0001f) { done = true; act = false; }
            if (act) {
                const float4 r2 = rec[3 * j + 2];  // r, g, b
                // C += c * alpha * T as fma(T, alpha*c, C) (reference SASS order)
                C0 = ffma(T, fmul(alpha, r2.x), C0);
                C1 = ffma(T, fmul(alpha, r2.y), C1);
                C2 = ffma(T, fmul(alpha, r2.z), C2);
                Dp = ffma(T, fmul(alpha, r1.z), Dp);
                T = test_T;
                last_contributor = (uint32_t)(b * kBatch + j + 1);
            }
            // n_touched: pixels that blended this Gaussian while still > 50 % transparent (forward.cu:368-370). The vote
            // is only parked here; counting and the shared-memory add are done for 32 instances at a time below.
            const uint32_t tv = __ballot_sync(0xffffffffu, act && test_T > 0.5f);
            if (lane == 0) s_tv[warp][k] = tv;
            if (__all_sync(0xffffffffu, done)) { ++k; break; }
        }
        __syncwarp();
        for (int i = (int)lane; i < k; i += 32) {
            const int c = __popc(s_tv[warp][i]);
            if (c) atomicAdd(&s_cnt[stage][s_list[warp][i]], c);
        }
        __syncthreads();   // every warp is done with this stage
        if (threadIdx.x < cnt) {
            const int c = s_cnt[stage][threadIdx.x];
            if (c) {
                atomicAdd(&n_touched[__float_as_int(rec[3 * threadIdx.x + 1].w)], c);
                s_cnt[stage][threadIdx.x] = 0;
            }
        }
        __syncthreads();   // ids read before the stage is overwritten
        if (threadIdx.x == 0 && b + 2 < nb) {
            const int cnt2 = min(kBatch, n - (b + 2) * kBatch);
            fence_proxy_async();
            bulk_load(&s_rec[stage][0], src + 3 * (size_t)(b + 2) * kBatch, (uint32_t)cnt2 * 48u, &s_bar[stage]);
        }
    }

    // an early stop leaves up to two bulk copies in flight: they must land before this CTA's shared memory
    // can be handed to another CTA
    if (threadIdx.x == 0) {
        for (int k = b; k < nb && k < b + 2; ++k) mbar_wait(&s_bar[k & 1], (uint32_t)((k >> 1) & 1));
    }

    if (inside) {
        final_T[pix_id] = T;
        n_contrib[pix_id] = last_contributor;
        const size_t HW = (size_t)H * W;
        out_color[pix_id] = ffma(bg[0], T, C0);
        out_color[HW + pix_id] = ffma(bg[1], T, C1);
        out_color[2 * HW + pix_id] = ffma(bg[2], T, C2);
        out_depth[pix_id] = Dp;
        out_opacity[pix_id] = fsub(1.0f, T);
    }
}

void launch_render_forward(int W, int H, int grid_x, int grid_y, const uint2* ranges, const float* records,
                           const float* bg, float* final_T, uint32_t* n_contrib, float* out_color, float* out_depth,
                           float* out_opacity, int* n_touched, cudaStream_t s) {
    render_forward_kernel<<<grid_x * grid_y, SPLAT_TILE_PIXELS, 0, s>>>(W, H, grid_x, ranges, records, bg, final_T,
                                                                        n_contrib, out_color, out_depth, out_opacity,
                                                                        n_touched);
}

// =====================================================================================================
// backward
// =====================================================================================================
// Recursive-halving all-lane reduction of 10 values: after the five exchange rounds the lane with
// bit0 == 0 and index k = 5*bit4 + 3*bit3 + 2*bit2 + bit1 (when valid) holds the warp total of value k.
__device__ __forceinline__ float warp_reduce10(const float (&v)[10], uint32_t lane) {
    const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4, h2 = lane & 2;
    float a[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const float send = h16 ? v[k] : v[k + 5];
        const float keep = h16 ? v[k + 5] : v[k];
        a[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
    float b[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float hi = (k + 3 < 5) ? a[k + 3] : 0.0f;
        const float send = h8 ? a[k] : hi;
        const float keep = h8 ? hi : a[k];
        b[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    float c[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const float hi = (k + 2 < 3) ? b[k + 2] : 0.0f;
        const float send = h4 ? b[k] : hi;
        const float keep = h4 ? hi : b[k];
        c[k] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    float d;
    {
        const float send = h2 ? c[0] : c[1];
        const float keep = h2 ? c[1] : c[0];
        d = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    }
    d += __shfl_xor_sync(0xffffffffu, d, 1);
    return d;
}

__global__ void __launch_bounds__(SPLAT_TILE_PIXELS)
render_backward_kernel(int W, int H, int grid_x, const uint2* __restrict__ ranges, const float* __restrict__ records,
                       const float* __restrict__ bg, const float* __restrict__ final_Ts,
                       const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels,
                       const float* __restrict__ dL_dpixels_depth, float* __restrict__ grad_accum) {
    __shared__ __align__(128) float4 s_rec[2][kBatch * 3];
    __shared__ __align__(8) uint64_t s_bar[2];
    __shared__ uint8_t s_list[SPLAT_TILE_PIXELS / 32][kBatch];

    const int tile = blockIdx.x;
    const int tile_x = tile % grid_x, tile_y = tile / grid_x;
    uint32_t px, py;
    pixel_of_thread(tile_x, tile_y, px, py);
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const uint32_t pix_id = (uint32_t)W * py + px;
    const float pixfx = (float)px, pixfy = (float)py;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float4 patch = patch_of_warp(tile_x, tile_y);

    const uint2 range = ranges[tile];
    const int n = (int)(range.y - range.x);
    const int nb = (n + kBatch - 1) / kBatch;
    const float4* src = reinterpret_cast<const float4*>(records) + 3 * (size_t)range.x;

    // batch b (counted from the BACK of the list) covers list indices [lo_b, hi_b)
    auto batch_lo = [&](int b) { return max(0, n - (b + 1) * kBatch); };
    auto batch_hi = [&](int b) { return n - b * kBatch; };

    if (threadIdx.x == 0) { mbar_init(&s_bar[0], 1); mbar_init(&s_bar[1], 1); fence_mbar_init(); }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int b = 0; b < 2 && b < nb; ++b) {
            const int lo = batch_lo(b), cnt = batch_hi(b) - lo;
            bulk_load(&s_rec[b][0], src + 3 * (size_t)lo, (uint32_t)cnt * 48u, &s_bar[b]);
        }
    }

    // which of the 10 reduced terms this lane ends up holding (see warp_reduce10), and where it goes in the
    // packed accumulator {sum u dx, sum u dy, sum u dx^2, sum u dx dy, sum u dy^2, sum u, r, g, b, depth, -, -}
    const int inner3 = ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
    const int inner5 = ((lane >> 3) & 1) * 3 + inner3;
    const int my_k = ((lane >> 4) & 1) * 5 + inner5;
    const bool holder = ((lane & 1) == 0) && inner3 < 3 && inner5 < 5;

    const float T_final = inside ? final_Ts[pix_id] : 0.0f;
    float T = T_final;
    const int last_contributor = inside ? (int)n_contrib[pix_id] : 0;
    const size_t HW = (size_t)H * W;
    float dLp0 = 0.0f, dLp1 = 0.0f, dLp2 = 0.0f, dLpd = 0.0f;
    if (inside) {
        dLp0 = dL_dpixels[pix_id]; dLp1 = dL_dpixels[HW + pix_id]; dLp2 = dL_dpixels[2 * HW + pix_id];
        dLpd = dL_dpixels_depth[pix_id];
    }
    // D = <what lies behind the Gaussian being processed, dL/dpixel>; behind the last one there is only the background
    float D = T_final * (bg[0] * dLp0 + bg[1] * dLp1 + bg[2] * dLp2);

    // nothing in this warp ever blended anything -> the warp only helps with barriers
    const int warp_last = __reduce_max_sync(0xffffffffu, last_contributor);

    for (int b = 0; b < nb; ++b) {
        const int stage = b & 1;
        const int lo = batch_lo(b), cnt = batch_hi(b) - lo;
        mbar_wait(&s_bar[stage], (uint32_t)((b >> 1) & 1));
        const float4* rec = &s_rec[stage][0];
        if (lo < warp_last) {
            // only list positions below the warp's deepest contributor matter; bin them against the warp's patch
            const int n_live = bin_batch_for_warp(rec, min(cnt, warp_last - lo), patch, s_list[warp], lane);
            for (int k = n_live - 1; k >= 0; --k) {
                const int j = s_list[warp][k];
                const int idx = lo + j;                        // position in the tile's list
                const float4 r0 = rec[3 * j];
                const float4 r1 = rec[3 * j + 1];
                const float dx = fsub(r0.x, pixfx), dy = fsub(r0.y, pixfy);
                const float power = gauss_power_ref(dx, dy, r0.z, r0.w, r1.x);
                // same expf / alpha as the forward so that the blend decisions replay exactly
                const float G0 = expf(power);
                const float alpha0 = fminf(0.99f, fmul(r1.y, G0));
                const bool skip = (idx >= last_contributor) || (power > 0.0f) || (alpha0 < 1.0f / 255.0f);
                if (__all_sync(0xffffffffu, skip)) continue;
                const float4 r2 = rec[3 * j + 2];
                // lanes that skip contribute through G = 0 and alpha = 0 (two selects instead of ten)
                const float G = skip ? 0.0f : G0;
                const float alpha = skip ? 0.0f : alpha0;
                float inv;
                asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(1.0f - alpha));
                // transmittance in front of this Gaussian (backward.cu:700)
                T = skip ? T : T * inv;
                // dL/dalpha (backward.cu:705-743) without the per-channel "colour behind" recursion: with
                // q = <colour/depth of this Gaussian, dL/dpixel> and D = <everything blended behind it (in absolute
                // weights alpha_j T_j) plus the background share T_final * bg, dL/dpixel>,
                //   sum_c (c - accum_rec_c) dL_c * T - T_final/(1-alpha) * <bg, dL>  =  T q - D / (1 - alpha),
                // and D grows by alpha T q. One scalar of state instead of nine, 7 operations instead of ~20.
                const float w = alpha * T;                                   // dchannel_dcolor
                const float q = r2.x * dLp0 + r2.y * dLp1 + r2.z * dLp2 + r1.z * dLpd;
                const float dL_dalpha = T * q - inv * D;
                D = w * q + D;
                const float dchannel_dcolor = w;

                // Everything that is constant per Gaussian (opacity, conic, the -0.5 and the NDC scale) is factored out
                // of the pixel sum: the warp reduces the moments of u = G * dL/dalpha over its pixels,
                //   sum u dx, sum u dy, sum u dx^2, sum u dx dy, sum u dy^2, sum u,
                // and the per-Gaussian backward turns them into dL/dmean2D, dL/dconic, dL/dopacity
                // (backward.cu:745-770; preprocess_bwd.cu: moments_to_gradients).
                const float u = G * dL_dalpha;
                const float ux = u * dx, uy = u * dy;
                float v[10];
                v[0] = ux;
                v[1] = uy;
                v[2] = ux * dx;
                v[3] = ux * dy;
                v[4] = uy * dy;
                v[5] = u;
                v[6] = dchannel_dcolor * dLp0;
                v[7] = dchannel_dcolor * dLp1;
                v[8] = dchannel_dcolor * dLp2;
                v[9] = dchannel_dcolor * dLpd;
                const float total = warp_reduce10(v, lane);
                if (holder && total != 0.0f) {
                    const uint32_t id = __float_as_uint(r1.w);
                    atomicAdd(&grad_accum[(size_t)id * SPLAT_GRAD_FLOATS + my_k], total);
                }
            }
        }
        __syncthreads();
        if (threadIdx.x == 0 && b + 2 < nb) {
            const int lo2 = batch_lo(b + 2), cnt2 = batch_hi(b + 2) - lo2;
            fence_proxy_async();
            bulk_load(&s_rec[stage][0], src + 3 * (size_t)lo2, (uint32_t)cnt2 * 48u, &s_bar[stage]);
        }
    }
}

void launch_render_backward(int W, int H, int grid_x, int grid_y, const uint2* ranges, const float* records,
                            const float* bg, const float* final_T, const uint32_t* n_contrib, const float* dL_dpix,
                            const float* dL_dpix_depth, float* grad_accum, cudaStream_t s) {
    render_backward_kernel<<<grid_x * grid_y, SPLAT_TILE_PIXELS, 0, s>>>(W, H, grid_x, ranges, records, bg, final_T,
                                                                         n_contrib, dL_dpix, dL_dpix_depth, grad_accum);
}

}  // namespace splat
