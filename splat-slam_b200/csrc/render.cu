// Tile renderers: forward front-to-back alpha compositing of colour/depth/opacity and the backward pass with
// back-to-front recomposition. Replaces DGR/cuda_rasterizer/forward.cu:263-392 and backward.cu:541-787.
//
// B200 design: one CTA per 16x16 tile (tile ids/ranges stay bit-compatible with the reference). Each warp owns
// an 8x4 pixel patch so whole-warp rejection of a Gaussian is likely. Per-instance 48-byte records (built in
// sorted order by the last pass of the instance sort) are streamed into a shared-memory ring with cp.async.bulk (TMA 1-D
// bulk copy, SASS UBLKCP) completing on an mbarrier; the consumer loops read them with broadcast LDS.128.
// The backward takes the per-Gaussian sums in a transposed second phase (lane = Gaussian) instead of a cross-lane
// reduction, and its warps walk the ring independently (no block barrier per batch); see render_backward_kernel.
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdlib>
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
// conservative footprint box stored in the record (preprocess_forward_kernel) and the survivors' batch indices are
// compacted, in order, into `list`. A (warp, instance) pair that cannot reach alpha >= 1/255 anywhere in the patch
// then costs ~0.5 instructions instead of a full power/exp evaluation by all 32 lanes. Comparisons are written so
// that NaN coordinates count as "inside" (the reference would still evaluate them). Returns the survivor count.
template <int kRecords = kBatch>
__device__ __forceinline__ int bin_batch_for_warp(const float4* rec, int cnt, float4 patch, uint8_t* list,
                                                  uint32_t lane) {
    int n = 0;
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < kRecords / 32; ++r) {
        const int j = r * 32 + (int)lane;
        bool live = false;
        if (j < cnt) {
            const float4 r0 = rec[3 * j];          // x, y, id, footprint box
            const uint32_t hw = __float_as_uint(r0.w);
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
// try_wait once; true when the phase with this parity has completed
__device__ __forceinline__ bool mbar_try(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}

constexpr int kFwdRecs = 64;     // records per ring stage
constexpr int kFwdStages = 8;    // stages of the record ring

// volatile read of a shared-memory word by its 32-bit shared address (a C++ volatile cast would go through a generic
// pointer and make ptxas address every shared array of the kernel through the generic window)
__device__ __forceinline__ int lds_volatile(uint32_t addr) {
    int v;
    asm volatile("ld.volatile.shared.s32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    return v;
}

// The warps of a tile see very different numbers of Gaussians (per-warp binning) and saturate at different depths, so
// they do not meet at block barriers: every warp walks the ring of record batches at its own pace, waiting only for
// "stage full" (mbarrier completed by the bulk copy). When a warp is done with a stage its lane 0 counts itself off
// and the LAST warp to do so re-arms the stage with the batch kFwdStages further on. A warp whose 32 pixels are all
// saturated (or outside the image) keeps counting itself off without looking at the records; when all eight are in
// that state nobody refills and everybody leaves (the reference's "whole tile done" exit, forward.cu:318-320).
//
// kLoss: the mapping loss that follows the rasterizer on every view of a mapping step
// (thirdparty/monogs/utils/slam_utils.py:71-105) is evaluated in the epilogue, from the pixel's colour and depth still
// in registers: dL/dimage and dL/ddepth are written next to the images, the loss sums are reduced per tile and
// finished by the last CTA (loss.cuh). One launch less per view and no re-read of the rendered images.
// lane 0: set bit `bit` of the id word of a record in GLOBAL memory when `on` (see the warp mask below)
__device__ __forceinline__ void mark_record(uint32_t* id_word, uint32_t bit, uint32_t lane, bool on) {
    asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.u32 q, %1, 0;\n\tsetp.eq.and.u32 p, %0, 0, q;\n\t@p red.global.or.b32 [%2], %3;\n\t}"
                 ::"r"(lane), "r"((uint32_t)on), "l"(id_word), "r"(bit) : "memory");
}

template <bool kLoss>
__global__ void __launch_bounds__(SPLAT_TILE_PIXELS)
render_forward_kernel(int W, int H, int grid_x, uint2* __restrict__ ranges, float* __restrict__ records, int warp_mask,
                      const float* __restrict__ bg, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                      float* __restrict__ out_color, float* __restrict__ out_depth, float* __restrict__ out_opacity,
                      int* __restrict__ n_touched, const uint32_t* __restrict__ sort_error, LossDesc loss) {
    // an instance sort whose look-back gave up must not go unnoticed: fail the stream loudly
    if (blockIdx.x == 0 && threadIdx.x == 0 && sort_error != nullptr && *sort_error != 0u) {
        printf("splat_raster: tile sort look-back timed out\n");
        __trap();
    }
    constexpr int kWarps = SPLAT_TILE_PIXELS / 32;
    __shared__ __align__(128) float4 s_rec[kFwdStages][kFwdRecs * 3];
    __shared__ __align__(8) uint64_t s_full[kFwdStages];
    __shared__ int s_done[kFwdStages];          // warps that are through with the stage's current batch
    __shared__ int s_saturated;                 // warps with nothing left to blend
    __shared__ int s_issued;                    // batches handed to the copy engine so far
    __shared__ uint8_t s_list[kWarps][kFwdRecs];
    __shared__ uint32_t s_tv[kWarps][kFwdRecs];
    __shared__ uint32_t s_av[kWarps][kFwdRecs];     // per listed instance: lanes that blended it (-> warp mask)

    const int tile = blockIdx.x;
    const int tile_x = tile % grid_x, tile_y = tile / grid_x;
    uint32_t px, py;
    pixel_of_thread(tile_x, tile_y, px, py);
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const uint32_t pix_id = (uint32_t)W * py + px;
    const float pixfx = (float)px, pixfy = (float)py;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float4 patch = patch_of_warp(tile_x, tile_y);
    uint8_t* my_list = s_list[warp];
    uint32_t* my_tv = s_tv[warp];
    uint32_t* my_av = s_av[warp];
    const uint32_t saturated_a = smem_u32(&s_saturated), issued_a = smem_u32(&s_issued);
    uint32_t tv_a = smem_u32(my_tv);
    uint32_t av_a = smem_u32(my_av);
    asm volatile("mov.u32 %0, %0;" : "+r"(av_a));
    asm volatile("mov.u32 %0, %0;" : "+r"(tv_a));     // opaque: keeps ptxas from re-deriving the address in the hot loop

    // The sort leaves (first position, ~(last position + 1)) per tile, both 0xFFFFFFFF for a tile without instances;
    // this CTA is the tile's first reader and rewrites the entry as [start, end) / (0, 0) for everybody after it
    // (the backward, inspection): the reference's identifyTileRanges result.
    uint2 range = ranges[tile];
    range.y = ~range.y;
    if (range.y <= range.x) range = make_uint2(0u, 0u);
    const int n = (int)(range.y - range.x);
    const int nb = (n + kFwdRecs - 1) / kFwdRecs;
    const float4* src = reinterpret_cast<const float4*>(records) + 3 * (size_t)range.x;
    auto fill = [&](int b) {   // one thread: bulk copy of batch b into its ring stage
        const int cnt = min(kFwdRecs, n - b * kFwdRecs);
        bulk_load(&s_rec[b % kFwdStages][0], src + 3 * (size_t)b * kFwdRecs, (uint32_t)cnt * 48u, &s_full[b % kFwdStages]);
    };

    if (threadIdx.x == 0) {
        for (int st = 0; st < kFwdStages; ++st) { mbar_init(&s_full[st], 1); s_done[st] = 0; }
        s_saturated = 0;
        s_issued = min(kFwdStages, nb);
        fence_mbar_init();
    }
    __syncthreads();                       // barriers armed; every thread has read the raw range entry
    if (threadIdx.x == 0) {
        ranges[tile] = range;
        for (int b = 0; b < kFwdStages && b < nb; ++b) fill(b);
    }

    bool done = !inside;
    float T = 1.0f, C0 = 0.0f, C1 = 0.0f, C2 = 0.0f, Dp = 0.0f;
    uint32_t last_contributor = 0;
    bool warp_done = __all_sync(0xffffffffu, done);
    if (warp_done && lane == 0) atomicAdd(&s_saturated, 1);

    int b = 0;
    for (; b < nb; ++b) {
        const int stage = b % kFwdStages;
        // wait for the batch - unless the whole tile has saturated in the meantime and nobody refills any more
        bool stop = false;
        for (;;) {
            const bool ok = mbar_try(&s_full[stage], (uint32_t)((b / kFwdStages) & 1));
            if (__all_sync(0xffffffffu, ok)) break;
            if (__any_sync(0xffffffffu, lds_volatile(saturated_a) == kWarps)) { stop = true; break; }
        }
        if (stop) break;
        if (!warp_done) {
            const int cnt = min(kFwdRecs, n - b * kFwdRecs);
            const float4* rec = &s_rec[stage][0];
            const int n_live = bin_batch_for_warp<kFwdRecs>(rec, cnt, patch, my_list, lane);
            // Warp mask for the backward: bit 24 + warp of a record's id word (ids stay below 2^24 when the mask is
            // in use) says "some pixel of this warp blended this instance". The backward then lists exactly the
            // instances that matter to the warp instead of re-testing footprints (a quarter of which pass the box
            // test without blending anything). Set in the GLOBAL copy of the record; the copy in shared memory that
            // this kernel reads ids from was loaded before and stays clean.
            uint32_t* id_words = reinterpret_cast<uint32_t*>(records) + 12 * ((size_t)range.x + (size_t)b * kFwdRecs) + 2;
            const uint32_t warp_bit = warp_mask ? (1u << (24 + warp)) : 0u;
            // n_touched counts pixels that blend a Gaussian while still > 50 % transparent (forward.cu:368-370). T only
            // falls, so once none of the warp's pixels is above 0.5 the vote, its store and the counting below are dead
            // weight for the rest of the tile (most of a deep tile): that case runs a loop without them.
            const bool counting = __any_sync(0xffffffffu, !done && T > 0.5f);
            int k = 0;
            if (counting) {
                for (; k < n_live; ++k) {
                    // survivors of the binning almost always touch some lane, so everything up to the blend is evaluated
                    // unconditionally and a single ballot per instance serves the n_touched count
                    const int j = my_list[k];
                    const float4 r0 = rec[3 * j];          // x, y, id, footprint box
                    const float4 r1 = rec[3 * j + 1];      // conic.x, conic.y, conic.z, opacity
                    const float dx = fsub(r0.x, pixfx), dy = fsub(r0.y, pixfy);
                    const float power = gauss_power_ref(dx, dy, r1.x, r1.y, r1.z);
                    // alpha = min(0.99, opacity * exp(power)); 1/255 cut; transmittance test (forward.cu:350-363)
                    const float alpha = fminf(0.99f, fmul(r1.w, expf(power)));
                    const float test_T = fmul(T, fsub(1.0f, alpha));
                    bool act = !done && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                    if (act && test_T < 0.0001f) { done = true; act = false; }
                    if (act) {
                        const float4 r2 = rec[3 * j + 2];  // r, g, b, depth
                        // C += c * alpha * T as fma(T, alpha*c, C) (reference SASS order)
                        C0 = ffma(T, fmul(alpha, r2.x), C0);
                        C1 = ffma(T, fmul(alpha, r2.y), C1);
                        C2 = ffma(T, fmul(alpha, r2.z), C2);
                        Dp = ffma(T, fmul(alpha, r2.w), Dp);
                        T = test_T;
                        last_contributor = (uint32_t)(b * kFwdRecs + j + 1);
                    }
                    // the vote is only parked here; it is counted for 32 instances at a time below
                    const uint32_t tv = __ballot_sync(0xffffffffu, act && test_T > 0.5f);
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.eq.u32 p, %0, 0;\n\t@p st.shared.u32 [%1], %2;\n\t}"
                                 ::"r"(lane), "r"(tv_a + 4u * (uint32_t)k), "r"(tv) : "memory");
                    const uint32_t av = __ballot_sync(0xffffffffu, act);
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.eq.u32 p, %0, 0;\n\t@p st.shared.u32 [%1], %2;\n\t}"
                                 ::"r"(lane), "r"(av_a + 4u * (uint32_t)k), "r"(av) : "memory");
                    if (__all_sync(0xffffffffu, done)) { ++k; break; }
                }
                __syncwarp();
                // one atomic per (warp, Gaussian) with votes instead of one per pixel
                for (int i = (int)lane; i < k; i += 32) {
                    const int c = __popc(my_tv[i]);
                    if (c) asm volatile("red.global.add.s32 [%0], %1;" ::"l"(n_touched + __float_as_int(rec[3 * my_list[i]].z)), "r"(c) : "memory");
                    if (warp_bit && my_av[i]) asm volatile("red.global.or.b32 [%0], %1;" ::"l"(id_words + 12 * my_list[i]), "r"(warp_bit) : "memory");
                }
            } else {
                for (; k < n_live; ++k) {
                    const int j = my_list[k];
                    const float4 r0 = rec[3 * j];
                    const float4 r1 = rec[3 * j + 1];
                    const float dx = fsub(r0.x, pixfx), dy = fsub(r0.y, pixfy);
                    const float power = gauss_power_ref(dx, dy, r1.x, r1.y, r1.z);
                    const float alpha = fminf(0.99f, fmul(r1.w, expf(power)));
                    const float test_T = fmul(T, fsub(1.0f, alpha));
                    bool act = !done && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                    if (act && test_T < 0.0001f) { done = true; act = false; }
                    if (__any_sync(0xffffffffu, act)) {
                        const float4 r2 = rec[3 * j + 2];
                        asm volatile("{\n\t.reg .pred p;\n\tsetp.eq.u32 p, %0, 0;\n\t@p st.shared.u32 [%1], %2;\n\t}"
                                     ::"r"(lane), "r"(av_a + 4u * (uint32_t)k), "r"(1u) : "memory");
                        // same arithmetic, predicated per lane instead of branched
                        const float c0 = ffma(T, fmul(alpha, r2.x), C0), c1 = ffma(T, fmul(alpha, r2.y), C1);
                        const float c2 = ffma(T, fmul(alpha, r2.z), C2), dp = ffma(T, fmul(alpha, r2.w), Dp);
                        C0 = act ? c0 : C0;
                        C1 = act ? c1 : C1;
                        C2 = act ? c2 : C2;
                        Dp = act ? dp : Dp;
                        T = act ? test_T : T;
                        last_contributor = act ? (uint32_t)(b * kFwdRecs + j + 1) : last_contributor;
                    } else {
                        asm volatile("{\n\t.reg .pred p;\n\tsetp.eq.u32 p, %0, 0;\n\t@p st.shared.u32 [%1], %2;\n\t}"
                                     ::"r"(lane), "r"(av_a + 4u * (uint32_t)k), "r"(0u) : "memory");
                        if (__all_sync(0xffffffffu, done)) { ++k; break; }
                    }
                }
                __syncwarp();
                if (warp_bit)
                    for (int i = (int)lane; i < k; i += 32)
                        if (my_av[i]) asm volatile("red.global.or.b32 [%0], %1;" ::"l"(id_words + 12 * my_list[i]), "r"(warp_bit) : "memory");
            }
            if (__all_sync(0xffffffffu, done)) {
                warp_done = true;
                if (lane == 0) atomicAdd(&s_saturated, 1);
            }
        }
        __syncwarp();                                  // every lane has read what it needs from this stage
        if (lane == 0) {
            __threadfence_block();
            if (atomicAdd(&s_done[stage], 1) == kWarps - 1) {
                s_done[stage] = 0;                     // published to the other warps by the refill's mbarrier
                if (b + kFwdStages < nb && lds_volatile(saturated_a) < kWarps) {
                    __threadfence_block();
                    fence_proxy_async();
                    fill(b + kFwdStages);
                    atomicMax(&s_issued, b + kFwdStages + 1);
                }
            }
        }
    }

    const size_t HW = (size_t)H * W;
    const float o0 = ffma(bg[0], T, C0), o1 = ffma(bg[1], T, C1), o2 = ffma(bg[2], T, C2);
    if (inside) {
        final_T[pix_id] = T;
        n_contrib[pix_id] = last_contributor;
        out_color[pix_id] = o0;
        out_color[HW + pix_id] = o1;
        out_color[2 * HW + pix_id] = o2;
        out_depth[pix_id] = Dp;
        out_opacity[pix_id] = fsub(1.0f, T);
    }

    // an early stop leaves bulk copies in flight that nobody waited for: they must land before this CTA's shared
    // memory can be handed to another CTA
    __syncthreads();
    if (threadIdx.x == 0) {
        const int issued = lds_volatile(issued_a);
        for (int k = b; k < issued; ++k) mbar_wait(&s_full[k % kFwdStages], (uint32_t)((k / kFwdStages) & 1));
    }

    if constexpr (kLoss) {
        const float ea = loss.initialization ? 1.0f : expf(loss.exposure_a[0]);
        const float eb = loss.initialization ? 0.0f : loss.exposure_b[0];
        const float w_rgb = loss.alpha / (3.0f * (float)HW), w_d = (1.0f - loss.alpha) / (float)HW;
        float v[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        if (inside) loss_pixel(loss, HW, (size_t)pix_id, o0, o1, o2, Dp, ea, eb, w_rgb, w_d, v);
        loss_finish<SPLAT_TILE_PIXELS>(v, loss.workspace, loss.out, w_rgb, w_d);
    }
}

void launch_render_forward(int W, int H, int grid_x, int grid_y, uint2* ranges, float* records, int warp_mask,
                           const float* bg, float* final_T, uint32_t* n_contrib, float* out_color, float* out_depth,
                           float* out_opacity, int* n_touched, const uint32_t* sort_error, const LossDesc* loss,
                           cudaStream_t s) {
    if (loss != nullptr)
        render_forward_kernel<true><<<grid_x * grid_y, SPLAT_TILE_PIXELS, 0, s>>>(
            W, H, grid_x, ranges, records, warp_mask, bg, final_T, n_contrib, out_color, out_depth, out_opacity,
            n_touched, sort_error, *loss);
    else
        render_forward_kernel<false><<<grid_x * grid_y, SPLAT_TILE_PIXELS, 0, s>>>(
            W, H, grid_x, ranges, records, warp_mask, bg, final_T, n_contrib, out_color, out_depth, out_opacity,
            n_touched, sort_error, LossDesc());
}

// =====================================================================================================
// backward
// =====================================================================================================
// Back-to-front over the same records and per-warp lists as the forward. The pixel-major pass replays the blend
// decisions exactly (same power / expf / alpha sequence) and produces, per (pixel, Gaussian), only the two scalars that
// all ten per-Gaussian gradient terms are linear in,
//     u = G * dL/dalpha        (the five geometric moments and dL/dopacity)
//     w = alpha * T            (colour and depth terms),
// with dL/dalpha = T q - D / (1 - alpha), q = <colour/depth of this Gaussian, dL/dpixel> and the single running scalar
// D = <everything blended behind it (absolute weights alpha_j T_j) plus the background share T_final * bg, dL/dpixel>,
// D += alpha T q (algebraically backward.cu:705-743 without the per-channel "colour behind" recursion).
// u and w are parked in shared memory, one row of 32 pixels per Gaussian. Once a warp has collected C Gaussians it
// switches roles: lane = (Gaussian g, 32/C-th part of the patch) walks over its pixels and accumulates the ten sums
// in registers. The pixel offsets inside an 8 x 4 patch are compile-time constants, so the moments are taken about
// the patch origin with immediates,
//     S0 = sum u,  Sx = sum u x,  Sxx = sum u x^2      per patch row (y is constant in a row),
// and shifted to the Gaussian's centre once per row (dx = gx - x, dy = gy - row). This replaces a cross-lane
// reduction of ten values per (warp, Gaussian) - 12 SHFL, 22 selects, 12 adds and the ten products feeding them in
// the round-1 kernel - by ~13 instructions, and the warp's ten totals of a Gaussian leave as three vector REDs
// (16 + 16 + 8 bytes, SASS REDG.F32x4 / F32x2) from one lane instead of ten scalar lanes.
// Everything that is constant per Gaussian (opacity, conic, the -0.5 and the NDC scale, backward.cu:745-770) is applied
// afterwards by the per-Gaussian backward (preprocess_bwd.cu) to the packed sums
//     {sum u dx, sum u dy, sum u dx^2, sum u dx dy | sum u dy^2, sum u, r, g | b, depth, -, -}.
template <int C>
struct BwdLayout {
    static constexpr int kRowF2 = 33;      // float2 per parked Gaussian: 32 pixels x (u, w) + 1 pad. 66 words = 2 mod 32:
                                           // the 16 rows a half-warp reads with 64-bit loads in the transposed phase
                                           // start in 16 different even banks; the pixel-major stores are contiguous
    static constexpr int kWarps = SPLAT_TILE_PIXELS / 32;
    static constexpr int kStages = 4;      // record ring: kStages batches of kRecs records in flight per CTA
    static constexpr int kRecs = 64;
    static constexpr size_t kRecOff = 0;                                         // float4 [kStages][kRecs * 3]
    static constexpr size_t kUwOff = kRecOff + kStages * kRecs * 48;             // float2 [kWarps][C][kRowF2]
    static constexpr size_t kInfoOff = kUwOff + (size_t)kWarps * C * kRowF2 * 8; // float4 [kWarps][C]  x, y, id, -
    static constexpr size_t kDlpOff = kInfoOff + (size_t)kWarps * C * 16;        // float4 [256] dL/dpixel
    static constexpr size_t kListOff = kDlpOff + SPLAT_TILE_PIXELS * 16;         // uint8  [kWarps][kRecs]
    static constexpr size_t kBarOff = kListOff + (size_t)kWarps * kRecs;         // uint64 [kStages] "stage is full"
    static constexpr size_t kDoneOff = kBarOff + kStages * 8;                    // int    [kStages] warps done with it
    static constexpr size_t kBytes = kDoneOff + kStages * 4;
};

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void red_add_v2(float* addr, float a, float b) {
    asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(addr), "f"(a), "f"(b) : "memory");
}

// Transposed phase: sums of one chunk of `count` (<= C) parked Gaussians over the warp's 32 pixels.
template <int C>
__device__ __forceinline__ void flush_chunk(const float2* __restrict__ uw, const float4* __restrict__ info,
                                            const float4* __restrict__ dlp, int count, float patch_x0, float patch_y0,
                                            uint32_t lane, float* __restrict__ grad_accum) {
    constexpr int kParts = 32 / C, kRows = 4 / kParts;
    const int g = (int)lane & (C - 1), part = (int)lane / C;
    const float4 inf = info[g];                                            // x, y, id
    const float gxr = inf.x - patch_x0, gyr = inf.y - patch_y0;
    float m0 = 0.f, mx = 0.f, my = 0.f, mxx = 0.f, mxy = 0.f, myy = 0.f, c0 = 0.f, c1 = 0.f, c2 = 0.f, cd = 0.f;
#pragma unroll
    for (int rr = 0; rr < kRows; ++rr) {
        const int row = part * kRows + rr;
        const float2* urow = uw + g * BwdLayout<C>::kRowF2 + row * 8;
        const float4* drow = dlp + row;        // dL/dpixel is stored column-major (x * 4 + row), see the kernel
        float S0 = 0.f, Sx = 0.f, Sxx = 0.f;
#pragma unroll
        for (int x = 0; x < 8; ++x) {
            const float2 v = urow[x];
            const float4 d = drow[4 * x];
            S0 += v.x;
            Sx = fmaf(v.x, (float)x, Sx);
            Sxx = fmaf(v.x, (float)(x * x), Sxx);
            c0 = fmaf(v.y, d.x, c0);
            c1 = fmaf(v.y, d.y, c1);
            c2 = fmaf(v.y, d.z, c2);
            cd = fmaf(v.y, d.w, cd);
        }
        // shift the row's moments from the patch origin to the Gaussian's centre: dx = gxr - x, dy = gyr - row
        const float dy = gyr - (float)row;
        const float rx = fmaf(gxr, S0, -Sx);                               // sum u dx
        const float rxx = fmaf(gxr, rx - Sx, Sxx);                         // sum u dx^2
        m0 += S0;
        mx += rx;
        my = fmaf(dy, S0, my);
        mxx += rxx;
        mxy = fmaf(dy, rx, mxy);
        myy = fmaf(dy * dy, S0, myy);
    }
#pragma unroll
    for (int o = 16; o >= C; o >>= 1) {
        m0 += __shfl_xor_sync(0xffffffffu, m0, o);
        mx += __shfl_xor_sync(0xffffffffu, mx, o);
        my += __shfl_xor_sync(0xffffffffu, my, o);
        mxx += __shfl_xor_sync(0xffffffffu, mxx, o);
        mxy += __shfl_xor_sync(0xffffffffu, mxy, o);
        myy += __shfl_xor_sync(0xffffffffu, myy, o);
        c0 += __shfl_xor_sync(0xffffffffu, c0, o);
        c1 += __shfl_xor_sync(0xffffffffu, c1, o);
        c2 += __shfl_xor_sync(0xffffffffu, c2, o);
        cd += __shfl_xor_sync(0xffffffffu, cd, o);
    }
    if (part == 0 && g < count) {
        float* dst = grad_accum + (size_t)__float_as_uint(inf.z) * SPLAT_GRAD_FLOATS;
        red_add_v4(dst, mx, my, mxx, mxy);
        red_add_v4(dst + 4, myy, m0, c0, c1);
        red_add_v2(dst + 8, c2, cd);
    }
}

// Listing of one staged batch for a warp from the forward's warp mask (bit 24 + warp of the id word): exactly the
// instances some pixel of the warp blended, in order. Same interface as bin_batch_for_warp.
template <int kRecords>
__device__ __forceinline__ int list_batch_by_mask(const float4* rec, int cnt, uint32_t warp, uint8_t* list, uint32_t lane) {
    int n = 0;
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < kRecords / 32; ++r) {
        const int j = r * 32 + (int)lane;
        bool live = false;
        if (j < cnt) live = (__float_as_uint(reinterpret_cast<const float*>(rec + 3 * j)[2]) >> (24 + warp)) & 1u;
        const uint32_t m = __ballot_sync(0xffffffffu, live);
        if (live) list[n + __popc(m & lt)] = (uint8_t)j;
        n += __popc(m);
    }
    __syncwarp();
    return n;
}

template <int C, bool kMask>
__global__ void __launch_bounds__(SPLAT_TILE_PIXELS, C == 8 ? 5 : 4)
render_backward_kernel(int W, int H, int grid_x, const uint2* __restrict__ ranges,
                       const float* __restrict__ records, const float* __restrict__ bg,
                       const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib,
                       const float* __restrict__ dL_dpixels, const float* __restrict__ dL_dpixels_depth,
                       float* __restrict__ grad_accum) {
    using L = BwdLayout<C>;
    constexpr int kRecs = L::kRecs, kStages = L::kStages;
    extern __shared__ __align__(128) unsigned char smem[];
    float4 (*s_rec)[kRecs * 3] = reinterpret_cast<float4 (*)[kRecs * 3]>(smem + L::kRecOff);
    uint64_t* s_full = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
    int* s_done = reinterpret_cast<int*>(smem + L::kDoneOff);

    const int tile = blockIdx.x;
    const int tile_x = tile % grid_x, tile_y = tile / grid_x;
    uint32_t px, py;
    pixel_of_thread(tile_x, tile_y, px, py);
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const uint32_t pix_id = (uint32_t)W * py + px;
    const float pixfx = (float)px, pixfy = (float)py;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float4 patch = patch_of_warp(tile_x, tile_y);
    uint8_t* s_list = smem + L::kListOff + warp * kRecs;
    const float2* s_uw = reinterpret_cast<const float2*>(smem + L::kUwOff) + (size_t)warp * C * L::kRowF2;
    const float4* s_info = reinterpret_cast<const float4*>(smem + L::kInfoOff) + warp * C;
    float4* s_dlp = reinterpret_cast<float4*>(smem + L::kDlpOff) + warp * 32;
    // 32-bit shared addresses for the stores of the hot loop: this lane's column of the next free row, and its info slot
    const uint32_t uw_first = smem_u32(s_uw) + lane * 8u;
    const uint32_t info_first = smem_u32(s_info), info_end = info_first + C * 16u;
    uint32_t uw_a = uw_first, info_a = info_first;

    const uint2 range = ranges[tile];
    const int n = (int)(range.y - range.x);
    const int nb = (n + kRecs - 1) / kRecs;
    const float4* src = reinterpret_cast<const float4*>(records) + 3 * (size_t)range.x;

    // batch b (counted from the BACK of the list) covers list indices [lo_b, hi_b)
    auto batch_lo = [&](int b) { return max(0, n - (b + 1) * kRecs); };
    auto batch_hi = [&](int b) { return n - b * kRecs; };
    auto fill = [&](int b) {   // one thread: bulk copy of batch b into its ring stage
        const int lo = batch_lo(b), cnt = batch_hi(b) - lo;
        bulk_load(&s_rec[b % kStages][0], src + 3 * (size_t)lo, (uint32_t)cnt * 48u, &s_full[b % kStages]);
    };

    if (threadIdx.x == 0) {
        for (int st = 0; st < kStages; ++st) { mbar_init(&s_full[st], 1); s_done[st] = 0; }
        fence_mbar_init();
    }
    __syncthreads();
    if (threadIdx.x == 0)
        for (int b = 0; b < kStages && b < nb; ++b) fill(b);

    const float T_final = inside ? final_Ts[pix_id] : 0.0f;
    float T = T_final;
    const int last_contributor = inside ? (int)n_contrib[pix_id] : 0;
    const size_t HW = (size_t)H * W;
    float dLp0 = 0.0f, dLp1 = 0.0f, dLp2 = 0.0f, dLpd = 0.0f;
    if (inside) {
        dLp0 = dL_dpixels[pix_id]; dLp1 = dL_dpixels[HW + pix_id]; dLp2 = dL_dpixels[2 * HW + pix_id];
        dLpd = dL_dpixels_depth[pix_id];
    }
    // read back by the other lanes of this warp only, in the transposed phase. Column-major (x * 4 + row): the lanes
    // that work on different patch rows in the same iteration then read neighbouring 16-byte words instead of words
    // 128 bytes apart in the same banks (4 -> 2 wavefronts per load)
    s_dlp[(lane & 7) * 4 + (lane >> 3)] = make_float4(dLp0, dLp1, dLp2, dLpd);
    // behind the last Gaussian there is only the background
    float D = T_final * (bg[0] * dLp0 + bg[1] * dLp1 + bg[2] * dLp2);
    // nothing in this warp ever blended anything -> the warp only keeps the ring going
    const int warp_last = __reduce_max_sync(0xffffffffu, last_contributor);

    // The warps of a tile see very different numbers of Gaussians (per-warp binning), so they do not meet at a block
    // barrier per batch: every warp walks the ring at its own pace, waiting only for "stage full" (mbarrier completed
    // by the bulk copy). When a warp is done with a stage its lane 0 counts itself off; the LAST warp to do so
    // re-arms the stage with the batch kStages further on. A fast warp can thus run up to kStages - 1 batches ahead.
    for (int b = 0; b < nb; ++b) {
        const int stage = b % kStages;
        const int lo = batch_lo(b), cnt = batch_hi(b) - lo;
        mbar_wait(&s_full[stage], (uint32_t)((b / kStages) & 1));
        const float4* rec = &s_rec[stage][0];
        if (lo < warp_last) {
            // only list positions below the warp's deepest contributor matter; bin them against the warp's patch
            const int n_live = kMask ? list_batch_by_mask<kRecs>(rec, min(cnt, warp_last - lo), warp, s_list, lane)
                                     : bin_batch_for_warp<kRecs>(rec, min(cnt, warp_last - lo), patch, s_list, lane);
            const int lc_rel = last_contributor - lo;      // the pixel's last contributor, relative to this batch
            for (int k = n_live - 1; k >= 0; --k) {
                const int j = s_list[k];
                const float4 r0 = rec[3 * j];              // x, y, id, footprint box
                const float4 r1 = rec[3 * j + 1];          // conic.x, conic.y, conic.z, opacity
                const float dx = fsub(r0.x, pixfx), dy = fsub(r0.y, pixfy);
                const float power = gauss_power_ref(dx, dy, r1.x, r1.y, r1.z);
                // same expf / alpha as the forward so that the blend decisions replay exactly
                const float G0 = expf(power);
                const float alpha0 = fminf(0.99f, fmul(r1.w, G0));
                const bool skip = (j >= lc_rel) || (power > 0.0f) || (alpha0 < 1.0f / 255.0f);
                if (!kMask) {        // with the mask every listed instance was blended by some lane
                    if (__all_sync(0xffffffffu, skip)) continue;
                }
                const float4 r2 = rec[3 * j + 2];          // r, g, b, depth
                // lanes that skip contribute through G = 0 and alpha = 0
                const float G = skip ? 0.0f : G0;
                const float alpha = skip ? 0.0f : alpha0;
                float inv;
                asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(1.0f - alpha));
                T = skip ? T : T * inv;                    // transmittance in front of this Gaussian (backward.cu:700)
                const float w = alpha * T;
                const float q = r2.x * dLp0 + r2.y * dLp1 + r2.z * dLp2 + r2.w * dLpd;
                const float dL_dalpha = T * q - inv * D;
                D = w * q + D;
                // park (u, w) of this pixel in the Gaussian's row; lane 0 adds the Gaussian's x, y, id
                asm volatile(
                    "{\n\t.reg .pred p;\n\t"
                    "st.shared.v2.f32 [%0], {%1, %2};\n\t"
                    "setp.eq.u32 p, %3, 0;\n\t"
                    "@p st.shared.v4.f32 [%4], {%5, %6, %7, %8};\n\t}"
                    ::"r"(uw_a), "f"(G * dL_dalpha), "f"(w), "r"(lane), "r"(info_a), "f"(r0.x), "f"(r0.y),
                      "f"(kMask ? __uint_as_float(__float_as_uint(r0.z) & 0x00FFFFFFu) : r0.z), "f"(r0.w)
                    : "memory");
                uw_a += L::kRowF2 * 8u;
                info_a += 16u;
                if (info_a == info_end) {
                    __syncwarp();
                    flush_chunk<C>(s_uw, s_info, s_dlp, C, patch.x, patch.y, lane, grad_accum);
                    __syncwarp();
                    uw_a = uw_first;
                    info_a = info_first;
                }
            }
        }
        __syncwarp();                                      // every lane has read what it needs from this stage
        if (lane == 0) {
            __threadfence_block();
            if (atomicAdd(&s_done[stage], 1) == L::kWarps - 1) {
                s_done[stage] = 0;                         // published to the other warps by the refill's mbarrier
                if (b + kStages < nb) {
                    __threadfence_block();
                    fence_proxy_async();
                    fill(b + kStages);
                }
            }
        }
    }
    if (info_a != info_first) {
        __syncwarp();
        flush_chunk<C>(s_uw, s_info, s_dlp, (int)((info_a - info_first) >> 4), patch.x, patch.y, lane, grad_accum);
    }
}

template <int C, bool kMask>
static void launch_bwd_chunked(int W, int H, int grid_x, int grid_y, const uint2* ranges, const float* records,
                               const float* bg, const float* final_T, const uint32_t* n_contrib, const float* dL_dpix,
                               const float* dL_dpix_depth, float* grad_accum, cudaStream_t s) {
    // opt in to more than 48 KB of shared memory once per device
    static bool configured[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        cudaFuncSetAttribute(render_backward_kernel<C, kMask>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)BwdLayout<C>::kBytes);
        configured[dev] = true;
    }
    render_backward_kernel<C, kMask><<<grid_x * grid_y, SPLAT_TILE_PIXELS, BwdLayout<C>::kBytes, s>>>(
        W, H, grid_x, ranges, records, bg, final_T, n_contrib, dL_dpix, dL_dpix_depth, grad_accum);
}

// Chunk size of the transposed phase: 8 (default) or 16 Gaussians per warp. Measured on the bench scene (B200):
// 16 is the faster kernel when it runs alone (0.441 vs 0.457 ms: fewer shuffles per Gaussian), 8 needs 35 KB of shared
// memory and 48 registers instead of 53 KB and 64, leaves room for the other views' kernels and wins when several
// views are in flight (906 vs 878 it/s with five view streams). SPLAT_BWD_CHUNK=16 selects the other one.
static int bwd_variant() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("SPLAT_BWD_CHUNK");
        v = (e && atoi(e) == 16) ? 16 : 8;
    }
    return v;
}

void launch_render_backward(int W, int H, int grid_x, int grid_y, const uint2* ranges, const float* records,
                            int warp_mask, const float* bg, const float* final_T, const uint32_t* n_contrib,
                            const float* dL_dpix, const float* dL_dpix_depth, float* grad_accum, cudaStream_t s) {
#define SPLAT_BWD(C, M) launch_bwd_chunked<C, M>(W, H, grid_x, grid_y, ranges, records, bg, final_T, n_contrib, dL_dpix, \
                                                 dL_dpix_depth, grad_accum, s)
    if (bwd_variant() == 8) { if (warp_mask) SPLAT_BWD(8, true); else SPLAT_BWD(8, false); }
    else { if (warp_mask) SPLAT_BWD(16, true); else SPLAT_BWD(16, false); }
#undef SPLAT_BWD
}

}  // namespace splat
