// Fused mapping loss, forward + backward in one pass over the image (the step that follows
// the rasterizer on every view). Replaces the ~10 elementwise/reduction kernels of
// thirdparty/monogs/utils/slam_utils.py:71-105 (get_loss_mapping / get_loss_mapping_rgbd, ssim_loss = False):
//   image_ab   = exp(a) * image + b                      (skipped when `initialization`)
//   rgb_mask   = gt_image.sum(0) > rgb_boundary_threshold
//   l1_rgb     = |image_ab * rgb_mask - gt_image * rgb_mask|
//   depth_mask = gt_depth > 0.01
//   l1_depth   = |depth * depth_mask - gt_depth * depth_mask|
//   loss       = alpha * mean(l1_rgb) + (1 - alpha) * mean(l1_depth)
// One thread per pixel reads the 8 input floats once and writes dL/dimage (3) and dL/ddepth (1); the four sums
// (l1_rgb, l1_depth, dL/da, dL/db) are block-reduced into per-block partials; the block that takes the last ticket adds
// them up in a fixed order (deterministic, no float atomics) and writes loss + sums, so the whole loss is ONE launch.
// HBM-bound: 48 bytes per pixel.
#include "common.cuh"
#include "kernels.h"

namespace splat {

constexpr int kLossBlock = 256;

__device__ __forceinline__ float sign0(float x) { return (x > 0.0f) ? 1.0f : ((x < 0.0f) ? -1.0f : 0.0f); }

__global__ void __launch_bounds__(kLossBlock)
mapping_loss_kernel(int N, const float* __restrict__ image, const float* __restrict__ depth,
                    const float* __restrict__ gt_image, const float* __restrict__ gt_depth,
                    const float* __restrict__ exposure_a, const float* __restrict__ exposure_b, float alpha,
                    float rgb_boundary_threshold, int initialization, float* __restrict__ workspace,
                    float* __restrict__ out, float* __restrict__ dL_dimage, float* __restrict__ dL_ddepth) {
    __shared__ float s_red[kLossBlock / 32][4];
    __shared__ bool s_last;
    unsigned int* ticket = reinterpret_cast<unsigned int*>(workspace);
    float* partials = workspace + 4;
    const float ea = initialization ? 1.0f : expf(exposure_a[0]);
    const float eb = initialization ? 0.0f : exposure_b[0];
    const float w_rgb = alpha / (3.0f * (float)N), w_d = (1.0f - alpha) / (float)N;
    float l1 = 0.0f, l1d = 0.0f, da = 0.0f, db = 0.0f;
    for (int i = blockIdx.x * kLossBlock + threadIdx.x; i < N; i += gridDim.x * kLossBlock) {
        const float g0 = gt_image[i], g1 = gt_image[N + i], g2 = gt_image[2 * (size_t)N + i];
        const float m = (__fadd_rn(__fadd_rn(g0, g1), g2) > rgb_boundary_threshold) ? 1.0f : 0.0f;
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            const float x = image[ch * (size_t)N + i];
            const float gt = ch == 0 ? g0 : (ch == 1 ? g1 : g2);
            // every op rounded on its own, as the reference's chain of separate torch kernels does (no FMA contraction:
            // the sign of a near-zero difference decides the gradient)
            const float diff = __fsub_rn(__fmul_rn(__fadd_rn(__fmul_rn(ea, x), eb), m), __fmul_rn(gt, m));
            l1 += fabsf(diff);
            const float g = w_rgb * sign0(diff) * m;     // dL / d image_ab
            dL_dimage[ch * (size_t)N + i] = g * ea;
            da += g * ea * x;
            db += g;
        }
        const float gd = gt_depth[i], d = depth[i];
        const float md = (gd > 0.01f) ? 1.0f : 0.0f;
        const float dd = __fsub_rn(__fmul_rn(d, md), __fmul_rn(gd, md));
        l1d += fabsf(dd);
        dL_ddepth[i] = w_d * sign0(dd) * md;
    }
    float v[4] = {l1, l1d, da, db};
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
        if (lane == 0) s_red[warp][k] = v[k];
    }
    __syncthreads();
    if (threadIdx.x < 4) {
        float s = 0.0f;
#pragma unroll
        for (int w = 0; w < kLossBlock / 32; ++w) s += s_red[w][threadIdx.x];
        partials[4 * blockIdx.x + threadIdx.x] = s;
    }
    // last-block-done: publish the partials, take a ticket; whoever draws the last one reduces them
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    float t[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    for (int blk = threadIdx.x; blk < (int)gridDim.x; blk += kLossBlock) {
        const float4 p = __ldcg(reinterpret_cast<const float4*>(partials) + blk);
        t[0] += p.x; t[1] += p.y; t[2] += p.z; t[3] += p.w;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t[k] += __shfl_xor_sync(0xffffffffu, t[k], o);
        if (lane == 0) s_red[warp][k] = t[k];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        float sum[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        for (int w = 0; w < kLossBlock / 32; ++w)
            for (int k = 0; k < 4; ++k) sum[k] += s_red[w][k];
        out[0] = sum[0] * w_rgb + sum[1] * w_d;
        out[1] = sum[0]; out[2] = sum[1]; out[3] = sum[2]; out[4] = sum[3];
        *ticket = 0u;     // leave the workspace ready for the next call on this stream
    }
}

int mapping_loss_blocks(int N) {
    const int need = (N + kLossBlock - 1) / kLossBlock;
    return need < 148 * 8 ? need : 148 * 8;     // one resident wave of 8 CTAs per SM, grid-stride beyond
}

void launch_mapping_loss(int N, const float* image, const float* depth, const float* gt_image, const float* gt_depth,
                         const float* exposure_a, const float* exposure_b, float alpha, float rgb_boundary_threshold,
                         int initialization, float* workspace, float* out, float* dL_dimage, float* dL_ddepth,
                         cudaStream_t s) {
    mapping_loss_kernel<<<mapping_loss_blocks(N), kLossBlock, 0, s>>>(N, image, depth, gt_image, gt_depth, exposure_a,
                                                                     exposure_b, alpha, rgb_boundary_threshold,
                                                                     initialization, workspace, out, dL_dimage,
                                                                     dL_ddepth);
}

}  // namespace splat
