// Per-pixel arithmetic and the final reduction of the mapping loss, shared by the stand-alone loss kernel
// (mapping_loss.cu) and the epilogue of the forward renderer (render.cu). Reference:
// thirdparty/monogs/utils/slam_utils.py:71-105 (get_loss_mapping / get_loss_mapping_rgbd, ssim_loss = False).
#pragma once
#include "common.cuh"

namespace splat {

struct LossDesc {
    const float* gt_image = nullptr;      // [3*N]
    const float* gt_depth = nullptr;      // [N]
    const float* exposure_a = nullptr;    // device scalar (unused with `initialization`)
    const float* exposure_b = nullptr;
    float alpha = 0.95f, rgb_boundary_threshold = 0.01f;
    int initialization = 0;
    float* workspace = nullptr;           // [4 + 4 * partial slots]: ticket word, then per-block partial sums
    float* out = nullptr;                 // [5] loss, sum l1_rgb, sum l1_depth, dL/da, dL/db
    float* dL_dimage = nullptr;           // [3*N]
    float* dL_ddepth = nullptr;           // [N]
};

__device__ __forceinline__ float loss_sign0(float x) { return (x > 0.0f) ? 1.0f : ((x < 0.0f) ? -1.0f : 0.0f); }

// One pixel: accumulates (l1_rgb, l1_depth, dL/da, dL/db) into v and writes the pixel's dL/dimage, dL/ddepth.
// Every op is rounded on its own, as the reference's chain of separate torch kernels does (no FMA contraction: the
// sign of a near-zero difference decides the gradient).
__device__ __forceinline__ void loss_pixel(const LossDesc& L, size_t N, size_t i, float x0, float x1, float x2, float d,
                                           float ea, float eb, float w_rgb, float w_d, float (&v)[4]) {
    const float g0 = L.gt_image[i], g1 = L.gt_image[N + i], g2 = L.gt_image[2 * N + i];
    const float m = (__fadd_rn(__fadd_rn(g0, g1), g2) > L.rgb_boundary_threshold) ? 1.0f : 0.0f;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
        const float x = ch == 0 ? x0 : (ch == 1 ? x1 : x2);
        const float gt = ch == 0 ? g0 : (ch == 1 ? g1 : g2);
        const float diff = __fsub_rn(__fmul_rn(__fadd_rn(__fmul_rn(ea, x), eb), m), __fmul_rn(gt, m));
        v[0] += fabsf(diff);
        const float g = w_rgb * loss_sign0(diff) * m;     // dL / d image_ab
        L.dL_dimage[ch * N + i] = g * ea;
        v[2] += g * ea * x;
        v[3] += g;
    }
    const float gd = L.gt_depth[i];
    const float md = (gd > 0.01f) ? 1.0f : 0.0f;
    const float dd = __fsub_rn(__fmul_rn(d, md), __fmul_rn(gd, md));
    v[1] += fabsf(dd);
    L.dL_ddepth[i] = w_d * loss_sign0(dd) * md;
}

// Block-wide: reduce the threads' four sums into partials[blockIdx.x]; the block that takes the last ticket adds all
// partials up in a fixed order (deterministic, no float atomics) and writes loss + sums. kThreads = block size.
template <int kThreads>
__device__ __forceinline__ void loss_finish(float (&v)[4], float* workspace, float* out, float w_rgb, float w_d) {
    __shared__ float s_red[kThreads / 32][4];
    __shared__ bool s_last;
    unsigned int* ticket = reinterpret_cast<unsigned int*>(workspace);
    float* partials = workspace + 4;
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
        for (int w = 0; w < kThreads / 32; ++w) s += s_red[w][threadIdx.x];
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
    for (int blk = threadIdx.x; blk < (int)gridDim.x; blk += kThreads) {
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
        for (int w = 0; w < kThreads / 32; ++w)
            for (int k = 0; k < 4; ++k) sum[k] += s_red[w][k];
        out[0] = sum[0] * w_rgb + sum[1] * w_d;
        out[1] = sum[0]; out[2] = sum[1]; out[3] = sum[2]; out[4] = sum[3];
        *ticket = 0u;     // leave the workspace ready for the next call on this stream
    }
}

}  // namespace splat
