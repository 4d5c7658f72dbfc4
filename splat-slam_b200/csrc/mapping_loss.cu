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
#include "loss.cuh"

namespace splat {

constexpr int kLossBlock = 256;

__global__ void __launch_bounds__(kLossBlock)
mapping_loss_kernel(int N, const float* __restrict__ image, const float* __restrict__ depth, LossDesc L) {
    const float ea = L.initialization ? 1.0f : expf(L.exposure_a[0]);
    const float eb = L.initialization ? 0.0f : L.exposure_b[0];
    const float w_rgb = L.alpha / (3.0f * (float)N), w_d = (1.0f - L.alpha) / (float)N;
    float v[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    for (int i = blockIdx.x * kLossBlock + threadIdx.x; i < N; i += gridDim.x * kLossBlock)
        loss_pixel(L, (size_t)N, (size_t)i, image[i], image[(size_t)N + i], image[2 * (size_t)N + i], depth[i], ea, eb,
                   w_rgb, w_d, v);
    loss_finish<kLossBlock>(v, L.workspace, L.out, w_rgb, w_d);
}

int mapping_loss_blocks(int N) {
    const int need = (N + kLossBlock - 1) / kLossBlock;
    return need < 148 * 8 ? need : 148 * 8;     // one resident wave of 8 CTAs per SM, grid-stride beyond
}

void launch_mapping_loss(int N, const float* image, const float* depth, const LossDesc& loss, cudaStream_t s) {
    mapping_loss_kernel<<<mapping_loss_blocks(N), kLossBlock, 0, s>>>(N, image, depth, loss);
}

}  // namespace splat
