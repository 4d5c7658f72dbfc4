// Fused parameter activations of the Gaussian model, forward and backward (first half):
// thirdparty/gaussian_splatting/scene/gaussian_model.py:53-61,76-94 applies, on every render call,
//   get_scaling = exp(_scaling)      get_rotation = normalize(_rotation)      get_opacity = sigmoid(_opacity)
// as ~6 separate torch kernels forward and ~10 backward. Here: one kernel each way, one thread per Gaussian,
// 32 B read + 32 B written forward; the backward recomputes the activations from the raw parameters instead of
// saving them. HBM-bound streaming kernels.
#include "common.cuh"
#include "kernels.h"

namespace splat {

__global__ void __launch_bounds__(256)
activations_forward_kernel(int P, const float* __restrict__ raw_scaling, const float* __restrict__ raw_rotation,
                           const float* __restrict__ raw_opacity, float* __restrict__ scales,
                           float* __restrict__ rotations, float* __restrict__ opacities) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
#pragma unroll
    for (int a = 0; a < 3; ++a) scales[3 * (size_t)i + a] = expf(raw_scaling[3 * (size_t)i + a]);
    const float4 q = reinterpret_cast<const float4*>(raw_rotation)[i];
    // torch.nn.functional.normalize: x / max(||x||_2, 1e-12)
    const float n = fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);
    reinterpret_cast<float4*>(rotations)[i] = make_float4(q.x / n, q.y / n, q.z / n, q.w / n);
    opacities[i] = 1.0f / (1.0f + expf(-raw_opacity[i]));
}

__global__ void __launch_bounds__(256)
activations_backward_kernel(int P, const float* __restrict__ raw_scaling, const float* __restrict__ raw_rotation,
                            const float* __restrict__ raw_opacity, const float* __restrict__ g_scales,
                            const float* __restrict__ g_rotations, const float* __restrict__ g_opacities,
                            float* __restrict__ d_scaling, float* __restrict__ d_rotation,
                            float* __restrict__ d_opacity) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    if (d_scaling) {
#pragma unroll
        for (int a = 0; a < 3; ++a)
            d_scaling[3 * (size_t)i + a] = g_scales ? g_scales[3 * (size_t)i + a] * expf(raw_scaling[3 * (size_t)i + a]) : 0.0f;
    }
    if (d_rotation) {
        float4 d = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (g_rotations) {
            const float4 q = reinterpret_cast<const float4*>(raw_rotation)[i];
            const float4 g = reinterpret_cast<const float4*>(g_rotations)[i];
            const float norm = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
            if (norm > 1e-12f) {          // y = q / n:  dq = (g - y (y . g)) / n
                const float4 y = make_float4(q.x / norm, q.y / norm, q.z / norm, q.w / norm);
                const float yg = y.x * g.x + y.y * g.y + y.z * g.z + y.w * g.w;
                d = make_float4((g.x - y.x * yg) / norm, (g.y - y.y * yg) / norm, (g.z - y.z * yg) / norm,
                                (g.w - y.w * yg) / norm);
            } else {                      // clamped denominator is a constant there
                d = make_float4(g.x / 1e-12f, g.y / 1e-12f, g.z / 1e-12f, g.w / 1e-12f);
            }
        }
        reinterpret_cast<float4*>(d_rotation)[i] = d;
    }
    if (d_opacity) {
        float v = 0.0f;
        if (g_opacities) {
            const float s = 1.0f / (1.0f + expf(-raw_opacity[i]));
            v = g_opacities[i] * s * (1.0f - s);
        }
        d_opacity[i] = v;
    }
}

void launch_activations_forward(int P, const float* raw_scaling, const float* raw_rotation, const float* raw_opacity,
                                float* scales, float* rotations, float* opacities, cudaStream_t s) {
    activations_forward_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, raw_scaling, raw_rotation, raw_opacity, scales,
                                                              rotations, opacities);
}

void launch_activations_backward(int P, const float* raw_scaling, const float* raw_rotation, const float* raw_opacity,
                                 const float* g_scales, const float* g_rotations, const float* g_opacities,
                                 float* d_scaling, float* d_rotation, float* d_opacity, cudaStream_t s) {
    activations_backward_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, raw_scaling, raw_rotation, raw_opacity, g_scales,
                                                               g_rotations, g_opacities, d_scaling, d_rotation,
                                                               d_opacity);
}

}  // namespace splat
