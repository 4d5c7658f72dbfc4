// Densification bookkeeping after backward, all views of a mapping iteration in one launch.
// Replaces, per rendered view k (src/mapper.py:519-529 and gaussian_model.py:738-742, boolean-mask indexing = a
// nonzero() + host sync + gather/scatter each):
//   vis                      = radii_k > 0
//   max_radii2D[vis]         = max(max_radii2D[vis], radii_k[vis])
//   xyz_gradient_accum[vis] += || viewspace_grad_k[vis, :2] ||
//   denom[vis]              += 1
// and the occlusion-aware visibility of the window views (src/mapper.py:494-499, 507-509):
//   visibility_k = (n_touched_k > 0).long();   n_obs = sum_k visibility_k
// One thread per Gaussian walks the K views in order, so the float accumulation order is the reference's.
// HBM-bound: 12 B per Gaussian per view (radius + 2 gradient floats; the third gradient float shares the sector)
// + 24 B for the three running statistics.
#include "common.cuh"
#include "kernels.h"

namespace splat {

struct DensifyViews {
    const int* radii[kDensifyMaxViews];
    const float* grad[kDensifyMaxViews];
    int count;
};

__global__ void __launch_bounds__(256)
densify_stats_kernel(int P, const __grid_constant__ DensifyViews views, int grad_stride, float* __restrict__ max_radii2D,
                     float* __restrict__ xyz_gradient_accum, float* __restrict__ denom) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    float mr = max_radii2D[i], acc = xyz_gradient_accum[i], dn = denom[i];
    bool touched = false;
    for (int k = 0; k < views.count; ++k) {
        const int r = __ldg(views.radii[k] + i);
        if (r <= 0) continue;
        const float gx = __ldg(views.grad[k] + (size_t)i * grad_stride);
        const float gy = __ldg(views.grad[k] + (size_t)i * grad_stride + 1);
        mr = fmaxf(mr, (float)r);
        // torch.norm over two elements: sqrt(x*x + y*y), products accumulated left to right
        acc = __fadd_rn(acc, __fsqrt_rn(__fmaf_rn(gy, gy, __fmul_rn(gx, gx))));
        dn = __fadd_rn(dn, 1.0f);
        touched = true;
    }
    if (touched) {
        max_radii2D[i] = mr;
        xyz_gradient_accum[i] = acc;
        denom[i] = dn;
    }
}

struct VisibilityViews {
    const int* n_touched[kDensifyMaxViews];
    long long* visibility[kDensifyMaxViews];
    int count;
};

__global__ void __launch_bounds__(256)
visibility_count_kernel(int P, const __grid_constant__ VisibilityViews views, int* __restrict__ n_obs, int accumulate) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    int seen = 0;
    for (int k = 0; k < views.count; ++k) {
        const int v = __ldg(views.n_touched[k] + i) > 0 ? 1 : 0;
        if (views.visibility[k]) views.visibility[k][i] = v;
        seen += v;
    }
    if (n_obs) n_obs[i] = accumulate ? n_obs[i] + seen : seen;
}

int launch_densify_stats(int P, int K, const int* const* radii, const float* const* grad, int grad_stride,
                         float* max_radii2D, float* xyz_gradient_accum, float* denom, cudaStream_t s) {
    int launches = 0;
    for (int at = 0; at < K; at += kDensifyMaxViews) {      // later chunks continue the same per-Gaussian sequence
        DensifyViews v;
        v.count = (K - at < kDensifyMaxViews) ? K - at : kDensifyMaxViews;
        for (int k = 0; k < v.count; ++k) { v.radii[k] = radii[at + k]; v.grad[k] = grad[at + k]; }
        densify_stats_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, v, grad_stride, max_radii2D, xyz_gradient_accum, denom);
        ++launches;
    }
    return launches;
}

int launch_visibility_count(int P, int K, const int* const* n_touched, long long* const* visibility, int* n_obs,
                            cudaStream_t s) {
    int launches = 0;
    for (int at = 0; at < K; at += kDensifyMaxViews) {
        VisibilityViews v;
        v.count = (K - at < kDensifyMaxViews) ? K - at : kDensifyMaxViews;
        for (int k = 0; k < v.count; ++k) {
            v.n_touched[k] = n_touched[at + k];
            v.visibility[k] = visibility ? visibility[at + k] : nullptr;
        }
        visibility_count_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, v, n_obs, at > 0 ? 1 : 0);
        ++launches;
    }
    return launches;
}

}  // namespace splat
