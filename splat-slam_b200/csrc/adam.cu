// Fused Adam step for the Gaussian parameter groups and the per-keyframe pose/exposure parameters (SURVEY.md §8f
// rank 2: the step right after backward, thirdparty/gaussian_splatting/scene/gaussian_model.py:264-302 builds the
// groups, src/mapper.py:557-562 steps them). The reference runs torch.optim.Adam(lr per group, eps = 1e-15), whose
// multi-tensor path issues, per step and per group list:
//   exp_avg.lerp_(grad, 1 - beta1); exp_avg_sq.mul_(beta2); exp_avg_sq.addcmul_(grad, grad, 1 - beta2);
//   denom = sqrt(exp_avg_sq) / sqrt(1 - beta2^t) + eps;  param.addcdiv_(exp_avg, denom, -lr / (1 - beta1^t))
// i.e. seven passes and 72 bytes per parameter. Here: ONE launch for up to kAdamMaxSegments tensors, 32 bytes per
// parameter (read p, g, m, v; write p, m, v and, on request, g = 0 for the next accumulation), each operation rounded
// where the chain of separate kernels rounds it, so the result tracks torch's to the last bit or two.
// HBM-bound streaming kernel.
#include "common.cuh"
#include "kernels.h"

namespace splat {

constexpr int kAdamBlock = 256;
constexpr int kAdamPerThread = 8;                                 // two float4 per stream per thread
constexpr int kAdamChunk = kAdamBlock * kAdamPerThread;           // elements per CTA

struct AdamBatch {
    AdamSegment seg[kAdamMaxSegments];
    int first_block[kAdamMaxSegments + 1];
    int count;
    int zero_grads;
};

__device__ __forceinline__ void adam_update(float& p, float g, float& m, float& v, const AdamSegment& s) {
    m = __fmaf_rn(s.w1, __fsub_rn(g, m), m);                          // lerp, weight < 0.5 branch
    v = __fmaf_rn(s.w2, __fmul_rn(g, g), __fmul_rn(v, s.beta2));      // mul_ then addcmul_
    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), s.bc2_sqrt), s.eps);
    p = __fmaf_rn(s.neg_step, __fdiv_rn(m, denom), p);                // addcdiv_
}

__global__ void __launch_bounds__(kAdamBlock)
adam_step_kernel(const __grid_constant__ AdamBatch batch) {
    int si = 0;
    while (si + 1 < batch.count && (int)blockIdx.x >= batch.first_block[si + 1]) ++si;
    const AdamSegment& s = batch.seg[si];
    const long long base = (long long)(blockIdx.x - batch.first_block[si]) * kAdamChunk;
    const bool zero = batch.zero_grads != 0;
    if (s.vec4 && base + kAdamChunk <= s.n) {
        float4* p4 = reinterpret_cast<float4*>(s.p + base);
        float4* g4 = reinterpret_cast<float4*>(s.g + base);
        float4* m4 = reinterpret_cast<float4*>(s.m + base);
        float4* v4 = reinterpret_cast<float4*>(s.v + base);
#pragma unroll
        for (int k = 0; k < kAdamPerThread / 4; ++k) {
            const int i = k * kAdamBlock + threadIdx.x;
            float4 p = p4[i], g = __ldcs(g4 + i), m = m4[i], v = v4[i];
            adam_update(p.x, g.x, m.x, v.x, s);
            adam_update(p.y, g.y, m.y, v.y, s);
            adam_update(p.z, g.z, m.z, v.z, s);
            adam_update(p.w, g.w, m.w, v.w, s);
            p4[i] = p; m4[i] = m; v4[i] = v;
            if (zero) g4[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        }
    } else {
        for (int k = 0; k < kAdamPerThread; ++k) {
            const long long i = base + k * kAdamBlock + threadIdx.x;
            if (i >= s.n) break;
            float p = s.p[i], m = s.m[i], v = s.v[i];
            adam_update(p, s.g[i], m, v, s);
            s.p[i] = p; s.m[i] = m; s.v[i] = v;
            if (zero) s.g[i] = 0.0f;
        }
    }
}

int launch_adam_step(int count, const AdamSegment* segments, int zero_grads, cudaStream_t stream) {
    int launches = 0;
    for (int at = 0; at < count; at += kAdamMaxSegments) {
        AdamBatch b;
        b.count = 0;
        b.zero_grads = zero_grads;
        int blocks = 0;
        for (int k = at; k < count && b.count < kAdamMaxSegments; ++k) {
            if (segments[k].n <= 0) continue;
            b.first_block[b.count] = blocks;
            b.seg[b.count] = segments[k];
            const uintptr_t bits = reinterpret_cast<uintptr_t>(segments[k].p) | reinterpret_cast<uintptr_t>(segments[k].g) |
                                   reinterpret_cast<uintptr_t>(segments[k].m) | reinterpret_cast<uintptr_t>(segments[k].v);
            b.seg[b.count].vec4 = (bits & 15u) == 0 ? 1 : 0;
            blocks += (int)((segments[k].n + kAdamChunk - 1) / kAdamChunk);
            ++b.count;
        }
        if (b.count == 0) continue;
        b.first_block[b.count] = blocks;
        adam_step_kernel<<<blocks, kAdamBlock, 0, stream>>>(b);
        ++launches;
    }
    return launches;
}

}  // namespace splat
