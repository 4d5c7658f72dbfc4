// Shared device helpers for the B200 (sm_100a) Gaussian rasterizer.
//
// Numerics contract: the discrete outputs (depth bits in the sort key, radii, tile rects, n_touched,
// n_contrib) must be bit-identical to the reference extension as nvcc/ptxas compile it for sm_100
// (DGR = thirdparty/diff-gaussian-rasterization-w-pose).  nvcc contracts a*b+c into FMA and ptxas fuses
// further mul/add pairs, so the exact rounding sequence is a property of the reference's *SASS*, not of its
// source.  Every bit-critical expression below is therefore written with explicit round-to-nearest
// intrinsics (__fmaf_rn/__fmul_rn/__fadd_rn are never re-fused) in the order the reference's sm_100 SASS
// evaluates them; the comments name the reference expression each sequence reproduces.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#define SPLAT_TILE 16                 // DGR/cuda_rasterizer/config.h:16-17 (BLOCK_X = BLOCK_Y = 16)
#define SPLAT_TILE_PIXELS 256
#define SPLAT_NEAR 0.001f             // as-installed near plane (Dockerfile:20 patches auxiliary.h:154)
#define SPLAT_REC_FLOATS 12           // per-instance render record: 48 B
#define SPLAT_GRAD_FLOATS 12          // per-Gaussian packed render-gradient accumulator: 48 B

namespace splat {

__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float ffma(float a, float b, float c) { return __fmaf_rn(a, b, c); }

// (p0 + p1) + p2 with p_i = u_i*v_i as the reference compiles every glm 3-term product sum:
// fma(u2, v2, fma(u0, v0, u1*v1))   (glm mat3*mat3, DGR/third_party/glm/glm/detail/type_mat3x3.inl:486-520)
__device__ __forceinline__ float dot3_ref(float u0, float v0, float u1, float v1, float u2, float v2) {
    return ffma(u2, v2, ffma(u0, v0, fmul(u1, v1)));
}

// transformPoint4x3 / 4x4 row: m0*x + m4*y + m8*z + m12  ->  (fma(z,m8, fma(x,m0, y*m4))) + m12
// (DGR/cuda_rasterizer/auxiliary.h:58-77)
__device__ __forceinline__ float affine_row_ref(float x, float y, float z, float m0, float m4, float m8, float m12) {
    return fadd(m12, ffma(z, m8, ffma(x, m0, fmul(y, m4))));
}

// ndc2Pix evaluated in double then narrowed: ((v + 1.0) * S - 1.0) * 0.5 (auxiliary.h:41-44); the reference
// SASS is DADD, DFMA(.., S, -1), DMUL 0.5, F2F.
__device__ __forceinline__ float ndc2pix_ref(float v, int S) {
    double d = __dadd_rn((double)v, 1.0);
    d = __fma_rn(d, (double)S, -1.0);
    d = __dmul_rn(d, 0.5);
    return (float)d;
}

// getRect (auxiliary.h:46-56): truncating float->int of (p -/+ r [+15]) / 16, clamped to the tile grid.
__device__ __forceinline__ void get_rect_ref(float px, float py, int radius, int gx, int gy,
                                             uint32_t& x0, uint32_t& y0, uint32_t& x1, uint32_t& y1) {
    const float r = (float)radius;
    const int ax = (int)fmul(fsub(px, r), 0.0625f);
    const int ay = (int)fmul(fsub(py, r), 0.0625f);
    const int bx = (int)fmul(fadd(fadd(fadd(px, r), 16.0f), -1.0f), 0.0625f);
    const int by = (int)fmul(fadd(fadd(fadd(py, r), 16.0f), -1.0f), 0.0625f);
    x0 = min((uint32_t)gx, (uint32_t)max(0, ax));
    y0 = min((uint32_t)gy, (uint32_t)max(0, ay));
    x1 = min((uint32_t)gx, (uint32_t)max(0, bx));
    y1 = min((uint32_t)gy, (uint32_t)max(0, by));
}

// Gaussian falloff exponent for pixel offset d = mean2D - pixel (forward.cu:344-349, backward.cu:681-684):
// power = -0.5f*(A*dx*dx + C*dy*dy) - B*dx*dy, evaluated as the reference SASS does:
//   t1 = dy*(dy*C); s = fma(dx, dx*A, t1); power = fma(s, -0.5, -(dy*(dx*B)))
__device__ __forceinline__ float gauss_power_ref(float dx, float dy, float A, float B, float C) {
    const float t1 = fmul(dy, fmul(dy, C));
    const float s = ffma(dx, fmul(dx, A), t1);
    const float t3 = fmul(dy, fmul(dx, B));
    return ffma(s, -0.5f, -t3);
}

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

// Conservative axis-aligned half extents (hx, hy), packed as two round-up fp16, of the region where a Gaussian with
// conic (A, B, C) and opacity o can pass the render kernels' alpha >= 1/255 test: alpha = o*exp(-q(d)) >= 1/255
// needs q(d) <= tau = ln(255*o), and {q <= tau} with q(d) = 0.5*(A dx^2 + C dy^2) + B dx dy is an ellipse whose
// bounding box has half widths sqrt(2 tau C/det), sqrt(2 tau A/det). Margins (+0.02 on tau, an anisotropy-scaled
// widening, x1.001, +0.05 px, fp16 round-up) cover the float rounding of power/expf, so a pixel the reference would
// blend is never outside the box. Degenerate conics -> "everywhere"; opacity < 1/255 -> "nowhere".
__device__ __forceinline__ uint32_t footprint_half_extents(float A, float B, float C, float o) {
    float hx = 65504.0f, hy = 65504.0f;
    const float tau0 = logf(255.0f * o) + 0.02f;
    const float det = A * C - B * B;
    if (tau0 < 0.0f) {
        hx = hy = -65504.0f;
    } else if (det > 0.0f && A > 0.0f && C > 0.0f) {
        // the terms of `power` reach 2*tau*A*C/det near the ellipse, so its float rounding error scales with the
        // anisotropy A*C/det: widen tau accordingly (negligible for ordinary splats, safe for needle-like ones)
        const float tau = tau0 * (1.0f + 4e-6f * (A * C / det));
        const float ex = sqrtf(2.0f * tau * C / det) * 1.001f + 0.05f;
        const float ey = sqrtf(2.0f * tau * A / det) * 1.001f + 0.05f;
        if (ex < 65504.0f) hx = ex;
        if (ey < 65504.0f) hy = ey;
    }
    const __half2 h = __halves2half2(__float2half_ru(hx), __float2half_ru(hy));
    return *reinterpret_cast<const uint32_t*>(&h);
}

}  // namespace splat

// ---- status codes of the C ABI ------------------------------------------------------------------------
#define SPLAT_OK 0
#define SPLAT_ERR_ARG (-1)
#define SPLAT_ERR_SIZE (-2)
#define SPLAT_CUDA_CHECK_LAUNCH()                                  \
    do {                                                           \
        cudaError_t e__ = cudaGetLastError();                      \
        if (e__ != cudaSuccess) return (int)e__;                   \
    } while (0)
