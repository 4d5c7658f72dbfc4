// Projection math shared by the forward and the backward per-Gaussian kernels, written in the exact rounding
// sequence of the reference's sm_100 SASS (see common.cuh). The backward re-derives a, b, c and the Jacobian rows
// with the SAME function, so the quantities it differentiates are bit-identical to what the forward used — the
// conic -> covariance gradient cancels catastrophically for needle-like splats and any other rounding of a*c - b*b
// would change those gradients by far more than the float-atomic noise of the reference itself.
#pragma once
#include "common.cuh"

namespace splat {

// Sigma = (S R)^T (S R) from scale and the UN-normalised quaternion (r,x,y,z) (DGR forward.cu:120-154).
// Rounding sequence reproduces the reference SASS: six cross terms are fma(u,v,+-(w*t)) with the rounded
// product being r*z, x*z and r*x respectively; squares are y*y, z*z rounded and x*x fused.
__device__ __forceinline__ void cov3d_ref(float sx, float sy, float sz, float mod, float4 q, float* cov) {
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    const float yy = fmul(y, y), zz = fmul(z, z);
    const float rz = fmul(r, z), xz = fmul(x, z), rx = fmul(r, x);
    const float xy_m_rz = ffma(x, y, -rz), xy_p_rz = ffma(x, y, rz);
    const float xz_p_ry = ffma(r, y, xz), xz_m_ry = ffma(-r, y, xz);
    const float yz_m_rx = ffma(y, z, -rx), yz_p_rx = ffma(y, z, rx);
    const float xx_zz = ffma(x, x, zz), xx_yy = ffma(x, x, yy), yy_zz = fadd(yy, zz);
    // rotation rows in the order the reference passes them to glm::mat3 (as columns)
    const float R00 = fsub(1.0f, fadd(yy_zz, yy_zz)), R01 = fadd(xy_m_rz, xy_m_rz), R02 = fadd(xz_p_ry, xz_p_ry);
    const float R10 = fadd(xy_p_rz, xy_p_rz), R11 = fsub(1.0f, fadd(xx_zz, xx_zz)), R12 = fadd(yz_m_rx, yz_m_rx);
    const float R20 = fadd(xz_m_ry, xz_m_ry), R21 = fadd(yz_p_rx, yz_p_rx), R22 = fsub(1.0f, fadd(xx_yy, xx_yy));
    const float s0 = fmul(mod, sx), s1 = fmul(mod, sy), s2 = fmul(mod, sz);
    // M (glm) column c = s (.) R-row c
    const float m00 = fmul(s0, R00), m01 = fmul(s1, R01), m02 = fmul(s2, R02);
    const float m10 = fmul(s0, R10), m11 = fmul(s1, R11), m12 = fmul(s2, R12);
    const float m20 = fmul(s0, R20), m21 = fmul(s1, R21), m22 = fmul(s2, R22);
    cov[0] = dot3_ref(m00, m00, m01, m01, m02, m02);
    cov[1] = dot3_ref(m10, m00, m11, m01, m12, m02);
    cov[2] = dot3_ref(m20, m00, m21, m01, m22, m02);
    cov[3] = dot3_ref(m10, m10, m11, m11, m12, m12);
    cov[4] = dot3_ref(m20, m10, m21, m11, m22, m12);
    cov[5] = dot3_ref(m20, m20, m21, m21, m22, m22);
}

struct EwaRef {
    float a, b, c;                 // 2D covariance with the 0.3 low-pass: (cov00 + 0.3, cov01, cov11 + 0.3)
    float J00, J02, J11, J12;      // non-zero Jacobian entries
    float T0[3], T1[3];            // rows of A = J W (glm's T[0][k], T[1][k])
    float txc, tyc;                // guard-band clamped view-space x, y
    float txtz, tytz, limx, limy;
};

// EWA projection of the 3D covariance with the 1.3*tanfov guard band and the 0.3 low-pass
// (DGR forward.cu:76-115; recomputed identically in backward.cu:171-206). t = view-space mean.
__device__ __forceinline__ EwaRef ewa_ref(float tx, float ty, float tz, float fx, float fy, float tanfovx,
                                          float tanfovy, const float* cov, const float* __restrict__ V) {
    EwaRef e;
    const float limx = fmul(tanfovx, 1.3f), limy = fmul(tanfovy, 1.3f);
    const float txtz = __fdiv_rn(tx, tz), tytz = __fdiv_rn(ty, tz);
    const float cxl = fminf(limx, fmaxf(-limx, txtz));
    const float cyl = fminf(limy, fmaxf(-limy, tytz));
    const float tz2 = fmul(tz, tz);
    const float J00 = __fdiv_rn(fx, tz);
    const float J02 = __fdiv_rn(fmul(fx, fmul(cxl, -tz)), tz2);
    const float J11 = __fdiv_rn(fy, tz);
    const float J12 = __fdiv_rn(fmul(fy, fmul(cyl, -tz)), tz2);
    float T0[3], T1[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        T0[r] = ffma(V[4 * r + 2], J02, fmul(V[4 * r], J00));
        T1[r] = ffma(V[4 * r + 2], J12, fmul(J11, V[4 * r + 1]));
        e.T0[r] = T0[r];
        e.T1[r] = T1[r];
    }
    e.J00 = J00; e.J02 = J02; e.J11 = J11; e.J12 = J12;
    e.txc = fmul(cxl, tz); e.tyc = fmul(cyl, tz);
    e.txtz = txtz; e.tytz = tytz; e.limx = limx; e.limy = limy;
    const float v0[3] = {cov[0], cov[1], cov[2]}, v1[3] = {cov[1], cov[3], cov[4]}, v2[3] = {cov[2], cov[4], cov[5]};
    const float X00 = dot3_ref(T0[0], v0[0], T0[1], v0[1], T0[2], v0[2]);
    const float X01 = dot3_ref(T1[0], v0[0], T1[1], v0[1], T1[2], v0[2]);
    const float X10 = dot3_ref(T0[0], v1[0], T0[1], v1[1], T0[2], v1[2]);
    const float X11 = dot3_ref(T1[0], v1[0], T1[1], v1[1], T1[2], v1[2]);
    const float X20 = dot3_ref(T0[0], v2[0], T0[1], v2[1], T0[2], v2[2]);
    const float X21 = dot3_ref(T1[0], v2[0], T1[1], v2[1], T1[2], v2[2]);
    const float c00 = dot3_ref(T0[0], X00, T0[1], X10, T0[2], X20);
    const float c01 = dot3_ref(T0[0], X01, T0[1], X11, T0[2], X21);
    const float c11 = dot3_ref(T1[0], X01, T1[1], X11, T1[2], X21);
    e.a = fadd(c00, 0.3f);
    e.b = c01;
    e.c = fadd(c11, 0.3f);
    return e;
}

}  // namespace splat
