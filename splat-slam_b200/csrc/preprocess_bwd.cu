// Per-Gaussian backward stage, one pass: conic -> cov2D -> (cov3D, mean, pose), mean2D -> (mean3D, pose),
// depth -> (mean3D, pose), SH backward, cov3D -> (scale, rotation), plus the block-level reduction of the
// camera-pose gradient. Fuses DGR/cuda_rasterizer/backward.cu:150-346 (computeCov2DCUDA), :418-539
// (preprocessCUDA), :21-145 (computeColorFromSH), :350-413 (computeCov3D) and the P x 6 `dL_dtau` reduction
// that the reference leaves to torch.sum (DGR/diff_gaussian_rasterization/__init__.py:152).
//
// Every output row is written exactly once (zeros for culled Gaussians), so none of the ten output tensors
// needs the zero-fill the reference pays for (rasterize_points.cu:160-170); only the 48-byte packed
// render-gradient accumulator, which backward render fills with atomics, is cleared beforehand.
#include "common.cuh"
#include "kernels.h"
#include "project.cuh"

namespace splat {

__device__ const float kB_SH_C0 = 0.28209479177387814f;
__device__ const float kB_SH_C1 = 0.4886025119029199f;
__device__ const float kB_SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                      -1.0925484305920792f, 0.5462742152960396f};
__device__ const float kB_SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                      0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                      -0.5900435899266435f};

struct f3 { float x, y, z; };
__device__ __forceinline__ f3 mk(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ f3 operator*(float s, f3 a) { return mk(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ f3 operator+(f3 a, f3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ float dot(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ f3 cross(f3 a, f3 b) {
    return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}

// d(normalize(v))/dv applied to dv (auxiliary.h:108-119)
__device__ __forceinline__ f3 dnormvdv(f3 v, f3 dv) {
    const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
    const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    f3 r;
    r.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
    r.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
    r.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
    return r;
}

// SH backward (backward.cu:21-145): writes dL_dsh rows [0, M), returns dL/dmean through the view direction.
// Everything that READS the coefficients (the direction derivatives dRGB/dx,y,z) comes before the first write of a
// gradient row. `dL_dsh_out` is either the global row or the caller's shared-memory staging row.
template <bool kHigherDegrees>
__device__ f3 sh_backward(int deg, int M, const float* shs, f3 pos, f3 campos, uint8_t clamp_bits, f3 dL_dRGB,
                          float* dL_dsh_out) {
    if (!kHigherDegrees) deg = 0;        // the SH degree 0 build carries none of the higher-degree code
    const f3 dir_orig = mk(pos.x - campos.x, pos.y - campos.y, pos.z - campos.z);
    const float inv_len = 1.0f / sqrtf(dot(dir_orig, dir_orig));
    const float x = dir_orig.x * inv_len, y = dir_orig.y * inv_len, z = dir_orig.z * inv_len;
    // coefficient row read with 128-bit loads when it is 16-byte aligned (M a multiple of 4), a degree at a time
    const bool vec = ((M & 3) == 0) && ((reinterpret_cast<uintptr_t>(shs) & 15u) == 0);
    const float4* sh4 = reinterpret_cast<const float4*>(shs);
    float f[48];
#define SH_LOAD(first4, count4)                                                        \
    _Pragma("unroll") for (int i__ = (first4); i__ < (first4) + (count4); ++i__) {     \
        if (vec) {                                                                     \
            const float4 v__ = __ldg(sh4 + i__);                                       \
            f[4 * i__] = v__.x; f[4 * i__ + 1] = v__.y; f[4 * i__ + 2] = v__.z; f[4 * i__ + 3] = v__.w; \
        } else {                                                                       \
            _Pragma("unroll") for (int j__ = 0; j__ < 4; ++j__)                        \
                f[4 * i__ + j__] = (4 * i__ + j__ < 3 * M) ? shs[4 * i__ + j__] : 0.0f; \
        }                                                                              \
    }
#define SHC(k) mk(f[3 * (k)], f[3 * (k) + 1], f[3 * (k) + 2])
    f3* out = reinterpret_cast<f3*>(dL_dsh_out);
    if (clamp_bits & 1) dL_dRGB.x = 0.0f;
    if (clamp_bits & 2) dL_dRGB.y = 0.0f;
    if (clamp_bits & 4) dL_dRGB.z = 0.0f;
    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
    f3 dRGBdx = mk(0, 0, 0), dRGBdy = mk(0, 0, 0), dRGBdz = mk(0, 0, 0);
    if (deg > 0) {
        SH_LOAD(0, 3)
        dRGBdx = (-kB_SH_C1) * SHC(3);
        dRGBdy = (-kB_SH_C1) * SHC(1);
        dRGBdz = kB_SH_C1 * SHC(2);
        if (deg > 1) {
            SH_LOAD(3, 4)
            dRGBdx = dRGBdx + (kB_SH_C2[0] * y) * SHC(4) + (kB_SH_C2[2] * 2.f * -x) * SHC(6) + (kB_SH_C2[3] * z) * SHC(7) +
                     (kB_SH_C2[4] * 2.f * x) * SHC(8);
            dRGBdy = dRGBdy + (kB_SH_C2[0] * x) * SHC(4) + (kB_SH_C2[1] * z) * SHC(5) + (kB_SH_C2[2] * 2.f * -y) * SHC(6) +
                     (kB_SH_C2[4] * 2.f * -y) * SHC(8);
            dRGBdz = dRGBdz + (kB_SH_C2[1] * y) * SHC(5) + (kB_SH_C2[2] * 2.f * 2.f * z) * SHC(6) + (kB_SH_C2[3] * x) * SHC(7);
            if (deg > 2) {
                SH_LOAD(7, 5)
                dRGBdx = dRGBdx + (kB_SH_C3[0] * 3.f * 2.f * xy) * SHC(9) + (kB_SH_C3[1] * yz) * SHC(10) +
                         (kB_SH_C3[2] * -2.f * xy) * SHC(11) + (kB_SH_C3[3] * -3.f * 2.f * xz) * SHC(12) +
                         (kB_SH_C3[4] * (-3.f * xx + 4.f * zz - yy)) * SHC(13) + (kB_SH_C3[5] * 2.f * xz) * SHC(14) +
                         (kB_SH_C3[6] * 3.f * (xx - yy)) * SHC(15);
                dRGBdy = dRGBdy + (kB_SH_C3[0] * 3.f * (xx - yy)) * SHC(9) + (kB_SH_C3[1] * xz) * SHC(10) +
                         (kB_SH_C3[2] * (-3.f * yy + 4.f * zz - xx)) * SHC(11) + (kB_SH_C3[3] * -3.f * 2.f * yz) * SHC(12) +
                         (kB_SH_C3[4] * -2.f * xy) * SHC(13) + (kB_SH_C3[5] * -2.f * yz) * SHC(14) +
                         (kB_SH_C3[6] * -3.f * 2.f * xy) * SHC(15);
                dRGBdz = dRGBdz + (kB_SH_C3[1] * xy) * SHC(10) + (kB_SH_C3[2] * 4.f * 2.f * yz) * SHC(11) +
                         (kB_SH_C3[3] * 3.f * (2.f * zz - xx - yy)) * SHC(12) + (kB_SH_C3[4] * 4.f * 2.f * xz) * SHC(13) +
                         (kB_SH_C3[5] * (xx - yy)) * SHC(14);
            }
        }
    }
    int written = 1;
    out[0] = kB_SH_C0 * dL_dRGB;
    if (deg > 0) {
        out[1] = (-kB_SH_C1 * y) * dL_dRGB;
        out[2] = (kB_SH_C1 * z) * dL_dRGB;
        out[3] = (-kB_SH_C1 * x) * dL_dRGB;
        written = 4;
        if (deg > 1) {
            out[4] = (kB_SH_C2[0] * xy) * dL_dRGB;
            out[5] = (kB_SH_C2[1] * yz) * dL_dRGB;
            out[6] = (kB_SH_C2[2] * (2.f * zz - xx - yy)) * dL_dRGB;
            out[7] = (kB_SH_C2[3] * xz) * dL_dRGB;
            out[8] = (kB_SH_C2[4] * (xx - yy)) * dL_dRGB;
            written = 9;
            if (deg > 2) {
                out[9] = (kB_SH_C3[0] * y * (3.f * xx - yy)) * dL_dRGB;
                out[10] = (kB_SH_C3[1] * xy * z) * dL_dRGB;
                out[11] = (kB_SH_C3[2] * y * (4.f * zz - xx - yy)) * dL_dRGB;
                out[12] = (kB_SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy)) * dL_dRGB;
                out[13] = (kB_SH_C3[4] * x * (4.f * zz - xx - yy)) * dL_dRGB;
                out[14] = (kB_SH_C3[5] * z * (xx - yy)) * dL_dRGB;
                out[15] = (kB_SH_C3[6] * x * (xx - 3.f * yy)) * dL_dRGB;
                written = 16;
            }
        }
    }
    for (int k = written; k < M; ++k) out[k] = mk(0, 0, 0);
    const f3 dL_ddir = mk(dot(dRGBdx, dL_dRGB), dot(dRGBdy, dL_dRGB), dot(dRGBdz, dL_dRGB));
    return dnormvdv(dir_orig, dL_ddir);
#undef SHC
#undef SH_LOAD
}

// Row stride (floats) of the shared-memory staging of SH rows: odd, so that the 256 threads of a block reading or
// writing element k of their own row hit 32 different banks.
__host__ __device__ inline int sh_stage_stride(int M) { return (3 * M) | 1; }

template <bool kHigherDegrees>
__global__ void __launch_bounds__(kPreBlock, 4)
preprocess_backward_kernel(BwdParams p, GeomState g, float* __restrict__ tau_partial, int stage_sh) {
    __shared__ float sV[16], sPm[16];
    __shared__ float s_tau[kPreBlock / 32][6];
    extern __shared__ float s_sh[];      // [kPreBlock][sh_stage_stride(M)] when the SH rows are staged (stage_sh)
    if (threadIdx.x < 16) { sV[threadIdx.x] = p.viewmatrix[threadIdx.x]; sPm[threadIdx.x] = p.projmatrix[threadIdx.x]; }
    // With more than one SH coefficient a thread's gradient row is 12*M bytes: per-thread stores would touch 32
    // different sectors per instruction (4 bytes of each). Instead every thread writes its row into shared memory
    // (odd row stride: no bank conflicts) and the block copies its rows out as one contiguous range of float4.
    // (Staging the coefficient READS the same way was measured and is slower: the per-thread loads overlap with the
    // covariance math, a separate load phase does not.)
    const int row = 3 * p.M, stride = sh_stage_stride(p.M);
    const int block_first = blockIdx.x * kPreBlock;
    const int nrows = min(kPreBlock, p.P - block_first);
    __syncthreads();
    const int idx = blockIdx.x * kPreBlock + threadIdx.x;
    float tau[6] = {0, 0, 0, 0, 0, 0};
    if (idx < p.P) {
        const bool vis = p.radii[idx] > 0;
        f3 dmean = mk(0, 0, 0);
        float dcov[6] = {0, 0, 0, 0, 0, 0};
        float4 acc0 = make_float4(0, 0, 0, 0), acc1 = acc0, acc2 = acc0;
        if (vis) {
            const float4* acc = reinterpret_cast<const float4*>(p.grad_accum) + 3 * (size_t)idx;
            acc0 = acc[0];   // moments of u = G dL/dalpha over the pixels: sum u dx, sum u dy, sum u dx^2, sum u dx dy
            acc1 = acc[1];   // sum u dy^2, sum u (= dL/dopacity), r, g
            acc2 = acc[2];   // b, depth, -, -
            {   // moments -> dL/dmean2D (NDC units), dL/dconic: the per-Gaussian factors the render backward left out
                // (backward.cu:745-770: dL_dG = opacity * dL_dalpha, dG_ddel = -G (conic . d), ddel_dx = 0.5 W)
                const float4 co = g.conic_opacity[idx];
                const float sx = acc0.x, sy = acc0.y, ho = -0.5f * co.w;
                acc0.x = ho * (float)p.W * (co.x * sx + co.y * sy);
                acc0.y = ho * (float)p.H * (co.z * sy + co.y * sx);
                acc0.z *= ho;
                acc0.w *= ho;
                acc1.x *= ho;
            }
            const f3 m = mk(p.means3D[3 * idx], p.means3D[3 * idx + 1], p.means3D[3 * idx + 2]);
            const float* cov3D = p.cov3D_precomp ? p.cov3D_precomp + 6 * (size_t)idx : g.cov3D + 6 * (size_t)idx;

            // ---------------- conic -> cov2D -> cov3D / t / W  (backward.cu:150-346) ----------------
            // view-space mean, guard-band clamp, Jacobian, A = J W and (a, b, c): the forward's own bit-exact sequence
            f3 t = mk(affine_row_ref(m.x, m.y, m.z, sV[0], sV[4], sV[8], sV[12]),
                      affine_row_ref(m.x, m.y, m.z, sV[1], sV[5], sV[9], sV[13]),
                      affine_row_ref(m.x, m.y, m.z, sV[2], sV[6], sV[10], sV[14]));
            const f3 p_C = t;   // unclamped camera-frame point, used by the mean2D/depth pose terms below
            const float hx = p.focal_x, hy = p.focal_y;
            const EwaRef e = ewa_ref(t.x, t.y, t.z, hx, hy, p.tanfovx, p.tanfovy, cov3D, sV);
            t.x = e.txc;
            t.y = e.tyc;
            const float x_grad_mul = (e.txtz < -e.limx || e.txtz > e.limx) ? 0.0f : 1.0f;
            const float y_grad_mul = (e.tytz < -e.limy || e.tytz > e.limy) ? 0.0f : 1.0f;
            const float J00 = e.J00, J02 = e.J02, J11 = e.J11, J12 = e.J12;
            // rows of the world->camera rotation: R(i,k) = view[4k+i]
            const f3 R0 = mk(sV[0], sV[4], sV[8]), R1 = mk(sV[1], sV[5], sV[9]), R2 = mk(sV[2], sV[6], sV[10]);
            const f3 A0 = mk(e.T0[0], e.T0[1], e.T0[2]);     // A = J R, rows 0 and 1
            const f3 A1 = mk(e.T1[0], e.T1[1], e.T1[2]);
            const f3 V0 = mk(cov3D[0], cov3D[1], cov3D[2]), V1 = mk(cov3D[1], cov3D[3], cov3D[4]),
                     V2 = mk(cov3D[2], cov3D[4], cov3D[5]);
            const f3 VA0 = mk(dot(V0, A0), dot(V1, A0), dot(V2, A0));
            const f3 VA1 = mk(dot(V0, A1), dot(V1, A1), dot(V2, A1));
            const float a = e.a, b = e.b, c = e.c;
            const float denom = ffma(a, c, -fmul(b, b));
            const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
            const float gx = acc0.z, gy = acc0.w, gz = acc1.x;   // dL_dconic slots [0],[1],[3]
            float dL_da = 0, dL_db = 0, dL_dc = 0;
            if (denom2inv != 0.0f) {
                dL_da = denom2inv * (-c * c * gx + 2 * b * c * gy + (denom - a * c) * gz);
                dL_dc = denom2inv * (-a * a * gz + 2 * a * b * gy + (denom - a * c) * gx);
                dL_db = denom2inv * 2 * (b * c * gx - (denom + 2 * b * b) * gy + a * b * gz);
                dcov[0] = A0.x * A0.x * dL_da + A0.x * A1.x * dL_db + A1.x * A1.x * dL_dc;
                dcov[3] = A0.y * A0.y * dL_da + A0.y * A1.y * dL_db + A1.y * A1.y * dL_dc;
                dcov[5] = A0.z * A0.z * dL_da + A0.z * A1.z * dL_db + A1.z * A1.z * dL_dc;
                dcov[1] = 2 * A0.x * A0.y * dL_da + (A0.x * A1.y + A0.y * A1.x) * dL_db + 2 * A1.x * A1.y * dL_dc;
                dcov[2] = 2 * A0.x * A0.z * dL_da + (A0.x * A1.z + A0.z * A1.x) * dL_db + 2 * A1.x * A1.z * dL_dc;
                dcov[4] = 2 * A0.z * A0.y * dL_da + (A0.y * A1.z + A0.z * A1.y) * dL_db + 2 * A1.y * A1.z * dL_dc;
            }
            const f3 dA0 = (2 * dL_da) * VA0 + dL_db * VA1;    // dL/dA(0,:)
            const f3 dA1 = (2 * dL_dc) * VA1 + dL_db * VA0;    // dL/dA(1,:)
            const float dJ00 = dot(R0, dA0), dJ02 = dot(R2, dA0), dJ11 = dot(R1, dA1), dJ12 = dot(R2, dA1);
            const float tz = 1.0f / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
            const f3 dt = mk(x_grad_mul * -hx * tz2 * dJ02, y_grad_mul * -hy * tz2 * dJ12,
                             -hx * tz2 * dJ00 - hy * tz2 * dJ11 + (2 * hx * t.x) * tz3 * dJ02 + (2 * hy * t.y) * tz3 * dJ12);
            // pose: d p_C / d rho = I, d p_C / d theta = -skew(t) with the CLAMPED t (backward.cu:273-288)
            const f3 th_t = cross(t, dt);
            tau[0] += dt.x; tau[1] += dt.y; tau[2] += dt.z;
            tau[3] += th_t.x; tau[4] += th_t.y; tau[5] += th_t.z;
            // mean through t: R^T dt (backward.cu:292-297)
            dmean = mk(sV[0] * dt.x + sV[1] * dt.y + sV[2] * dt.z, sV[4] * dt.x + sV[5] * dt.y + sV[6] * dt.z,
                       sV[8] * dt.x + sV[9] * dt.y + sV[10] * dt.z);
            // rotation columns through W (backward.cu:299-343): theta += sum_k col_k(R) x dL/dR(:,k)
            const f3 dW0 = J00 * dA0, dW1 = J11 * dA1, dW2 = J02 * dA0 + J12 * dA1;   // rows m of dL/dR(m, k)
            {
                const f3 c0 = mk(sV[0], sV[1], sV[2]), c1 = mk(sV[4], sV[5], sV[6]), c2 = mk(sV[8], sV[9], sV[10]);
                const f3 w0 = cross(c0, mk(dW0.x, dW1.x, dW2.x));
                const f3 w1 = cross(c1, mk(dW0.y, dW1.y, dW2.y));
                const f3 w2 = cross(c2, mk(dW0.z, dW1.z, dW2.z));
                tau[3] += w0.x + w1.x + w2.x; tau[4] += w0.y + w1.y + w2.y; tau[5] += w0.z + w1.z + w2.z;
            }

            // ---------------- mean2D -> mean3D / pose (backward.cu:446-512) ----------------
            const float m_hx = sPm[0] * m.x + sPm[4] * m.y + sPm[8] * m.z + sPm[12];
            const float m_hy = sPm[1] * m.x + sPm[5] * m.y + sPm[9] * m.z + sPm[13];
            const float m_hw = sPm[3] * m.x + sPm[7] * m.y + sPm[11] * m.z + sPm[15];
            const float m_w = 1.0f / (m_hw + 0.0000001f);
            const float mul1 = m_hx * m_w * m_w, mul2 = m_hy * m_w * m_w;
            const float g2x = acc0.x, g2y = acc0.y;
            dmean.x += (sPm[0] * m_w - sPm[3] * mul1) * g2x + (sPm[1] * m_w - sPm[3] * mul2) * g2y;
            dmean.y += (sPm[4] * m_w - sPm[7] * mul1) * g2x + (sPm[5] * m_w - sPm[7] * mul2) * g2y;
            dmean.z += (sPm[8] * m_w - sPm[11] * mul1) * g2x + (sPm[9] * m_w - sPm[11] * mul2) * g2y;
            {
                const float alpha = m_w, beta = -m_hx * m_w * m_w, gamma = -m_hy * m_w * m_w;
                const float pa = p.projmatrix_raw[0], pb = p.projmatrix_raw[5], pe = p.projmatrix_raw[11];
                const f3 v1 = mk(alpha * pa, 0.0f, beta * pe), v2 = mk(0.0f, alpha * pb, gamma * pe);
                const f3 th1 = cross(p_C, v1), th2 = cross(p_C, v2);
                tau[0] += g2x * v1.x + g2y * v2.x; tau[1] += g2x * v1.y + g2y * v2.y; tau[2] += g2x * v1.z + g2y * v2.z;
                tau[3] += g2x * th1.x + g2y * th2.x; tau[4] += g2x * th1.y + g2y * th2.y; tau[5] += g2x * th1.z + g2y * th2.z;
            }
            // ---------------- depth -> mean3D / pose (backward.cu:514-528) ----------------
            const float gd = acc2.y;
            dmean.x += gd * sV[2]; dmean.y += gd * sV[6]; dmean.z += gd * sV[10];
            tau[2] += gd; tau[3] += gd * p_C.y; tau[4] += -gd * p_C.x;

            // ---------------- SH (backward.cu:532-534) ----------------
            if (p.shs != nullptr && p.dL_dsh != nullptr) {
                float* my_row = s_sh + threadIdx.x * stride;
                const f3 dm = sh_backward<kHigherDegrees>(p.D, p.M, p.shs + (size_t)idx * row, m,
                                          mk(p.campos[0], p.campos[1], p.campos[2]), g.clamped[idx],
                                          mk(acc1.z, acc1.w, acc2.x), stage_sh ? my_row : p.dL_dsh + (size_t)idx * row);
                dmean = dmean + dm;
                tau[0] -= dm.x; tau[1] -= dm.y; tau[2] -= dm.z;
            }
            // ---------------- cov3D -> scale / rotation (backward.cu:350-413) ----------------
            if (p.scales != nullptr && p.dL_dscale != nullptr) {
                const float4 q4 = reinterpret_cast<const float4*>(p.rotations)[idx];
                const float r = q4.x, x = q4.y, y = q4.z, z = q4.w;
                // rows of Rq; glm's R has them as columns, M = S * R  ->  M(col c) = s (.) Rq-row c
                const f3 q0 = mk(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y));
                const f3 q1 = mk(2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x));
                const f3 q2 = mk(2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
                const f3 s = mk(p.scale_modifier * p.scales[3 * idx], p.scale_modifier * p.scales[3 * idx + 1],
                                p.scale_modifier * p.scales[3 * idx + 2]);
                // symmetric dL/dSigma with halved off-diagonals
                const f3 S0 = mk(dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]), S1 = mk(0.5f * dcov[1], dcov[3], 0.5f * dcov[4]),
                         S2 = mk(0.5f * dcov[2], 0.5f * dcov[4], dcov[5]);
                // Sigma_ij = sum_k s_k^2 Rq(i,k) Rq(j,k). With G = dL/dSigma (symmetric):
                //   dL/ds_k = 2 s_k * sum_ij G_ij Rq(i,k) Rq(j,k);  dL/dRq(i,k) = 2 s_k^2 * sum_j G_ij Rq(j,k)
                const f3 col0 = mk(q0.x, q1.x, q2.x), col1 = mk(q0.y, q1.y, q2.y), col2 = mk(q0.z, q1.z, q2.z);
                const f3 Gc0 = mk(dot(S0, col0), dot(S1, col0), dot(S2, col0));
                const f3 Gc1 = mk(dot(S0, col1), dot(S1, col1), dot(S2, col1));
                const f3 Gc2 = mk(dot(S0, col2), dot(S1, col2), dot(S2, col2));
                p.dL_dscale[3 * idx] = 2.f * s.x * dot(col0, Gc0);
                p.dL_dscale[3 * idx + 1] = 2.f * s.y * dot(col1, Gc1);
                p.dL_dscale[3 * idx + 2] = 2.f * s.z * dot(col2, Gc2);
                // dL/dRq(i,k): D[i][k]
                const f3 D0 = mk(2.f * s.x * s.x * Gc0.x, 2.f * s.y * s.y * Gc1.x, 2.f * s.z * s.z * Gc2.x);
                const f3 D1 = mk(2.f * s.x * s.x * Gc0.y, 2.f * s.y * s.y * Gc1.y, 2.f * s.z * s.z * Gc2.y);
                const f3 D2 = mk(2.f * s.x * s.x * Gc0.z, 2.f * s.y * s.y * Gc1.z, 2.f * s.z * s.z * Gc2.z);
                // chain through Rq(q) for the raw quaternion (no normalisation Jacobian, backward.cu:410-412)
                float4 dq;
                dq.x = 2.f * (-z * D0.y + y * D0.z + z * D1.x - x * D1.z - y * D2.x + x * D2.y);
                dq.y = 2.f * (y * D0.y + z * D0.z + y * D1.x - 2.f * x * D1.y - r * D1.z + z * D2.x + r * D2.y - 2.f * x * D2.z);
                dq.z = 2.f * (-2.f * y * D0.x + x * D0.y + r * D0.z + x * D1.x + z * D1.z - r * D2.x + z * D2.y - 2.f * y * D2.z);
                dq.w = 2.f * (-2.f * z * D0.x - r * D0.y + x * D0.z + r * D1.x - 2.f * z * D1.y + y * D1.z + x * D2.x + y * D2.y);
                reinterpret_cast<float4*>(p.dL_drot)[idx] = dq;
            }
        } else {
            if (p.shs != nullptr && p.dL_dsh != nullptr) {
                float* dst = stage_sh ? s_sh + threadIdx.x * stride : p.dL_dsh + (size_t)idx * row;
                for (int k = 0; k < row; ++k) dst[k] = 0.0f;
            }
            if (p.scales != nullptr && p.dL_dscale != nullptr) {
                p.dL_dscale[3 * idx] = 0.0f; p.dL_dscale[3 * idx + 1] = 0.0f; p.dL_dscale[3 * idx + 2] = 0.0f;
                reinterpret_cast<float4*>(p.dL_drot)[idx] = make_float4(0, 0, 0, 0);
            }
        }
        // outputs the caller does not want (NULL) are skipped: a pose-only tracking step needs none of them
        if (p.dL_dmean2D) {
            p.dL_dmean2D[3 * idx] = acc0.x; p.dL_dmean2D[3 * idx + 1] = acc0.y; p.dL_dmean2D[3 * idx + 2] = 0.0f;
        }
        if (p.dL_dopacity) p.dL_dopacity[idx] = acc1.y;
        if (p.dL_dcolors) {
            p.dL_dcolors[3 * idx] = acc1.z; p.dL_dcolors[3 * idx + 1] = acc1.w; p.dL_dcolors[3 * idx + 2] = acc2.x;
        }
        if (p.dL_dmean3D) {
            p.dL_dmean3D[3 * idx] = dmean.x; p.dL_dmean3D[3 * idx + 1] = dmean.y; p.dL_dmean3D[3 * idx + 2] = dmean.z;
        }
        if (p.dL_dcov3D) {
#pragma unroll
            for (int k = 0; k < 6; ++k) p.dL_dcov3D[6 * (size_t)idx + k] = dcov[k];
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (stage_sh) {                      // gradient rows: shared -> global; the block's rows are one contiguous range
        __syncthreads();
        float* dst = p.dL_dsh + (size_t)block_first * row;
        if ((row & 3) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15u) == 0) {
            const int q = row >> 2, total = nrows * q;          // float4 chunks per row / in the block
            for (int f = threadIdx.x; f < total; f += kPreBlock) {
                const int r = f / q, c = f - r * q;
                const float* src = s_sh + r * stride + 4 * c;
                reinterpret_cast<float4*>(dst)[f] = make_float4(src[0], src[1], src[2], src[3]);
            }
        } else {
            for (int r = warp; r < nrows; r += kPreBlock / 32)
                for (int k = lane; k < row; k += 32) dst[(size_t)r * row + k] = s_sh[r * stride + k];
        }
    }
    // camera-pose gradient: block-reduce the six components, one partial row per block
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        float v = tau[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) s_tau[warp][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        float s = 0.0f;
#pragma unroll
        for (int w = 0; w < kPreBlock / 32; ++w) s += s_tau[w][threadIdx.x];
        tau_partial[6 * (size_t)blockIdx.x + threadIdx.x] = s;
    }
}

void launch_preprocess_backward(const BwdParams& p, const GeomState& g, cudaStream_t s) {
    const int nblk = (p.P + kPreBlock - 1) / kPreBlock;
    // stage the SH rows through shared memory when they are longer than one coefficient (see the kernel)
    int stage_sh = (p.shs != nullptr && p.dL_dsh != nullptr && p.M > 1) ? 1 : 0;
    size_t smem = stage_sh ? sizeof(float) * (size_t)kPreBlock * sh_stage_stride(p.M) : 0;
    constexpr size_t kOptIn = 64 * 1024;   // enough for M <= 21 coefficients (degree 3 has 16)
    if (smem > 48 * 1024 && smem <= kOptIn) {
        static int raised[64] = {};        // per device: 1 = opted in to kOptIn of dynamic shared memory, -1 = refused
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev >= 0 && dev < 64 && raised[dev] == 0) {
            const cudaError_t e1 = cudaFuncSetAttribute(preprocess_backward_kernel<true>,
                                                        cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kOptIn);
            const cudaError_t e2 = cudaFuncSetAttribute(preprocess_backward_kernel<false>,
                                                        cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kOptIn);
            raised[dev] = (e1 == cudaSuccess && e2 == cudaSuccess) ? 1 : -1;
            if (raised[dev] < 0) cudaGetLastError();   // handled below, not an error of this call
        }
        if (dev < 0 || dev >= 64 || raised[dev] < 0) smem = kOptIn + 1;
    }
    if (smem > kOptIn) {
        // a padded SH layout with more rows than the staging buffer can hold (or no opt-in): every thread stores its
        // own row directly, slower but correct for any M
        stage_sh = 0;
        smem = 0;
    }
    if (p.shs != nullptr && p.D > 0)
        preprocess_backward_kernel<true><<<nblk, kPreBlock, smem, s>>>(p, g, p.dL_dtau, stage_sh);
    else
        preprocess_backward_kernel<false><<<nblk, kPreBlock, smem, s>>>(p, g, p.dL_dtau, stage_sh);
}

}  // namespace splat
