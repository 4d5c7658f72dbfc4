// Per-Gaussian forward stage: near cull, 3D covariance, EWA 2D covariance, conic, radius, tile rectangle,
// SH -> RGB, and the per-block tile count that feeds the offset scan.
// Replaces DGR/cuda_rasterizer/forward.cu:158-258 (preprocessCUDA) and rasterizer_impl.cu:54-66 (checkFrustum).
#include <cstdio>
#include "common.cuh"
#include "kernels.h"
#include "project.cuh"

namespace splat {

// SH basis constants (DGR/cuda_rasterizer/auxiliary.h:20-39)
__device__ const float kSH_C0 = 0.28209479177387814f;
__device__ const float kSH_C1 = 0.4886025119029199f;
__device__ const float kSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                    -1.0925484305920792f, 0.5462742152960396f};
__device__ const float kSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                    0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                    -0.5900435899266435f};

// SH -> RGB with +0.5 and clamp at 0 (DGR forward.cu:22-73). Colours only need 1e-5, so plain float code.
// A row is 12*M bytes; when it is 16-byte aligned (M a multiple of 4, i.e. the SH degree 1 and 3 layouts) it is read
// with 128-bit loads, a degree at a time — a quarter of the load instructions and of the L1 sector traffic of 4-byte
// loads, which is what bounded this kernel at SH degree 3.
template <bool kHigherDegrees>
__device__ __forceinline__ float3 sh_to_rgb(int deg, int M, const float* __restrict__ sh, float3 pos, float3 campos,
                                            uint8_t& clamp_bits) {
    float3 dir = make_float3(pos.x - campos.x, pos.y - campos.y, pos.z - campos.z);
    const float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
    dir.x /= len; dir.y /= len; dir.z /= len;
    const bool vec = ((M & 3) == 0) && ((reinterpret_cast<uintptr_t>(sh) & 15u) == 0);
    const float4* sh4 = reinterpret_cast<const float4*>(sh);
    float f[48];
#define SH_LOAD(first4, count4)                                                        \
    _Pragma("unroll") for (int i__ = (first4); i__ < (first4) + (count4); ++i__) {     \
        if (vec) {                                                                     \
            const float4 v__ = __ldg(sh4 + i__);                                       \
            f[4 * i__] = v__.x; f[4 * i__ + 1] = v__.y; f[4 * i__ + 2] = v__.z; f[4 * i__ + 3] = v__.w; \
        } else {                                                                       \
            _Pragma("unroll") for (int j__ = 0; j__ < 4; ++j__)                        \
                f[4 * i__ + j__] = (4 * i__ + j__ < 3 * M) ? sh[4 * i__ + j__] : 0.0f; \
        }                                                                              \
    }
#define SH_ACC(w, k) { const float w__ = (w); r.x += w__ * f[3 * (k)]; r.y += w__ * f[3 * (k) + 1]; r.z += w__ * f[3 * (k) + 2]; }
    float3 r;
    if (!kHigherDegrees || deg == 0) {   // 3 floats: not worth a 16-byte load, and M may be 1
        r = make_float3(kSH_C0 * sh[0], kSH_C0 * sh[1], kSH_C0 * sh[2]);
    } else {
        SH_LOAD(0, 3)                    // coefficients 0..3
        r = make_float3(kSH_C0 * f[0], kSH_C0 * f[1], kSH_C0 * f[2]);
        const float x = dir.x, y = dir.y, z = dir.z;
        SH_ACC(-kSH_C1 * y, 1) SH_ACC(kSH_C1 * z, 2) SH_ACC(-kSH_C1 * x, 3)
        if (deg > 1) {
            SH_LOAD(3, 4)                // floats 12..27: coefficients 4..8 (+ the first float of 9)
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            SH_ACC(kSH_C2[0] * xy, 4) SH_ACC(kSH_C2[1] * yz, 5) SH_ACC(kSH_C2[2] * (2.0f * zz - xx - yy), 6)
            SH_ACC(kSH_C2[3] * xz, 7) SH_ACC(kSH_C2[4] * (xx - yy), 8)
            if (deg > 2) {
                SH_LOAD(7, 5)            // floats 28..47: coefficients 9..15
                SH_ACC(kSH_C3[0] * y * (3.0f * xx - yy), 9) SH_ACC(kSH_C3[1] * xy * z, 10)
                SH_ACC(kSH_C3[2] * y * (4.0f * zz - xx - yy), 11)
                SH_ACC(kSH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy), 12)
                SH_ACC(kSH_C3[4] * x * (4.0f * zz - xx - yy), 13) SH_ACC(kSH_C3[5] * z * (xx - yy), 14)
                SH_ACC(kSH_C3[6] * x * (xx - 3.0f * yy), 15)
            }
        }
    }
#undef SH_ACC
#undef SH_LOAD
    r.x += 0.5f; r.y += 0.5f; r.z += 0.5f;
    clamp_bits = (r.x < 0.0f ? 1 : 0) | (r.y < 0.0f ? 2 : 0) | (r.z < 0.0f ? 4 : 0);
    return make_float3(fmaxf(r.x, 0.0f), fmaxf(r.y, 0.0f), fmaxf(r.z, 0.0f));
}

// kHigherDegrees = false is the SH degree 0 build (Splat-SLAM's own setting): without the higher-degree code it needs
// fewer registers and keeps its occupancy.
template <bool kHigherDegrees>
__global__ void __launch_bounds__(kPreBlock)
preprocess_forward_kernel(int P, int D, int M,
                          const float* __restrict__ means3D, const float* __restrict__ scales, float scale_modifier,
                          const float* __restrict__ rotations, const float* __restrict__ opacities,
                          const float* __restrict__ shs, const float* __restrict__ cov3D_precomp,
                          const float* __restrict__ colors_precomp, const float* __restrict__ viewmatrix,
                          const float* __restrict__ projmatrix, const float* __restrict__ campos,
                          int W, int H, float tanfovx, float tanfovy, float focal_x, float focal_y,
                          int grid_x, int grid_y, int prefiltered,
                          int* __restrict__ radii, int* __restrict__ n_touched, GeomState g) {
    __shared__ float sV[16], sPm[16];
    if (threadIdx.x < 16) { sV[threadIdx.x] = viewmatrix[threadIdx.x]; sPm[threadIdx.x] = projmatrix[threadIdx.x]; }
    __syncthreads();
    const int idx = blockIdx.x * kPreBlock + threadIdx.x;
    uint32_t tiles = 0;
    if (idx < P) {
        int radius_out = 0;
        float tz_out = 0.0f;
        do {
            const float px = means3D[3 * idx], py = means3D[3 * idx + 1], pz = means3D[3 * idx + 2];
            // in_frustum (auxiliary.h:139-164): cull on view-space z only
            const float tz = affine_row_ref(px, py, pz, sV[2], sV[6], sV[10], sV[14]);
            if (tz <= SPLAT_NEAR) {
                if (prefiltered) { printf("Point is filtered although prefiltered is set. This shouldn't happen!"); __trap(); }
                break;
            }
            const float tx = affine_row_ref(px, py, pz, sV[0], sV[4], sV[8], sV[12]);
            const float ty = affine_row_ref(px, py, pz, sV[1], sV[5], sV[9], sV[13]);
            const float hx = affine_row_ref(px, py, pz, sPm[0], sPm[4], sPm[8], sPm[12]);
            const float hy = affine_row_ref(px, py, pz, sPm[1], sPm[5], sPm[9], sPm[13]);
            const float hw = affine_row_ref(px, py, pz, sPm[3], sPm[7], sPm[11], sPm[15]);
            const float p_w = __frcp_rn(fadd(hw, 0.0000001f));
            const float projx = fmul(hx, p_w), projy = fmul(hy, p_w);
            float cov[6];
            if (cov3D_precomp != nullptr) {
#pragma unroll
                for (int k = 0; k < 6; ++k) cov[k] = cov3D_precomp[6 * idx + k];
            } else {
                const float4 q = reinterpret_cast<const float4*>(rotations)[idx];
                cov3d_ref(scales[3 * idx], scales[3 * idx + 1], scales[3 * idx + 2], scale_modifier, q, cov);
#pragma unroll
                for (int k = 0; k < 6; ++k) g.cov3D[6 * idx + k] = cov[k];
            }
            const EwaRef ewa = ewa_ref(tx, ty, tz, focal_x, focal_y, tanfovx, tanfovy, cov, sV);
            const float3 abc = make_float3(ewa.a, ewa.b, ewa.c);
            // det = a*c - b*b as fma(a, c, -(b*b)); conic = inverse (forward.cu:216-221)
            const float det = ffma(abc.x, abc.z, -fmul(abc.y, abc.y));
            if (det == 0.0f) break;
            const float det_inv = __frcp_rn(det);
            const float4 con_o = make_float4(fmul(abc.z, det_inv), fmul(-abc.y, det_inv), fmul(abc.x, det_inv),
                                             opacities[idx]);
            // radius from the larger eigenvalue (forward.cu:227-231)
            const float mid = fmul(fadd(abc.x, abc.z), 0.5f);
            const float sq = __fsqrt_rn(fmaxf(ffma(mid, mid, -det), 0.1f));
            const float lam = fmaxf(fadd(mid, sq), fsub(mid, sq));
            const int my_radius = __float2int_ru(fmul(__fsqrt_rn(lam), 3.0f));
            const float2 pim = make_float2(ndc2pix_ref(projx, W), ndc2pix_ref(projy, H));
            uint32_t x0, y0, x1, y1;
            get_rect_ref(pim.x, pim.y, my_radius, grid_x, grid_y, x0, y0, x1, y1);
            const uint32_t area = (x1 - x0) * (y1 - y0);
            if (area == 0) break;
            float3 c;
            if (colors_precomp == nullptr) {
                uint8_t cb;
                c = sh_to_rgb<kHigherDegrees>(D, M, shs + (size_t)idx * M * 3, make_float3(px, py, pz),
                              make_float3(campos[0], campos[1], campos[2]), cb);
                g.rgb[3 * idx] = c.x; g.rgb[3 * idx + 1] = c.y; g.rgb[3 * idx + 2] = c.z;
                g.clamped[idx] = cb;
            } else {
                c = make_float3(colors_precomp[3 * idx], colors_precomp[3 * idx + 1], colors_precomp[3 * idx + 2]);
            }
            g.depths[idx] = tz;
            g.means2D[idx] = pim;
            g.conic_opacity[idx] = con_o;
            // everything the render kernels need for this Gaussian, as one 48-byte record
            // {x, y, id, footprint box | conic.x, conic.y, conic.z, opacity | r, g, b, depth}: the first quarter is all the
            // per-warp binning reads, the first two decide whether a pixel blends, the last is only read if one does
            g.rec[3 * (size_t)idx + 0] = make_float4(
                pim.x, pim.y, __int_as_float(idx),
                __uint_as_float(footprint_half_extents(con_o.x, con_o.y, con_o.z, con_o.w)));
            g.rec[3 * (size_t)idx + 1] = con_o;
            g.rec[3 * (size_t)idx + 2] = make_float4(c.x, c.y, c.z, tz);
            radius_out = my_radius;
            tiles = area;
            tz_out = tz;
        } while (false);
        radii[idx] = radius_out;
        n_touched[idx] = 0;
        g.tiles_touched[idx] = tiles;
        // input of the per-Gaussian depth sort (radix_sort.cu): culled Gaussians go to the end
        g.dkey[idx] = tiles ? __float_as_uint(tz_out) : 0xFFFFFFFFu;
        g.order[idx] = (uint32_t)idx;
    }
    // R = total instances, available as soon as this kernel ends: the host reads it back while the depth sort runs
    // instead of after it (the reference learns it only after its scan, rasterizer_impl.cu:280-284)
    __shared__ uint32_t s_tiles[kPreBlock / 32];
    uint32_t sum = tiles;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if ((threadIdx.x & 31) == 0) s_tiles[threadIdx.x >> 5] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t total = 0;
#pragma unroll
        for (int w = 0; w < kPreBlock / 32; ++w) total += s_tiles[w];
        if (total) atomicAdd(g.num_rendered + 1, total);
    }
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ viewmatrix,
                                    bool* __restrict__ present) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float tz = affine_row_ref(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2], viewmatrix[2],
                                    viewmatrix[6], viewmatrix[10], viewmatrix[14]);
    present[idx] = !(tz <= SPLAT_NEAR);
}

void launch_preprocess_forward(const FwdParams& p, const GeomState& g, cudaStream_t s) {
    const int nblk = (p.P + kPreBlock - 1) / kPreBlock;
    auto kernel = (p.shs != nullptr && p.D > 0) ? preprocess_forward_kernel<true> : preprocess_forward_kernel<false>;
    kernel<<<nblk, kPreBlock, 0, s>>>(
        p.P, p.D, p.M, p.means3D, p.scales, p.scale_modifier, p.rotations, p.opacities, p.shs, p.cov3D_precomp,
        p.colors_precomp, p.viewmatrix, p.projmatrix, p.campos, p.W, p.H, p.tanfovx, p.tanfovy, p.focal_x, p.focal_y,
        p.grid_x, p.grid_y, p.prefiltered, p.radii, p.n_touched, g);
}

void launch_mark_visible(int P, const float* means3D, const float* viewmatrix, bool* present, cudaStream_t s) {
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, viewmatrix, present);
}

}  // namespace splat
