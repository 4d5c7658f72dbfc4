// Map deformation after a pose / depth update of keyframes: moves the Gaussians anchored to a
// keyframe with that keyframe. Replaces Mapper.update_mapping_points (src/mapper.py:154-255), which per keyframe runs
// ~40 small torch kernels, three boolean-mask index ops (each a nonzero() + host sync) and P-sized clones:
//   rigid:      mean' = T [mean,1],  T = inv(inv(w2c_old) @ w2c_new);   rot' = quat(T) (x) normalize(rot)
//   non-rigid:  p = w2c_old [mean,1];  (u,v) = trunc((K p).xy / (K p).z) clamped to the image;
//               s = 1 + (depth_new[v,u] - depth_old[v,u]) / p.z;  s = 1 where a depth is 0 or s <= 0;
//               mean' = T [inv(w2c_old) [s p, 1], 1];  rot' as above;  log-scale' = log-scale + log(s)
// and normalises EVERY rotation (the reference writes get_rotation back for all Gaussians, mapper.py:176,245).
// Two launches for any number of keyframes: a setup kernel derives T, inv(w2c_old) and quat(T) per keyframe (4x4
// adjugate inverses in double, on the device — nothing is read back), then one pass over the Gaussians looks each
// one's keyframe up in that table. HBM-bound: 4 B id + 40 B read + 40 B written per moved Gaussian, 4 + 32 B for the
// others (id + rotation).
#include "common.cuh"
#include "kernels.h"

namespace splat {

// per keyframe, written by the setup kernel: [0..11] w2c_old rows 0-2, [12..23] inv(w2c_old) rows 0-2,
// [24..35] T rows 0-2, [36..39] quat(T) (w,x,y,z)
constexpr int kDeformTableFloats = 40;

struct DeformKeyframes {
    const float* w2c_old[kDeformMaxKeyframes];
    const float* w2c_new[kDeformMaxKeyframes];
    const float* depth_new[kDeformMaxKeyframes];
    const float* depth_old[kDeformMaxKeyframes];
    int frame_idx[kDeformMaxKeyframes];
    int rigid[kDeformMaxKeyframes];
    int count;
};

__device__ bool invert4x4(const double* m, double* inv) {
    // adjugate / determinant, row-major
    inv[0] = m[5] * m[10] * m[15] - m[5] * m[11] * m[14] - m[9] * m[6] * m[15] + m[9] * m[7] * m[14] + m[13] * m[6] * m[11] - m[13] * m[7] * m[10];
    inv[4] = -m[4] * m[10] * m[15] + m[4] * m[11] * m[14] + m[8] * m[6] * m[15] - m[8] * m[7] * m[14] - m[12] * m[6] * m[11] + m[12] * m[7] * m[10];
    inv[8] = m[4] * m[9] * m[15] - m[4] * m[11] * m[13] - m[8] * m[5] * m[15] + m[8] * m[7] * m[13] + m[12] * m[5] * m[11] - m[12] * m[7] * m[9];
    inv[12] = -m[4] * m[9] * m[14] + m[4] * m[10] * m[13] + m[8] * m[5] * m[14] - m[8] * m[6] * m[13] - m[12] * m[5] * m[10] + m[12] * m[6] * m[9];
    inv[1] = -m[1] * m[10] * m[15] + m[1] * m[11] * m[14] + m[9] * m[2] * m[15] - m[9] * m[3] * m[14] - m[13] * m[2] * m[11] + m[13] * m[3] * m[10];
    inv[5] = m[0] * m[10] * m[15] - m[0] * m[11] * m[14] - m[8] * m[2] * m[15] + m[8] * m[3] * m[14] + m[12] * m[2] * m[11] - m[12] * m[3] * m[10];
    inv[9] = -m[0] * m[9] * m[15] + m[0] * m[11] * m[13] + m[8] * m[1] * m[15] - m[8] * m[3] * m[13] - m[12] * m[1] * m[11] + m[12] * m[3] * m[9];
    inv[13] = m[0] * m[9] * m[14] - m[0] * m[10] * m[13] - m[8] * m[1] * m[14] + m[8] * m[2] * m[13] + m[12] * m[1] * m[10] - m[12] * m[2] * m[9];
    inv[2] = m[1] * m[6] * m[15] - m[1] * m[7] * m[14] - m[5] * m[2] * m[15] + m[5] * m[3] * m[14] + m[13] * m[2] * m[7] - m[13] * m[3] * m[6];
    inv[6] = -m[0] * m[6] * m[15] + m[0] * m[7] * m[14] + m[4] * m[2] * m[15] - m[4] * m[3] * m[14] - m[12] * m[2] * m[7] + m[12] * m[3] * m[6];
    inv[10] = m[0] * m[5] * m[15] - m[0] * m[7] * m[13] - m[4] * m[1] * m[15] + m[4] * m[3] * m[13] + m[12] * m[1] * m[7] - m[12] * m[3] * m[5];
    inv[14] = -m[0] * m[5] * m[14] + m[0] * m[6] * m[13] + m[4] * m[1] * m[14] - m[4] * m[2] * m[13] - m[12] * m[1] * m[6] + m[12] * m[2] * m[5];
    inv[3] = -m[1] * m[6] * m[11] + m[1] * m[7] * m[10] + m[5] * m[2] * m[11] - m[5] * m[3] * m[10] - m[9] * m[2] * m[7] + m[9] * m[3] * m[6];
    inv[7] = m[0] * m[6] * m[11] - m[0] * m[7] * m[10] - m[4] * m[2] * m[11] + m[4] * m[3] * m[10] + m[8] * m[2] * m[7] - m[8] * m[3] * m[6];
    inv[11] = -m[0] * m[5] * m[11] + m[0] * m[7] * m[9] + m[4] * m[1] * m[11] - m[4] * m[3] * m[9] - m[8] * m[1] * m[7] + m[8] * m[3] * m[5];
    inv[15] = m[0] * m[5] * m[10] - m[0] * m[6] * m[9] - m[4] * m[1] * m[10] + m[4] * m[2] * m[9] + m[8] * m[1] * m[6] - m[8] * m[2] * m[5];
    const double det = m[0] * inv[0] + m[1] * inv[4] + m[2] * inv[8] + m[3] * inv[12];
    if (det == 0.0) return false;
    const double r = 1.0 / det;
    for (int i = 0; i < 16; ++i) inv[i] *= r;
    return true;
}

__global__ void deform_setup_kernel(const __grid_constant__ DeformKeyframes kfs, float* __restrict__ table) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= kfs.count) return;
    double a[16], b[16], ai[16], bi[16], t[16];
    for (int i = 0; i < 16; ++i) { a[i] = kfs.w2c_old[k][i]; b[i] = kfs.w2c_new[k][i]; }
    float* out = table + (size_t)k * kDeformTableFloats;
    const bool ok = invert4x4(a, ai) && invert4x4(b, bi);
    // T = inv(inv(A) B) = inv(B) A                                   (mapper.py:166,229)
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c) {
            double s = 0.0;
            for (int j = 0; j < 4; ++j) s += bi[4 * r + j] * a[4 * j + c];
            t[4 * r + c] = ok ? s : (r == c ? 1.0 : 0.0);             // singular input: leave the points where they are
        }
    for (int i = 0; i < 12; ++i) {
        out[i] = (float)a[i];
        out[12 + i] = ok ? (float)ai[i] : (float)(i % 5 == 0);
        out[24 + i] = (float)t[i];
    }
    // rotation_matrix_to_quaternion (general_utils.py:138-162) on the float32 matrix, as the reference feeds it
    const float r00 = out[24], r01 = out[25], r02 = out[26], r10 = out[28], r11 = out[29], r12 = out[30], r20 = out[32],
                r21 = out[33], r22 = out[34];
    float qw = sqrtf(fmaxf(0.0f, 1.0f + r00 + r11 + r22)) / 2.0f;
    float qx = sqrtf(fmaxf(0.0f, 1.0f + r00 - r11 - r22)) / 2.0f;
    float qy = sqrtf(fmaxf(0.0f, 1.0f - r00 + r11 - r22)) / 2.0f;
    float qz = sqrtf(fmaxf(0.0f, 1.0f - r00 - r11 + r22)) / 2.0f;
    auto sgn = [](float v) { return v > 0.0f ? 1.0f : (v < 0.0f ? -1.0f : 0.0f); };
    qx *= sgn(qx * (r21 - r12));
    qy *= sgn(qy * (r02 - r20));
    qz *= sgn(qz * (r10 - r01));
    out[36] = qw; out[37] = qx; out[38] = qy; out[39] = qz;
}

__device__ __forceinline__ float3 affine_rows(const float* m, float3 p) {     // rows 0-2 of a 4x4 times [p,1]
    return make_float3(fmaf(m[2], p.z, fmaf(m[1], p.y, fmaf(m[0], p.x, m[3]))),
                       fmaf(m[6], p.z, fmaf(m[5], p.y, fmaf(m[4], p.x, m[7]))),
                       fmaf(m[10], p.z, fmaf(m[9], p.y, fmaf(m[8], p.x, m[11]))));
}

__global__ void __launch_bounds__(256)
deform_points_kernel(int P, const __grid_constant__ DeformKeyframes kfs, const float* __restrict__ table,
                     const int* __restrict__ kf_ids, const float* __restrict__ intrinsics, int W, int H,
                     int normalize_all, float* __restrict__ xyz, float* __restrict__ rotation,
                     float* __restrict__ scaling) {
    __shared__ float s_tab[kDeformMaxKeyframes * kDeformTableFloats];
    __shared__ float s_K[9];
    for (int i = threadIdx.x; i < kfs.count * kDeformTableFloats; i += blockDim.x) s_tab[i] = table[i];
    if (threadIdx.x < 9) s_K[threadIdx.x] = intrinsics ? intrinsics[threadIdx.x] : 0.0f;
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const int id = kf_ids[i];
    int k = -1;
    for (int j = 0; j < kfs.count; ++j)
        if (kfs.frame_idx[j] == id) { k = j; break; }
    if (k < 0 && !normalize_all) return;
    // every rotation is written back normalised: x / max(||x||, 1e-12)     (torch.nn.functional.normalize)
    float4 q = reinterpret_cast<float4*>(rotation)[i];
    const float n = fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);
    q = make_float4(q.x / n, q.y / n, q.z / n, q.w / n);
    if (k < 0) {
        reinterpret_cast<float4*>(rotation)[i] = q;
        return;
    }
    const float* tab = s_tab + k * kDeformTableFloats;
    float3 mean = make_float3(xyz[3 * (size_t)i], xyz[3 * (size_t)i + 1], xyz[3 * (size_t)i + 2]);
    if (!kfs.rigid[k]) {
        float3 pc = affine_rows(tab, mean);                                        // old camera frame
        const float px = fmaf(s_K[2], pc.z, fmaf(s_K[1], pc.y, s_K[0] * pc.x));
        const float py = fmaf(s_K[5], pc.z, fmaf(s_K[4], pc.y, s_K[3] * pc.x));
        const float pz = fmaf(s_K[8], pc.z, fmaf(s_K[7], pc.y, s_K[6] * pc.x));
        // .long() truncates toward zero; NaN / inf (point on the camera plane) land on a border pixel
        const float fu = px / pz, fv = py / pz;
        int u = (fu >= (float)(W - 1)) ? W - 1 : ((fu > 0.0f) ? (int)fu : 0);
        int v = (fv >= (float)(H - 1)) ? H - 1 : ((fv > 0.0f) ? (int)fv : 0);
        const float dn = kfs.depth_new[k][(size_t)v * W + u], dold = kfs.depth_old[k][(size_t)v * W + u];
        float s = 1.0f + (1.0f / pc.z) * (dn - dold);
        if (dn == 0.0f || dold == 0.0f) s = 1.0f;
        if (s <= 0.0f) s = 1.0f;
        pc = make_float3(s * pc.x, s * pc.y, s * pc.z);
        mean = affine_rows(tab + 12, pc);                                          // back to the world
        const float ls = logf(s);
        scaling[3 * (size_t)i] += ls;
        scaling[3 * (size_t)i + 1] += ls;
        scaling[3 * (size_t)i + 2] += ls;
    }
    mean = affine_rows(tab + 24, mean);
    xyz[3 * (size_t)i] = mean.x;
    xyz[3 * (size_t)i + 1] = mean.y;
    xyz[3 * (size_t)i + 2] = mean.z;
    // quaternion_multiply(quat(T), q)                                         (general_utils.py:164-176)
    const float w1 = tab[36], x1 = tab[37], y1 = tab[38], z1 = tab[39];
    const float w2 = q.x, x2 = q.y, y2 = q.z, z2 = q.w;
    reinterpret_cast<float4*>(rotation)[i] = make_float4(w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                                                         w1 * y2 + y1 * w2 + z1 * x2 - x1 * z2, w1 * z2 + z1 * w2 + x1 * y2 - y1 * x2);
}

int launch_deform_points(int P, int K, const DeformKeyframe* kfs, const int* kf_ids, const float* intrinsics, int W, int H,
                         float* table, float* xyz, float* rotation, float* scaling, cudaStream_t s) {
    int launches = 0;
    for (int at = 0; at < K; at += kDeformMaxKeyframes) {
        DeformKeyframes b;
        b.count = (K - at < kDeformMaxKeyframes) ? K - at : kDeformMaxKeyframes;
        for (int k = 0; k < b.count; ++k) {
            const DeformKeyframe& in = kfs[at + k];
            b.w2c_old[k] = in.w2c_old; b.w2c_new[k] = in.w2c_new;
            b.depth_new[k] = in.depth_new; b.depth_old[k] = in.depth_old;
            b.frame_idx[k] = in.frame_idx; b.rigid[k] = in.rigid;
        }
        deform_setup_kernel<<<1, kDeformMaxKeyframes, 0, s>>>(b, table);
        // the rotations of ALL Gaussians are normalised once, by the first chunk
        deform_points_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, b, table, kf_ids, intrinsics, W, H, at == 0 ? 1 : 0, xyz,
                                                            rotation, scaling);
        launches += 2;
    }
    return launches;
}

}  // namespace splat
