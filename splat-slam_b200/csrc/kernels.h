// Internal launch interface between api.cu and the kernel translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "state.h"
#include "loss.cuh"

namespace splat {

struct FwdParams {
    int P, D, M, W, H, grid_x, grid_y, prefiltered;
    float scale_modifier, tanfovx, tanfovy, focal_x, focal_y;
    const float *means3D, *scales, *rotations, *opacities, *shs, *cov3D_precomp, *colors_precomp;
    const float *viewmatrix, *projmatrix, *campos;
    int* radii;
    int* n_touched;
};

struct BwdParams {
    int P, D, M, R, W, H, grid_x, grid_y;
    float scale_modifier, tanfovx, tanfovy, focal_x, focal_y;
    const float *bg, *means3D, *shs, *colors_precomp, *scales, *rotations, *cov3D_precomp;
    const float *viewmatrix, *projmatrix, *projmatrix_raw, *campos;
    const int* radii;
    const float *dL_dpix, *dL_dpix_depth;
    float* grad_accum;  // [P * SPLAT_GRAD_FLOATS], zeroed by the caller-visible entry point
    float *dL_dmean2D, *dL_dcolors, *dL_dopacity, *dL_dmean3D, *dL_dcov3D, *dL_dsh, *dL_dscale, *dL_drot, *dL_dtau;
};

// preprocess.cu
void launch_preprocess_forward(const FwdParams& p, const GeomState& g, cudaStream_t s);
void launch_mark_visible(int P, const float* means3D, const float* viewmatrix, bool* present, cudaStream_t s);
// preprocess_bwd.cu
void launch_preprocess_backward(const BwdParams& p, const GeomState& g, cudaStream_t s);
// binning.cu
void launch_depth_sort_and_offsets(int P, const GeomState& g, cudaStream_t s);
void launch_duplicate_with_keys(int P, int grid_x, int grid_y, const int* radii, const GeomState& g,
                                uint32_t* tkeys_out, uint32_t* vals_out, uint2* ranges, cudaStream_t s);
void launch_reset_ranges(int tiles, uint2* ranges, cudaStream_t s);
// radix_sort.cu
int tile_sort_passes(int grid_x, int grid_y);
// where the last pass of the instance sort puts its result (see radix_onesweep_kernel<true>)
struct RecordSink {
    const float4* gauss_rec = nullptr;   // [3P] per-Gaussian render records
    float4* records = nullptr;           // [3R] per-instance records in sorted order
    uint2* ranges = nullptr;             // [tiles] preset to 0xFFFFFFFF (launch_reset_ranges)
    uint64_t* keys64 = nullptr;          // [R] optional: the reference's sorted 64-bit keys, for inspection
    uint32_t* point_list = nullptr;      // [R] optional: sorted Gaussian ids, for inspection
};
void launch_sort_pairs(uint32_t* k0, uint32_t* v0, uint32_t* k1, uint32_t* v1, int n, int passes, uint32_t* scratch,
                       cudaStream_t s, const RecordSink* sink = nullptr);
// mapping_loss.cu
int mapping_loss_blocks(int N);
void launch_mapping_loss(int N, const float* image, const float* depth, const LossDesc& loss, cudaStream_t s);
// adam.cu
constexpr int kAdamMaxSegments = 32;      // tensors per launch (kernel-parameter space)
struct AdamSegment {
    float* p; float* g; float* m; float* v;
    long long n;
    float w1, beta2, w2, bc2_sqrt, eps, neg_step;
    int vec4;
};
int launch_adam_step(int count, const AdamSegment* segments, int zero_grads, cudaStream_t stream);   // -> launches
// densify.cu
constexpr int kDensifyMaxViews = 16;      // views per launch (kernel-parameter space)
int launch_densify_stats(int P, int K, const int* const* radii, const float* const* grad, int grad_stride,
                         float* max_radii2D, float* xyz_gradient_accum, float* denom, cudaStream_t s);   // -> launches
int launch_visibility_count(int P, int K, const int* const* n_touched, long long* const* visibility, int* n_obs,
                            cudaStream_t s);                                                             // -> launches
// knn.cu
size_t knn_scratch_bytes(size_t P);
int launch_knn_mean_dist2(int P, const float* points, void* scratch, float* mean_dist2, cudaStream_t s);   // -> launches
// deform.cu
constexpr int kDeformMaxKeyframes = 64;   // keyframes per launch pair (kernel-parameter space)
struct DeformKeyframe {
    const float* w2c_old; const float* w2c_new; const float* depth_new; const float* depth_old;
    int frame_idx; int rigid;
};
int launch_deform_points(int P, int K, const DeformKeyframe* kfs, const int* kf_ids, const float* intrinsics, int W, int H,
                         float* table, float* xyz, float* rotation, float* scaling, cudaStream_t s);     // -> launches
// activations.cu
void launch_activations_forward(int P, const float* raw_scaling, const float* raw_rotation, const float* raw_opacity,
                                float* scales, float* rotations, float* opacities, cudaStream_t s);
void launch_activations_backward(int P, const float* raw_scaling, const float* raw_rotation, const float* raw_opacity,
                                 const float* g_scales, const float* g_rotations, const float* g_opacities,
                                 float* d_scaling, float* d_rotation, float* d_opacity, cudaStream_t s);
// allreduce.cu
constexpr int kMaxPeerRanks = 8;          // GPUs of one NVSwitch node
int launch_allreduce_peer(float* const* buffers, int world, int rank, size_t count, cudaStream_t s);      // -> launches, < 0 unsupported
int launch_allreduce_multimem(float* multicast, int world, int rank, size_t count, cudaStream_t s);       // -> launches
// render.cu
// warp_mask: the forward marks, in bits 24..31 of every record's id word, which warps of the tile blended the instance,
// and the backward lists instances by that mask (needs ids < 2^24; 0 = footprint test in both)
constexpr int kWarpMaskMaxGaussians = 1 << 24;
void launch_render_forward(int W, int H, int grid_x, int grid_y, uint2* ranges, float* records, int warp_mask,
                           const float* bg, float* final_T, uint32_t* n_contrib, float* out_color, float* out_depth,
                           float* out_opacity, int* n_touched, const uint32_t* sort_error, const LossDesc* loss,
                           cudaStream_t s);
void launch_render_backward(int W, int H, int grid_x, int grid_y, const uint2* ranges, const float* records,
                            int warp_mask, const float* bg, const float* final_T, const uint32_t* n_contrib, const float* dL_dpix,
                            const float* dL_dpix_depth, float* grad_accum, cudaStream_t s);

}  // namespace splat
