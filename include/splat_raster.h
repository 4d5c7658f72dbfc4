/* splat_raster.h — C ABI of the B200-native differentiable Gaussian-splatting rasterizer with pose gradients.
 *
 * Drop-in boundary for the path Splat-SLAM reaches through `diff_gaussian_rasterization._C`
 * (DGR = thirdparty/diff-gaussian-rasterization-w-pose in the reference tree):
 *   _C.rasterize_gaussians            DGR/rasterize_points.cu:35-122   -> CudaRasterizer::Rasterizer::forward  (DGR/cuda_rasterizer/rasterizer.h:33-59,  rasterizer_impl.cu:198-344)
 *   _C.rasterize_gaussians_backward   DGR/rasterize_points.cu:124-211  -> CudaRasterizer::Rasterizer::backward (DGR/cuda_rasterizer/rasterizer.h:61-91,  rasterizer_impl.cu:348-455)
 *   _C.mark_visible                   DGR/rasterize_points.cu:213-232  -> CudaRasterizer::Rasterizer::markVisible (DGR/cuda_rasterizer/rasterizer.h:24-31, rasterizer_impl.cu:141-153)
 *
 * Conventions
 *   - plain C: device pointers, ints/floats, a cudaStream_t passed as void*; no torch types, no exceptions.
 *   - every function returns 0 on success, a positive cudaError_t value if a launch/runtime call failed,
 *     or a negative SPLAT_ERR_* code for bad arguments. Kernel faults surface asynchronously, as in the reference
 *     (its CHECK_CUDA only synchronises when debug=true, DGR/cuda_rasterizer/auxiliary.h:166-173).
 *   - the library never allocates or frees device memory: the caller owns inputs, outputs and the three opaque
 *     state buffers (geometry / binning / image), sized by the *_bytes queries below. This replaces the
 *     reference's std::function resize callbacks (DGR/rasterize_points.cu:27-33, rasterizer_impl.cu:228-243,286-288).
 *   - all float tensors are contiguous float32, matrices are the 16 floats the reference receives
 *     (viewmatrix = W2C^T i.e. column-major W2C, projmatrix = viewmatrix @ projmatrix_raw).
 *   - optional inputs are passed as NULL exactly where the reference passes an empty tensor
 *     (shs XOR colors_precomp; (scales, rotations) XOR cov3D_precomp).
 *   - work is enqueued on `stream`; nothing synchronises the device.
 */
#ifndef SPLAT_RASTER_H_
#define SPLAT_RASTER_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPLAT_OK 0
#define SPLAT_ERR_ARG (-1)   /* NULL / inconsistent argument */
#define SPLAT_ERR_SIZE (-2)  /* negative or overflowing size */

/* ABI version of this header (bumped on any signature change). */
int splat_raster_abi_version(void);

/* Sizes of the caller-allocated opaque state buffers.
 * Replace `required<GeometryState/ImageState/BinningState>` (DGR/cuda_rasterizer/rasterizer_impl.h:65-72). */
size_t splat_raster_geom_bytes(int P);
size_t splat_raster_image_bytes(int W, int H);
size_t splat_raster_binning_bytes(int R);

/* Forward, stage 1 of 2: per-Gaussian preprocess + instance-offset scan.
 * Replaces FORWARD::preprocess + cub::DeviceScan::InclusiveSum (DGR/cuda_rasterizer/rasterizer_impl.cu:250-280,
 * forward.cu:158-258). Writes radii[P] (int32), zeroes n_touched[P] (int32), fills `geom`, and leaves the number of
 * tile instances R = num_rendered both in `geom` and, if `num_rendered_host` is not NULL, in that (pinned) host
 * int via an async copy on `stream` — the caller synchronises the stream, reads R and sizes the binning buffer,
 * which is the same single device->host hand-off the reference has (rasterizer_impl.cu:284). */
int splat_raster_forward_preprocess(
    int P, int D, int M, int W, int H,
    const float* means3D, const float* scales, float scale_modifier, const float* rotations,
    const float* opacities, const float* shs, const float* cov3D_precomp, const float* colors_precomp,
    const float* viewmatrix, const float* projmatrix, const float* campos,
    float tan_fovx, float tan_fovy, int prefiltered,
    int* radii, int* n_touched, void* geom, int* num_rendered_host, void* stream);

/* Blocks the calling host thread until the num_rendered written by its last splat_raster_forward_preprocess call on
 * the current device has reached `num_rendered_host` (which must then be pinned host memory). The preprocess call
 * hands the count over right after its first kernel, so this returns while the per-Gaussian depth sort it also
 * enqueued is still running; a plain stream synchronisation gives the same value later. The reference blocks in a
 * 4-byte cudaMemcpy after its scan (DGR/cuda_rasterizer/rasterizer_impl.cu:284). */
int splat_raster_forward_wait(void);

/* Forward, stage 2 of 2: key generation, 64-bit (tile|depth) radix sort, tile ranges, render.
 * Replaces duplicateWithKeys, cub::DeviceRadixSort::SortPairs, identifyTileRanges and FORWARD::render
 * (DGR/cuda_rasterizer/rasterizer_impl.cu:286-341, forward.cu:263-392).
 * out_color[3*H*W], out_depth[H*W], out_opacity[H*W] are fully written; n_touched[P] is accumulated into. */
int splat_raster_forward_render(
    int P, int R, int W, int H, const float* background, const float* colors_precomp, const int* radii,
    void* geom, void* binning, void* image,
    float* out_color, float* out_depth, float* out_opacity, int* n_touched, void* stream);

/* Backward. Replaces CudaRasterizer::Rasterizer::backward (DGR/cuda_rasterizer/rasterizer_impl.cu:348-455):
 * BACKWARD::render (backward.cu:563-787) then computeCov2DCUDA + preprocessCUDA (backward.cu:150-346,418-539).
 *   grad_scratch : P * 12 floats of scratch, cleared by this call (packed per-Gaussian render gradients)
 *   dL_dtau_partial : ceil(P/256) * 6 floats; the camera-pose gradient is the column sum
 *                     (rho = [0:3], theta = [3:6]) — the reference returns P x 6 and sums in Python
 *                     (DGR/diff_gaussian_rasterization/__init__.py:152-154)
 * All other outputs are fully written, zeros for culled Gaussians: dL_dmeans2D[P*3], dL_dcolors[P*3],
 * dL_dopacity[P], dL_dmeans3D[P*3], dL_dcov3D[P*6], dL_dsh[P*M*3] (if shs), dL_dscales[P*3] and
 * dL_drotations[P*4] (if scales). Any of these eight may be NULL when the caller does not need it (the reference
 * always computes all of them, DGR/diff_gaussian_rasterization/__init__.py:107-171 has no needs_input_grad check);
 * a pose-only tracking step passes NULL for all of them. Exceptions: dL_dscales/dL_drotations come as a pair, and
 * dL_dsh is required when D > 0 because the SH direction term also feeds the pose gradient (backward.cu:141-143). */
int splat_raster_backward(
    int P, int D, int M, int R, int W, int H,
    const float* background, const float* means3D, const float* shs, const float* colors_precomp,
    const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
    const float* viewmatrix, const float* projmatrix, const float* projmatrix_raw, const float* campos,
    float tan_fovx, float tan_fovy, const int* radii,
    void* geom, void* binning, void* image,
    const float* dL_dout_color, const float* dL_dout_depth,
    float* grad_scratch,
    float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D,
    float* dL_dsh, float* dL_dscales, float* dL_drotations, float* dL_dtau_partial, void* stream);

/* Backward for hosts that have no way to obtain device memory at call time. (No counterpart in the reference: its
 * backward allocates eleven tensors per call through torch, DGR/rasterize_points.cu:160-170.)
 *
 * splat_raster_backward needs `grad_scratch`, P*12 floats that somebody has to allocate. Some users of this library
 * cannot do that where the call happens: a pose-tracking loop that must keep a fixed frame time (cudaFree waits for
 * the whole device to go idle and cudaMalloc may, which also holds up the mapper process sharing the GPU), a
 * host that handed all of its device memory to fixed pools at start-up, or a thin binding in a language runtime
 * that only forwards the tensors it was given. For them this entry point keeps the scratch inside the library, in a
 * fixed device array of SPLAT_NOALLOC_MAX_GAUSSIANS * 12 floats (12 MiB per GPU, held from the moment the library
 * is loaded on that GPU until the process ends). It calls no cudaMalloc/cudaFree and no allocating library routine;
 * the caller still passes every input and output, exactly as for splat_raster_backward minus `grad_scratch`
 * (per-Gaussian outputs may be NULL when not wanted, e.g. all of them for pose-only tracking; dL_dtau_partial is
 * always required).
 *
 *   - P above SPLAT_NOALLOC_MAX_GAUSSIANS is refused with SPLAT_ERR_CAPACITY before anything is launched; use
 *     splat_raster_backward with your own scratch for larger clouds.
 *   - The first call on each device looks up where the array lives on that device (cudaGetSymbolAddress, with that
 *     device current) and remembers it. Make that first call before starting any stream capture.
 *   - All calls on one GPU share the one array. It still holds the sums of the previous call, so each call first
 *     clears the P*12 floats it is about to accumulate into. Two calls whose GPU work overlaps (different streams,
 *     or different host threads) would clear and add to the same floats at the same time and corrupt each other:
 *     keep calls on one device ordered on one stream, or serialise them yourself.
 * Results equal those of splat_raster_backward up to the summation order of float atomics. */
#define SPLAT_NOALLOC_MAX_GAUSSIANS 262144
#define SPLAT_ERR_CAPACITY (-3) /* more Gaussians than the built-in scratch of the *_noalloc entry point can hold */
int splat_raster_backward_noalloc(
    int P, int D, int M, int R, int W, int H,
    const float* background, const float* means3D, const float* shs, const float* colors_precomp,
    const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
    const float* viewmatrix, const float* projmatrix, const float* projmatrix_raw, const float* campos,
    float tan_fovx, float tan_fovy, const int* radii,
    void* geom, void* binning, void* image,
    const float* dL_dout_color, const float* dL_dout_depth,
    float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D,
    float* dL_dsh, float* dL_dscales, float* dL_drotations, float* dL_dtau_partial, void* stream);

/* present[i] = Gaussian i passes the near-plane test. Replaces Rasterizer::markVisible / checkFrustum
 * (DGR/cuda_rasterizer/rasterizer_impl.cu:54-66,141-153). `present` is P bytes (bool). */
int splat_raster_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                              unsigned char* present, void* stream);

/* Fused mapping loss, forward + backward in one pass (the step after the rasterizer on every view).
 * Replaces get_loss_mapping / get_loss_mapping_rgbd with ssim_loss = False
 * (thirdparty/monogs/utils/slam_utils.py:71-105):
 *   loss = alpha * mean|(exp(a)*image + b)*m_rgb - gt_image*m_rgb| + (1-alpha) * mean|depth*m_d - gt_depth*m_d|,
 *   m_rgb = gt_image.sum(0) > rgb_boundary_threshold, m_d = gt_depth > 0.01; `initialization` skips the exposure map.
 * image/gt_image/dL_dimage are [3*H*W], depth/gt_depth/dL_ddepth [H*W], exposure_a/b one device float each.
 * workspace: 4 + 4 * splat_raster_mapping_loss_blocks(W,H) floats, 16-byte aligned; its first 32-bit word is a ticket
 *   counter that must be 0 on entry and is 0 again when the call has finished (zero it once, then reuse it for calls
 *   that are ordered on one stream; concurrent streams need one workspace each).
 * out: 5 floats = (loss, sum l1_rgb, sum l1_depth, dL/da, dL/db), summed in a fixed order (run-to-run deterministic).
 * Gradients are for dL/dloss = 1. */
int splat_raster_mapping_loss_blocks(int W, int H);

/* The same loss evaluated in the epilogue of the forward renderer, from the pixel's colour and depth still in
 * registers: splat_raster_forward_render_loss = splat_raster_forward_render + this descriptor. dL/dimage and dL/ddepth
 * are written next to the images, `out` as for splat_raster_mapping_loss (the sums run over tiles instead of blocks
 * of 256 pixels, so `out` agrees with the stand-alone kernel to rounding, not bit for bit; the gradients are
 * identical). One launch and one pass over the images less per view (slam_utils.py:71-105 follows every render of a
 * mapping step, reference src/mapper.py:426-490). workspace: 4 + 4 * splat_raster_loss_partials(W, H) floats, ticket
 * word zero on entry as above. */
typedef struct splat_raster_loss_desc {
    const float* gt_image;      /* [3*H*W] */
    const float* gt_depth;      /* [H*W] */
    const float* exposure_a;    /* one device float; NULL with initialization */
    const float* exposure_b;
    float alpha;
    float rgb_boundary_threshold;
    int initialization;
    float* workspace;
    float* out;                 /* [5] */
    float* dL_dimage;           /* [3*H*W] */
    float* dL_ddepth;           /* [H*W] */
} splat_raster_loss_desc;
int splat_raster_loss_partials(int W, int H);
int splat_raster_forward_render_loss(
    int P, int R, int W, int H, const float* background, const float* colors_precomp, const int* radii,
    void* geom, void* binning, void* image,
    float* out_color, float* out_depth, float* out_opacity, int* n_touched,
    const splat_raster_loss_desc* loss, void* stream);

int splat_raster_mapping_loss(int W, int H, const float* image, const float* depth, const float* gt_image,
                              const float* gt_depth, const float* exposure_a, const float* exposure_b, float alpha,
                              float rgb_boundary_threshold, int initialization, float* workspace, float* out,
                              float* dL_dimage, float* dL_ddepth, void* stream);

/* Parameter activations of the Gaussian model in one launch each way. Replaces the properties
 * get_scaling / get_rotation / get_opacity (thirdparty/gaussian_splatting/scene/gaussian_model.py:53-61,76-94):
 *   scales = exp(raw_scaling) [P*3];  rotations = raw_rotation / max(||raw_rotation||, 1e-12) [P*4];
 *   opacities = sigmoid(raw_opacity) [P].
 * The backward takes the raw parameters again (nothing is saved) and the gradients w.r.t. the three outputs; any g_* may
 * be NULL (treated as zero) and any d_* may be NULL (not wanted). rotation arrays must be 16-byte aligned. */
int splat_raster_activations_forward(int P, const float* raw_scaling, const float* raw_rotation,
                                     const float* raw_opacity, float* scales, float* rotations, float* opacities,
                                     void* stream);
int splat_raster_activations_backward(int P, const float* raw_scaling, const float* raw_rotation,
                                      const float* raw_opacity, const float* g_scales, const float* g_rotations,
                                      const float* g_opacities, float* d_scaling, float* d_rotation, float* d_opacity,
                                      void* stream);

/* Fused Adam step over a list of tensors in one launch per 32 tensors (the step after backward).
 * Replaces `self.gaussians.optimizer.step()` / `self.keyframe_optimizers.step()` (src/mapper.py:557-562), i.e.
 * torch.optim.Adam as configured in thirdparty/gaussian_splatting/scene/gaussian_model.py:264-302 (one lr per group,
 * eps = 1e-15, betas 0.9/0.999, no weight decay, no amsgrad) and src/mapper.py:1111:
 *   m = lerp(m, g, 1-beta1); v = v*beta2 + (1-beta2)*g*g; p -= lr/(1-beta1^step) * m / (sqrt(v)/sqrt(1-beta2^step) + eps)
 * `step` is the 1-based count of this update. All pointers are device float arrays of n elements; `segments` itself is
 * a HOST array. zero_grads != 0 also clears each grad (for callers that accumulate into a persistent buffer instead
 * of `zero_grad(set_to_none=True)`). */
typedef struct splat_raster_adam_segment {
    float* param;
    float* grad;
    float* exp_avg;
    float* exp_avg_sq;
    long long n;
    double lr, beta1, beta2, eps;
    long long step;
} splat_raster_adam_segment;
int splat_raster_adam_step(int num_segments, const splat_raster_adam_segment* segments, int zero_grads, void* stream);

/* Densification bookkeeping for the K views of one mapping iteration, one launch per 16 views.
 * Replaces the per-view loop of src/mapper.py:519-529 + GaussianModel.add_densification_stats
 * (thirdparty/gaussian_splatting/scene/gaussian_model.py:738-742): for every view k in order and every Gaussian with
 * radii_k > 0:  max_radii2D = max(max_radii2D, radii_k);  xyz_gradient_accum += hypot of viewspace_grad_k[:2];
 * denom += 1.  `radii` / `viewspace_grad` are HOST arrays of K device pointers ([P] int, [P*grad_stride] float with
 * grad_stride >= 2, 3 for the rasterizer's dL_dmeans2D); the three statistics are [P] device floats updated in place. */
int splat_raster_densify_stats(int P, int K, const int* const* radii, const float* const* viewspace_grad,
                               int grad_stride, float* max_radii2D, float* xyz_gradient_accum, float* denom,
                               void* stream);

/* Occlusion-aware visibility of the window views (src/mapper.py:494-499,507-509): visibility[k][i] =
 * n_touched[k][i] > 0 as int64 (the reference's `.long()`), n_obs[i] = number of views that touched Gaussian i.
 * `n_touched` / `visibility` are HOST arrays of K device pointers; `visibility` (or single entries) and `n_obs` may
 * be NULL. */
int splat_raster_visibility_count(int P, int K, const int* const* n_touched, long long* const* visibility,
                                  int* n_obs, void* stream);

/* mean_dist2[i] = mean of the squared distances from point i to its three nearest neighbours.
 * Replaces simple_knn._C.distCUDA2 (thirdparty/simple-knn/spatial.cu:15-26 -> SimpleKNN::knn, simple_knn.cu:174-216),
 * which gaussian_model.py:194-200 calls at every keyframe insert. points is [P*3] device floats, mean_dist2 [P];
 * scratch is splat_raster_knn_scratch_bytes(P) bytes of device memory with no state between calls. Like the
 * reference, fewer than four points leave unfilled neighbour slots at FLT_MAX (the mean overflows to +inf), and
 * coincident points count with distance 0. No host synchronisation. */
size_t splat_raster_knn_scratch_bytes(int P);
int splat_raster_knn_mean_dist2(int P, const float* points, void* scratch, float* mean_dist2, void* stream);

/* Map deformation: move the Gaussians anchored to updated keyframes with their keyframe.
 * Replaces Mapper.update_mapping_points (src/mapper.py:154-255), for any number of keyframes at once:
 *   rigid != 0 : mean' = T [mean,1] with T = inv(inv(w2c_old) @ w2c_new); rot' = quat(T) (x) normalize(rot)
 *   rigid == 0 : additionally rescale the point along its old camera ray by s = 1 + (depth_new - depth_old) / z_cam at
 *                the pixel it projects to (clamped to the image; s = 1 where a depth is 0 or s <= 0) before moving
 *                it, and add log(s) to its three log-scales.
 * Every Gaussian's rotation is written back normalised, as the reference does (mapper.py:176,245).
 * keyframes is a HOST array; its matrices are DEVICE pointers to 16 row-major floats, depth maps DEVICE [H*W] floats
 * (unused when rigid). kf_ids [P] is the device copy of GaussianModel.unique_kfIDs (int32), intrinsics 9 device
 * floats (may be NULL when every keyframe is rigid). table: SPLAT_DEFORM_TABLE_FLOATS device floats of scratch.
 * xyz [P*3], rotation [P*4] (16-byte aligned), scaling [P*3] are updated in place. No host synchronisation. */
#define SPLAT_DEFORM_TABLE_FLOATS 2560
typedef struct splat_raster_deform_keyframe {
    const float* w2c_old;
    const float* w2c_new;
    const float* depth_new;
    const float* depth_old;
    int frame_idx;
    int rigid;
} splat_raster_deform_keyframe;
int splat_raster_deform_points(int P, int K, const splat_raster_deform_keyframe* keyframes, const int* kf_ids,
                               const float* intrinsics, int W, int H, float* table, float* xyz, float* rotation,
                               float* scaling, void* stream);

/* Gradient exchange of the view-sharded mapping window inside one NVSwitch node, in place, fp32 sum (SURVEY.md §8e;
 * the reference is single-GPU: no counterpart). Every rank holds its copy of the flat gradient buffer in memory that
 * all ranks have mapped; `buffers[r]` is rank r's copy as seen from THIS process (a peer pointer), `multicast` the
 * NVSwitch multicast address of the same allocation. One launch per rank does reduce-scatter + all-gather: rank
 * `rank` sums its 1/world slice over all copies and writes the total into every copy, so all ranks end up with
 * bit-identical sums. `count` floats, a multiple of 4; pointers 16-byte aligned; world <= 8.
 * The kernel never waits for another GPU: the CALLER orders it between two cross-rank barriers on `stream` (before:
 * every rank's gradients are complete; after: every slice has landed everywhere). */
int splat_raster_allreduce_peer(float* const* buffers, int world, int rank, size_t count, void* stream);
int splat_raster_allreduce_multimem(float* multicast, int world, int rank, size_t count, void* stream);

/* Scheduling option, off by default: run the short latency-bound kernels of a forward (everything except the two
 * render kernels) on a highest-priority side stream, ordered against the caller's stream with events, so that with
 * several views in flight on different streams they are not queued behind another view's render CTAs.
 * enable: 1 on, 0 off, < 0 query only. Returns the previous setting. No counterpart in the reference. */
int splat_raster_priority_lanes(int enable);

/* Test/diagnostic view into the opaque buffers (device pointers into the caller's buffers; layout is private).
 * Mirrors what the parity harness carves out of the reference's buffers (rasterizer_impl.cu:155-194).
 * `keys` and `point_list` (the reference's sorted 64-bit keys and point list) are not needed by the render kernels,
 * which read per-instance records; a forward only materialises them while splat_raster_keep_sort_output is on
 * (enable: 1 on, 0 off, < 0 query; returns the previous setting; library-wide, off by default). The 8 bytes per
 * instance for the keys are part of splat_raster_binning_bytes only while it is on: query the size and run the forward
 * under the same setting. */
int splat_raster_keep_sort_output(int enable);
typedef struct splat_raster_views {
    const float* depths;              /* [P] */
    const float* means2D;             /* [P*2] */
    const float* conic_opacity;       /* [P*4] */
    const float* rgb;                 /* [P*3] */
    const float* cov3D;               /* [P*6] */
    const unsigned int* tiles_touched;/* [P] */
    const unsigned long long* keys;   /* [R] sorted keys */
    const unsigned int* point_list;   /* [R] sorted Gaussian ids */
    const unsigned int* ranges;       /* [tiles*2] */
    const float* final_T;             /* [H*W] */
    const unsigned int* n_contrib;    /* [H*W] */
} splat_raster_views;
int splat_raster_debug_views(int P, int R, int W, int H, void* geom, void* binning, void* image,
                             splat_raster_views* out);

/* ---- diagnostics (measurement aids for bench.py; no counterpart in the reference, whose only aid is
 * debug=true -> cudaDeviceSynchronize after each stage, DGR/cuda_rasterizer/auxiliary.h:166-173) ---- */

/* Debug mode, the equivalent of the reference's `debug` flag (CHECK_CUDA(..., debug), DGR/cuda_rasterizer/
 * auxiliary.h:166-173: cudaDeviceSynchronize after every stage, throw with the place of the failure). While enabled,
 * every stage of splat_raster_forward_preprocess / _forward_render / _backward is followed by a wait for its stream and
 * an error check; the first failing stage ends the call, which returns the CUDA error code, and
 * splat_raster_last_failed_stage() names the stage (per host thread; "" if none; `reset` clears it).
 * `fail_stage` >= 0 (an index into splat_raster_stage_name) makes that stage REPORT a failure without faulting
 * anything on the device: a hook for testing the caller's error path. Pass -1 otherwise. Library-wide switch;
 * returns the previous setting. splat_raster_debug_sync_count counts the per-stage waits done so far. */
int splat_raster_set_debug(int enable, int fail_stage);
const char* splat_raster_last_failed_stage(int reset);
unsigned long long splat_raster_debug_sync_count(int reset);

/* Number of kernels this library has launched in this process (optionally reset). */
unsigned long long splat_raster_launch_count(int reset);

/* Per-stage timing with CUDA events recorded on the launching stream. Off by default; when on, each stage is
 * bracketed by two event records and harvested lazily. Returns the previous setting. A measurement aid for ONE host
 * thread driving ONE device: the event slots are process-wide and not synchronised (the launch counter is atomic). */
int splat_raster_profile(int enable);
int splat_raster_stage_count(void);
const char* splat_raster_stage_name(int stage);
/* ms_out[stage] = accumulated milliseconds, calls_out[stage] = number of timed executions (arrays of
 * splat_raster_stage_count() entries). Synchronises on outstanding events. */
int splat_raster_profile_read(float* ms_out, int* calls_out, int reset);

#ifdef __cplusplus
}
#endif
#endif /* SPLAT_RASTER_H_ */
