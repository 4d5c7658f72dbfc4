// extern "C" entry points declared in include/splat_raster.h: argument checks, state carving, launch sequences.
#include "../../include/splat_raster.h"
#include "common.cuh"
#include "kernels.h"

#include <atomic>
#include <cmath>
#include <cstdlib>
#include <map>
#include <mutex>
#include <utility>
#include <vector>

using namespace splat;

namespace {
inline int grid_dim(int v) { return (v + SPLAT_TILE - 1) / SPLAT_TILE; }

// One event per host thread and device marks the point where num_rendered has reached pinned host memory
// (splat_raster_forward_preprocess records it, splat_raster_forward_wait waits on it).
constexpr int kMaxDevices = 64;
thread_local cudaEvent_t t_handoff[kMaxDevices];
thread_local bool t_handoff_recorded[kMaxDevices];
cudaEvent_t handoff_event() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return nullptr;
    if (!t_handoff_recorded[dev]) {
        if (cudaEventCreateWithFlags(&t_handoff[dev], cudaEventDisableTiming) != cudaSuccess) return nullptr;
        t_handoff_recorded[dev] = true;
    }
    return t_handoff[dev];
}

// Optional high-priority side lane for the short, latency-bound kernels of a forward (preprocess, the sort passes, key
// emission, record gather). With several views in flight on different streams the two render kernels keep every SM
// busy for hundreds of microseconds; CTAs of a higher-priority stream are scheduled first whenever a slot frees up, so
// a view's ~20 small kernels slip through another view's render instead of queueing behind its remaining CTAs.
// Ordering with the caller's stream is kept by an event pair (fork at entry, join before the render kernel / return).
struct Lane {
    cudaStream_t hp = nullptr;
    cudaEvent_t fork = nullptr, join = nullptr;
};
int g_priority_lanes = 0;
bool profiling_enabled();
std::mutex g_lane_mutex;
std::map<std::pair<int, cudaStream_t>, Lane> g_lanes;

cudaStream_t lane_begin(cudaStream_t s) {
    if (!g_priority_lanes || profiling_enabled()) return s;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return s;
    Lane lane;
    {
        std::lock_guard<std::mutex> lock(g_lane_mutex);
        Lane& l = g_lanes[std::make_pair(dev, s)];
        if (!l.hp) {
            int lo = 0, hi = 0;
            cudaDeviceGetStreamPriorityRange(&lo, &hi);       // hi is the numerically lowest = greatest priority
            if (cudaStreamCreateWithPriority(&l.hp, cudaStreamNonBlocking, hi) != cudaSuccess ||
                cudaEventCreateWithFlags(&l.fork, cudaEventDisableTiming) != cudaSuccess ||
                cudaEventCreateWithFlags(&l.join, cudaEventDisableTiming) != cudaSuccess) {
                l = Lane();
                return s;
            }
        }
        lane = l;
    }
    if (cudaEventRecord(lane.fork, s) != cudaSuccess || cudaStreamWaitEvent(lane.hp, lane.fork, 0) != cudaSuccess) return s;
    return lane.hp;
}

void lane_end(cudaStream_t s, cudaStream_t used) {
    if (used == s) return;
    int dev = 0;
    cudaGetDevice(&dev);
    Lane lane;
    {
        std::lock_guard<std::mutex> lock(g_lane_mutex);
        lane = g_lanes[std::make_pair(dev, s)];
    }
    cudaEventRecord(lane.join, used);
    cudaStreamWaitEvent(s, lane.join, 0);
}

// ---- diagnostics: kernel-launch counter and optional per-stage CUDA-event timing ------------------------
enum Stage { ST_PREPROCESS = 0, ST_SCAN, ST_DUP, ST_SORT, ST_RENDER_FWD, ST_ZERO, ST_RENDER_BWD,
             ST_PREPROCESS_BWD, ST_COUNT };
const char* kStageNames[ST_COUNT] = {"preprocess_forward", "depth_sort_and_offsets", "duplicate_with_keys",
                                     "tile_sort_records_ranges", "render_forward", "zero_grad_scratch",
                                     "render_backward", "preprocess_backward"};
int g_keep_sort_output = 0;
// The forward-to-backward warp mask lives in bits 24..31 of the records' id words: clouds of 2^24 Gaussians or more
// run without it (footprint test in the backward). SPLAT_NO_WARP_MASK=1 forces that path for any size (tests).
int use_warp_mask(int P) {
    static const int off = [] { const char* e = getenv("SPLAT_NO_WARP_MASK"); return (e && atoi(e) != 0) ? 1 : 0; }();
    return (!off && P < kWarpMaskMaxGaussians) ? 1 : 0;
}
std::atomic<unsigned long long> g_launches{0};
int g_prof = 0;
bool profiling_enabled() { return g_prof != 0; }
int g_prof_dev = -1;
cudaEvent_t g_ev[ST_COUNT][2];
bool g_pending[ST_COUNT];
float g_ms[ST_COUNT];
int g_calls[ST_COUNT];

void prof_harvest(int st) {
    if (!g_pending[st]) return;
    float ms = 0.0f;
    if (cudaEventSynchronize(g_ev[st][1]) == cudaSuccess && cudaEventElapsedTime(&ms, g_ev[st][0], g_ev[st][1]) == cudaSuccess) {
        g_ms[st] += ms;
        g_calls[st] += 1;
    }
    g_pending[st] = false;
}
void prof_begin(int st, cudaStream_t s) {
    if (!g_prof) return;
    int dev = 0;
    cudaGetDevice(&dev);
    if (g_prof_dev != dev) {   // events belong to a device: (re)create them for the current one
        for (int i = 0; i < ST_COUNT; ++i) {
            cudaEventCreate(&g_ev[i][0]);
            cudaEventCreate(&g_ev[i][1]);
            g_pending[i] = false;
        }
        g_prof_dev = dev;
    }
    prof_harvest(st);
    cudaEventRecord(g_ev[st][0], s);
}
// ---- debug mode: the reference's CHECK_CUDA(..., debug) (DGR/cuda_rasterizer/auxiliary.h:166-173) synchronises after
// every stage and throws with the place of the failure. Here: wait for the stage's stream, pick up any error, remember
// which stage it was; the status goes back to the caller, who raises.
int g_debug = 0;
int g_debug_fail_stage = -1;                 // test hook: this stage reports a failure (nothing faults on the device)
std::atomic<unsigned long long> g_debug_syncs{0};
thread_local int t_failed_stage = -1;

int prof_end(int st, cudaStream_t s, int kernels) {
    g_launches += (unsigned long long)kernels;
    if (g_prof) {
        cudaEventRecord(g_ev[st][1], s);
        g_pending[st] = true;
    }
    if (!g_debug) return SPLAT_OK;
    cudaError_t e = cudaStreamSynchronize(s);
    g_debug_syncs += 1;
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e == cudaSuccess && g_debug_fail_stage == st) e = cudaErrorUnknown;
    if (e != cudaSuccess) t_failed_stage = st;
    return (int)e;
}
#define SPLAT_STAGE_END(st, s, kernels)              \
    do {                                             \
        const int rc__ = prof_end((st), (s), (kernels)); \
        if (rc__ != SPLAT_OK) return rc__;           \
    } while (0)
}  // namespace

extern "C" {

int splat_raster_abi_version(void) { return 1; }

size_t splat_raster_geom_bytes(int P) {
    if (P < 0) return 0;
    size_t b = 0;
    GeomState::carve(nullptr, (size_t)P, &b);
    return b;
}

size_t splat_raster_image_bytes(int W, int H) {
    if (W <= 0 || H <= 0) return 0;
    size_t b = 0;
    ImageState::carve(nullptr, (size_t)W * H, (size_t)grid_dim(W) * grid_dim(H), &b);
    return b;
}

size_t splat_raster_binning_bytes(int R) {
    if (R < 0) return 0;
    size_t b = 0;
    BinningState::carve(nullptr, (size_t)R, &b, g_keep_sort_output != 0);
    return b;
}

int splat_raster_forward_preprocess(int P, int D, int M, int W, int H, const float* means3D, const float* scales,
                                    float scale_modifier, const float* rotations, const float* opacities,
                                    const float* shs, const float* cov3D_precomp, const float* colors_precomp,
                                    const float* viewmatrix, const float* projmatrix, const float* campos,
                                    float tan_fovx, float tan_fovy, int prefiltered, int* radii, int* n_touched,
                                    void* geom, int* num_rendered_host, void* stream) {
    if (P < 0 || W <= 0 || H <= 0 || D < 0 || D > 3) return SPLAT_ERR_SIZE;
    if (!geom || !viewmatrix || !projmatrix) return SPLAT_ERR_ARG;
    const cudaStream_t caller = (cudaStream_t)stream;
    cudaStream_t s = caller;
    GeomState g = GeomState::carve(geom, (size_t)P, nullptr);
    if (P == 0) {
        cudaMemsetAsync(g.num_rendered, 0, sizeof(uint32_t), s);
        if (num_rendered_host) *num_rendered_host = 0;
        return SPLAT_OK;
    }
    if (!means3D || !opacities || !radii || !n_touched) return SPLAT_ERR_ARG;
    if ((shs == nullptr) == (colors_precomp == nullptr)) return SPLAT_ERR_ARG;            // exactly one colour source
    if (((scales == nullptr) || (rotations == nullptr)) == (cov3D_precomp == nullptr)) return SPLAT_ERR_ARG;
    if (shs && (M < (D + 1) * (D + 1) || !campos)) return SPLAT_ERR_ARG;
    FwdParams p;
    p.P = P; p.D = D; p.M = M; p.W = W; p.H = H;
    p.grid_x = grid_dim(W); p.grid_y = grid_dim(H);
    p.prefiltered = prefiltered;
    p.scale_modifier = scale_modifier; p.tanfovx = tan_fovx; p.tanfovy = tan_fovy;
    // focal lengths exactly as the reference derives them on the host (rasterizer_impl.cu:225-226)
    p.focal_y = H / (2.0f * tan_fovy);
    p.focal_x = W / (2.0f * tan_fovx);
    p.means3D = means3D; p.scales = scales; p.rotations = rotations; p.opacities = opacities; p.shs = shs;
    p.cov3D_precomp = cov3D_precomp; p.colors_precomp = colors_precomp;
    p.viewmatrix = viewmatrix; p.projmatrix = projmatrix; p.campos = campos;
    p.radii = radii; p.n_touched = n_touched;
    s = lane_begin(caller);
    cudaMemsetAsync(g.num_rendered, 0, 4 * sizeof(uint32_t), s);     // word 1: instance total summed by preprocess
    prof_begin(ST_PREPROCESS, s);
    launch_preprocess_forward(p, g, s);
    SPLAT_STAGE_END(ST_PREPROCESS, s, 1);
    SPLAT_CUDA_CHECK_LAUNCH();
    if (num_rendered_host) {
        // hand R to the host now; the depth sort below runs while the caller sizes its binning buffer
        cudaError_t e = cudaMemcpyAsync(num_rendered_host, g.num_rendered + 1, sizeof(int), cudaMemcpyDeviceToHost, s);
        if (e != cudaSuccess) return (int)e;
        cudaEvent_t ev = handoff_event();
        if (ev) {
            e = cudaEventRecord(ev, s);
            if (e != cudaSuccess) return (int)e;
        }
    }
    prof_begin(ST_SCAN, s);
    launch_depth_sort_and_offsets(P, g, s);   // histogram kernel + 4 sort passes + block sums (with scan)
    SPLAT_STAGE_END(ST_SCAN, s, 6);
    lane_end(caller, s);
    SPLAT_CUDA_CHECK_LAUNCH();
    return SPLAT_OK;
}

int splat_raster_priority_lanes(int enable) {
    const int was = g_priority_lanes;
    if (enable >= 0) g_priority_lanes = enable ? 1 : 0;
    if (enable == 0) {
        // Switching the lanes off also releases them: one side stream and two events per (device, caller stream) that
        // ever ran a forward. Work still queued on a destroyed stream or event completes first (CUDA defers the release).
        std::lock_guard<std::mutex> lock(g_lane_mutex);
        int cur = 0;
        cudaGetDevice(&cur);
        for (auto& kv : g_lanes) {
            Lane& l = kv.second;
            if (!l.hp) continue;
            cudaSetDevice(kv.first.first);
            cudaStreamDestroy(l.hp);
            cudaEventDestroy(l.fork);
            cudaEventDestroy(l.join);
        }
        g_lanes.clear();
        cudaSetDevice(cur);
    }
    return was;
}

int splat_raster_forward_wait(void) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return SPLAT_ERR_ARG;
    if (!t_handoff_recorded[dev]) return SPLAT_OK;
    const cudaError_t e = cudaEventSynchronize(t_handoff[dev]);
    return e == cudaSuccess ? SPLAT_OK : (int)e;
}

}  // extern "C"

namespace {
int loss_desc_from_abi(int W, int H, const splat_raster_loss_desc* in, LossDesc* out) {
    if (W <= 0 || H <= 0) return SPLAT_ERR_SIZE;
    if (!in->gt_image || !in->gt_depth || !in->workspace || !in->out || !in->dL_dimage || !in->dL_ddepth)
        return SPLAT_ERR_ARG;
    if ((reinterpret_cast<uintptr_t>(in->workspace) & 15u) != 0) return SPLAT_ERR_ARG;
    if (!in->initialization && (!in->exposure_a || !in->exposure_b)) return SPLAT_ERR_ARG;
    out->gt_image = in->gt_image; out->gt_depth = in->gt_depth;
    out->exposure_a = in->exposure_a; out->exposure_b = in->exposure_b;
    out->alpha = in->alpha; out->rgb_boundary_threshold = in->rgb_boundary_threshold;
    out->initialization = in->initialization;
    out->workspace = in->workspace; out->out = in->out; out->dL_dimage = in->dL_dimage; out->dL_ddepth = in->dL_ddepth;
    return SPLAT_OK;
}

int forward_render_impl(int P, int R, int W, int H, const float* background, const float* colors_precomp,
                        const int* radii, void* geom, void* binning, void* image, float* out_color,
                        float* out_depth, float* out_opacity, int* n_touched, const LossDesc* loss, void* stream) {
    if (P < 0 || R < 0 || W <= 0 || H <= 0) return SPLAT_ERR_SIZE;
    if (!geom || !binning || !image || !background || !out_color || !out_depth || !out_opacity) return SPLAT_ERR_ARG;
    if (P > 0 && (!radii || !n_touched)) return SPLAT_ERR_ARG;
    const cudaStream_t caller = (cudaStream_t)stream;
    cudaStream_t s = lane_begin(caller);     // key emission, tile sort and record gather on the side lane, if enabled
    const int gx = grid_dim(W), gy = grid_dim(H);
    GeomState g = GeomState::carve(geom, (size_t)P, nullptr);
    BinningState b = BinningState::carve(binning, (size_t)R, nullptr, g_keep_sort_output != 0);
    ImageState im = ImageState::carve(image, (size_t)W * H, (size_t)gx * gy, nullptr);
    const int passes = tile_sort_passes(gx, gy);
    if (R <= 0) launch_reset_ranges(gx * gy, im.ranges, s);     // otherwise preset by the key-emission kernel
    if (R > 0) {
        // the ping-pong sort reads its input from the first buffer pair; the last pass writes records, not pairs
        prof_begin(ST_DUP, s);
        launch_duplicate_with_keys(P, gx, gy, radii, g, b.tkey, b.point_list, im.ranges, s);
        SPLAT_STAGE_END(ST_DUP, s, 1);
        SPLAT_CUDA_CHECK_LAUNCH();
        RecordSink sink;
        sink.gauss_rec = g.rec;
        sink.records = reinterpret_cast<float4*>(b.records);
        sink.ranges = im.ranges;
        if (g_keep_sort_output) {      // inspection: also materialise the reference's sorted keys and point list
            sink.keys64 = b.keys;
            sink.point_list = (passes & 1) ? b.vals_alt : b.point_list;   // not the buffer the last pass reads
        }
        prof_begin(ST_SORT, s);
        launch_sort_pairs(b.tkey, b.point_list, b.tkey_alt, b.vals_alt, R, passes, b.sort_scratch, s, &sink);
        SPLAT_STAGE_END(ST_SORT, s, passes + 1);
        SPLAT_CUDA_CHECK_LAUNCH();
    }
    (void)colors_precomp;   // colours were folded into the per-Gaussian record by the preprocess kernel
    lane_end(caller, s);
    s = caller;
    SPLAT_CUDA_CHECK_LAUNCH();
    prof_begin(ST_RENDER_FWD, s);
    launch_render_forward(W, H, gx, gy, im.ranges, b.records, use_warp_mask(P), background, im.final_T, im.n_contrib, out_color,
                          out_depth, out_opacity, n_touched, R > 0 ? sort_error_flag(b.sort_scratch) : nullptr, loss, s);
    SPLAT_STAGE_END(ST_RENDER_FWD, s, 1);
    SPLAT_CUDA_CHECK_LAUNCH();
    return SPLAT_OK;
}
}  // namespace

extern "C" {

int splat_raster_forward_render(int P, int R, int W, int H, const float* background, const float* colors_precomp,
                                const int* radii, void* geom, void* binning, void* image, float* out_color,
                                float* out_depth, float* out_opacity, int* n_touched, void* stream) {
    return forward_render_impl(P, R, W, H, background, colors_precomp, radii, geom, binning, image, out_color, out_depth,
                               out_opacity, n_touched, nullptr, stream);
}

int splat_raster_forward_render_loss(int P, int R, int W, int H, const float* background, const float* colors_precomp,
                                     const int* radii, void* geom, void* binning, void* image, float* out_color,
                                     float* out_depth, float* out_opacity, int* n_touched,
                                     const splat_raster_loss_desc* loss, void* stream) {
    if (!loss) return SPLAT_ERR_ARG;
    LossDesc L;
    const int rc = loss_desc_from_abi(W, H, loss, &L);
    if (rc != SPLAT_OK) return rc;
    return forward_render_impl(P, R, W, H, background, colors_precomp, radii, geom, binning, image, out_color, out_depth,
                               out_opacity, n_touched, &L, stream);
}

int splat_raster_loss_partials(int W, int H) {
    if (W <= 0 || H <= 0) return 0;
    const int tiles = grid_dim(W) * grid_dim(H), blocks = mapping_loss_blocks(W * H);
    return tiles > blocks ? tiles : blocks;     // one workspace serves both the stand-alone and the fused loss
}

int splat_raster_backward(int P, int D, int M, int R, int W, int H, const float* background, const float* means3D,
                          const float* shs, const float* colors_precomp, const float* scales, float scale_modifier,
                          const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                          const float* projmatrix, const float* projmatrix_raw, const float* campos, float tan_fovx,
                          float tan_fovy, const int* radii, void* geom, void* binning, void* image,
                          const float* dL_dout_color, const float* dL_dout_depth, float* grad_scratch,
                          float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D,
                          float* dL_dcov3D, float* dL_dsh, float* dL_dscales, float* dL_drotations,
                          float* dL_dtau_partial, void* stream) {
    if (P < 0 || R < 0 || W <= 0 || H <= 0 || D < 0 || D > 3) return SPLAT_ERR_SIZE;
    if (P == 0) return SPLAT_OK;
    if (!geom || !binning || !image || !background || !means3D || !radii || !viewmatrix || !projmatrix ||
        !projmatrix_raw || !dL_dout_color || !dL_dout_depth || !grad_scratch || !dL_dtau_partial)
        return SPLAT_ERR_ARG;
    if (shs && !campos) return SPLAT_ERR_ARG;
    if (shs && D > 0 && !dL_dsh) return SPLAT_ERR_ARG;        // the SH direction term also feeds the pose gradient
    if (scales && (!rotations || ((dL_dscales == nullptr) != (dL_drotations == nullptr)))) return SPLAT_ERR_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    const int gx = grid_dim(W), gy = grid_dim(H);
    GeomState g = GeomState::carve(geom, (size_t)P, nullptr);
    BinningState b = BinningState::carve(binning, (size_t)R, nullptr, g_keep_sort_output != 0);
    ImageState im = ImageState::carve(image, (size_t)W * H, (size_t)gx * gy, nullptr);
    prof_begin(ST_ZERO, s);
    cudaError_t e = cudaMemsetAsync(grad_scratch, 0, sizeof(float) * SPLAT_GRAD_FLOATS * (size_t)P, s);
    SPLAT_STAGE_END(ST_ZERO, s, 0);
    if (e != cudaSuccess) return (int)e;
    if (R > 0) {
        prof_begin(ST_RENDER_BWD, s);
        launch_render_backward(W, H, gx, gy, im.ranges, b.records, use_warp_mask(P), background, im.final_T, im.n_contrib, dL_dout_color,
                               dL_dout_depth, grad_scratch, s);
        SPLAT_STAGE_END(ST_RENDER_BWD, s, 1);
        SPLAT_CUDA_CHECK_LAUNCH();
    }
    BwdParams p;
    p.P = P; p.D = D; p.M = M; p.R = R; p.W = W; p.H = H; p.grid_x = gx; p.grid_y = gy;
    p.scale_modifier = scale_modifier; p.tanfovx = tan_fovx; p.tanfovy = tan_fovy;
    p.focal_y = H / (2.0f * tan_fovy);
    p.focal_x = W / (2.0f * tan_fovx);
    p.bg = background; p.means3D = means3D; p.shs = shs; p.colors_precomp = colors_precomp; p.scales = scales;
    p.rotations = rotations; p.cov3D_precomp = cov3D_precomp; p.viewmatrix = viewmatrix; p.projmatrix = projmatrix;
    p.projmatrix_raw = projmatrix_raw; p.campos = campos; p.radii = radii;
    p.dL_dpix = dL_dout_color; p.dL_dpix_depth = dL_dout_depth; p.grad_accum = grad_scratch;
    p.dL_dmean2D = dL_dmeans2D; p.dL_dcolors = dL_dcolors; p.dL_dopacity = dL_dopacity; p.dL_dmean3D = dL_dmeans3D;
    p.dL_dcov3D = dL_dcov3D; p.dL_dsh = dL_dsh; p.dL_dscale = dL_dscales; p.dL_drot = dL_drotations;
    p.dL_dtau = dL_dtau_partial;
    prof_begin(ST_PREPROCESS_BWD, s);
    launch_preprocess_backward(p, g, s);
    SPLAT_STAGE_END(ST_PREPROCESS_BWD, s, 1);
    SPLAT_CUDA_CHECK_LAUNCH();
    return SPLAT_OK;
}

int splat_raster_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                              unsigned char* present, void* stream) {
    (void)projmatrix;   // the reference's test only uses the view-space depth (auxiliary.h:154)
    if (P < 0) return SPLAT_ERR_SIZE;
    if (P == 0) return SPLAT_OK;
    if (!means3D || !viewmatrix || !present) return SPLAT_ERR_ARG;
    launch_mark_visible(P, means3D, viewmatrix, reinterpret_cast<bool*>(present), (cudaStream_t)stream);
    g_launches += 1;
    SPLAT_CUDA_CHECK_LAUNCH();
    return SPLAT_OK;
}

int splat_raster_mapping_loss_blocks(int W, int H) {
    if (W <= 0 || H <= 0) return 0;
    return mapping_loss_blocks(W * H);
}

int splat_raster_mapping_loss(int W, int H, const float* image, const float* depth, const float* gt_image,
                              const float* gt_depth, const float* exposure_a, const float* exposure_b, float alpha,
                              float rgb_boundary_threshold, int initialization, float* workspace, float* out,
                              float* dL_dimage, float* dL_ddepth, void* stream) {
    if (W <= 0 || H <= 0) return SPLAT_ERR_SIZE;
    if (!image || !depth || !gt_image || !gt_depth || !workspace || !out || !dL_dimage || !dL_ddepth)
        return SPLAT_ERR_ARG;
    if ((reinterpret_cast<uintptr_t>(workspace) & 15u) != 0) return SPLAT_ERR_ARG;
    if (!initialization && (!exposure_a || !exposure_b)) return SPLAT_ERR_ARG;
    LossDesc L;
    L.gt_image = gt_image; L.gt_depth = gt_depth; L.exposure_a = exposure_a; L.exposure_b = exposure_b;
    L.alpha = alpha; L.rgb_boundary_threshold = rgb_boundary_threshold; L.initialization = initialization;
    L.workspace = workspace; L.out = out; L.dL_dimage = dL_dimage; L.dL_ddepth = dL_ddepth;
    launch_mapping_loss(W * H, image, depth, L, (cudaStream_t)stream);
    g_launches += 1;
    SPLAT_CUDA_CHECK_LAUNCH();
    return SPLAT_OK;
}

int splat_raster_activations_forward(int P, const float* raw_scaling, const float* raw_rotation,
                                     const float* raw_opacity, float* scales, float* rotations, float* opacities,
                                     void* stream) {
    if (P < 0) return SPLAT_ERR_SIZE;
    if (P == 0) return SPLAT_OK;
    if (!raw_scaling || !raw_rotation || !raw_opacity || !scales || !rotations || !opacities) return SPLAT_ERR_ARG;
    if (((reinterpret_cast<uintptr_t>(raw_rotation) | reinterpret_cast<uintptr_t>(rotations)) & 15u) != 0) return SPLAT_ERR_ARG;
    launch_activations_forward(P, raw_scaling, raw_rotation, raw_opacity, scales, rotations, opacities,
                               (cudaStream_t)stream);
    g_launches += 1;
    SPLAT_CUDA_CHECK_LAUNCH();
    return SPLAT_OK;
}

int splat_raster_activations_backward(int P, const float* raw_scaling, const float* raw_rotation,
                                      const float* raw_opacity, const float* g_scales, const float* g_rotations,
                                      const float* g_opacities, float* d_scaling, float* d_rotation, float* d_opacity,
                                      void* stream) {
    if (P < 0) return SPLAT_ERR_SIZE;
    if (P == 0) return SPLAT_OK;
    if (!raw_scaling || !raw_rotation || !raw_opacity) return SPLAT_ERR_ARG;
    uintptr_t bits = reinterpret_cast<uintptr_t>(raw_rotation) | reinterpret_cast<uintptr_t>(g_rotations) |
                     reinterpret_cast<uintptr_t>(d_rotation);
    if ((bits & 15u) != 0) return SPLAT_ERR_ARG;
    launch_activations_backward(P, raw_scaling, raw_rotation, raw_opacity, g_scales, g_rotations, g_opacities, d_scaling,
                                d_rotation, d_opacity, (cudaStream_t)stream);
    g_launches += 1;
    SPLAT_CUDA_CHECK_LAUNCH();
    return SPLAT_OK;
}

int splat_raster_adam_step(int num_segments, const splat_raster_adam_segment* segments, int zero_grads, void* stream) {
    if (num_segments < 0 || (num_segments > 0 && !segments)) return SPLAT_ERR_ARG;
    std::vector<AdamSegment> segs;
    segs.reserve((size_t)num_segments);
    for (int i = 0; i < num_segments; ++i) {
        const splat_raster_adam_segment& in = segments[i];
        if (in.n < 0 || in.step < 1) return SPLAT_ERR_ARG;
        if (in.n == 0) continue;
        if (!in.param || !in.grad || !in.exp_avg || !in.exp_avg_sq) return SPLAT_ERR_ARG;
        // scalars in double, then narrowed, exactly as torch hands Python floats to its float kernels
        const double bc1 = 1.0 - std::pow(in.beta1, (double)in.step);
        const double bc2 = 1.0 - std::pow(in.beta2, (double)in.step);
        AdamSegment s;
        s.p = in.param; s.g = in.grad; s.m = in.exp_avg; s.v = in.exp_avg_sq; s.n = in.n;
        s.w1 = (float)(1.0 - in.beta1);
        s.beta2 = (float)in.beta2;
        s.w2 = (float)(1.0 - in.beta2);
        s.bc2_sqrt = (float)std::sqrt(bc2);
        s.eps = (float)in.eps;
        s.neg_step = (float)(-(in.lr / bc1));
        s.vec4 = 0;
        segs.push_back(s);
    }
    if (segs.empty()) return SPLAT_OK;
    g_launches += launch_adam_step((int)segs.size(), segs.data(), zero_grads, (cudaStream_t)stream);
    SPLAT_CUDA_CHECK_LAUNCH();
    return SPLAT_OK;
}

int splat_raster_densify_stats(int P, int K, const int* const* radii, const float* const* viewspace_grad,
                               int grad_stride, float* max_radii2D, float* xyz_gradient_accum, float* denom,
                               void* stream) {
    if (P < 0 || K < 0 || grad_stride < 2) return SPLAT_ERR_ARG;
    if (P == 0 || K == 0) return SPLAT_OK;
    if (!radii || !viewspace_grad || !max_radii2D || !xyz_gradient_accum || !denom) return SPLAT_ERR_ARG;
    for (int k = 0; k < K; ++k)
        if (!radii[k] || !viewspace_grad[k]) return SPLAT_ERR_ARG;
    g_launches += launch_densify_stats(P, K, radii, viewspace_grad, grad_stride, max_radii2D, xyz_gradient_accum, denom,
                                       (cudaStream_t)stream);
    SPLAT_CUDA_CHECK_LAUNCH();
    return SPLAT_OK;
}

int splat_raster_visibility_count(int P, int K, const int* const* n_touched, long long* const* visibility,
                                  int* n_obs, void* stream) {
    if (P < 0 || K < 0) return SPLAT_ERR_ARG;
    if (P == 0 || K == 0) return SPLAT_OK;
    if (!n_touched) return SPLAT_ERR_ARG;
    for (int k = 0; k < K; ++k)
        if (!n_touched[k]) return SPLAT_ERR_ARG;
    g_launches += launch_visibility_count(P, K, n_touched, visibility, n_obs, (cudaStream_t)stream);
    SPLAT_CUDA_CHECK_LAUNCH();
    return SPLAT_OK;
}

size_t splat_raster_knn_scratch_bytes(int P) { return P > 0 ? knn_scratch_bytes((size_t)P) : 256; }

int splat_raster_knn_mean_dist2(int P, const float* points, void* scratch, float* mean_dist2, void* stream) {
    if (P < 0) return SPLAT_ERR_SIZE;
    if (P == 0) return SPLAT_OK;
    if (!points || !scratch || !mean_dist2) return SPLAT_ERR_ARG;
    if ((reinterpret_cast<uintptr_t>(scratch) & 255u) != 0) return SPLAT_ERR_ARG;
    g_launches += launch_knn_mean_dist2(P, points, scratch, mean_dist2, (cudaStream_t)stream);
    SPLAT_CUDA_CHECK_LAUNCH();
    return SPLAT_OK;
}

int splat_raster_deform_points(int P, int K, const splat_raster_deform_keyframe* keyframes, const int* kf_ids,
                               const float* intrinsics, int W, int H, float* table, float* xyz, float* rotation,
                               float* scaling, void* stream) {
    static_assert(SPLAT_DEFORM_TABLE_FLOATS == kDeformMaxKeyframes * 40, "table size");
    if (P < 0 || K < 0) return SPLAT_ERR_SIZE;
    if (P == 0 || K == 0) return SPLAT_OK;
    if (!keyframes || !kf_ids || !table || !xyz || !rotation || !scaling) return SPLAT_ERR_ARG;
    if ((reinterpret_cast<uintptr_t>(rotation) & 15u) != 0) return SPLAT_ERR_ARG;
    std::vector<DeformKeyframe> kfs((size_t)K);
    for (int k = 0; k < K; ++k) {
        const splat_raster_deform_keyframe& in = keyframes[k];
        if (!in.w2c_old || !in.w2c_new) return SPLAT_ERR_ARG;
        if (!in.rigid && (!in.depth_new || !in.depth_old || !intrinsics || W <= 0 || H <= 0)) return SPLAT_ERR_ARG;
        kfs[(size_t)k] = DeformKeyframe{in.w2c_old, in.w2c_new, in.depth_new, in.depth_old, in.frame_idx, in.rigid};
    }
    g_launches += launch_deform_points(P, K, kfs.data(), kf_ids, intrinsics, W, H, table, xyz, rotation, scaling,
                                       (cudaStream_t)stream);
    SPLAT_CUDA_CHECK_LAUNCH();
    return SPLAT_OK;
}

int splat_raster_allreduce_peer(float* const* buffers, int world, int rank, size_t count, void* stream) {
    if (!buffers || world < 1 || rank < 0 || rank >= world || (count & 3u) != 0) return SPLAT_ERR_ARG;
    if (world > kMaxPeerRanks) return SPLAT_ERR_SIZE;
    for (int r = 0; r < world; ++r)
        if (!buffers[r] || (reinterpret_cast<uintptr_t>(buffers[r]) & 15u) != 0) return SPLAT_ERR_ARG;
    if (world == 1 || count == 0) return SPLAT_OK;
    const int k = launch_allreduce_peer(buffers, world, rank, count, (cudaStream_t)stream);
    if (k < 0) return SPLAT_ERR_SIZE;
    g_launches += (unsigned long long)k;
    SPLAT_CUDA_CHECK_LAUNCH();
    return SPLAT_OK;
}

int splat_raster_allreduce_multimem(float* multicast, int world, int rank, size_t count, void* stream) {
    if (!multicast || world < 1 || rank < 0 || rank >= world || (count & 3u) != 0) return SPLAT_ERR_ARG;
    if ((reinterpret_cast<uintptr_t>(multicast) & 15u) != 0) return SPLAT_ERR_ARG;
    if (world == 1 || count == 0) return SPLAT_OK;
    g_launches += (unsigned long long)launch_allreduce_multimem(multicast, world, rank, count, (cudaStream_t)stream);
    SPLAT_CUDA_CHECK_LAUNCH();
    return SPLAT_OK;
}

int splat_raster_keep_sort_output(int enable) {
    const int was = g_keep_sort_output;
    if (enable >= 0) g_keep_sort_output = enable ? 1 : 0;
    return was;
}

int splat_raster_debug_views(int P, int R, int W, int H, void* geom, void* binning, void* image,
                             splat_raster_views* out) {
    if (!out || P < 0 || R < 0 || W <= 0 || H <= 0) return SPLAT_ERR_ARG;
    const int gx = grid_dim(W), gy = grid_dim(H);
    GeomState g = GeomState::carve(geom, (size_t)P, nullptr);
    BinningState b = BinningState::carve(binning, (size_t)R, nullptr, true);   // keys: only there if the forward ran with keep_sort_output
    ImageState im = ImageState::carve(image, (size_t)W * H, (size_t)gx * gy, nullptr);
    out->depths = g.depths;
    out->means2D = reinterpret_cast<const float*>(g.means2D);
    out->conic_opacity = reinterpret_cast<const float*>(g.conic_opacity);
    out->rgb = g.rgb;
    out->cov3D = g.cov3D;
    out->tiles_touched = g.tiles_touched;
    out->keys = reinterpret_cast<const unsigned long long*>(b.keys);
    out->point_list = (tile_sort_passes(gx, gy) & 1) ? b.vals_alt : b.point_list;
    out->ranges = reinterpret_cast<const unsigned int*>(im.ranges);
    out->final_T = im.final_T;
    out->n_contrib = im.n_contrib;
    return SPLAT_OK;
}

}  // extern "C"

// ---- backward without any allocation ---------------------------------------------------------------------
// Scratch for splat_raster_backward_noalloc: the packed per-Gaussian render gradients (12 floats each) that the
// regular entry point receives from its caller. 262144 Gaussians x 48 B = 12 MiB, resident on every GPU that loads
// this library until the process exits. Device code may name the array directly; HOST code must not use the name as
// an address (it is only a host-side handle for the symbol) and goes through cudaGetSymbolAddress below.
static __device__ __align__(256) float g_noalloc_grad_scratch[(size_t)SPLAT_NOALLOC_MAX_GAUSSIANS * SPLAT_GRAD_FLOATS];

namespace {
constexpr int kNoAllocMaxDevices = 64;
float* g_noalloc_ptr[kNoAllocMaxDevices];   // per device: where that device's copy of the array lives
}  // namespace

extern "C" {

int splat_raster_backward_noalloc(int P, int D, int M, int R, int W, int H, const float* background,
                                  const float* means3D, const float* shs, const float* colors_precomp,
                                  const float* scales, float scale_modifier, const float* rotations,
                                  const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                                  const float* projmatrix_raw, const float* campos, float tan_fovx, float tan_fovy,
                                  const int* radii, void* geom, void* binning, void* image,
                                  const float* dL_dout_color, const float* dL_dout_depth, float* dL_dmeans2D,
                                  float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D,
                                  float* dL_dsh, float* dL_dscales, float* dL_drotations, float* dL_dtau_partial,
                                  void* stream) {
    if (P < 0 || R < 0 || W <= 0 || H <= 0 || D < 0 || D > 3) return SPLAT_ERR_SIZE;
    if (P > SPLAT_NOALLOC_MAX_GAUSSIANS) return SPLAT_ERR_CAPACITY;   // would not fit the built-in scratch
    if (P == 0) return SPLAT_OK;
    // same contract as splat_raster_backward: a NULL per-Gaussian output means "not wanted" (pose-only tracking passes
    // NULL for all of them); only the pose partials are mandatory
    if (!geom || !binning || !image || !background || !means3D || !radii || !viewmatrix || !projmatrix ||
        !projmatrix_raw || !dL_dout_color || !dL_dout_depth || !dL_dtau_partial)
        return SPLAT_ERR_ARG;
    if (shs && !campos) return SPLAT_ERR_ARG;
    if (shs && D > 0 && !dL_dsh) return SPLAT_ERR_ARG;        // the SH direction term also feeds the pose gradient
    if (scales && (!rotations || ((dL_dscales == nullptr) != (dL_drotations == nullptr)))) return SPLAT_ERR_ARG;

    // One-time set-up per device, done by the first call made with that device current: ask the runtime where this
    // device's copy of the array is. (Keep that first call outside of any stream capture.)
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= kNoAllocMaxDevices) return SPLAT_ERR_ARG;
    if (g_noalloc_ptr[dev] == nullptr) {
        void* addr = nullptr;
        e = cudaGetSymbolAddress(&addr, g_noalloc_grad_scratch);
        if (e != cudaSuccess) return (int)e;
        if (addr == nullptr) return SPLAT_ERR_ARG;
        g_noalloc_ptr[dev] = static_cast<float*>(addr);
    }
    float* scratch = g_noalloc_ptr[dev];

    cudaStream_t s = (cudaStream_t)stream;
    const int gx = grid_dim(W), gy = grid_dim(H);
    GeomState g = GeomState::carve(geom, (size_t)P, nullptr);
    BinningState b = BinningState::carve(binning, (size_t)R, nullptr, g_keep_sort_output != 0);
    ImageState im = ImageState::carve(image, (size_t)W * H, (size_t)gx * gy, nullptr);
    // the array still holds the previous call's sums: the part this call accumulates into starts from zero
    e = cudaMemsetAsync(scratch, 0, sizeof(float) * SPLAT_GRAD_FLOATS * (size_t)P, s);
    if (e != cudaSuccess) return (int)e;
    if (R > 0) {
        launch_render_backward(W, H, gx, gy, im.ranges, b.records, use_warp_mask(P), background, im.final_T, im.n_contrib, dL_dout_color,
                               dL_dout_depth, scratch, s);
        g_launches += 1;
        SPLAT_CUDA_CHECK_LAUNCH();
    }
    BwdParams p;
    p.P = P; p.D = D; p.M = M; p.R = R; p.W = W; p.H = H; p.grid_x = gx; p.grid_y = gy;
    p.scale_modifier = scale_modifier; p.tanfovx = tan_fovx; p.tanfovy = tan_fovy;
    p.focal_y = H / (2.0f * tan_fovy);
    p.focal_x = W / (2.0f * tan_fovx);
    p.bg = background; p.means3D = means3D; p.shs = shs; p.colors_precomp = colors_precomp; p.scales = scales;
    p.rotations = rotations; p.cov3D_precomp = cov3D_precomp; p.viewmatrix = viewmatrix; p.projmatrix = projmatrix;
    p.projmatrix_raw = projmatrix_raw; p.campos = campos; p.radii = radii;
    p.dL_dpix = dL_dout_color; p.dL_dpix_depth = dL_dout_depth; p.grad_accum = scratch;
    p.dL_dmean2D = dL_dmeans2D; p.dL_dcolors = dL_dcolors; p.dL_dopacity = dL_dopacity; p.dL_dmean3D = dL_dmeans3D;
    p.dL_dcov3D = dL_dcov3D; p.dL_dsh = dL_dsh; p.dL_dscale = dL_dscales; p.dL_drot = dL_drotations;
    p.dL_dtau = dL_dtau_partial;
    launch_preprocess_backward(p, g, s);
    g_launches += 1;
    SPLAT_CUDA_CHECK_LAUNCH();
    return SPLAT_OK;
}

unsigned long long splat_raster_launch_count(int reset) {
    return reset ? g_launches.exchange(0) : g_launches.load();
}

int splat_raster_set_debug(int enable, int fail_stage) {
    const int was = g_debug;
    g_debug = enable ? 1 : 0;
    g_debug_fail_stage = enable ? fail_stage : -1;
    return was;
}

const char* splat_raster_last_failed_stage(int reset) {
    const int st = t_failed_stage;
    if (reset) t_failed_stage = -1;
    return (st >= 0 && st < ST_COUNT) ? kStageNames[st] : "";
}

unsigned long long splat_raster_debug_sync_count(int reset) {
    return reset ? g_debug_syncs.exchange(0) : g_debug_syncs.load();
}

int splat_raster_profile(int enable) {
    const int old = g_prof;
    g_prof = enable ? 1 : 0;
    return old;
}

int splat_raster_stage_count(void) { return ST_COUNT; }

const char* splat_raster_stage_name(int stage) { return (stage >= 0 && stage < ST_COUNT) ? kStageNames[stage] : ""; }

int splat_raster_profile_read(float* ms_out, int* calls_out, int reset) {
    if (!ms_out || !calls_out) return SPLAT_ERR_ARG;
    for (int i = 0; i < ST_COUNT; ++i) {
        if (g_prof_dev >= 0) prof_harvest(i);
        ms_out[i] = g_ms[i];
        calls_out[i] = g_calls[i];
        if (reset) { g_ms[i] = 0.0f; g_calls[i] = 0; }
    }
    return SPLAT_OK;
}

}  // extern "C"
