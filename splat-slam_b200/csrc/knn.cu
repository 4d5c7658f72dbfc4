// Mean squared distance to the three nearest neighbours of every point: the initial scale of
// new Gaussians at every keyframe insert (thirdparty/gaussian_splatting/scene/gaussian_model.py:194-200 calls
// simple_knn._C.distCUDA2). Replaces thirdparty/simple-knn/simple_knn.cu:63-216 (SimpleKNN::knn: two CUB reductions
// with blocking copies of the bounds to the host, coord2Morton, CUB radix sort, boxMinMax over 1024-point boxes,
// boxMeanDist scanning every box that intersects the running 3rd-best radius, through indirect point loads).
//
// The result is an exact 3-NN over squared distances d = dx*dx + dy*dy + dz*dz, so it does not depend on the search
// structure; only the distance expression's rounding matters, and that is kept (one rounded product, then two fused
// multiply-adds, in the order nvcc contracts the reference's expression). What is different here:
//   * nothing goes to the host: bounds are block partials that the Morton kernel folds itself;
//   * the Morton sort is this library's own one-sweep radix sort (radix_sort.cu), four 8-bit passes over 30-bit codes;
//   * the sorted points are gathered once into a padded float4 array, so the search reads coalesced / broadcast 16-byte
//     records instead of points[indices[i]];
//   * two box levels (1024 points, and 32 = one warp) instead of one, tested warp-uniformly with a vote; the search
//     starts at the query's own warp and walks the 1024-boxes outward in Morton order so the bound tightens early.
// Instruction-bound (distance evaluations), not HBM-bound: 16 B read per point per visiting warp, from L1/L2.
#include "common.cuh"
#include "kernels.h"
#include "state.h"

#include <cfloat>

namespace splat {

constexpr int kKnnBox = 1024;        // level-1 box = one gather CTA
constexpr int kKnnSub = 32;          // level-2 box = one warp
constexpr int kKnnBoundBlocks = 296; // 148 SMs x 2

struct KnnScratch {
    uint32_t *key0, *val0, *key1, *val1, *sort_scratch;
    float4* sorted;      // [nbox * 1024] (x, y, z, original index bits), padded with +inf
    float4* sub_box;     // [nbox * 32][2] (min, max)
    float4* box;         // [nbox][2]
    float* partials;     // [kKnnBoundBlocks][6]
    static KnnScratch carve(void* buf, size_t P, size_t* bytes) {
        Carver c(buf);
        KnnScratch k;
        const size_t nbox = (P + kKnnBox - 1) / kKnnBox;
        k.key0 = c.take<uint32_t>(P);
        k.val0 = c.take<uint32_t>(P);
        k.key1 = c.take<uint32_t>(P);
        k.val1 = c.take<uint32_t>(P);
        k.sort_scratch = c.take<uint32_t>(sort_scratch_words(P));
        k.sorted = c.take<float4>(nbox * kKnnBox);
        k.sub_box = c.take<float4>(nbox * (kKnnBox / kKnnSub) * 2);
        k.box = c.take<float4>(nbox * 2);
        k.partials = c.take<float>(kKnnBoundBlocks * 6);
        if (bytes) *bytes = c.off + 256;
        return k;
    }
};

size_t knn_scratch_bytes(size_t P) {
    size_t bytes = 0;
    KnnScratch::carve(nullptr, P, &bytes);
    return bytes;
}

__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__global__ void __launch_bounds__(256)
knn_bounds_kernel(int P, const float* __restrict__ points, float* __restrict__ partials) {
    __shared__ float s[8][6];
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = blockIdx.x * 256 + threadIdx.x; i < P; i += gridDim.x * 256) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float v = points[3 * (size_t)i + a];
            lo[a] = fminf(lo[a], v);
            hi[a] = fmaxf(hi[a], v);
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float l = warp_min(lo[a]), h = warp_max(hi[a]);
        if (lane == 0) { s[warp][a] = l; s[warp][3 + a] = h; }
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        float v = s[0][threadIdx.x];
        for (int w = 1; w < 8; ++w) v = threadIdx.x < 3 ? fminf(v, s[w][threadIdx.x]) : fmaxf(v, s[w][threadIdx.x]);
        partials[6 * blockIdx.x + threadIdx.x] = v;
    }
}

__device__ __forceinline__ uint32_t spread10(uint32_t x) {       // 10 bits -> every third bit
    x = (x | (x << 16)) & 0x030000FFu;
    x = (x | (x << 8)) & 0x0300F00Fu;
    x = (x | (x << 4)) & 0x030C30C3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}

__global__ void __launch_bounds__(256)
knn_morton_kernel(int P, const float* __restrict__ points, const float* __restrict__ partials, int nparts,
                  uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    __shared__ float s_b[6];
    if (threadIdx.x < 6) {
        float v = partials[threadIdx.x];
        for (int b = 1; b < nparts; ++b)
            v = threadIdx.x < 3 ? fminf(v, partials[6 * b + threadIdx.x]) : fmaxf(v, partials[6 * b + threadIdx.x]);
        s_b[threadIdx.x] = v;
    }
    __syncthreads();
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= P) return;
    uint32_t code = 0;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float ext = s_b[3 + a] - s_b[a];
        float q = ext > 0.0f ? (points[3 * (size_t)i + a] - s_b[a]) / ext * 1023.0f : 0.0f;
        q = fminf(fmaxf(q, 0.0f), 1023.0f);                    // also maps NaN to 0
        code |= spread10((uint32_t)q) << a;
    }
    keys[i] = code;
    vals[i] = (uint32_t)i;
}

// One CTA per 1024 Morton-sorted points: gather them into the padded record array, reduce the warp's and the CTA's
// bounding boxes.
__global__ void __launch_bounds__(kKnnBox)
knn_gather_boxes_kernel(int P, const float* __restrict__ points, const uint32_t* __restrict__ order,
                        float4* __restrict__ sorted, float4* __restrict__ sub_box, float4* __restrict__ box) {
    __shared__ float s[kKnnBox / 32][6];
    const int i = blockIdx.x * kKnnBox + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float4 rec = make_float4(INFINITY, INFINITY, INFINITY, __uint_as_float(0xFFFFFFFFu));
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (i < P) {
        const uint32_t id = order[i];
        rec = make_float4(points[3 * (size_t)id], points[3 * (size_t)id + 1], points[3 * (size_t)id + 2],
                          __uint_as_float(id));
        lo[0] = hi[0] = rec.x; lo[1] = hi[1] = rec.y; lo[2] = hi[2] = rec.z;
    }
    sorted[i] = rec;
#pragma unroll
    for (int a = 0; a < 3; ++a) { lo[a] = warp_min(lo[a]); hi[a] = warp_max(hi[a]); }
    if (lane == 0) {
        const size_t sb = (size_t)blockIdx.x * (kKnnBox / 32) + warp;
        sub_box[2 * sb] = make_float4(lo[0], lo[1], lo[2], 0.0f);
        sub_box[2 * sb + 1] = make_float4(hi[0], hi[1], hi[2], 0.0f);
#pragma unroll
        for (int a = 0; a < 3; ++a) { s[warp][a] = lo[a]; s[warp][3 + a] = hi[a]; }
    }
    __syncthreads();
    if (warp == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { lo[a] = warp_min(s[lane][a]); hi[a] = warp_max(s[lane][3 + a]); }
        if (lane == 0) {
            box[2 * blockIdx.x] = make_float4(lo[0], lo[1], lo[2], 0.0f);
            box[2 * blockIdx.x + 1] = make_float4(hi[0], hi[1], hi[2], 0.0f);
        }
    }
}

__device__ __forceinline__ float dist2_ref(const float4& a, const float4& b) {
    const float dx = __fsub_rn(a.x, b.x), dy = __fsub_rn(a.y, b.y), dz = __fsub_rn(a.z, b.z);
    // simple_knn.cu:132-133 as nvcc contracts `dx*dx + dy*dy + dz*dz` (SASS of the sm_100 build: the y product is the
    // one that is rounded on its own, x and z are fused)
    return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
}

// Lower bound of dist2_ref to any point inside the box: per axis the float distance to the box is <= the float distance
// to any point in it (subtraction rounds monotonically), and the sum of squares above is monotone in each term.
__device__ __forceinline__ float box_dist2(const float4& lo, const float4& hi, const float4& p) {
    const float dx = fmaxf(fmaxf(__fsub_rn(lo.x, p.x), __fsub_rn(p.x, hi.x)), 0.0f);
    const float dy = fmaxf(fmaxf(__fsub_rn(lo.y, p.y), __fsub_rn(p.y, hi.y)), 0.0f);
    const float dz = fmaxf(fmaxf(__fsub_rn(lo.z, p.z), __fsub_rn(p.z, hi.z)), 0.0f);
    return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
}

__device__ __forceinline__ void keep3(float d, float& b0, float& b1, float& b2) {      // simple_knn.cu:134-142
    if (b0 > d) { const float t = b0; b0 = d; d = t; }
    if (b1 > d) { const float t = b1; b1 = d; d = t; }
    if (b2 > d) { b2 = d; }
}

__device__ __forceinline__ void scan_sub_box(const float4* __restrict__ sorted, int first, int self, const float4& me,
                                             float& b0, float& b1, float& b2) {
#pragma unroll 8
    for (int j = 0; j < kKnnSub; ++j) {
        const float4 o = __ldg(sorted + first + j);            // same address across the warp: one broadcast load
        const float d = dist2_ref(o, me);
        if (first + j != self) keep3(d, b0, b1, b2);
    }
}

__global__ void __launch_bounds__(256)
knn_search_kernel(int P, int nbox, const float4* __restrict__ sorted, const float4* __restrict__ sub_box,
                  const float4* __restrict__ box, float* __restrict__ mean_dist2) {
    const int q = blockIdx.x * 256 + threadIdx.x;               // padded positions take part in the votes
    const float4 me = sorted[q];
    const bool live = q < P;
    float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;
    const int own_sub = q / kKnnSub, own_box = q / kKnnBox;
    scan_sub_box(sorted, own_sub * kKnnSub, q, me, b0, b1, b2);
    for (int step = 0; step < 2 * nbox; ++step) {                // own box, then alternately one further down / up
        const int b = own_box + ((step & 1) ? -((step + 1) >> 1) : (step >> 1));
        if (b < 0 || b >= nbox) continue;
        const bool want = live && box_dist2(__ldg(box + 2 * b), __ldg(box + 2 * b + 1), me) < b2;
        if (!__any_sync(0xffffffffu, want)) continue;
        for (int sb = b * (kKnnBox / kKnnSub); sb < (b + 1) * (kKnnBox / kKnnSub); ++sb) {
            if (sb == own_sub) continue;
            const bool want_sub = live && box_dist2(__ldg(sub_box + 2 * sb), __ldg(sub_box + 2 * sb + 1), me) < b2;
            if (!__any_sync(0xffffffffu, want_sub)) continue;
            scan_sub_box(sorted, sb * kKnnSub, q, me, b0, b1, b2);   // every lane may as well use the candidates
        }
    }
    if (live) mean_dist2[__float_as_uint(me.w)] = __fdiv_rn(__fadd_rn(__fadd_rn(b0, b1), b2), 3.0f);
}

int launch_knn_mean_dist2(int P, const float* points, void* scratch, float* mean_dist2, cudaStream_t s) {
    if (P <= 0) return 0;
    KnnScratch k = KnnScratch::carve(scratch, (size_t)P, nullptr);
    const int nbox = (P + kKnnBox - 1) / kKnnBox;
    int nparts = (P + 255) / 256;
    if (nparts > kKnnBoundBlocks) nparts = kKnnBoundBlocks;
    knn_bounds_kernel<<<nparts, 256, 0, s>>>(P, points, k.partials);
    knn_morton_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, points, k.partials, nparts, k.key0, k.val0);
    launch_sort_pairs(k.key0, k.val0, k.key1, k.val1, P, 4, k.sort_scratch, s);     // 30-bit codes, result in (key0,val0)
    knn_gather_boxes_kernel<<<nbox, kKnnBox, 0, s>>>(P, points, k.val0, k.sorted, k.sub_box, k.box);
    knn_search_kernel<<<nbox * (kKnnBox / 256), 256, 0, s>>>(P, nbox, k.sorted, k.sub_box, k.box, mean_dist2);
    return 4 + 5;      // + histogram and four one-sweep passes inside launch_sort_pairs
}

}  // namespace splat
