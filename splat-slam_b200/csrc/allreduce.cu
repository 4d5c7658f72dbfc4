// Gradient exchange of the view-sharded mapping window over NVLink peer memory (the reference is
// single-GPU and has no counterpart). Every rank holds the flat Gaussian-gradient buffer in memory that all ranks of
// the node have mapped (torch symmetric memory hands out the peer pointers / the NVSwitch multicast address).
// One kernel per rank does reduce-scatter and all-gather in a single pass: rank r owns the r-th slice, sums it over
// all ranks' buffers and writes the total back into every buffer. Each element is summed by exactly one rank, in rank
// order, so all ranks end up with bit-identical gradients (NCCL's ring gives that too; a per-rank local sum would not).
//   * peer variant: plain loads/stores on the mapped peer pointers (world loads in flight per element, world stores);
//   * multimem variant: the NVSwitch reduces (multimem.ld_reduce) and broadcasts (multimem.st) in the fabric, so a rank
//     moves 1/world of the buffer each way instead of (world-1)/world.
// The caller brackets the call with two cross-rank barriers (before: every rank's gradients are complete; after:
// every rank's slice has landed everywhere); the kernel itself never waits for another GPU.
#include "common.cuh"
#include "kernels.h"

namespace splat {

struct PeerBuffers {
    float* p[kMaxPeerRanks];
};

__device__ __forceinline__ float4 ld_sys_v4(const float* p) {
    float4 v;
    asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_sys_v4(float* p, float4 v) {
    asm volatile("st.volatile.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

template <int W>
__global__ void __launch_bounds__(512)
allreduce_peer_kernel(PeerBuffers bufs, size_t first4, size_t end4) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = first4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < end4; i += stride) {
        float4 v[W];
#pragma unroll
        for (int r = 0; r < W; ++r) v[r] = ld_sys_v4(bufs.p[r] + 4 * i);
        float4 s = v[0];
#pragma unroll
        for (int r = 1; r < W; ++r) { s.x += v[r].x; s.y += v[r].y; s.z += v[r].z; s.w += v[r].w; }
#pragma unroll
        for (int r = 0; r < W; ++r) st_sys_v4(bufs.p[r] + 4 * i, s);
    }
}

__global__ void __launch_bounds__(512)
allreduce_multimem_kernel(float* mc, size_t first4, size_t end4) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = first4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < end4; i += stride) {
        float4 s;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(s.x), "=f"(s.y), "=f"(s.z), "=f"(s.w)
                     : "l"(mc + 4 * i)
                     : "memory");
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc + 4 * i), "f"(s.x),
                     "f"(s.y), "f"(s.z), "f"(s.w)
                     : "memory");
    }
}

// slice of rank `rank` in units of float4: [first, end)
static void slice_of(size_t count4, int world, int rank, size_t& first, size_t& end) {
    const size_t per = (count4 + (size_t)world - 1) / (size_t)world;
    first = per * (size_t)rank < count4 ? per * (size_t)rank : count4;
    end = first + per < count4 ? first + per : count4;
}

int launch_allreduce_peer(float* const* buffers, int world, int rank, size_t count, cudaStream_t s) {
    PeerBuffers b;
    for (int r = 0; r < kMaxPeerRanks; ++r) b.p[r] = r < world ? buffers[r] : nullptr;
    size_t first, end;
    slice_of(count / 4, world, rank, first, end);
    if (end <= first) return 0;
    const size_t n = end - first;
    int grid = (int)((n + 511) / 512);
    if (grid > 148 * 4) grid = 148 * 4;
    switch (world) {
        case 2: allreduce_peer_kernel<2><<<grid, 512, 0, s>>>(b, first, end); break;
        case 3: allreduce_peer_kernel<3><<<grid, 512, 0, s>>>(b, first, end); break;
        case 4: allreduce_peer_kernel<4><<<grid, 512, 0, s>>>(b, first, end); break;
        case 5: allreduce_peer_kernel<5><<<grid, 512, 0, s>>>(b, first, end); break;
        case 6: allreduce_peer_kernel<6><<<grid, 512, 0, s>>>(b, first, end); break;
        case 7: allreduce_peer_kernel<7><<<grid, 512, 0, s>>>(b, first, end); break;
        case 8: allreduce_peer_kernel<8><<<grid, 512, 0, s>>>(b, first, end); break;
        default: return -1;
    }
    return 1;
}

int launch_allreduce_multimem(float* multicast, int world, int rank, size_t count, cudaStream_t s) {
    size_t first, end;
    slice_of(count / 4, world, rank, first, end);
    if (end <= first) return 0;
    const size_t n = end - first;
    int grid = (int)((n + 511) / 512);
    if (grid > 148 * 4) grid = 148 * 4;
    allreduce_multimem_kernel<<<grid, 512, 0, s>>>(multicast, first, end);
    return 1;
}

}  // namespace splat
