// The bar for the instance sort: what the reference runs per view for binning, timed alone on the same GPU:
//   cub::DeviceScan::InclusiveSum over P tile counts            (DGR/cuda_rasterizer/rasterizer_impl.cu:280)
//   cub::DeviceRadixSort::SortPairs<uint64, uint32> over R keys, bits [0, 32 + ceil(log2(tiles)))   (:303-311)
// with P = 1 M, R = 3.45 M, 3225 tiles (the bench scene). Compare with this library's stage_ms
// depth_sort_and_offsets + duplicate_with_keys + tile_sort_records_ranges (which also build the render records).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a tests/perf/cub_sort_bench.cu -o tests/perf/_build/cub_sort_bench
#include <cub/cub.cuh>
#include <cstdio>
#include <cstdint>
#include <vector>
#include <random>

int main(int argc, char** argv) {
    const int P = argc > 1 ? atoi(argv[1]) : 1000000;
    const int R = argc > 2 ? atoi(argv[2]) : 3451769;
    const int tiles = argc > 3 ? atoi(argv[3]) : 3225;
    int bits = 0;
    while ((tiles >> bits) != 0) ++bits;
    std::mt19937 rng(1);
    std::vector<uint64_t> keys(R);
    std::vector<uint32_t> vals(R), counts(P);
    for (int i = 0; i < R; ++i) {
        const float depth = 0.5f + 5.5f * (rng() / 4294967296.0f);
        uint32_t db;
        memcpy(&db, &depth, 4);
        keys[i] = ((uint64_t)(rng() % tiles) << 32) | db;
        vals[i] = rng() % P;
    }
    for (int i = 0; i < P; ++i) counts[i] = rng() % 8;
    uint64_t *k0, *k1;
    uint32_t *v0, *v1, *c0, *c1;
    cudaMalloc(&k0, 8ull * R); cudaMalloc(&k1, 8ull * R); cudaMalloc(&v0, 4ull * R); cudaMalloc(&v1, 4ull * R);
    cudaMalloc(&c0, 4ull * P); cudaMalloc(&c1, 4ull * P);
    cudaMemcpy(k0, keys.data(), 8ull * R, cudaMemcpyHostToDevice);
    cudaMemcpy(v0, vals.data(), 4ull * R, cudaMemcpyHostToDevice);
    cudaMemcpy(c0, counts.data(), 4ull * P, cudaMemcpyHostToDevice);
    size_t tmp_sort = 0, tmp_scan = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmp_sort, k0, k1, v0, v1, R, 0, 32 + bits);
    cub::DeviceScan::InclusiveSum(nullptr, tmp_scan, c0, c1, P);
    void *t_sort, *t_scan;
    cudaMalloc(&t_sort, tmp_sort); cudaMalloc(&t_scan, tmp_scan);
    cudaEvent_t e0, e1, e2;
    cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&e2);
    float ms_scan = 0, ms_sort = 0;
    const int iters = 50;
    for (int it = -5; it < iters; ++it) {
        cudaEventRecord(e0);
        cub::DeviceScan::InclusiveSum(t_scan, tmp_scan, c0, c1, P);
        cudaEventRecord(e1);
        cub::DeviceRadixSort::SortPairs(t_sort, tmp_sort, k0, k1, v0, v1, R, 0, 32 + bits);
        cudaEventRecord(e2);
        cudaEventSynchronize(e2);
        if (it >= 0) {
            float a, b;
            cudaEventElapsedTime(&a, e0, e1); cudaEventElapsedTime(&b, e1, e2);
            ms_scan += a; ms_sort += b;
        }
    }
    printf("{\"P\": %d, \"R\": %d, \"tiles\": %d, \"sort_bits\": %d, \"cub_inclusive_sum_ms\": %.4f, \"cub_sort_pairs_u64_u32_ms\": %.4f, \"error\": \"%s\"}\n",
           P, R, tiles, 32 + bits, ms_scan / iters, ms_sort / iters, cudaGetErrorString(cudaGetLastError()));
    return 0;
}
