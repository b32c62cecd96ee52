// Measurement only (not part of the library): how long does the radix-sort formulation's FIRST step take on B200?
// north_star names "a warp-cooperative hash or radix-sort pass over packed vertex keys"; for an incoherent corner stream
// the sort would have to order 50 M (key record, corner id) pairs before anything else can run coherently.  This times
// the vendor's tuned LSD radix sort (cub::DeviceRadixSort::SortPairs, 32-bit keys limited to the bits the key needs) on
// that many random pairs; the whole record schedule on the permuted 16.8 M-triangle mesh takes 1.99 ms, its insert pass
// 0.93 ms.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o sort_probe sort_probe.cu && ./sort_probe
#include <cub/cub.cuh>
#include <cstdio>
#include <cstdint>
#include <vector>

__global__ void fill(uint32_t* k, uint32_t* v, uint32_t n, uint32_t range) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        uint32_t x = i * 2654435761u;
        x ^= x >> 15; x *= 0x2c1b3c6du; x ^= x >> 12; x *= 0x297a2d39u; x ^= x >> 15;
        k[i] = (uint32_t)(((uint64_t)x * range) >> 32);
        v[i] = i;
    }
}

int main() {
    const uint32_t n = 50320896u, range = 8392609u;  // corners and key records of the 16.8 M-triangle grid
    uint32_t *k0, *k1, *v0, *v1;
    cudaMalloc(&k0, 4ull * n); cudaMalloc(&k1, 4ull * n); cudaMalloc(&v0, 4ull * n); cudaMalloc(&v1, 4ull * n);
    fill<<<148 * 8, 256>>>(k0, v0, n, range);
    size_t tmp_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, k0, k1, v0, v1, n, 0, 24);
    void* tmp;
    cudaMalloc(&tmp, tmp_bytes);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int bits : {24}) {  // 2^24 > 8 392 609 key records
        cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, k0, k1, v0, v1, n, 0, bits);  // warm-up
        cudaEventRecord(e0);
        const int reps = 10;
        for (int r = 0; r < reps; ++r) cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, k0, k1, v0, v1, n, 0, bits);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        printf("{\"pairs\": %u, \"key_bits\": %d, \"ms_per_sort\": %.4f, \"gpairs_per_s\": %.2f}\n", n, bits, ms / reps, n / (ms / reps) / 1e6);
    }
    return cudaGetLastError() != cudaSuccess;
}
