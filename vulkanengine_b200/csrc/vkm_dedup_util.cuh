// Geometry::DeduplicateVertices (reference: engine/src/resources/geometry/geometry.cpp:84-126) on the GPU,
// with the reference's LITERAL semantics:
//   for i in order:  found_i  <=>  exists j < ucount_i with vertices[i] == vertices[j]   (epsilon ==, all 16 floats,
//                    compared against the INPUT array, first such j wins), ucount_{i+1} = ucount_i + !found_i;
//   a duplicate's indices become j, a kept vertex's indices become its rank among kept vertices.
// Decomposition: (1) m_i = smallest j < i that is epsilon-equal to i  -- data parallel (cell hash over the
// quantised position, 27 neighbour cells); (2) the recurrence found_i = (m_i < ucount_i) -- inherently
// sequential but one integer compare per vertex: a single warp walks it; (3) emit -- data parallel.
#pragma once

#include "vkm_device.cuh"

namespace vkm {

constexpr float kEps = 1.192092896e-07f;  // FLOAT_EPSILON, defines.hpp:67
constexpr uint32_t kNoMatch = 0xFFFFFFFFu;

__device__ __forceinline__ bool eps_equal16(const float4* __restrict__ a, const float4* __restrict__ b) {
    bool eq = true;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const float4 x = a[q], y = b[q];
        eq &= fabsf(__fsub_rn(x.x, y.x)) < kEps;
        eq &= fabsf(__fsub_rn(x.y, y.y)) < kEps;
        eq &= fabsf(__fsub_rn(x.z, y.z)) < kEps;
        eq &= fabsf(__fsub_rn(x.w, y.w)) < kEps;
    }
    return eq;  // NaN / Inf anywhere -> false, as in glm::epsilonEqual
}

// Cell coordinate of one position component.  Below 1024 the cell width is epsilon, so epsilon-equal values sit
// in the same or an adjacent cell; from 1024 up, distinct floats are further apart than epsilon and the cell is
// the value itself (flagged with bit 62 so the two ranges never meet).
__device__ __forceinline__ int64_t cell_of(float a, bool* wide) {
    if (fabsf(a) < 1024.0f) {
        *wide = true;
        return (int64_t)floor((double)a * 8388608.0);  // a / 2^-23, exact in double
    }
    *wide = false;
    return (int64_t)__float_as_uint(a) | (1ll << 62);
}

__device__ __forceinline__ uint32_t cell_bucket(int64_t x, int64_t y, int64_t z, uint32_t mask) {
    uint64_t h = mix64((uint64_t)x * 0x9E3779B97F4A7C15ull + 1);
    h = mix64(h ^ (uint64_t)y);
    h = mix64(h ^ (uint64_t)z);
    return (uint32_t)h & mask;
}

__device__ __forceinline__ bool finite3(const float4 p) {
    return !nonfinite_bits(__float_as_uint(p.x)) && !nonfinite_bits(__float_as_uint(p.y)) &&
           !nonfinite_bits(__float_as_uint(p.z));
}

__global__ void __launch_bounds__(kThreads) k_dd_build(const float4* __restrict__ vtx, uint32_t n, uint32_t* head,
                                                       uint32_t* next, uint32_t mask) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 p = vtx[4ull * i];
        next[i] = kNoMatch;
        if (!finite3(p)) continue;  // never equal to anything, itself included
        bool w;
        const uint32_t b = cell_bucket(cell_of(p.x, &w), cell_of(p.y, &w), cell_of(p.z, &w), mask);
        next[i] = atomicExch(head + b, i);
    }
}

__global__ void __launch_bounds__(kThreads) k_dd_match(const float4* __restrict__ vtx, uint32_t n,
                                                       const uint32_t* __restrict__ head,
                                                       const uint32_t* __restrict__ next, uint32_t mask, uint32_t* m) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 p = vtx[4ull * i];
        uint32_t best = kNoMatch;
        if (finite3(p)) {
            bool wx, wy, wz;
            const int64_t cx = cell_of(p.x, &wx), cy = cell_of(p.y, &wy), cz = cell_of(p.z, &wz);
            for (int dx = wx ? -1 : 0; dx <= (wx ? 1 : 0); ++dx)
                for (int dy = wy ? -1 : 0; dy <= (wy ? 1 : 0); ++dy)
                    for (int dz = wz ? -1 : 0; dz <= (wz ? 1 : 0); ++dz) {
                        const uint32_t b = cell_bucket(cx + dx, cy + dy, cz + dz, mask);
                        for (uint32_t j = head[b]; j != kNoMatch; j = next[j])
                            if (j < i && j < best && eps_equal16(vtx + 4ull * i, vtx + 4ull * j)) best = j;
                    }
        }
        m[i] = best;
    }
}

// One warp walks ucount_{i+1} = ucount_i + !(m_i < ucount_i); rank[i] = ucount_i, found[i] in bit 31.
// The recurrence is sequential, but for a chunk of 32 vertices that starts at count u0 almost every lane is decided
// without it: m < u0 is found whatever happens inside the chunk (the count only grows), m >= u0 + lane is not found
// whatever happens (the count grows by at most one per vertex).  Only when some lane falls in between does the warp
// replay that chunk one vertex at a time.  Four chunks of m are loaded ahead, so the walk is not paced by load latency.
__global__ void k_dd_sequence(const uint32_t* __restrict__ m, uint32_t n, uint32_t* rank_found, uint32_t* out_count) {
    const uint32_t lane = threadIdx.x;
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t u = 0;
    for (uint32_t base0 = 0; base0 < n; base0 += 128) {
        uint32_t mm[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint32_t i = base0 + 32 * q + lane;
            mm[q] = (i < n) ? m[i] : kNoMatch;
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint32_t base = base0 + 32 * q;
            if (base >= n) break;
            const uint32_t i = base + lane, mi = mm[q];
            const bool live = i < n;
            const bool surely_found = live && mi < u;
            const bool surely_new = live && (mi == kNoMatch || mi >= u + lane);
            uint32_t mine;
            if (__all_sync(0xFFFFFFFFu, !live || surely_found || surely_new)) {
                const uint32_t news = __ballot_sync(0xFFFFFFFFu, surely_new);
                mine = (u + __popc(news & lt)) | (surely_found ? 0x80000000u : 0u);
                u += __popc(news);
            } else {
                mine = 0;
                const uint32_t cnt = min(32u, n - base);
                for (uint32_t k = 0; k < cnt; ++k) {
                    const uint32_t mk = __shfl_sync(0xFFFFFFFFu, mi, k);
                    const bool found = mk < u;
                    if (lane == k) mine = u | (found ? 0x80000000u : 0u);
                    u += found ? 0u : 1u;
                }
            }
            if (live) rank_found[i] = mine;
        }
    }
    if (lane == 0) *out_count = u;
}

__global__ void __launch_bounds__(kThreads) k_dd_emit(const float4* __restrict__ vtx, uint32_t n,
                                                      const uint32_t* __restrict__ rank_found, float4* out) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t rf = rank_found[i];
        if (rf & 0x80000000u) continue;
#pragma unroll
        for (int q = 0; q < 4; ++q) out[4ull * rf + q] = vtx[4ull * i + q];
    }
}

__global__ void __launch_bounds__(kThreads) k_dd_remap(uint32_t* indices, uint32_t index_count, uint32_t n,
                                                       const uint32_t* __restrict__ rank_found,
                                                       const uint32_t* __restrict__ m,
                                                       const uint32_t* __restrict__ unique_count) {
    const uint32_t removed = n - *unique_count;
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < index_count; k += gridDim.x * blockDim.x) {
        const uint32_t x = indices[k];
        if (x < n) {
            const uint32_t rf = rank_found[x];
            indices[k] = (rf & 0x80000000u) ? m[x] : rf;  // duplicate -> matched INPUT index j (ReassignIndex to = j)
        } else {
            indices[k] = x - removed;  // beyond the array: only ever decremented, once per removed duplicate
        }
    }
}

}  // namespace vkm
