// Renderer handoff as a headless device buffer ("next" row f2).
//
// The reference hands finished geometry to the renderer through GeometrySystem::AcquireGeometryFromConfig ->
// VulkanRendererBackend::CreateGeometry -> UploadDataRange (geometry_system.cpp:136-176, vulkan.cpp:15-27,
// 691-725): every geometry gets a first-fit byte range (Freelist::AllocateBlock, freelist.cpp:144-159) in ONE
// global vertex buffer and ONE global index buffer, and is later drawn with vkCmdBindVertexBuffers(offset) /
// vkCmdBindIndexBuffer(offset, UINT32) (vulkan.cpp:664-672).  Here the two global buffers are plain device
// allocations; an import's shapes are copied device-to-device into first-fit ranges and the (offset, size) pairs
// are returned for the draw calls.  The Vulkan side (external-memory import of these buffers) is unavailable on
// this box and not built.
#include "../../include/vkmesh_b200.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <vector>

struct vkm_rangealloc {
    uint64_t total = 0;
    std::vector<std::pair<uint64_t, uint64_t>> free_ranges;  // (offset, size), sorted by offset, never adjacent
};

struct vkm_geometry_pool {
    int device = 0;
    void* d_vertices = nullptr;
    void* d_indices = nullptr;
    vkm_rangealloc* va = nullptr;
    vkm_rangealloc* ia = nullptr;
};

extern "C" {

vkm_rangealloc* vkm_rangealloc_create(uint64_t total_bytes) {
    vkm_rangealloc* a = new vkm_rangealloc();
    a->total = total_bytes;
    if (total_bytes) a->free_ranges.push_back({0, total_bytes});
    return a;
}
void vkm_rangealloc_destroy(vkm_rangealloc* a) { delete a; }

// First fit, lowest offset first -- the order Freelist::AllocateBlock walks its node list.  UINT64_MAX = no room.
uint64_t vkm_rangealloc_alloc(vkm_rangealloc* a, uint64_t size) {
    if (!a) return UINT64_MAX;
    for (size_t i = 0; i < a->free_ranges.size(); ++i) {
        auto& r = a->free_ranges[i];
        if (r.second >= size) {
            const uint64_t off = r.first;
            r.first += size;
            r.second -= size;
            if (r.second == 0) a->free_ranges.erase(a->free_ranges.begin() + i);
            return off;
        }
    }
    return UINT64_MAX;
}

// Returns a range and merges it with free neighbours (FreelistNode::FreeBlock, freelist.cpp:36-72).
int vkm_rangealloc_free(vkm_rangealloc* a, uint64_t offset, uint64_t size) {
    if (!a || offset + size > a->total) return VKM_ERR_INVALID_ARG;
    if (size == 0) return VKM_OK;
    auto it = std::lower_bound(a->free_ranges.begin(), a->free_ranges.end(), std::make_pair(offset, (uint64_t)0));
    if (it != a->free_ranges.end() && offset + size > it->first) return VKM_ERR_INVALID_ARG;  // overlaps a free range
    if (it != a->free_ranges.begin() && (it - 1)->first + (it - 1)->second > offset) return VKM_ERR_INVALID_ARG;
    it = a->free_ranges.insert(it, {offset, size});
    if (it + 1 != a->free_ranges.end() && it->first + it->second == (it + 1)->first) {
        it->second += (it + 1)->second;
        a->free_ranges.erase(it + 1);
    }
    if (it != a->free_ranges.begin() && (it - 1)->first + (it - 1)->second == it->first) {
        (it - 1)->second += it->second;
        a->free_ranges.erase(it);
    }
    return VKM_OK;
}

uint64_t vkm_rangealloc_free_space(const vkm_rangealloc* a) {
    uint64_t s = 0;
    if (a) for (auto& r : a->free_ranges) s += r.second;
    return s;
}

int vkm_pool_create(int device, uint64_t vertex_bytes, uint64_t index_bytes, vkm_geometry_pool** out) {
    if (!out) return VKM_ERR_INVALID_ARG;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || device >= n) {
        cudaGetLastError();
        return VKM_ERR_NO_DEVICE;
    }
    vkm_geometry_pool* p = new vkm_geometry_pool();
    p->device = device;
    cudaSetDevice(device);
    if (cudaMalloc(&p->d_vertices, std::max<uint64_t>(vertex_bytes, 1)) != cudaSuccess ||
        cudaMalloc(&p->d_indices, std::max<uint64_t>(index_bytes, 1)) != cudaSuccess) {
        cudaGetLastError();
        if (p->d_vertices) cudaFree(p->d_vertices);
        delete p;
        return VKM_ERR_OOM;
    }
    p->va = vkm_rangealloc_create(vertex_bytes);
    p->ia = vkm_rangealloc_create(index_bytes);
    *out = p;
    return VKM_OK;
}

void vkm_pool_destroy(vkm_geometry_pool* p) {
    if (!p) return;
    cudaSetDevice(p->device);
    cudaFree(p->d_vertices);
    cudaFree(p->d_indices);
    vkm_rangealloc_destroy(p->va);
    vkm_rangealloc_destroy(p->ia);
    delete p;
}

const void* vkm_pool_vertex_buffer(const vkm_geometry_pool* p) { return p ? p->d_vertices : nullptr; }
const void* vkm_pool_index_buffer(const vkm_geometry_pool* p) { return p ? p->d_indices : nullptr; }

// CreateGeometry for shape `shape` of a finished import: first-fit ranges + device-to-device copies on `stream`.
int vkm_pool_acquire(vkm_geometry_pool* p, const vkm_plan_result* r, const uint64_t* shape_corner_offsets, uint32_t shape,
                     vkm_pool_geometry* out, void* stream) {
    if (!p || !r || !shape_corner_offsets || !out) return VKM_ERR_INVALID_ARG;
    const uint64_t v0 = r->shape_vertex_offsets[shape], v1 = r->shape_vertex_offsets[shape + 1];
    const uint64_t c0 = shape_corner_offsets[shape], c1 = shape_corner_offsets[shape + 1];
    out->vertex_count = (uint32_t)(v1 - v0);
    out->index_count = (uint32_t)(c1 - c0);
    out->vertex_bytes = (v1 - v0) * 64;
    out->index_bytes = (c1 - c0) * 4;
    out->vertex_offset = vkm_rangealloc_alloc(p->va, out->vertex_bytes);
    if (out->vertex_offset == UINT64_MAX) return VKM_ERR_CAPACITY;
    out->index_offset = vkm_rangealloc_alloc(p->ia, out->index_bytes);
    if (out->index_offset == UINT64_MAX) {
        vkm_rangealloc_free(p->va, out->vertex_offset, out->vertex_bytes);
        return VKM_ERR_CAPACITY;
    }
    cudaSetDevice(p->device);
    cudaError_t e = cudaSuccess;
    if (out->vertex_bytes)
        e = cudaMemcpyAsync((char*)p->d_vertices + out->vertex_offset, (const char*)r->d_vertices + v0 * 64,
                            out->vertex_bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream);
    if (e == cudaSuccess && out->index_bytes)
        e = cudaMemcpyAsync((char*)p->d_indices + out->index_offset, (const char*)r->d_indices + c0 * 4, out->index_bytes,
                            cudaMemcpyDeviceToDevice, (cudaStream_t)stream);
    return e == cudaSuccess ? VKM_OK : VKM_ERR_CUDA;
}

int vkm_pool_release(vkm_geometry_pool* p, const vkm_pool_geometry* g) {
    if (!p || !g) return VKM_ERR_INVALID_ARG;
    int rc = vkm_rangealloc_free(p->va, g->vertex_offset, g->vertex_bytes);
    if (rc == VKM_OK) rc = vkm_rangealloc_free(p->ia, g->index_offset, g->index_bytes);
    return rc;
}

}  // extern "C"
