// Procedural generators ("next" row f4): GeometrySystem::GeneratePlainConfig and
// GeometrySystem::GenerateCubeConfig -- engine/src/systems/geometry/geometry_system.cpp:182-294 and :296-439.
// The plane is filled by one thread per segment with the reference's float expressions (no contraction); the
// cube's 24 vertices are the reference's closed-form table followed by GenerateNormals + GenerateTangents on the
// GPU (geometry_system.cpp:420-421).  The plane's TODO at :227 ("we can always deduplicate them later") names
// Geometry::DeduplicateVertices as the intended follow-up: `deduplicate != 0` runs it.
#include "../../include/vkmesh_b200.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>

extern "C" void vkm_internal_set_error(const char* msg);
extern "C" int vkm_internal_device(int* sm);  // vkmesh.cu

namespace {

__global__ void k_plane(float4* vtx, uint32_t* idx, uint32_t xs, uint32_t ys, float width, float height, float tile_x,
                        float tile_y) {
    const uint32_t n = xs * ys;
    const float seg_w = __fdiv_rn(width, (float)xs), seg_h = __fdiv_rn(height, (float)ys);
    const float half_w = __fmul_rn(width, 0.5f), half_h = __fmul_rn(height, 0.5f);
    for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < n; s += gridDim.x * blockDim.x) {
        const uint32_t x = s % xs, y = s / xs;
        const float min_x = __fsub_rn(__fmul_rn((float)x, seg_w), half_w);
        const float min_y = __fsub_rn(__fmul_rn((float)y, seg_h), half_h);
        const float max_x = __fadd_rn(min_x, seg_w), max_y = __fadd_rn(min_y, seg_h);
        const float min_u = __fmul_rn(__fdiv_rn((float)x, (float)xs), tile_x);
        const float min_v = __fmul_rn(__fdiv_rn((float)y, (float)ys), tile_y);
        const float max_u = __fmul_rn(__fdiv_rn((float)(x + 1), (float)xs), tile_x);
        const float max_v = __fmul_rn(__fdiv_rn((float)(y + 1), (float)ys), tile_y);
        const float px[4] = {min_x, max_x, min_x, max_x}, py[4] = {min_y, max_y, max_y, min_y};
        const float pu[4] = {min_u, max_u, min_u, max_u}, pv[4] = {min_v, max_v, max_v, min_v};
        const uint32_t v0 = s * 4;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float4* d = vtx + 4ull * (v0 + k);
            d[0] = make_float4(px[k], py[k], 0.f, 0.f);   // position.xyz, normal.x
            d[1] = make_float4(0.f, 0.f, pu[k], pv[k]);   // normal.yz, texcoord
            d[2] = make_float4(0.f, 0.f, 0.f, 0.f);       // color
            d[3] = make_float4(0.f, 0.f, 0.f, 0.f);       // tangent
        }
        uint32_t* o = idx + 6ull * s;
        o[0] = v0 + 0; o[1] = v0 + 1; o[2] = v0 + 2; o[3] = v0 + 0; o[4] = v0 + 3; o[5] = v0 + 1;
    }
}

void init_out(vkm_shape_out* out) {
    memset(out, 0, sizeof *out);
    out->vertex_size = 64;
    out->index_size = 4;
}

}  // namespace

extern "C" {

int vkm_generate_plane_config(float width, float height, uint32_t xs, uint32_t ys, float tile_x, float tile_y,
                              int deduplicate, vkm_shape_out* out) {
    if (!out) return VKM_ERR_INVALID_ARG;
    init_out(out);
    if (width == 0) width = 1.0f;   // the reference's defaults (geometry_system.cpp:187-211)
    if (height == 0) height = 1.0f;
    if (xs < 1) xs = 1;
    if (ys < 1) ys = 1;
    if (tile_x == 0) tile_x = 1.0f;
    if (tile_y == 0) tile_y = 1.0f;
    const uint64_t nseg = (uint64_t)xs * ys;
    if (nseg * 4 > 0x3FFFFFFFull) { vkm_internal_set_error("vkm_generate_plane_config: too many segments"); return VKM_ERR_INVALID_ARG; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        vkm_internal_set_error("no CUDA device (there is no CPU path)");
        return VKM_ERR_NO_DEVICE;
    }
    int sms = 148;
    const int dev = vkm_internal_device(&sms);  // the device selected with vkm_init()
    if (dev >= ndev || cudaSetDevice(dev) != cudaSuccess) {
        cudaGetLastError();
        vkm_internal_set_error("vkm_generate_plane_config: the device chosen with vkm_init is not available");
        return VKM_ERR_NO_DEVICE;
    }
    float4* d_v = nullptr;
    uint32_t* d_i = nullptr;
    int rc = VKM_OK;
    void* hv = malloc(nseg * 4 * 64);
    void* hi = malloc(nseg * 6 * 4);
    if (!hv || !hi) rc = VKM_ERR_OOM;
    if (rc == VKM_OK && (cudaMalloc(&d_v, nseg * 4 * 64) != cudaSuccess || cudaMalloc(&d_i, nseg * 6 * 4) != cudaSuccess)) rc = VKM_ERR_OOM;
    if (rc == VKM_OK) {
        k_plane<<<(unsigned)std::min<uint64_t>((nseg + 255) / 256, (uint64_t)sms * 8), 256>>>(d_v, d_i, xs, ys, width, height, tile_x, tile_y);
        if (cudaMemcpy(hv, d_v, nseg * 4 * 64, cudaMemcpyDeviceToHost) != cudaSuccess ||
            cudaMemcpy(hi, d_i, nseg * 6 * 4, cudaMemcpyDeviceToHost) != cudaSuccess) {
            vkm_internal_set_error(cudaGetErrorString(cudaGetLastError()));
            rc = VKM_ERR_CUDA;
        }
    }
    cudaFree(d_v);
    cudaFree(d_i);
    if (rc != VKM_OK) {
        free(hv);
        free(hi);
        return rc;
    }
    out->vertex_count = (uint32_t)(nseg * 4);
    out->index_count = (uint32_t)(nseg * 6);
    out->vertices = hv;
    out->indices = hi;
    if (deduplicate) {
        void* uniq = nullptr;
        const uint32_t n = vkm_deduplicate_vertices(out->vertex_count, out->vertices, out->index_count, (uint32_t*)out->indices, &uniq);
        if (n == 0xFFFFFFFFu) {
            vkm_free_shape(out);
            return VKM_ERR_CUDA;
        }
        free(out->vertices);
        out->vertices = uniq;
        out->vertex_count = n;
    }
    return VKM_OK;
}

int vkm_generate_cube_config(float width, float height, float depth, float tile_x, float tile_y, vkm_shape_out* out) {
    if (!out) return VKM_ERR_INVALID_ARG;
    init_out(out);
    if (width == 0) width = 1.0f;
    if (height == 0) height = 1.0f;
    if (depth == 0) depth = 1.0f;
    if (tile_x == 0) tile_x = 1.0f;
    if (tile_y == 0) tile_y = 1.0f;
    const float hw = width * 0.5f, hh = height * 0.5f, hd = depth * 0.5f;
    const float mnx = -hw, mny = -hh, mnz = -hd, mxx = hw, mxy = hh, mxz = hd;
    // four corners per face in the reference's order (geometry_system.cpp:347-404)
    const float P[6][4][3] = {
        {{mnx, mny, mxz}, {mxx, mxy, mxz}, {mnx, mxy, mxz}, {mxx, mny, mxz}},  // front
        {{mxx, mny, mnz}, {mnx, mxy, mnz}, {mxx, mxy, mnz}, {mnx, mny, mnz}},  // back
        {{mnx, mny, mnz}, {mnx, mxy, mxz}, {mnx, mxy, mnz}, {mnx, mny, mxz}},  // left
        {{mxx, mny, mxz}, {mxx, mxy, mnz}, {mxx, mxy, mxz}, {mxx, mny, mnz}},  // right
        {{mxx, mny, mxz}, {mnx, mny, mnz}, {mxx, mny, mnz}, {mnx, mny, mxz}},  // bottom
        {{mnx, mxy, mxz}, {mxx, mxy, mnz}, {mnx, mxy, mnz}, {mxx, mxy, mxz}},  // top
    };
    const float UV[4][2] = {{0.f, 0.f}, {tile_x, tile_y}, {0.f, tile_y}, {tile_x, 0.f}};
    vkm_vertex3d* v = (vkm_vertex3d*)calloc(24, sizeof(vkm_vertex3d));
    uint32_t* idx = (uint32_t*)malloc(36 * 4);
    if (!v || !idx) {
        free(v);
        free(idx);
        return VKM_ERR_OOM;
    }
    for (int f = 0; f < 6; ++f) {
        for (int k = 0; k < 4; ++k) {
            memcpy(v[4 * f + k].position, P[f][k], 12);
            memcpy(v[4 * f + k].texcoord, UV[k], 8);
        }
        const uint32_t o = 4 * f, t[6] = {o, o + 1, o + 2, o, o + 3, o + 1};
        memcpy(idx + 6 * f, t, 24);
    }
    int rc = vkm_generate_normals(24, v, 36, idx);
    if (rc == VKM_OK) rc = vkm_generate_tangents(24, v, 36, idx);
    if (rc != VKM_OK) {
        free(v);
        free(idx);
        return rc;
    }
    out->vertex_count = 24;
    out->index_count = 36;
    out->vertices = v;
    out->indices = idx;
    return VKM_OK;
}

}  // extern "C"
