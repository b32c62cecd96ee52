// E3DM import cache (host-side file format code; no device work).
// Replaces MeshLoader::WriteToE3DM / LoadE3DM -- engine/src/systems/resource/loaders/mesh/mesh_loader.cpp:63-126 /
// :128-198.  Layout, little-endian and unpadded: u16 version = 1; u32 name_len; name; u64 n_configs; then per
// config: u32 vertex_count; u32 vertex_size; vertex bytes; u32 index_count; u32 index_size; index bytes;
// u32 len; config name; u32 len; material name.  The extent is not stored (the reference leaves it
// uninitialised after LoadE3DM; here it is zeroed).
#include "../../include/vkmesh_b200.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace {
bool put(FILE* f, const void* p, size_t n) { return n == 0 || fwrite(p, 1, n, f) == n; }
bool get(FILE* f, void* p, size_t n) { return n == 0 || fread(p, 1, n, f) == n; }
char* get_string(FILE* f) {
    uint32_t len = 0;
    if (!get(f, &len, 4)) return nullptr;
    char* s = (char*)malloc((size_t)len + 1);
    if (!s || !get(f, s, len)) {
        free(s);
        return nullptr;
    }
    s[len] = 0;
    return s;
}
}  // namespace

extern "C" {

int vkm_e3dm_write(const char* path, const char* mesh_name, const vkm_e3dm_config* configs, uint64_t count) {
    if (!path || !mesh_name || (count && !configs)) return VKM_ERR_INVALID_ARG;
    FILE* f = fopen(path, "wb");
    if (!f) return VKM_ERR_INVALID_ARG;
    bool ok = true;
    const uint16_t version = 0x0001;
    const uint32_t name_len = (uint32_t)strlen(mesh_name);
    ok &= put(f, &version, 2) && put(f, &name_len, 4) && put(f, mesh_name, name_len) && put(f, &count, 8);
    for (uint64_t i = 0; ok && i < count; ++i) {
        const vkm_shape_out& s = configs[i].shape;
        const char* cn = configs[i].name ? configs[i].name : "";
        const char* mn = configs[i].material_name ? configs[i].material_name : "";
        const uint32_t cl = (uint32_t)strlen(cn), ml = (uint32_t)strlen(mn);
        ok &= put(f, &s.vertex_count, 4) && put(f, &s.vertex_size, 4) &&
              put(f, s.vertices, (size_t)s.vertex_count * s.vertex_size) && put(f, &s.index_count, 4) &&
              put(f, &s.index_size, 4) && put(f, s.indices, (size_t)s.index_count * s.index_size) && put(f, &cl, 4) &&
              put(f, cn, cl) && put(f, &ml, 4) && put(f, mn, ml);
    }
    ok &= (fclose(f) == 0);
    return ok ? VKM_OK : VKM_ERR_INVALID_ARG;
}

void vkm_e3dm_free(vkm_e3dm_config* configs, uint64_t count) {
    if (!configs) return;
    for (uint64_t i = 0; i < count; ++i) {
        free(configs[i].shape.vertices);
        free(configs[i].shape.indices);
        free((void*)configs[i].name);
        free((void*)configs[i].material_name);
    }
    free(configs);
}

int vkm_e3dm_read(const char* path, char* mesh_name, int mesh_name_cap, vkm_e3dm_config** out_configs,
                  uint64_t* out_count) {
    if (!path || !out_configs || !out_count) return VKM_ERR_INVALID_ARG;
    *out_configs = nullptr;
    *out_count = 0;
    FILE* f = fopen(path, "rb");
    if (!f) return VKM_ERR_INVALID_ARG;
    uint16_t version = 0;
    uint64_t count = 0;
    bool ok = get(f, &version, 2);
    char* name = ok ? get_string(f) : nullptr;
    ok = ok && name && get(f, &count, 8) && count < (1u << 24);
    if (ok && mesh_name && mesh_name_cap > 0) {
        strncpy(mesh_name, name, (size_t)mesh_name_cap - 1);
        mesh_name[mesh_name_cap - 1] = 0;
    }
    free(name);
    vkm_e3dm_config* cfg = ok ? (vkm_e3dm_config*)calloc(count ? count : 1, sizeof(vkm_e3dm_config)) : nullptr;
    ok = ok && cfg;
    for (uint64_t i = 0; ok && i < count; ++i) {
        vkm_shape_out& s = cfg[i].shape;
        ok = get(f, &s.vertex_count, 4) && get(f, &s.vertex_size, 4);
        const size_t vb = ok ? (size_t)s.vertex_count * s.vertex_size : 0;
        s.vertices = ok ? malloc(vb ? vb : 1) : nullptr;
        ok = ok && s.vertices && get(f, s.vertices, vb) && get(f, &s.index_count, 4) && get(f, &s.index_size, 4);
        const size_t ib = ok ? (size_t)s.index_count * s.index_size : 0;
        s.indices = ok ? malloc(ib ? ib : 1) : nullptr;
        ok = ok && s.indices && get(f, s.indices, ib);
        if (ok) ok = (cfg[i].name = get_string(f)) != nullptr;
        if (ok) ok = (cfg[i].material_name = get_string(f)) != nullptr;
    }
    fclose(f);
    if (!ok) {
        vkm_e3dm_free(cfg, count);
        return VKM_ERR_INVALID_ARG;
    }
    *out_configs = cfg;
    *out_count = count;
    return VKM_OK;
}

}  // extern "C"
