// TEST INFRASTRUCTURE ONLY (oracle). Not part of the shipped product path.
//
// Plain-C driver around the UNMODIFIED reference mesh-import sources.  The
// reference translation units are compiled where they lie under
// /root/reference by oracle/Makefile; nothing from them is copied here.  This
// file only (a) exposes their entry points through `extern "C"` so that tests
// and bench.py's cpu_baseline leg can call them through ctypes, and (b) lets
// the reference's own MeshLoader::LoadOBJ loop
// (engine/src/systems/resource/loaders/mesh/mesh_loader.cpp:200-328) run on
// in-memory tinyobj-shaped arrays: the library is linked with
// `--wrap=<mangled tinyobj::LoadObj>`, and the wrapper below serves file names
// that start with "mem:" from a thread-local injected mesh instead of parsing
// OBJ text.  Every other file name is forwarded to the real tinyobj parser.
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "resources/geometry/geometry.hpp"
#include "systems/geometry/geometry_system.hpp"
#include "core/utils/freelist.hpp"
#include "systems/resource/loaders/mesh/mesh_loader.hpp"
#include "systems/resource/resources/mesh/mesh_resource.hpp"
#include "vendor/tinyobj/tinyobj.hpp"

namespace {

struct InjectedMesh {
    const float* pos; uint64_t np;
    const float* nrm; uint64_t nn;
    const float* uv;  uint64_t nt;
    const int32_t* idx;              // 3 ints per corner: v, n, t (tinyobj::index_t order)
    const uint64_t* shape_offsets;   // nshapes + 1 corner offsets
    uint32_t nshapes;
};

thread_local const InjectedMesh* g_injected = nullptr;

// MeshLoader keeps LoadOBJ / LoadE3DM / WriteToE3DM protected (mesh_loader.hpp:45-47).
struct OpenLoader : Engine::MeshLoader {
    OpenLoader() : Engine::MeshLoader(0u, std::string(".")) {}
    using Engine::MeshLoader::LoadE3DM;
    using Engine::MeshLoader::LoadOBJ;
    using Engine::MeshLoader::WriteToE3DM;
};

struct ParsedObj {
    tinyobj::attrib_t attrib;
    std::vector<tinyobj::shape_t> shapes;
    std::vector<tinyobj::material_t> materials;
};

void copy_str(const std::string& s, char* out, int cap) {
    if (!out || cap <= 0) return;
    int n = (int)s.size() < cap - 1 ? (int)s.size() : cap - 1;
    std::memcpy(out, s.data(), n);
    out[n] = 0;
}

}  // namespace

// ---- tinyobj::LoadObj interposer (see oracle/Makefile: -Wl,--wrap=...) ----
extern "C" bool VKREF_REAL_LOADOBJ(tinyobj::attrib_t*, std::vector<tinyobj::shape_t>*,
                                   std::vector<tinyobj::material_t>*, std::string*, std::string*,
                                   const char*, const char*, bool, bool);

extern "C" bool VKREF_WRAP_LOADOBJ(tinyobj::attrib_t* attrib, std::vector<tinyobj::shape_t>* shapes,
                                   std::vector<tinyobj::material_t>* materials, std::string* warn,
                                   std::string* err, const char* filename, const char* mtl_basedir,
                                   bool triangulate, bool default_vcols_fallback) {
    if (filename && std::strncmp(filename, "mem:", 4) == 0 && g_injected) {
        const InjectedMesh& m = *g_injected;
        attrib->vertices.assign(m.pos, m.pos + 3 * m.np);
        attrib->normals.assign(m.nrm, m.nrm + 3 * m.nn);
        attrib->texcoords.assign(m.uv, m.uv + 2 * m.nt);
        shapes->clear();
        shapes->resize(m.nshapes);
        static_assert(sizeof(tinyobj::index_t) == 12, "index_t is three ints");
        for (uint32_t s = 0; s < m.nshapes; ++s) {
            uint64_t c0 = m.shape_offsets[s], c1 = m.shape_offsets[s + 1];
            tinyobj::shape_t& sh = (*shapes)[s];
            sh.name = "shape" + std::to_string(s);
            const tinyobj::index_t* first = reinterpret_cast<const tinyobj::index_t*>(m.idx) + c0;
            sh.mesh.indices.assign(first, first + (c1 - c0));
            sh.mesh.num_face_vertices.assign((c1 - c0) / 3, 3u);
            sh.mesh.material_ids.assign((c1 - c0) / 3 ? (c1 - c0) / 3 : 1, -1);
        }
        return true;
    }
    return VKREF_REAL_LOADOBJ(attrib, shapes, materials, warn, err, filename, mtl_basedir, triangulate,
                              default_vcols_fallback);
}

extern "C" {

// ---------------- OBJ text -> tinyobj arrays (the hot path's input) ----------------
void* ref_parse_obj(const char* path, const char* mtl_dir) {
    ParsedObj* p = new ParsedObj();
    std::string warn, err;
    if (!tinyobj::LoadObj(&p->attrib, &p->shapes, &p->materials, &warn, &err, path, mtl_dir)) {
        delete p;
        return nullptr;
    }
    return p;
}
void ref_parsed_sizes(void* h, uint64_t out[4]) {
    ParsedObj* p = (ParsedObj*)h;
    out[0] = p->attrib.vertices.size() / 3;
    out[1] = p->attrib.normals.size() / 3;
    out[2] = p->attrib.texcoords.size() / 2;
    out[3] = p->shapes.size();
}
void ref_parsed_attribs(void* h, float* pos, float* nrm, float* uv) {
    ParsedObj* p = (ParsedObj*)h;
    std::memcpy(pos, p->attrib.vertices.data(), p->attrib.vertices.size() * 4);
    std::memcpy(nrm, p->attrib.normals.data(), p->attrib.normals.size() * 4);
    std::memcpy(uv, p->attrib.texcoords.data(), p->attrib.texcoords.size() * 4);
}
uint64_t ref_parsed_shape_info(void* h, uint32_t i, char* name, int name_cap, int* material0, char* mat,
                               int mat_cap) {
    ParsedObj* p = (ParsedObj*)h;
    const tinyobj::shape_t& s = p->shapes[i];
    copy_str(s.name, name, name_cap);
    int m0 = s.mesh.material_ids.empty() ? -1 : s.mesh.material_ids[0];
    if (material0) *material0 = m0;
    copy_str(m0 >= 0 && (size_t)m0 < p->materials.size() ? p->materials[m0].name : std::string(), mat, mat_cap);
    return s.mesh.indices.size();
}
void ref_parsed_shape_indices(void* h, uint32_t i, int32_t* idx_vnt) {
    ParsedObj* p = (ParsedObj*)h;
    const auto& v = p->shapes[i].mesh.indices;
    std::memcpy(idx_vnt, v.data(), v.size() * sizeof(tinyobj::index_t));
}
void ref_parsed_free(void* h) { delete (ParsedObj*)h; }

// ---------------- the reference import path ----------------
void* ref_load_obj(const char* path, const char* name) {
    OpenLoader l;
    return l.LoadOBJ(path, name);
}

void* ref_import_arrays(const float* pos, uint64_t np, const float* nrm, uint64_t nn, const float* uv,
                        uint64_t nt, const int32_t* idx_vnt, const uint64_t* shape_offsets, uint32_t nshapes) {
    InjectedMesh m{pos, np, nrm, nn, uv, nt, idx_vnt, shape_offsets, nshapes};
    g_injected = &m;
    OpenLoader l;
    // The path needs a '/': GetFullDirectoryFromPath (core/utils/string.cpp:208-215) never terminates without one.
    Engine::MeshResource* r = l.LoadOBJ("mem:/injected", "injected");
    g_injected = nullptr;
    return r;
}

uint64_t ref_mesh_config_count(void* h) { return ((Engine::MeshResource*)h)->GetConfigs().size(); }

void ref_mesh_config(void* h, uint64_t i, uint32_t* vertex_count, uint32_t* index_count, void** vertices,
                     void** indices, float extent9[9], char* name, int name_cap, char* mat, int mat_cap) {
    Engine::GeometryConfig& c = ((Engine::MeshResource*)h)->GetConfigs()[i];
    *vertex_count = c.vertex_count;
    *index_count = c.index_count;
    *vertices = c.vertices;
    *indices = c.indices;
    std::memcpy(extent9, &c.extent, 36);
    copy_str(c.name, name, name_cap);
    copy_str(c.material_name, mat, mat_cap);
}
void ref_mesh_free(void* h) { delete (Engine::MeshResource*)h; }

int ref_write_e3dm(void* h, const char* path, const char* name) {
    OpenLoader l;
    return l.WriteToE3DM(path, name, ((Engine::MeshResource*)h)->GetConfigs()) ? 1 : 0;
}
void* ref_load_e3dm(const char* path, const char* name) {
    OpenLoader l;
    return l.LoadE3DM(path, name);
}

// ---------------- geometry utilities (geometry.cpp:20-126) ----------------
void ref_generate_tangents(uint32_t vc, void* v, uint32_t ic, uint32_t* idx) {
    Engine::Geometry::GenerateTangents(vc, (Engine::Vertex3D*)v, ic, idx);
}
void ref_generate_normals(uint32_t vc, void* v, uint32_t ic, uint32_t* idx) {
    Engine::Geometry::GenerateNormals(vc, (Engine::Vertex3D*)v, ic, idx);
}
uint32_t ref_deduplicate_vertices(uint32_t vc, void* v, uint32_t ic, uint32_t* idx, void** out_v) {
    Engine::Vertex3D* out = nullptr;
    uint32_t n = Engine::Geometry::DeduplicateVertices(vc, (Engine::Vertex3D*)v, ic, idx, &out);
    *out_v = out;
    return n;
}
void ref_free(void* p) { std::free(p); }

// ---------------- procedural generators (geometry_system.cpp:182-439) ----------------
// GeneratePlainConfig / GenerateCubeConfig are non-static members that never touch the object; constructing a real
// GeometrySystem would call into the renderer, so they are invoked on raw, suitably aligned storage.
static Engine::GeometrySystem* fake_geometry_system() {
    alignas(Engine::GeometrySystem) static unsigned char storage[sizeof(Engine::GeometrySystem)];
    return reinterpret_cast<Engine::GeometrySystem*>(storage);
}
static void hand_over(Engine::GeometryConfig& c, uint32_t* vc, uint32_t* ic, void** v, void** idx) {
    *vc = c.vertex_count;
    *ic = c.index_count;
    *v = c.vertices;   // malloc()ed by Platform::AllocMemory; release with ref_free
    *idx = c.indices;
}
void ref_generate_plane(float width, float height, uint32_t xs, uint32_t ys, float tile_x, float tile_y, uint32_t* vc,
                        uint32_t* ic, void** v, void** idx) {
    Engine::GeometryConfig c = fake_geometry_system()->GeneratePlainConfig(width, height, xs, ys, tile_x, tile_y, "p", "m");
    hand_over(c, vc, ic, v, idx);
}
void ref_generate_cube(float width, float height, float depth, float tile_x, float tile_y, uint32_t* vc, uint32_t* ic,
                       void** v, void** idx) {
    Engine::GeometryConfig c = fake_geometry_system()->GenerateCubeConfig(width, height, depth, tile_x, tile_y, "c", "m");
    hand_over(c, vc, ic, v, idx);
}

// ---------------- Freelist (core/utils/freelist.cpp): the first-fit allocator behind the renderer's buffers ----------------
void* ref_freelist_create(uint64_t total) { return new Engine::Freelist(total); }
void ref_freelist_destroy(void* f) { delete (Engine::Freelist*)f; }
uint64_t ref_freelist_alloc(void* f, uint64_t size) {  // Freelist::AllocateBlock (:144-164): offset, or UINT64_MAX
    Engine::FreelistNode* n = ((Engine::Freelist*)f)->AllocateBlock(size);
    return n ? n->GetMemoryOffset() : UINT64_MAX;
}
int ref_freelist_free(void* f, uint64_t offset) { return ((Engine::Freelist*)f)->FreeByOffset(offset) ? 1 : 0; }
uint64_t ref_freelist_free_space(void* f) { return ((Engine::Freelist*)f)->FreeSpace(); }

// ---------------- std::hash<Vertex3D> / operator== (renderer_types.hpp:20-27,72-86) ----------------
uint64_t ref_hash_vertex(const void* v64) { return std::hash<Engine::Vertex3D>()(*(const Engine::Vertex3D*)v64); }
int ref_vertex_equal(const void* a, const void* b) {
    return (*(const Engine::Vertex3D*)a == *(const Engine::Vertex3D*)b) ? 1 : 0;
}
uint32_t ref_sizeof_vertex(void) { return (uint32_t)sizeof(Engine::Vertex3D); }

}  // extern "C"
