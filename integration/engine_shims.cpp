// Engine-side bindings for libvkmesh_b200: the reference's own signatures, forwarding to the C ABI.
// This file is compiled INSIDE the engine tree (it includes the engine's headers, nothing is copied from them):
//
//   Engine::Geometry::GenerateTangents / GenerateNormals / DeduplicateVertices
//       replace the bodies in engine/src/resources/geometry/geometry.cpp:20-126
//   Engine::MeshLoader::LoadOBJ
//       replaces the body in engine/src/systems/resource/loaders/mesh/mesh_loader.cpp:200-328; tinyobj still
//       parses the text, the per-shape loop (:228-325) becomes one vkm_import_shapes call.
//
// oracle/Makefile (target `dropin`) builds a demo that links this file in front of the unmodified reference
// objects: the engine's own MeshLoader::Load -> (this) LoadOBJ -> CUDA -> the engine's own WriteToE3DM.
#include <cstring>
#include <vector>

#include "core/logger/logger.hpp"
#include "core/utils/string.hpp"
#include "resources/geometry/geometry.hpp"
#include "systems/resource/loaders/mesh/mesh_loader.hpp"
#include "systems/resource/resources/mesh/mesh_resource.hpp"
#include "vendor/tinyobj/tinyobj.hpp"
#include "vkmesh_b200.h"

namespace Engine {

static_assert(sizeof(Vertex3D) == sizeof(vkm_vertex3d), "Vertex3D is the 64-byte interleaved layout");
static_assert(sizeof(GeometryExtent) == sizeof(vkm_extent), "GeometryExtent is 9 packed floats");

void Geometry::GenerateTangents(u32 vertex_count, Vertex3D* vertices, u32 index_count, u32* indices) {
    if (vkm_generate_tangents(vertex_count, vertices, index_count, indices) != VKM_OK) {
        ERROR("Geometry::GenerateTangents - %s", vkm_last_error());
    }
}

void Geometry::GenerateNormals(u32 vertex_count, Vertex3D* vertices, u32 index_count, u32* indices) {
    if (vkm_generate_normals(vertex_count, vertices, index_count, indices) != VKM_OK) {
        ERROR("Geometry::GenerateNormals - %s", vkm_last_error());
    }
}

u32 Geometry::DeduplicateVertices(u32 vertex_count, Vertex3D* vertices, u32 index_count, u32* indices,
                                  Vertex3D** out_vertices) {
    void* out = nullptr;  // malloc()ed, like Platform::AllocMemory
    const u32 n = vkm_deduplicate_vertices(vertex_count, vertices, index_count, indices, &out);
    *out_vertices = static_cast<Vertex3D*>(out);
    return n;
}

MeshResource* MeshLoader::LoadOBJ(const std::string& file_path, const std::string& name) {
    tinyobj::attrib_t attributes;
    std::vector<tinyobj::shape_t> shapes;
    std::vector<tinyobj::material_t> materials;
    std::string warnings, errors;
    const std::string directory = GetFullDirectoryFromPath(file_path);
    if (!tinyobj::LoadObj(&attributes, &shapes, &materials, &warnings, &errors, file_path.c_str(), directory.c_str())) {
        ERROR("MeshLoader::LoadOBJ - failed to load '%s'", file_path.c_str());
        return nullptr;
    }

    static_assert(sizeof(tinyobj::index_t) == 3 * sizeof(int32_t), "index_t is (vertex, normal, texcoord)");
    std::vector<tinyobj::index_t> corners;
    std::vector<uint64_t> offsets{0};
    for (const auto& shape : shapes) {
        corners.insert(corners.end(), shape.mesh.indices.begin(), shape.mesh.indices.end());
        offsets.push_back(corners.size());
    }
    std::vector<vkm_shape_out> out(shapes.size());
    const int rc = vkm_import_shapes(attributes.vertices.data(), attributes.vertices.size() / 3, attributes.normals.data(),
                                     attributes.normals.size() / 3, attributes.texcoords.data(),
                                     attributes.texcoords.size() / 2, reinterpret_cast<const int32_t*>(corners.data()),
                                     offsets.data(), static_cast<uint32_t>(shapes.size()), out.data());
    if (rc != VKM_OK) {
        ERROR("MeshLoader::LoadOBJ - %s", vkm_last_error());
        return nullptr;
    }

    GeometryConfigs configs;
    configs.reserve(shapes.size());
    for (u32 i = 0; i < shapes.size(); ++i) {
        GeometryConfig config{};
        config.vertex_size = out[i].vertex_size;
        config.vertex_count = out[i].vertex_count;
        config.vertices = out[i].vertices;  // malloc()ed: ~MeshResource frees them
        config.index_size = out[i].index_size;
        config.index_count = out[i].index_count;
        config.indices = out[i].indices;
        std::memcpy(&config.extent, &out[i].extent, sizeof(GeometryExtent));
        const u32 current_material =
            shapes[i].mesh.material_ids.empty() ? INVALID_ID : static_cast<u32>(shapes[i].mesh.material_ids[0]);
        if (current_material != INVALID_ID) config.material_name = materials[current_material].name;
        config.name = shapes[i].name;
        configs.push_back(config);
    }
    return new MeshResource(id, name, file_path, configs);
}

}  // namespace Engine
