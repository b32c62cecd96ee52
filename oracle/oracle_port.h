/* TEST INFRASTRUCTURE ONLY (oracle).  Never linked into, imported by or executed from the
 * product path; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may use it, and only as the checker.
 *
 * Plain-C restatement of the reference's mesh-import algorithm.  Parity status: PINNED --
 * checked against the reference-shipped known-answer file assets/models/falcon.e3dm
 * (through tests/golden/) and against oracle/_ref/libvkref.so (the reference's own sources
 * compiled unmodified) on every fixture and seeded synthetic mesh in tests/.
 */
#ifndef ORACLE_PORT_H
#define ORACLE_PORT_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Engine::Vertex3D, renderer/renderer_types.hpp:13-19: 16 packed floats, 64 bytes. */
typedef struct orc_vertex {
    float position[3];
    float normal[3];
    float texcoord[2];
    float color[4];
    float tangent[4];
} orc_vertex;

/* Engine::GeometryExtent, mesh_resource.hpp:8-12. */
typedef struct orc_extent {
    float center[3], min_extents[3], max_extents[3];
} orc_extent;

typedef struct orc_shape {
    uint32_t vertex_count;
    uint32_t index_count;
    orc_vertex* vertices; /* malloc'd */
    uint32_t* indices;    /* malloc'd */
    orc_extent extent;
} orc_shape;

/* std::hash<Vertex3D>, renderer_types.hpp:72-86, over libstdc++'s std::hash<float>. */
uint64_t orc_hash_vertex(const orc_vertex* v);
/* Vertex3D::operator==, renderer_types.hpp:20-27 (glm::epsilonEqual, FLOAT_EPSILON). */
int orc_vertex_equal(const orc_vertex* a, const orc_vertex* b);

/* Per-shape loop of MeshLoader::LoadOBJ, mesh_loader.cpp:228-325.
 * idx_vnt: 3 ints per corner in tinyobj::index_t order (vertex, normal, texcoord).
 * shape_offsets: nshapes+1 corner offsets. Returns 0, or -1 on an out-of-range index
 * (undefined behaviour in the reference). */
int orc_import_shapes(const float* pos, uint64_t np, const float* nrm, uint64_t nn, const float* uv, uint64_t nt,
                      const int32_t* idx_vnt, const uint64_t* shape_offsets, uint32_t nshapes, orc_shape* out);
void orc_free_shape(orc_shape* s);

/* Geometry::GenerateTangents / GenerateNormals / DeduplicateVertices, geometry.cpp:20-126. */
void orc_generate_tangents(uint32_t vertex_count, orc_vertex* vertices, uint32_t index_count, const uint32_t* indices);
void orc_generate_normals(uint32_t vertex_count, orc_vertex* vertices, uint32_t index_count, const uint32_t* indices);
uint32_t orc_deduplicate_vertices(uint32_t vertex_count, const orc_vertex* vertices, uint32_t index_count,
                                  uint32_t* indices, orc_vertex** out_vertices);
void orc_free(void* p);

/* GeometrySystem::GeneratePlainConfig / GenerateCubeConfig, geometry_system.cpp:182-294 / :296-439 (buffers only).
 * vertices: 4*xs*ys (plane) or 24 (cube) orc_vertex; indices: 6*xs*ys or 36 u32; caller allocates. */
void orc_generate_plane(float width, float height, uint32_t xs, uint32_t ys, float tile_x, float tile_y,
                        orc_vertex* vertices, uint32_t* indices);
void orc_generate_cube(float width, float height, float depth, float tile_x, float tile_y, orc_vertex* vertices,
                       uint32_t* indices);

#ifdef __cplusplus
}
#endif
#endif
