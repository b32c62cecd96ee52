/*
 * vkmesh_b200 -- B200-native (CUDA sm_100a) implementation of everdoze/VulkanEngine's mesh-import
 * geometry path, behind a C ABI.  No CPU fallback: every entry point fails with VKM_ERR_NO_DEVICE
 * when there is no CUDA device.
 *
 * Each entry point cites the reference interface it replaces (paths relative to the reference
 * root).  INTEGRATION.md shows the engine-side binding.
 */
#ifndef VKMESH_B200_H
#define VKMESH_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- status codes (the reference signals failure with nullptr / false + a log line,
 *      mesh_loader.cpp:55-58, 219-222; the shims in INTEGRATION.md map non-zero to that) ---- */
#define VKM_OK 0
#define VKM_ERR_INVALID_ARG 1 /* null pointer, corner count not a multiple of 3, sizes beyond 32-bit limits */
#define VKM_ERR_BAD_INDEX 2   /* a v/vn/vt index is negative or out of range (undefined behaviour in the reference) */
#define VKM_ERR_OOM 3         /* host or device allocation failed */
#define VKM_ERR_CUDA 4        /* a CUDA call failed; see vkm_last_error() */
#define VKM_ERR_NO_DEVICE 5   /* no CUDA device: there is no CPU path */
#define VKM_ERR_CAPACITY 6    /* a plan's fixed capacity was exceeded */
#define VKM_ERR_UNSUPPORTED 7 /* OBJ construct that is left to the caller's CPU parser */

/* Engine::Vertex3D -- engine/src/renderer/renderer_types.hpp:13-19 (64 bytes, packed floats). */
typedef struct vkm_vertex3d {
    float position[3];
    float normal[3];
    float texcoord[2];
    float color[4];
    float tangent[4];
} vkm_vertex3d;

/* Engine::GeometryExtent -- engine/src/systems/resource/resources/mesh/mesh_resource.hpp:8-12. */
typedef struct vkm_extent {
    float center[3];
    float min_extents[3];
    float max_extents[3];
} vkm_extent;

/* The buffer part of Engine::GeometryConfig -- mesh_resource.hpp:14-24.  `vertices` and `indices` are
 * malloc() memory, as the reference's are (Platform::AllocMemory, mesh_loader.cpp:303-304), so the
 * engine's ~MeshResource / DisposeConfig can free() them (mesh_resource.cpp:13-18). */
typedef struct vkm_shape_out {
    uint32_t vertex_size;  /* 64 */
    uint32_t vertex_count;
    void* vertices;
    uint32_t index_size;   /* 4 */
    uint32_t index_count;
    void* indices;
    vkm_extent extent;
} vkm_shape_out;

/* ---- library state ---- */
int vkm_device_count(void);
/* Selects the CUDA devices the host entry points may use (default: device 0).  Optional. */
int vkm_init(const int* device_ordinals, int count);
void vkm_shutdown(void);
/* Thread-local message of the last failing call on this thread. */
const char* vkm_last_error(void);
const char* vkm_version(void);

/* ---- host entry points: drop-in for the reference's CPU functions (host pointers in, host pointers out) ---- */

/* Replaces static void Geometry::GenerateTangents(u32 vertex_count, Vertex3D* vertices, u32 index_count,
 * u32* indices) -- engine/src/resources/geometry/geometry.cpp:20-54 (declared geometry.hpp:49).
 * In place: writes vertices[i].tangent for every vertex some triangle references; the triangle with the
 * highest number wins, as in the reference's sequential loop. */
int vkm_generate_tangents(uint32_t vertex_count, void* vertices, uint32_t index_count, const uint32_t* indices);

/* Replaces static void Geometry::GenerateNormals(...) -- geometry.cpp:56-82 (geometry.hpp:50).  Flat face
 * normals, last triangle wins. */
int vkm_generate_normals(uint32_t vertex_count, void* vertices, uint32_t index_count, const uint32_t* indices);

/* Replaces static u32 Geometry::DeduplicateVertices(u32 vertex_count, Vertex3D* vertices, u32 index_count,
 * u32* indices, Vertex3D** out_vertices) -- geometry.cpp:94-126 (geometry.hpp:51).  Rewrites `indices` in
 * place, returns the unique count and a malloc()ed *out_vertices.  Returns UINT32_MAX on failure. */
uint32_t vkm_deduplicate_vertices(uint32_t vertex_count, const void* vertices, uint32_t index_count,
                                  uint32_t* indices, void** out_vertices);

/* Replaces the per-shape loop of MeshResource* MeshLoader::LoadOBJ(const std::string&, const std::string&)
 * -- engine/src/systems/resource/loaders/mesh/mesh_loader.cpp:228-325: face expansion, first-occurrence
 * vertex dedup, bounding box and tangent generation for every shape tinyobj produced.
 *   pos / nrm / uv  : tinyobj::attrib_t::vertices / normals / texcoords (tinyobj.hpp:387-398)
 *   idx_vnt         : concatenated tinyobj::index_t streams, 3 ints per corner in (vertex, normal, texcoord)
 *                     order (tinyobj.hpp:347-351)
 *   shape_corner_offsets : nshapes + 1 offsets into idx_vnt, in corners (each a multiple of 3)
 *   out             : nshapes entries, filled with malloc()ed buffers (or see vkm_set_result_memory); release
 *                     with vkm_free_shape().
 * The input arrays may be ordinary pageable memory (staged through the library's page-locked ring) or come from
 * vkm_host_alloc() (copied directly).  Device scratch is cached between calls and sized from the inputs; when a
 * mesh has more unique vertices than its attribute arrays suggest the import is repeated once with the exact
 * size. */
int vkm_import_shapes(const float* pos, uint64_t np, const float* nrm, uint64_t nn, const float* uv, uint64_t nt,
                      const int32_t* idx_vnt, const uint64_t* shape_corner_offsets, uint32_t nshapes,
                      vkm_shape_out* out);
void vkm_free_shape(vkm_shape_out* shape);

/* Where the host entry points (vkm_import_shapes, vkm_import_batch, vkm_import_obj_text) put their results.
 *   0 (default): malloc()ed buffers, one pair per shape -- what GeometryConfig owns in the reference
 *                (Platform::AllocMemory == malloc at mesh_loader.cpp:303-311, released with
 *                Platform::FrMemory == free at mesh_resource.cpp:13-18).  Costs one host copy out of the
 *                page-locked staging.
 *   1          : page-locked blocks owned by the library and recycled: one block per call, the device writes
 *                the results into it directly (no staging copy, no page faults).  The buffers MUST then be
 *                released with vkm_free_shape(), never free(); idle blocks are kept for reuse up to
 *                VKM_RESULT_POOL_MAX_MB (environment; default a quarter of host memory, at least 8192) and dropped
 *                by vkm_shutdown() or mode 0.
 * vkm_free_shape() accepts buffers of either kind, so the mode can change while results are outstanding.
 * No counterpart in the reference (its loader has no device). */
int vkm_set_result_memory(int mode);

/* A batch of independent meshes (each what one vkm_import_shapes call takes) over every GPU selected with
 * vkm_init(): the multi-GPU form of the same reference loop (each mesh is an independent dedup domain,
 * mesh_loader.cpp:322-324 clears all state between shapes; the reference itself has no multi-mesh entry).  Meshes
 * are partitioned over the devices by size (longest first); each device runs two lanes so that one mesh's copies
 * overlap another's kernels.  Small meshes are grouped: a group's arrays are laid end to end on the device and run
 * as one segmented import (every mesh's indices are checked against its own record counts first), so thousands of
 * small assets do not pay thousands of launches and synchronisations; results and statuses stay per mesh.  outs[i] must hold meshes[i].nshapes entries; statuses (optional) gets one code per
 * mesh.  Returns the first failure's code, VKM_OK if every mesh imported. */
typedef struct vkm_mesh_in {
    const float* pos;  uint64_t np;
    const float* nrm;  uint64_t nn;
    const float* uv;   uint64_t nt;
    const int32_t* idx_vnt;
    const uint64_t* shape_corner_offsets; /* nshapes + 1 */
    uint32_t nshapes;
} vkm_mesh_in;
int vkm_import_batch(const vkm_mesh_in* meshes, uint32_t count, vkm_shape_out* const* outs, int* statuses);

/* ---- OBJ text -> tinyobj-shaped streams on the GPU.  Replaces the call tinyobj::LoadObj(&attributes, &shapes,
 *      &materials, ...) in MeshLoader::LoadOBJ (mesh_loader.cpp:219; parser tinyobj.hpp:2557-3115) for `v`/`vn`/`vt`
 *      records, faces of 3 to 255 corners (quads by the shorter diagonal, larger polygons by the parser's own ear
 *      clipping, tinyobj.hpp:1714-1935), `o`/`g` shapes, `usemtl` and `mtllib` -- same floats (the parser's decimal
 *      routine is reproduced), same index_t streams in the same triangle order, same shape split, names and
 *      material names.  VKM_ERR_UNSUPPORTED is left for l / p / vw / t records, polygons beyond 255 corners and
 *      quads / polygons that name a vertex beyond the file's `v` records (tinyobj's treatment of those depends on
 *      where in the file the record appears); a zero or unresolvable index is VKM_ERR_BAD_INDEX (tinyobj fails the
 *      load). ---- */
typedef struct vkm_obj_streams {
    float* pos;   uint64_t np;  /* attrib_t::vertices  (malloc) */
    float* nrm;   uint64_t nn;  /* attrib_t::normals */
    float* uv;    uint64_t nt;  /* attrib_t::texcoords */
    int32_t* idx_vnt;           /* concatenated index_t streams, 3 ints per corner (v, n, t), -1 = absent */
    uint64_t* shape_corner_offsets; /* nshapes + 1 */
    uint32_t nshapes;
    char** shape_names;         /* shape_t::name */
    char** usemtl_names;        /* the raw `usemtl` name in force at the shape's first face */
    char** material_names;      /* materials[shape.mesh.material_ids[0]].name, "" when the id is -1: what
                                   mesh_loader.cpp:231, 312-314 put into GeometryConfig::material_name */
} vkm_obj_streams;
/* A material file the OBJ text may name on a `mtllib` line (the text entry points do not touch the file system:
 * the caller reads the .mtl it finds next to the .obj, as tinyobj's MaterialFileReader would, tinyobj.hpp:2442-2480). */
typedef struct vkm_mtl_source {
    const char* filename;  /* as written after `mtllib`, e.g. "falcon.mtl" */
    const char* text;
    uint64_t len;
} vkm_mtl_source;
int vkm_parse_obj(int device, const char* text, uint64_t len, vkm_obj_streams* out);  /* no .mtl: material_names all "" */
int vkm_parse_obj_mtl(int device, const char* text, uint64_t len, const vkm_mtl_source* mtls, uint32_t nmtl,
                      vkm_obj_streams* out);
void vkm_free_obj_streams(vkm_obj_streams* streams);

/* The whole of MeshLoader::LoadOBJ (mesh_loader.cpp:200-328) from OBJ text: parse on the GPU, run the geometry
 * stage on the device-resident streams (they never visit the host), return GeometryConfig-shaped buffers
 * (malloc()ed) with shape names.  Same restrictions as vkm_parse_obj. */
typedef struct vkm_obj_import {
    uint32_t nshapes;
    vkm_shape_out* shapes;
    char** shape_names;
    char** usemtl_names;
    char** material_names;      /* GeometryConfig::material_name (see vkm_obj_streams) */
} vkm_obj_import;
int vkm_import_obj_text(int device, const char* text, uint64_t len, vkm_obj_import* out);
int vkm_import_obj_text_mtl(int device, const char* text, uint64_t len, const vkm_mtl_source* mtls, uint32_t nmtl,
                            vkm_obj_import* out);
void vkm_free_obj_import(vkm_obj_import* result);

/* ---- E3DM import cache: replaces b8 MeshLoader::WriteToE3DM(const std::string&, const std::string&,
 *      GeometryConfigs&) and MeshResource* MeshLoader::LoadE3DM(const std::string&, const std::string&)
 *      -- mesh_loader.cpp:63-126 and :128-198.  Host-side file code; byte-compatible with the reference's files. ---- */
typedef struct vkm_e3dm_config {
    vkm_shape_out shape;        /* extent is not part of the format: zero after a read */
    const char* name;           /* GeometryConfig::name */
    const char* material_name;  /* GeometryConfig::material_name */
} vkm_e3dm_config;
int vkm_e3dm_write(const char* path, const char* mesh_name, const vkm_e3dm_config* configs, uint64_t count);
/* Reads a file into malloc()ed configs (vertices / indices / strings); release with vkm_e3dm_free(). */
int vkm_e3dm_read(const char* path, char* mesh_name, int mesh_name_cap, vkm_e3dm_config** out_configs,
                  uint64_t* out_count);
void vkm_e3dm_free(vkm_e3dm_config* configs, uint64_t count);

/* ---- procedural generators.  Replace GeometryConfig GeometrySystem::GeneratePlainConfig(f32 width, f32 height,
 *      u32 x_segment_count, u32 y_segment_count, f32 tile_x, f32 tile_y, name, material_name) and
 *      GeometrySystem::GenerateCubeConfig(f32 width, f32 height, f32 depth, f32 tile_x, f32 tile_y, name,
 *      material_name) -- engine/src/systems/geometry/geometry_system.cpp:182-294 and :296-439 (buffers only; the
 *      names stay with the caller).  `deduplicate` runs Geometry::DeduplicateVertices on the plane, the follow-up
 *      the reference's TODO at :227 names.  Buffers are malloc()ed; release with vkm_free_shape(). ---- */
int vkm_generate_plane_config(float width, float height, uint32_t x_segment_count, uint32_t y_segment_count,
                              float tile_x, float tile_y, int deduplicate, vkm_shape_out* out);
int vkm_generate_cube_config(float width, float height, float depth, float tile_x, float tile_y, vkm_shape_out* out);

/* Pinned host memory for callers that want full-speed PCIe transfers through the host entry points. */
void* vkm_host_alloc(uint64_t bytes);
void vkm_host_free(void* p);

/* ---- device-resident entry points (inputs and outputs stay in HBM; used by bench.py and by a renderer
 *      that consumes the buffers on the GPU).  They replace the same reference loop as vkm_import_shapes;
 *      the reference has no device-resident counterpart (its only device boundary is the Vulkan staging
 *      copy, engine/src/renderer/backend/vulkan/vulkan.cpp:15-27). ---- */
typedef struct vkm_plan vkm_plan;

typedef struct vkm_plan_desc {
    int device;                /* CUDA ordinal */
    uint64_t max_corners;      /* capacity: corners per import (3 x triangles) */
    uint64_t max_vertices;     /* capacity: unique vertices per import; 0 = worst case (= max_corners) */
    uint64_t max_attrib;       /* largest of Np, Nn, Nt (sizes the hash tables; 0 = derive from max_corners) */
    uint32_t max_shapes;
} vkm_plan_desc;

/* Per-import summary, valid after vkm_plan_finish(). */
typedef struct vkm_plan_result {
    uint64_t total_vertices;       /* sum over shapes */
    uint64_t total_indices;
    const void* d_vertices;        /* device: total_vertices x 64 B, shapes back to back */
    const uint32_t* d_indices;     /* device: total_indices x u32, per-shape local indices */
    const uint32_t* shape_vertex_offsets; /* host, nshapes + 1 */
    const vkm_extent* shape_extents;      /* host, nshapes */
    float phase_ms[8];             /* device time per pipeline phase when timing was requested, else 0 */
} vkm_plan_result;

int vkm_plan_create(const vkm_plan_desc* desc, vkm_plan** out);
void vkm_plan_destroy(vkm_plan* plan);
/* Enqueues the whole import on `stream` (a cudaStream_t; NULL = the plan's own stream).  All pointers are
 * device pointers except shape_corner_offsets (host).  Asynchronous.
 *   - Alignment (checked, VKM_ERR_INVALID_ARG): d_idx_vnt 16 bytes, d_uv 8 bytes, d_pos / d_nrm 4 bytes -- the kernels
 *     read them with 128-/64-bit loads.  cudaMalloc results qualify; a sub-range of a packed arena may not.
 *   - A plan holds ONE set of outputs and ONE status block.  Several imports may be queued on one stream before
 *     vkm_plan_finish (bench.py does, to time back-to-back imports), but only the LAST one's results and status
 *     are observable; an error or a failed speculation in an earlier queued import goes unnoticed.  Callers that
 *     consume every result call vkm_plan_finish after every vkm_plan_import.
 *   - Schedules.  While the plan expects a "clean" mesh (no two distinct (v,vn,vt) triplets with equal values, no
 *     NaN/Inf attribute: true for nearly every mesh, assumed for a new plan) it launches the short record
 *     schedule, which also proves the assumption; if the proof fails, vkm_plan_finish discards that run and
 *     launches the general schedule once (vkm_plan_stats.mis_speculated).  After a non-clean import the general
 *     schedule runs first.  All schedules give identical results. */
int vkm_plan_import(vkm_plan* plan, const float* d_pos, uint64_t np, const float* d_nrm, uint64_t nn,
                    const float* d_uv, uint64_t nt, const int32_t* d_idx_vnt, const uint64_t* shape_corner_offsets,
                    uint32_t nshapes, void* stream);
/* Waits for the import, checks the device status word and fills `result`. */
int vkm_plan_finish(vkm_plan* plan, vkm_plan_result* result);
/* What vkm_plan_finish had to do for the last import (no reference counterpart: the reference has one code path). */
typedef struct vkm_plan_stats {
    uint32_t runs;            /* schedules launched (1 = the first one stood) */
    uint32_t schedule;        /* the one whose results stood: 0 record (clean speculation), 1 short general, 2 full general */
    uint32_t mis_speculated;  /* 1 = a clean-speculating run was discarded because the mesh was not clean */
    uint32_t table_overflows; /* re-runs because a hash table was too small or its keys clustered */
    uint32_t reseeds;         /* hash generation changes among those re-runs */
    uint32_t hash_seed;       /* hash generation in force */
} vkm_plan_stats;
int vkm_plan_get_stats(const vkm_plan* plan, vkm_plan_stats* out);
/* The same for the plan behind the host entry points (vkm_import_shapes and friends; the calling thread's last call). */
int vkm_last_import_stats(vkm_plan_stats* out);
/* When enabled, vkm_plan_import records CUDA events between phases (costs a few microseconds). */
void vkm_plan_set_timing(vkm_plan* plan, int enabled);
/* Names of the phases reported in vkm_plan_result.phase_ms (NULL past the last). */
const char* vkm_plan_phase_name(int phase);

/* ---- renderer handoff as a headless device buffer.  Replaces the bookkeeping of
 *      GeometrySystem::AcquireGeometryFromConfig -> VulkanRendererBackend::CreateGeometry -> UploadDataRange
 *      (geometry_system.cpp:136-176; vulkan.cpp:15-27, 691-725): one global vertex buffer and one global index
 *      buffer, first-fit byte ranges (Freelist::AllocateBlock, core/utils/freelist.cpp:144-159), and the
 *      (offset, size) pair per geometry that the draw call binds with (vulkan.cpp:664-672).  The buffers are plain
 *      device memory; the Vulkan external-memory import is unavailable on this box. ---- */
typedef struct vkm_rangealloc vkm_rangealloc;       /* first-fit range allocator (host bookkeeping only) */
vkm_rangealloc* vkm_rangealloc_create(uint64_t total_bytes);
void vkm_rangealloc_destroy(vkm_rangealloc* a);
uint64_t vkm_rangealloc_alloc(vkm_rangealloc* a, uint64_t size);  /* lowest fitting offset, UINT64_MAX if none */
int vkm_rangealloc_free(vkm_rangealloc* a, uint64_t offset, uint64_t size);
uint64_t vkm_rangealloc_free_space(const vkm_rangealloc* a);

typedef struct vkm_geometry_pool vkm_geometry_pool;
typedef struct vkm_pool_geometry {
    uint64_t vertex_offset, vertex_bytes; /* byte range in the global vertex buffer */
    uint64_t index_offset, index_bytes;   /* byte range in the global index buffer (u32 indices) */
    uint32_t vertex_count, index_count;
} vkm_pool_geometry;
int vkm_pool_create(int device, uint64_t vertex_bytes, uint64_t index_bytes, vkm_geometry_pool** out);
void vkm_pool_destroy(vkm_geometry_pool* pool);
const void* vkm_pool_vertex_buffer(const vkm_geometry_pool* pool);
const void* vkm_pool_index_buffer(const vkm_geometry_pool* pool);
/* Places shape `shape` of a finished import (vkm_plan_finish) into the pool: device-to-device copies on `stream`. */
int vkm_pool_acquire(vkm_geometry_pool* pool, const vkm_plan_result* result, const uint64_t* shape_corner_offsets,
                     uint32_t shape, vkm_pool_geometry* out, void* stream);
int vkm_pool_release(vkm_geometry_pool* pool, const vkm_pool_geometry* geometry);

/* Device-resident Geometry::GenerateTangents / GenerateNormals (geometry.cpp:20-82) on caller-owned device
 * buffers; `d_scratch` must hold vertex_count u32 (it is overwritten). */
int vkm_generate_tangents_dev(uint32_t vertex_count, void* d_vertices, uint32_t index_count,
                              const uint32_t* d_indices, uint32_t* d_scratch, void* stream);
int vkm_generate_normals_dev(uint32_t vertex_count, void* d_vertices, uint32_t index_count,
                             const uint32_t* d_indices, uint32_t* d_scratch, void* stream);

/* Tangent generation for callers that must not allocate device memory.  (This entry point has no counterpart in
 * the reference; it exists only because of how this library is deployed.)
 *
 * Who needs it: code that runs while the renderer is drawing -- a hot-reload hook, a procedural-geometry update
 * inside the frame loop -- where a cudaMalloc/cudaFree would synchronise the whole device and stall the frames in
 * flight, and engines that hand all remaining device memory to their own allocator at start-up, so that a later
 * cudaMalloc can simply fail.  Such a caller owns `d_vertices` and `d_indices` but has nowhere to put the
 * per-vertex "last triangle" words that vkm_generate_tangents_dev takes as `d_scratch`.
 *
 * Here that scratch is a fixed array inside the library's device image: VKM_NOALLOC_MAX_VERTICES words (4 MiB),
 * resident on every GPU that loads the library until the process exits.  The call makes no cudaMalloc/cudaFree
 * and uses no allocating library routine.  Same results as vkm_generate_tangents_dev, asynchronous on `stream`.
 *   - vertex_count > VKM_NOALLOC_MAX_VERTICES is rejected with VKM_ERR_CAPACITY (nothing is launched).
 *   - The first call on each device looks the array's device address up (once per device, with that device
 *     current).  Make that first call during set-up, before any stream capture begins.
 *   - All calls on a device share the one array and it keeps whatever the previous call left; every call first
 *     zeroes the words it is going to use.  Two calls whose GPU work overlaps (different streams, or different
 *     host threads on the same device) race on the array: serialise them. */
#define VKM_NOALLOC_MAX_VERTICES (1u << 20)
int vkm_generate_tangents_dev_noalloc(uint32_t vertex_count, void* d_vertices, uint32_t index_count,
                                      const uint32_t* d_indices, void* stream);

/* Thin device-memory helpers so that non-CUDA callers (ctypes, the bench) need no second runtime. */
int vkm_dev_alloc(int device, uint64_t bytes, void** d_ptr);
int vkm_dev_free(int device, void* d_ptr);
int vkm_dev_upload(int device, void* d_dst, const void* h_src, uint64_t bytes);
int vkm_dev_download(int device, void* h_dst, const void* d_src, uint64_t bytes);
int vkm_dev_sync(int device);
int vkm_dev_mem_info(int device, uint64_t* free_bytes, uint64_t* total_bytes);  /* cudaMemGetInfo */

#ifdef __cplusplus
}
#endif
#endif /* VKMESH_B200_H */
