/* TEST INFRASTRUCTURE ONLY (oracle) -- see oracle_port.h.  CPU restatement, written from the
 * reference's behaviour, of:
 *   - std::hash<Vertex3D> / Vertex3D::operator==     renderer/renderer_types.hpp:20-27, 50-88
 *   - the per-shape loop of MeshLoader::LoadOBJ      systems/resource/loaders/mesh/mesh_loader.cpp:228-325
 *   - Geometry::GenerateTangents / GenerateNormals   resources/geometry/geometry.cpp:20-82
 *   - Geometry::DeduplicateVertices + ReassignIndex  resources/geometry/geometry.cpp:84-126
 *
 * Third-party arithmetic that is NOT under /root/reference: the reference dedups through
 * std::unordered_map with std::hash<float>.  The oracle build links libstdc++ 13.3, whose
 * std::hash<float> is `x == 0 ? 0 : _Hash_bytes(&x, 4, 0xc70f6907)` (bits/functional_hash.h)
 * and whose 64-bit _Hash_bytes is the murmur-style mixer restated in hash_bytes4() below.
 * libstdc++ caches the hash code in each node (the user hash is not noexcept), so a lookup
 * matches a stored key iff the 64-bit hash codes are equal AND operator== holds.
 *
 * Build: gcc -O2 -ffp-contract=off (no FMA contraction; see oracle/Makefile).
 */
#include "oracle_port.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define ORC_FLOAT_EPSILON 1.192092896e-07f /* defines.hpp:67 */

/* ------------------------------------------------------------------------------------------
 * Hashing
 * ---------------------------------------------------------------------------------------- */
static uint64_t shift_mix(uint64_t v) { return v ^ (v >> 47); }

/* libstdc++ _Hash_bytes for len == 4 on a 64-bit target: no full 8-byte block, one 4-byte tail. */
static uint64_t hash_bytes4(uint32_t bits) {
    const uint64_t mul = 0xc6a4a7935bd1e995ull;
    uint64_t h = 0xc70f6907ull ^ (4ull * mul);
    h ^= (uint64_t)bits;
    h *= mul;
    h = shift_mix(h) * mul;
    h = shift_mix(h);
    return h;
}

static uint64_t hash_float(float x) {
    uint32_t b;
    if (x == 0.0f) return 0; /* +0 and -0 both hash to 0; NaN != 0 so NaN is hashed by bits */
    memcpy(&b, &x, 4);
    return hash_bytes4(b);
}

static uint64_t hash_vec(const float* v, int n) { /* renderer_types.hpp:51-69 */
    uint64_t h = 0;
    for (int i = 0; i < n; ++i) h ^= hash_float(v[i]) << i;
    return h;
}

uint64_t orc_hash_vertex(const orc_vertex* v) { /* renderer_types.hpp:72-86 */
    uint64_t h = 17;
    h = h * 31 + hash_vec(v->position, 3);
    h = h * 31 + hash_vec(v->normal, 3);
    h = h * 31 + hash_vec(v->texcoord, 2);
    h = h * 31 + hash_vec(v->color, 4);
    h = h * 31 + hash_vec(v->tangent, 4);
    return h;
}

static int eps_eq(const float* a, const float* b, int n) { /* glm/gtc/epsilon.inl:32-35: abs(x - y) < eps */
    for (int i = 0; i < n; ++i)
        if (!(fabsf(a[i] - b[i]) < ORC_FLOAT_EPSILON)) return 0;
    return 1;
}

int orc_vertex_equal(const orc_vertex* a, const orc_vertex* b) { /* renderer_types.hpp:20-27 */
    return eps_eq(a->position, b->position, 3) && eps_eq(a->normal, b->normal, 3) &&
           eps_eq(a->texcoord, b->texcoord, 2) && eps_eq(a->color, b->color, 4) && eps_eq(a->tangent, b->tangent, 4);
}

/* ------------------------------------------------------------------------------------------
 * Geometry utilities
 * ---------------------------------------------------------------------------------------- */
static void normalize3(float* v) { /* glm normalize = v * inversesqrt(dot(v,v)); dot3 = (x*x + y*y) + z*z */
    float d = (v[0] * v[0] + v[1] * v[1]) + v[2] * v[2];
    float inv = 1.0f / sqrtf(d);
    v[0] *= inv;
    v[1] *= inv;
    v[2] *= inv;
}

void orc_generate_tangents(uint32_t vertex_count, orc_vertex* vtx, uint32_t index_count, const uint32_t* indices) {
    (void)vertex_count; /* unused by the reference body too */
    for (uint32_t i = 0; i + 3 <= index_count; i += 3) { /* whole triangles; the reference assumes index_count % 3 == 0 */
        const uint32_t i0 = indices[i], i1 = indices[i + 1], i2 = indices[i + 2];
        float e1[3], e2[3], t[3];
        for (int k = 0; k < 3; ++k) {
            e1[k] = vtx[i1].position[k] - vtx[i0].position[k];
            e2[k] = vtx[i2].position[k] - vtx[i0].position[k];
        }
        const float du1 = vtx[i1].texcoord[0] - vtx[i0].texcoord[0];
        const float dv1 = vtx[i1].texcoord[1] - vtx[i0].texcoord[1];
        const float du2 = vtx[i2].texcoord[0] - vtx[i0].texcoord[0];
        const float dv2 = vtx[i2].texcoord[1] - vtx[i0].texcoord[1];
        const float dividend = du1 * dv2 - du2 * dv1;
        const float fc = 1.0f / dividend;
        for (int k = 0; k < 3; ++k) t[k] = fc * (dv2 * e1[k] - dv1 * e2[k]);
        normalize3(t);
        const float w = (dividend < 0.0f) ? -1.0f : 1.0f;
        const uint32_t tri[3] = {i0, i1, i2};
        for (int c = 0; c < 3; ++c) { /* assignment: the last triangle that references a vertex wins */
            float* out = vtx[tri[c]].tangent;
            out[0] = t[0];
            out[1] = t[1];
            out[2] = t[2];
            out[3] = w;
        }
    }
}

void orc_generate_normals(uint32_t vertex_count, orc_vertex* vtx, uint32_t index_count, const uint32_t* indices) {
    (void)vertex_count;
    for (uint32_t i = 0; i + 3 <= index_count; i += 3) { /* whole triangles; the reference assumes index_count % 3 == 0 */
        const uint32_t i0 = indices[i], i1 = indices[i + 1], i2 = indices[i + 2];
        float a[3], b[3], n[3];
        for (int k = 0; k < 3; ++k) {
            a[k] = vtx[i1].position[k] - vtx[i0].position[k];
            b[k] = vtx[i2].position[k] - vtx[i0].position[k];
        }
        /* glm/detail/func_geometric.inl:68-79 */
        n[0] = a[1] * b[2] - b[1] * a[2];
        n[1] = a[2] * b[0] - b[2] * a[0];
        n[2] = a[0] * b[1] - b[0] * a[1];
        normalize3(n);
        const uint32_t tri[3] = {i0, i1, i2};
        for (int c = 0; c < 3; ++c) memcpy(vtx[tri[c]].normal, n, 12);
    }
}

uint32_t orc_deduplicate_vertices(uint32_t vertex_count, const orc_vertex* vertices, uint32_t index_count,
                                  uint32_t* indices, orc_vertex** out_vertices) {
    /* Literal behaviour of geometry.cpp:94-126, including its quirk: candidate i is compared with the
     * INPUT entries [0, unique_count), not with the unique list, and duplicates are remapped through
     * ReassignIndex(from = i - found_count, to = j), which also shifts every larger index down by one. */
    orc_vertex* uniq = (orc_vertex*)malloc(sizeof(orc_vertex) * (vertex_count ? vertex_count : 1));
    uint32_t ucount = 0, found_count = 0;
    for (uint32_t i = 0; i < vertex_count; ++i) {
        int found = 0;
        for (uint32_t j = 0; j < ucount; ++j) {
            if (orc_vertex_equal(&vertices[i], &vertices[j])) {
                const uint32_t from = i - found_count, to = j;
                for (uint32_t k = 0; k < index_count; ++k) {
                    if (indices[k] == from) indices[k] = to;
                    else if (indices[k] > from) indices[k]--;
                }
                found = 1;
                found_count++;
                break;
            }
        }
        if (!found) uniq[ucount++] = vertices[i];
    }
    *out_vertices = (orc_vertex*)malloc(sizeof(orc_vertex) * (ucount ? ucount : 1));
    memcpy(*out_vertices, uniq, sizeof(orc_vertex) * ucount);
    free(uniq);
    return ucount;
}

void orc_free(void* p) { free(p); }

/* ------------------------------------------------------------------------------------------
 * Procedural generators
 * ---------------------------------------------------------------------------------------- */
void orc_generate_plane(float width, float height, uint32_t xs, uint32_t ys, float tile_x, float tile_y,
                        orc_vertex* v, uint32_t* indices) {
    if (width == 0) width = 1.0f;
    if (height == 0) height = 1.0f;
    if (xs < 1) xs = 1;
    if (ys < 1) ys = 1;
    if (tile_x == 0) tile_x = 1.0f;
    if (tile_y == 0) tile_y = 1.0f;
    memset(v, 0, sizeof(orc_vertex) * 4u * xs * ys);
    const float seg_w = width / xs, seg_h = height / ys, half_w = width * 0.5f, half_h = height * 0.5f;
    for (uint32_t y = 0; y < ys; ++y)
        for (uint32_t x = 0; x < xs; ++x) {
            const float min_x = (x * seg_w) - half_w, min_y = (y * seg_h) - half_h;
            const float max_x = min_x + seg_w, max_y = min_y + seg_h;
            const float min_u = (x / (float)xs) * tile_x, min_v = (y / (float)ys) * tile_y;
            const float max_u = ((x + 1) / (float)xs) * tile_x, max_v = ((y + 1) / (float)ys) * tile_y;
            const uint32_t o = (y * xs + x) * 4;
            const float px[4] = {min_x, max_x, min_x, max_x}, py[4] = {min_y, max_y, max_y, min_y};
            const float pu[4] = {min_u, max_u, min_u, max_u}, pv[4] = {min_v, max_v, max_v, min_v};
            for (int k = 0; k < 4; ++k) {
                v[o + k].position[0] = px[k];
                v[o + k].position[1] = py[k];
                v[o + k].texcoord[0] = pu[k];
                v[o + k].texcoord[1] = pv[k];
            }
            uint32_t* t = indices + (y * xs + x) * 6;
            t[0] = o; t[1] = o + 1; t[2] = o + 2; t[3] = o; t[4] = o + 3; t[5] = o + 1;
        }
}

void orc_generate_cube(float width, float height, float depth, float tile_x, float tile_y, orc_vertex* v,
                       uint32_t* indices) {
    if (width == 0) width = 1.0f;
    if (height == 0) height = 1.0f;
    if (depth == 0) depth = 1.0f;
    if (tile_x == 0) tile_x = 1.0f;
    if (tile_y == 0) tile_y = 1.0f;
    const float hw = width * 0.5f, hh = height * 0.5f, hd = depth * 0.5f;
    const float a = -hw, b = -hh, c = -hd, A = hw, B = hh, C = hd;
    const float P[6][4][3] = {
        {{a, b, C}, {A, B, C}, {a, B, C}, {A, b, C}}, {{A, b, c}, {a, B, c}, {A, B, c}, {a, b, c}},
        {{a, b, c}, {a, B, C}, {a, B, c}, {a, b, C}}, {{A, b, C}, {A, B, c}, {A, B, C}, {A, b, c}},
        {{A, b, C}, {a, b, c}, {A, b, c}, {a, b, C}}, {{a, B, C}, {A, B, c}, {a, B, c}, {A, B, C}},
    };
    const float UV[4][2] = {{0.f, 0.f}, {tile_x, tile_y}, {0.f, tile_y}, {tile_x, 0.f}};
    memset(v, 0, sizeof(orc_vertex) * 24);
    for (int f = 0; f < 6; ++f) {
        for (int k = 0; k < 4; ++k) {
            memcpy(v[4 * f + k].position, P[f][k], 12);
            memcpy(v[4 * f + k].texcoord, UV[k], 8);
        }
        const uint32_t o = 4 * f;
        uint32_t* t = indices + 6 * f;
        t[0] = o; t[1] = o + 1; t[2] = o + 2; t[3] = o; t[4] = o + 3; t[5] = o + 1;
    }
    orc_generate_normals(24, v, 36, indices);
    orc_generate_tangents(24, v, 36, indices);
}

/* ------------------------------------------------------------------------------------------
 * Import: expand + AABB + first-occurrence dedup + tangents, per shape
 * ---------------------------------------------------------------------------------------- */
typedef struct {
    uint32_t* head;  /* bucket -> node, 0xFFFFFFFF = empty */
    uint32_t* next;  /* node -> node */
    uint64_t* hash;  /* node -> cached hash code */
    uint32_t mask;
} vmap;

/* A key never compares equal to ITSELF when some component c has !(|c - c| < eps): c - c is NaN for
 * NaN and for +-Inf alike, so every non-finite component makes the corner unmatchable. */
static int never_matches(const orc_vertex* v) {
    const float* f = (const float*)v;
    for (int i = 0; i < 8; ++i)
        if (!(fabsf(f[i] - f[i]) < ORC_FLOAT_EPSILON)) return 1;
    return 0;
}

int orc_import_shapes(const float* pos, uint64_t np, const float* nrm, uint64_t nn, const float* uv, uint64_t nt,
                      const int32_t* idx, const uint64_t* shape_offsets, uint32_t nshapes, orc_shape* out) {
    for (uint32_t s = 0; s < nshapes; ++s) memset(&out[s], 0, sizeof(orc_shape));
    for (uint32_t s = 0; s < nshapes; ++s) {
        const uint64_t c0 = shape_offsets[s], c1 = shape_offsets[s + 1];
        const uint64_t ncorners = (c1 - c0) / 3 * 3; /* whole triangles only (fv = 3, mesh_loader.cpp:233-287) */
        orc_shape* o = &out[s];
        orc_vertex* verts = (orc_vertex*)malloc(sizeof(orc_vertex) * (ncorners ? ncorners : 1));
        uint32_t* indices = (uint32_t*)malloc(sizeof(uint32_t) * (ncorners ? ncorners : 1));
        uint32_t nverts = 0;

        uint32_t nb = 16;
        while (nb < 2 * ncorners && nb < (1u << 30)) nb <<= 1;
        vmap m;
        m.mask = nb - 1;
        m.head = (uint32_t*)malloc(sizeof(uint32_t) * nb);
        memset(m.head, 0xFF, sizeof(uint32_t) * nb);
        m.next = (uint32_t*)malloc(sizeof(uint32_t) * (ncorners ? ncorners : 1));
        m.hash = (uint64_t*)malloc(sizeof(uint64_t) * (ncorners ? ncorners : 1));

        orc_extent ext;
        memset(&ext, 0, sizeof ext); /* GeometryExtent extent{} : the box always contains the origin */

        for (uint64_t c = 0; c < ncorners; ++c) {
            const int32_t vi = idx[3 * (c0 + c) + 0], ni = idx[3 * (c0 + c) + 1], ti = idx[3 * (c0 + c) + 2];
            if (vi < 0 || (uint64_t)vi >= np || ni < 0 || (uint64_t)ni >= nn || ti < 0 || (uint64_t)ti >= nt) {
                free(verts); free(indices); free(m.head); free(m.next); free(m.hash);
                for (uint32_t k = 0; k < s; ++k) orc_free_shape(&out[k]);
                return -1;
            }
            orc_vertex v;
            memset(&v, 0, sizeof v); /* Vertex3D vertex{} : color and tangent stay zero */
            memcpy(v.position, pos + 3 * (uint64_t)vi, 12);
            memcpy(v.texcoord, uv + 2 * (uint64_t)ti, 8);
            memcpy(v.normal, nrm + 3 * (uint64_t)ni, 12);
            for (int k = 0; k < 3; ++k) { /* strict compares: NaN never updates, -0 never replaces +0 */
                if (v.position[k] < ext.min_extents[k]) ext.min_extents[k] = v.position[k];
                if (v.position[k] > ext.max_extents[k]) ext.max_extents[k] = v.position[k];
            }
            /* uniqueVertices.count(v) == 0 -> uniqueVertices[v] = vertices.size(); vertices.push_back(v);
             * indices.push_back(uniqueVertices[v]);                                    (mesh_loader.cpp:280-285)
             * A key that contains NaN or +-Inf never compares equal, itself included: count() is 0, the vertex is
             * appended, and the second operator[] default-inserts yet another node whose value is 0, so the
             * index written is 0.  Those nodes can never be found again, so they are not stored here. */
            if (never_matches(&v)) {
                verts[nverts++] = v;
                indices[c] = 0;
                continue;
            }
            const uint64_t h = orc_hash_vertex(&v);
            const uint32_t b = (uint32_t)(h ^ (h >> 32)) & m.mask;
            uint32_t found = 0xFFFFFFFFu;
            for (uint32_t n = m.head[b]; n != 0xFFFFFFFFu; n = m.next[n]) {
                if (m.hash[n] == h && orc_vertex_equal(&verts[n], &v)) { found = n; break; }
            }
            if (found == 0xFFFFFFFFu) {
                found = nverts;
                verts[nverts] = v;
                m.hash[nverts] = h;
                m.next[nverts] = m.head[b]; /* libstdc++ inserts a new node at the front of its bucket */
                m.head[b] = nverts;
                nverts++;
            }
            indices[c] = found;
        }
        for (int k = 0; k < 3; ++k) ext.center[k] = (ext.min_extents[k] + ext.max_extents[k]) / 2.0f;

        orc_generate_tangents(nverts, verts, (uint32_t)ncorners, indices);

        free(m.head); free(m.next); free(m.hash);
        o->vertex_count = nverts;
        o->index_count = (uint32_t)ncorners;
        o->vertices = (orc_vertex*)realloc(verts, sizeof(orc_vertex) * (nverts ? nverts : 1));
        o->indices = indices;
        o->extent = ext;
    }
    return 0;
}

void orc_free_shape(orc_shape* s) {
    free(s->vertices);
    free(s->indices);
    s->vertices = NULL;
    s->indices = NULL;
}
