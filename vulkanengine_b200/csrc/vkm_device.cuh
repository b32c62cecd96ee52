// Device-side building blocks shared by the kernels of the mesh-import path.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace vkm {

constexpr uint32_t kEmpty32 = 0xFFFFFFFFu;     // empty slot (both tables hold one u32 per slot)
constexpr uint32_t kUnmatchable = 0xFFFFFFFFu; // aux marker: key holds NaN/Inf and never matches anything
constexpr int kMaxProbe = 2048;

// Device status block, read back by the host after every import.
struct Status {
    uint32_t bad_index;          // some v/vn/vt index out of range
    uint32_t overflow1;          // triplet table too small
    uint32_t overflow2;          // value table too small
    uint32_t has_unmatchable;    // at least one NaN/Inf key was seen
    uint32_t capacity_exceeded;  // unique vertices > plan capacity
    uint32_t total_vertices;
    uint32_t has_value_dup;      // two distinct triplets carried the same value key (duplicated records, +-0 twins)
    uint32_t has_displaced;      // mode 0: some triplet sits outside its key record's own slot pair
};

// Everything a kernel of the import pipeline may touch.  Passed by value.
struct ImportArgs {
    // inputs (tinyobj streams, device resident)
    const float* pos;
    const float* nrm;
    const float* uv;
    const int32_t* idx;  // 3 ints per corner: v, n, t
    uint32_t np, nn, nt;
    uint32_t ncorners;
    const uint32_t* shape_off;  // nshapes + 1 corner offsets
    uint32_t nshapes;
    // scratch
    uint32_t* table1;  // triplet table: first (lowest) corner that carries the slot's (shape, v, n, t)
    uint32_t m1;
    uint32_t t1_mode;   // 0: locality-preserving home slot (key attribute index * scale), 1: scrambled hash
    uint32_t t1_sel;    // which of v / n / t is the key attribute (the largest attribute array)
    uint32_t t1_scale;  // slots per key-attribute record
    uint4* trip;        // [m1 / 2] mode 0: (v, n, t, shape) of the FIRST variant of key record p, written by the thread
                        // that claims slot 2p as its home slot; dense (every referenced record has one), never cleared
    uint32_t* table2;  // value table: (tag << cbits | first corner); 2^cbits > ncorners, so a word is never kEmpty32
    uint32_t m2;
    uint32_t cbits;    // bits that hold the corner id in a table2 word; the remaining high bits are a hash tag
    uint32_t* slot1;   // [ncorners] slot of every corner in table1
    uint32_t* aux1;    // [m1] per triplet: slot in table2, later the final local vertex index
    uint32_t* fbits;   // [nwords + 1] bit c set <=> corner c creates a vertex
    uint32_t* fprefix; // [nwords + 1] exclusive popcount prefix of fbits
    uint32_t* block_sums;
    uint32_t nwords;
    uint32_t* ext_enc;  // [nshapes * 6] order-encoded min xyz, max xyz
    uint32_t* vbase;    // [nshapes + 1] first global vertex id of every shape
    uint32_t* lastface; // [vertex capacity] 1 + highest face id that references the vertex
    // record schedule (vkm_records.cuh): one 16-byte record per key-attribute record + a small hashed side table
    uint32_t* rfirst;    // [nkey] first (lowest) corner of the key record's first variant; all-ones = unclaimed
    uint2* ident;        // [nkey] that variant's other two indices (a, b); all-ones = not published yet
    uint32_t nkey;       // records of the key attribute (t1_sel)
    uint32_t* sec_first; // [msec] side table: first corner of a triplet that is not its key record's first variant
    uint32_t* sec_last;  // [msec] its last corner
    uint32_t msec;
    uint32_t* vid_of;    // [nkey + msec] local vertex index of every slot (written by the vertex pass)
    uint32_t seed;       // hash generation of this run (see hash_triplet)
    uint32_t* hbox;      // [6] order-encoded min / max of ALL position records (only steers the fingerprint table's bucket choice)
    uint32_t cell_vpc;   // record schedule: vertices per spatial cell, in quarters (0 = scrambled bucket choice)
    // outputs
    uint32_t* out_idx;  // [ncorners] per-shape local indices
    float4* out_vtx;    // [vertex capacity * 4]
    uint32_t vertex_capacity;
    Status* status;
};

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 33;
    x *= 0xff51afd7ed558ccdull;
    x ^= x >> 33;
    x *= 0xc4ceb9fe1a85ec53ull;
    x ^= x >> 33;
    return x;
}

// `seed` is the plan's hash generation: 0 until a table overflows at a size that should have held the mesh (keys
// clustering in THIS hash), then bumped so that the re-run spreads them differently.
__device__ __forceinline__ uint64_t hash_triplet(uint32_t v, uint32_t n, uint32_t t, uint32_t shape, uint32_t seed) {
    const uint64_t a = (uint64_t)v | ((uint64_t)n << 32);
    const uint64_t b = (uint64_t)t | ((uint64_t)shape << 32);
    return mix64(a ^ mix64(b + 0x9e3779b97f4a7c15ull + 0xd6e8feb86659fd93ull * seed));
}

// Range reduction without a power-of-two table: floor(x * m / 2^32).
__device__ __forceinline__ uint32_t reduce32(uint32_t x, uint32_t m) { return __umulhi(x, m); }

__device__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16;
    x *= 0x7feb352du;
    x ^= x >> 15;
    x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}

// Home slot of a triplet.  Mode 0 keeps the order of the key attribute's indices, so that a mesh whose faces
// are stored coherently touches the table coherently (neighbouring corners -> neighbouring slots -> shared
// sectors): record k of the key attribute owns slots [k*scale, (k+1)*scale); its first variant takes the first
// one, further variants (same record, other n/t) probe on.  Mode 1 is a full scramble, used when mode 0
// clusters (many distinct triplets per key record).
__device__ __forceinline__ uint32_t home_slot(const ImportArgs& a, uint32_t v, uint32_t n, uint32_t t, uint32_t sh) {
    if (a.t1_mode == 0) {
        const uint32_t key = a.t1_sel == 0 ? v : (a.t1_sel == 1 ? n : t);
        return key * a.t1_scale;  // < np|nn|nt * scale <= m1 (host guarantees no 32-bit overflow)
    }
    return reduce32((uint32_t)hash_triplet(v, n, t, sh, a.seed), a.m1);
}

__device__ __forceinline__ uint64_t hash_key8(const uint32_t k[8], uint32_t shape, uint32_t seed) {
    uint64_t h = mix64(((uint64_t)k[0] | ((uint64_t)k[1] << 32)) + 0x9e3779b97f4a7c15ull * (shape + 1) + 0xd6e8feb86659fd93ull * seed);
    h = mix64(h ^ ((uint64_t)k[2] | ((uint64_t)k[3] << 32)));
    h = mix64(h ^ ((uint64_t)k[4] | ((uint64_t)k[5] << 32)));
    h = mix64(h ^ ((uint64_t)k[6] | ((uint64_t)k[7] << 32)));
    return h;
}

// Largest s with shape_off[s] <= c (empty shapes are skipped because the search is an upper bound).
__device__ __forceinline__ uint32_t find_shape(const uint32_t* __restrict__ shape_off, uint32_t nshapes, uint32_t c) {
    if (nshapes <= 1) return 0;
    uint32_t lo = 0, hi = nshapes;  // invariant: shape_off[lo] <= c < shape_off[hi]
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (__ldg(shape_off + mid) <= c) lo = mid; else hi = mid;
    }
    return lo;
}

// The same for a warp whose lanes mostly fall into one shape (neighbouring corners / slots): the first active lane
// searches, the others check its answer with two loads and search themselves only when it does not fit.  With
// thousands of shapes the plain search is 14 dependent loads per thread.
__device__ __forceinline__ uint32_t find_shape_near(const uint32_t* __restrict__ shape_off, uint32_t nshapes, uint32_t c) {
    if (nshapes <= 1) return 0;
    const uint32_t mask = __activemask();
    const int leader = __ffs(mask) - 1;
    uint32_t hint = 0;
    if ((int)(threadIdx.x & 31) == leader) hint = find_shape(shape_off, nshapes, c);
    hint = __shfl_sync(mask, hint, leader);
    if (__ldg(shape_off + hint) <= c && c < __ldg(shape_off + hint + 1)) return hint;
    return find_shape(shape_off, nshapes, c);
}

// Monotone float -> uint map (so that integer atomicMin/Max order floats).  Never fed NaN.
__device__ __forceinline__ uint32_t enc_float(float f) {
    const uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ __forceinline__ float dec_float(uint32_t e) {
    const uint32_t b = (e & 0x80000000u) ? (e & 0x7FFFFFFFu) : ~e;
#ifdef __CUDA_ARCH__
    return __uint_as_float(b);
#else
    float f;
    memcpy(&f, &b, 4);
    return f;
#endif
}

__device__ __forceinline__ bool nonfinite_bits(uint32_t b) { return (b & 0x7F800000u) == 0x7F800000u; }
__device__ __forceinline__ uint32_t canon_bits(uint32_t b) { return b == 0x80000000u ? 0u : b; }

// Programmatic dependent launch (the import's kernels are launched with the stream-serialisation attribute): the
// first statement of every import kernel lets the NEXT kernel's CTAs become resident as this grid's CTAs retire, and
// then waits until the PREVIOUS kernel has completed and its memory is visible.  Nothing the previous kernel wrote
// is touched before pdl_enter() returns.  Launched without the attribute, both instructions do nothing.
__device__ __forceinline__ void pdl_enter() {
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

// Global rank (number of vertex-creating corners before corner c).
__device__ __forceinline__ uint32_t rank_of(const ImportArgs& a, uint32_t c) {
    const uint32_t w = c >> 5;
    return a.fprefix[w] + __popc(a.fbits[w] & ((1u << (c & 31)) - 1u));
}

// A "clean" import has no value duplicates (two distinct triplets with the same canonical 8 floats) and no NaN/Inf
// keys.  Then every distinct triplet IS one vertex, created by the very triplet its corners carry, and:
//   - a corner's final index is the rank of its slot's first corner (no second-level lookup);
//   - tangents can be produced while the vertex record is written: the winning triangle's corners read
//     pos[v] / uv[t] of their own triplets, which are bit-identical to the stored vertices (a value duplicate
//     could differ from the stored first-seen bits in the sign of a zero; a NaN/Inf corner points at vertex 0).
// The flags are final once k_resolve_slots has finished.
__device__ __forceinline__ bool import_is_clean(const ImportArgs& a) {
    return a.status->has_value_dup == 0 && a.status->has_unmatchable == 0;
}

// The triplet stored in slot s (whose first corner is c).  In mode 0, while no triplet was displaced out of its own
// slot pair, an even slot 2p can only hold the first variant of key record p, and that variant's claimer wrote
// trip[p]: one aligned 16-byte read.  Every other case (second variants in odd slots, displaced entries, mode 1)
// reads the triplet back from the slot's first corner in the index stream -- always exact, just a gather.
__device__ __forceinline__ uint4 slot_triplet(const ImportArgs& a, uint32_t s, uint32_t c, bool fast) {
    if (fast && !(s & 1u)) return a.trip[s >> 1];
    const int32_t* o = a.idx + 3ull * min(c, a.ncorners - 1);
    return make_uint4((uint32_t)__ldg(o), (uint32_t)__ldg(o + 1), (uint32_t)__ldg(o + 2),
                      find_shape(a.shape_off, a.nshapes, c));
}
__device__ __forceinline__ bool trip_is_fast(const ImportArgs& a) {
    return a.t1_mode == 0 && a.status->has_displaced == 0;
}

// Loads the 8 live floats of a corner's expanded vertex (position, normal, texcoord) as raw bits.
// Indices are clamped so that a flagged bad index can never fault.
__device__ __forceinline__ void gather_key(const ImportArgs& a, uint32_t v, uint32_t n, uint32_t t, uint32_t k[8]) {
    v = min(v, a.np - 1);
    n = min(n, a.nn - 1);
    t = min(t, a.nt - 1);
    const uint32_t* p = reinterpret_cast<const uint32_t*>(a.pos) + 3ull * v;
    const uint32_t* q = reinterpret_cast<const uint32_t*>(a.nrm) + 3ull * n;
    const uint32_t* r = reinterpret_cast<const uint32_t*>(a.uv) + 2ull * t;
    k[0] = __ldg(p);
    k[1] = __ldg(p + 1);
    k[2] = __ldg(p + 2);
    k[3] = __ldg(q);
    k[4] = __ldg(q + 1);
    k[5] = __ldg(q + 2);
    const uint2 tt = __ldg(reinterpret_cast<const uint2*>(r));
    k[6] = tt.x;
    k[7] = tt.y;
}

// Box of a SAMPLE (every 8th block of 256 records) of the finite position words, referenced or not: it only steers where
// in a value table a key starts (store_cell_map below; positions outside it are clamped to the border cells), so it can
// cost locality, never correctness.  hbox[0..2] = min words, hbox[3..5] = max words (initialised by the clear kernels).
__device__ __forceinline__ void sample_position_box(const ImportArgs& a, uint32_t tid, uint32_t nth) {
    uint32_t mn[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, mx[3] = {0u, 0u, 0u};
    const uint32_t* pw = reinterpret_cast<const uint32_t*>(a.pos);
    // (blocks of 256 consecutive records, every 8th block: whole sectors are used)
    for (uint32_t r = ((tid >> 8) << 11) + (tid & 255u); r < a.np; r += (nth >> 8) << 11) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const uint32_t b = __ldg(pw + 3ull * r + k);
            if (!nonfinite_bits(b)) {
                const uint32_t e = enc_float(__uint_as_float(canon_bits(b)));
                mn[k] = min(mn[k], e);
                mx[k] = max(mx[k], e);
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        mn[k] = __reduce_min_sync(0xFFFFFFFFu, mn[k]);
        mx[k] = __reduce_max_sync(0xFFFFFFFFu, mx[k]);
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            if (mn[k] < a.hbox[k]) atomicMin(a.hbox + k, mn[k]);
            if (mx[k] > a.hbox[3 + k]) atomicMax(a.hbox + 3 + k, mx[k]);
        }
    }
}

__device__ __forceinline__ float pick3(uint32_t ax, float x, float y, float z) { return ax == 0 ? x : (ax == 1 ? y : z); }

__device__ __forceinline__ void store_cell_map(const ImportArgs& a) {
    uint32_t* o = a.hbox + 8;
    o[0] = 0;
    if (a.cell_vpc == 0 || a.seed != 0 || a.nshapes > 1) return;
    float mnx, mny, mnz, ex, ey, ez;
    {
        const uint32_t l0 = a.hbox[0], l1 = a.hbox[1], l2 = a.hbox[2], h0 = a.hbox[3], h1 = a.hbox[4], h2 = a.hbox[5];
        if (l0 > h0 || l1 > h1 || l2 > h2) return;  // no finite position at all
        mnx = dec_float(l0); mny = dec_float(l1); mnz = dec_float(l2);
        ex = __fsub_rn(dec_float(h0), mnx); ey = __fsub_rn(dec_float(h1), mny); ez = __fsub_rn(dec_float(h2), mnz);
    }
    // the two longest axes, longest first
    uint32_t s0 = 0, s1 = 1;
    if (ex >= ey && ex >= ez) { s0 = 0; s1 = ey >= ez ? 1u : 2u; }
    else if (ey >= ex && ey >= ez) { s0 = 1; s1 = ex >= ez ? 0u : 2u; }
    else { s0 = 2; s1 = ex >= ey ? 0u : 1u; }
    const float e0 = pick3(s0, ex, ey, ez), e1 = pick3(s1, ex, ey, ez);
    // as many cells as four times the buckets: four consecutive cells of a row share a bucket (cell_bucket), so the
    // lattice covers the table once and a bucket collects about (vertices / buckets) = 4 keys of its 8 words
    const float cells = fmaxf(a.cell_vpc == 4 ? 4.0f * (float)(a.m2 >> 3) : (float)(a.m2 >> 1), 1.0f);  // (general schedule: a cell per pair of words)
    const float area = __fmul_rn(e0, e1);
    if (!(area > 0.0f) || !(area < 3.0e38f)) return;  // a line, a point, or an overflowing box
    const float h = __fsqrt_rn(__fdiv_rn(area, cells));
    if (!(h > 0.0f)) return;
    const float inv_h = __fdiv_rn(1.0f, h);
    if (!(inv_h < 3.0e38f)) return;
    o[1] = s0 | (s1 << 2);
    o[2] = (uint32_t)fminf(__fmul_rn(e0, inv_h), 4.0e6f) + 2u;
    o[3] = __float_as_uint(pick3(s0, mnx, mny, mnz));
    o[4] = __float_as_uint(pick3(s1, mnx, mny, mnz));
    o[5] = __float_as_uint(inv_h);
    o[0] = 1;
}

// General schedule: where a value key starts in its single-word table -- the lattice cell of its position, two words per
// cell, row-major, so that neighbours in space (for a coherently stored mesh: in the slot domain) probe neighbouring
// sectors instead of random ones.  Equal keys have equal positions, hence equal homes.
__device__ __forceinline__ uint32_t cell_home(const uint32_t* __restrict__ cm, const uint32_t kc[8], uint32_t m2) {
    const uint32_t axes = __ldg(cm + 1), g0 = __ldg(cm + 2);
    const float min0 = __uint_as_float(__ldg(cm + 3)), min1 = __uint_as_float(__ldg(cm + 4)), inv_h = __uint_as_float(__ldg(cm + 5));
    const float p0 = pick3(axes & 3u, __uint_as_float(kc[0]), __uint_as_float(kc[1]), __uint_as_float(kc[2]));
    const float p1 = pick3(axes >> 2, __uint_as_float(kc[0]), __uint_as_float(kc[1]), __uint_as_float(kc[2]));
    const float c0 = fminf(fmaxf(__fmul_rn(__fsub_rn(p0, min0), inv_h), 0.0f), 4.0e6f);
    const float c1 = fminf(fmaxf(__fmul_rn(__fsub_rn(p1, min1), inv_h), 0.0f), 4.0e6f);
    const uint64_t id = (uint64_t)(uint32_t)c0 + (uint64_t)(uint32_t)c1 * g0;
    return (uint32_t)((2ull * id) % m2);
}

}  // namespace vkm
