// sm_100a kernels of the mesh-import path (all HBM/L2-bound integer + gather work; no tensor cores).
//
//   k_insert_triplets   A2/A3  read the (v,n,t) index stream with 128-bit loads, insert every corner into the
//                              triplet table with atomicMin(first corner)                     [C-scale]
//   k_resolve_slots     A2/A3/A4  per distinct triplet (slot domain): gather 8 floats, bounding box, NaN/Inf
//                              screening, insert the VALUE key into table2                    [T-scale]
//   k_mark_unmatchable  A3     NaN/Inf keys: every occurrence creates a vertex (only runs if any seen)
//   k_redo_flags_*      A3     only if value duplicates exist: rebuild the flag bits from table2
//   k_scan_tiles/finish A3/A8  popcount prefix over the flag bitmap (stable first-occurrence order), per-shape
//                              vertex bases, lastface reset
//   k_final_index       A3     per triplet: final local vertex index
//   k_write_indices     A3/A5  per corner: index buffer + atomicMax(last face per vertex)   [C-scale]
//   k_write_vertices    A2/A8  stable compaction: 64-byte Vertex3D records
//   k_tangents/normals  A5/A6  per vertex: evaluate the winning (highest-numbered) triangle
//
// Row ids refer to SURVEY.md section 8(a).
#pragma once

#include "vkm_device.cuh"

namespace vkm {

constexpr int kThreads = 256;

// ---------------------------------------------------------------------------------------------
// Per-import reset of the scratch in ONE launch (six memset nodes cost more launch gaps than bytes on small
// meshes): tables -> empty, flag bitmap -> 0, boxes -> enc(+0) (GeometryExtent extent{}, mesh_loader.cpp:230),
// status -> 0.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) k_clear(ImportArgs a) {
    pdl_enter();
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    const uint4 ff = make_uint4(kEmpty32, kEmpty32, kEmpty32, kEmpty32);
    uint4* t1 = reinterpret_cast<uint4*>(a.table1);
    for (uint32_t i = tid; i < (a.m1 >> 2); i += nth) t1[i] = ff;
    for (uint32_t i = (a.m1 & ~3u) + tid; i < a.m1; i += nth) a.table1[i] = kEmpty32;
    uint4* t2 = reinterpret_cast<uint4*>(a.table2);
    for (uint32_t i = tid; i < (a.m2 >> 2); i += nth) t2[i] = ff;
    for (uint32_t i = (a.m2 & ~3u) + tid; i < a.m2; i += nth) a.table2[i] = kEmpty32;
    for (uint32_t i = tid; i <= a.nwords; i += nth) a.fbits[i] = 0;
    for (uint32_t i = tid; i < a.nshapes * 6u; i += nth) a.ext_enc[i] = 0x80000000u;
    if (tid < sizeof(Status) / 4) reinterpret_cast<uint32_t*>(a.status)[tid] = 0;
    if (a.cell_vpc) {  // position-ordered value table: sampled position box (k_sample_box), cell map off until k_cell_map
        if (tid < 6) a.hbox[tid] = tid < 3 ? 0xFFFFFFFFu : 0u;
        if (tid == 6) a.hbox[8] = 0;
    }
}

// Position-ordered homes in the value table (insert_value): sample the position box, then one thread derives the
// lattice (vkm_device.cuh).  Two short launches in front of the insert pass; only for single shapes with a table too
// large for L2, in hash generation 0.
__global__ void __launch_bounds__(kThreads) k_sample_box(ImportArgs a) {
    pdl_enter();
    sample_position_box(a, blockIdx.x * blockDim.x + threadIdx.x, gridDim.x * blockDim.x);
}
__global__ void k_cell_map(ImportArgs a) {
    pdl_enter();
    if (blockIdx.x == 0 && threadIdx.x == 0) store_cell_map(a);
}

// ---------------------------------------------------------------------------------------------
// K1: triplet insert.  Each thread owns 4 consecutive corners = 48 bytes = three 128-bit loads.
// The table holds one u32 per slot: the lowest corner id seen so far for the slot's triplet.  Equality
// is decided exactly, by reading the occupant's own triplet from the index stream (for coherent meshes
// the occupant is a recent corner, so that read hits L1/L2).  The four corners of a thread are inserted
// one after the other on purpose: corners of one quad repeat triplets, so the later ones hit L1 (a variant
// that kept four probes in flight per thread measured 1.6x slower on B200: more registers, no reuse).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ bool same_triplet(const int32_t* __restrict__ idx, uint32_t occ, uint32_t v, uint32_t n,
                                             uint32_t t, uint32_t sh_begin, uint32_t sh_end) {
    if (occ < sh_begin || occ >= sh_end) return false;
    const int32_t* o = idx + 3ull * occ;
    return (uint32_t)__ldg(o) == v && (uint32_t)__ldg(o + 1) == n && (uint32_t)__ldg(o + 2) == t;
}

// ADDR: 0/1/2 = mode 0 keyed on v / n / t, 3 = mode 1 (scrambled).  MULTI: more than one shape in the stream.
// Specialising removes the per-corner mode/shape branching; the kernel is instruction-issue bound.
template <int ADDR, bool MULTI>
__global__ void __launch_bounds__(kThreads) k_insert_triplets(ImportArgs a) {
    pdl_enter();
    const uint32_t ngroups = (a.ncorners + 3) >> 2;
    for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < ngroups; g += gridDim.x * blockDim.x) {
        const uint32_t c0 = g << 2;
        int32_t w[12];
        if (c0 + 4 <= a.ncorners) {
            const int4* src = reinterpret_cast<const int4*>(a.idx + 3ull * c0);
            const int4 x = __ldg(src), y = __ldg(src + 1), z = __ldg(src + 2);
            w[0] = x.x; w[1] = x.y; w[2] = x.z; w[3] = x.w;
            w[4] = y.x; w[5] = y.y; w[6] = y.z; w[7] = y.w;
            w[8] = z.x; w[9] = z.y; w[10] = z.z; w[11] = z.w;
        } else {
#pragma unroll
            for (int i = 0; i < 12; ++i) w[i] = (c0 + i / 3 < a.ncorners) ? __ldg(a.idx + 3ull * c0 + i) : 0;
        }
        uint32_t sh = 0, sh_begin = 0, sh_end = a.ncorners;
        if (MULTI) {
            sh = find_shape(a.shape_off, a.nshapes, min(c0, a.ncorners - 1));
            sh_begin = a.shape_off[sh];
            sh_end = a.shape_off[sh + 1];
        }
        uint32_t slots[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t c = c0 + i;
            slots[i] = 0;
            if (c >= a.ncorners) continue;
            if (MULTI && c >= sh_end) {
                sh = find_shape(a.shape_off, a.nshapes, c);
                sh_begin = a.shape_off[sh];
                sh_end = a.shape_off[sh + 1];
            }
            const uint32_t v = (uint32_t)w[3 * i], n = (uint32_t)w[3 * i + 1], t = (uint32_t)w[3 * i + 2];
            if (v >= a.np || n >= a.nn || t >= a.nt) {  // negative ints are huge as unsigned
                a.status->bad_index = 1;
                continue;
            }
            uint32_t s;
            if (ADDR == 0) s = v * 2;
            else if (ADDR == 1) s = n * 2;
            else if (ADDR == 2) s = t * 2;
            else s = reduce32((uint32_t)hash_triplet(v, n, t, sh, a.seed), a.m1);
            bool ok = false;
            for (int probe = 0; probe < kMaxProbe; ++probe) {
                uint32_t cur = __ldcg(a.table1 + s);
                if (cur == kEmpty32) {
                    const uint32_t prev = atomicCAS(a.table1 + s, kEmpty32, c);
                    if (prev == kEmpty32) {
                        if (ADDR != 3) {
                            // first variant of its key record: leave the triplet for the slot-domain passes
                            if (probe == 0) a.trip[s >> 1] = make_uint4(v, n, t, sh);
                            else if (probe >= 2) a.status->has_displaced = 1;  // landed outside its own slot pair
                        }
                        ok = true;
                        break;
                    }
                    cur = prev;
                }
                if (cur == c) { ok = true; break; }
                if (same_triplet(a.idx, cur, v, n, t, sh_begin, sh_end)) {
                    if (cur > c) atomicMin(a.table1 + s, c);
                    ok = true;
                    break;
                }
                s = (s + 1 == a.m1) ? 0 : s + 1;
            }
            if (!ok) { a.status->overflow1 = 1; s = 0; }
            slots[i] = s;
        }
        if (c0 + 4 <= a.ncorners) {
            *reinterpret_cast<uint4*>(a.slot1 + c0) = make_uint4(slots[0], slots[1], slots[2], slots[3]);
        } else {
            for (int i = 0; i < 4 && c0 + i < a.ncorners; ++i) a.slot1[c0 + i] = slots[i];
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Value-level dedup (second level).  Distinct triplets can still carry identical values (duplicated
// v / vn / vt records), and +0 / -0 must merge, so every distinct triplet inserts its canonical 8-float
// key into a tagged table with atomicMin(first corner).  Tags only skip gathers: equality is decided
// exactly on the occupant's own values.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ bool same_value_key(const ImportArgs& a, uint32_t occ, const uint32_t k[8]) {
    const int32_t* o = a.idx + 3ull * occ;
    uint32_t q[8];
    gather_key(a, (uint32_t)__ldg(o), (uint32_t)__ldg(o + 1), (uint32_t)__ldg(o + 2), q);
    bool eq = true;
#pragma unroll
    for (int i = 0; i < 8; ++i) eq &= (canon_bits(q[i]) == k[i]);
    return eq;
}

__device__ __forceinline__ uint32_t insert_value(const ImportArgs& a, uint32_t c, const uint32_t k[8], uint32_t sh,
                                                 uint32_t sh_begin, uint32_t sh_end, bool spatial) {
    const uint64_t h = hash_key8(k, sh, a.seed);
    const uint32_t cmask = (a.cbits >= 32) ? 0xFFFFFFFFu : ((1u << a.cbits) - 1u);
    const uint32_t tag = (a.cbits >= 32) ? 0u : ((uint32_t)(h >> 32) >> a.cbits) << a.cbits;  // high (32 - cbits) bits
    const uint32_t mine = tag | c;
    uint32_t s = spatial ? cell_home(a.hbox + 8, k, a.m2) : reduce32((uint32_t)h, a.m2);
    for (int probe = 0; probe < kMaxProbe; ++probe) {
        uint32_t cur = __ldcg(a.table2 + s);
        if (cur == kEmpty32) {
            const uint32_t prev = atomicCAS(a.table2 + s, kEmpty32, mine);
            if (prev == kEmpty32) return s;
            cur = prev;
        }
        if ((cur & ~cmask) == tag) {
            const uint32_t occ = cur & cmask;
            if (occ == c) return s;
            if (occ >= sh_begin && occ < sh_end && same_value_key(a, occ, k)) {
                if (occ > c) atomicMin(a.table2 + s, mine);  // same tag: orders by corner
                a.status->has_value_dup = 1;  // the later arrival of every duplicate pair lands here
                return s;
            }
        }
        s = (s + 1 == a.m2) ? 0 : s + 1;
    }
    a.status->overflow2 = 1;
    return 0;
}

__device__ __forceinline__ void box_atomics(uint32_t* e, const uint32_t mn[3], const uint32_t mx[3]) {
    // bounds converge quickly: read (possibly stale, monotone) values first, atomics only on improvement
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        if (mn[k] < e[k]) atomicMin(e + k, mn[k]);
        if (mx[k] > e[3 + k]) atomicMax(e + 3 + k, mx[k]);
    }
}

// ---------------------------------------------------------------------------------------------
// K2 (A2 + A4 + second level of A3), slot domain: one thread per triplet-table slot.  An occupied slot
// is one distinct (shape, v, n, t): gather its 8 floats once, feed the shape's bounding box (the box
// over distinct triplets equals the box over all corners), screen NaN/Inf, insert the value key.
// Slot order follows the key attribute's index order, so the gathers are as coherent as the mesh.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 8) k_resolve_slots(ImportArgs a, bool write_aux) {
    pdl_enter();
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t neutral = 0x80000000u;  // enc_float(+0): the reference's box always contains the origin
    // One thread per PAIR of slots: in mode 0 a key record owns slots (2k, 2k+1) and the odd one is only used by
    // a second variant of the record, so nearly every thread has exactly one triplet to resolve.
    const uint32_t npairs = a.m1 >> 1;
    const uint32_t padded = (npairs + 31) & ~31u;  // whole warps stay in the loop (warp reductions below)
    const bool fast = trip_is_fast(a);
    const bool spatial = a.cell_vpc == 1 && __ldg(a.hbox + 8) != 0;  // position-ordered homes in the value table
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < padded; p += gridDim.x * blockDim.x) {
        uint32_t mn[3] = {neutral, neutral, neutral}, mx[3] = {neutral, neutral, neutral};
        uint2 cc = make_uint2(kEmpty32, kEmpty32);
        if (p < npairs) cc = __ldcs(reinterpret_cast<const uint2*>(a.table1) + p);
        uint32_t sh = 0xFFFFFFFFu;  // shape of this thread's box contribution
        bool mixed = false;         // the two slots belong to different shapes: use per-slot atomics
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const uint32_t c = j ? cc.y : cc.x;
            if (c == kEmpty32) continue;
            const uint32_t s = 2 * p + j;
            const uint4 tr = slot_triplet(a, s, c, fast);
            uint32_t k[8];
            gather_key(a, tr.x, tr.y, tr.z, k);
            bool bad = false;
#pragma unroll
            for (int i = 0; i < 8; ++i) bad |= nonfinite_bits(k[i]);
            uint32_t pmn[3] = {neutral, neutral, neutral}, pmx[3] = {neutral, neutral, neutral};
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                // strict </> against NaN never updates the reference's box; +-Inf does
                if ((k[i] & 0x7FFFFFFFu) <= 0x7F800000u) pmn[i] = pmx[i] = enc_float(__uint_as_float(k[i]));
            }
            if (sh == 0xFFFFFFFFu || sh == tr.w) {
                sh = tr.w;
#pragma unroll
                for (int i = 0; i < 3; ++i) {
                    mn[i] = min(mn[i], pmn[i]);
                    mx[i] = max(mx[i], pmx[i]);
                }
            } else {
                mixed = true;
                box_atomics(a.ext_enc + 6u * tr.w, pmn, pmx);
            }
            // aux1 (value-table slot per triplet) only feeds the dirty-only kernels: the short schedule skips it
            if (bad) {
                if (write_aux) a.aux1[s] = kUnmatchable;
                a.status->has_unmatchable = 1;
            } else {
#pragma unroll
                for (int i = 0; i < 8; ++i) k[i] = canon_bits(k[i]);
                const uint32_t s2 = insert_value(a, c, k, tr.w, a.shape_off[tr.w], a.shape_off[tr.w + 1], spatial);
                if (write_aux) a.aux1[s] = s2;
                // Speculation: no two distinct triplets share a value key (true for almost every mesh), so the
                // first corner of every distinct triplet creates a vertex.  If a duplicate shows up anywhere,
                // has_value_dup is raised and k_redo_flags_* rebuild the flags from the value table.
                atomicOr(a.fbits + (c >> 5), 1u << (c & 31));
            }
        }
        // bounding box: one set of (improvement-checked) atomics per warp when its slots share a shape
        const bool have = (sh != 0xFFFFFFFFu);
        const uint32_t any = __ballot_sync(0xFFFFFFFFu, have);
        if (any) {
            const uint32_t sh0 = __shfl_sync(0xFFFFFFFFu, sh, __ffs(any) - 1);
            if (__all_sync(0xFFFFFFFFu, !have || (sh == sh0 && !mixed))) {
#pragma unroll
                for (int i = 0; i < 3; ++i) {
                    mn[i] = __reduce_min_sync(0xFFFFFFFFu, mn[i]);
                    mx[i] = __reduce_max_sync(0xFFFFFFFFu, mx[i]);
                }
                if (lane == 0) box_atomics(a.ext_enc + 6u * sh0, mn, mx);
            } else if (have) {
                box_atomics(a.ext_enc + 6u * sh, mn, mx);
            }
        }
    }
}

// Every occurrence of a NaN/Inf key creates a vertex (and gets index 0).
__global__ void __launch_bounds__(kThreads) k_mark_unmatchable(ImportArgs a) {
    pdl_enter();
    if (a.status->has_unmatchable == 0) return;
    for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < a.ncorners; c += gridDim.x * blockDim.x) {
        if (a.aux1[a.slot1[c]] == kUnmatchable) atomicOr(a.fbits + (c >> 5), 1u << (c & 31));
    }
}

// Only when value duplicates exist: the speculative flags are wrong for the later triplet of each duplicate
// group.  Clear them, then flag the lowest corner of every distinct value key from the value table.
__global__ void __launch_bounds__(kThreads) k_redo_flags_clear(ImportArgs a) {
    pdl_enter();
    if (a.status->has_value_dup == 0) return;
    for (uint32_t w = blockIdx.x * blockDim.x + threadIdx.x; w < a.nwords; w += gridDim.x * blockDim.x) a.fbits[w] = 0;
}

__global__ void __launch_bounds__(kThreads) k_redo_flags_mark(ImportArgs a) {
    pdl_enter();
    if (a.status->has_value_dup == 0) return;
    for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < a.m2; s += gridDim.x * blockDim.x) {
        const uint32_t cur = a.table2[s];
        if (cur != kEmpty32) {
            const uint32_t c = (a.cbits >= 32) ? cur : (cur & ((1u << a.cbits) - 1u));
            atomicOr(a.fbits + (c >> 5), 1u << (c & 31));
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Popcount prefix over the flag bitmap: three small kernels (the bitmap is C/8 bytes).
// ---------------------------------------------------------------------------------------------
constexpr int kScanThreads = 1024;
constexpr int kScanItems = 4;  // words per thread
constexpr int kScanTile = kScanThreads * kScanItems;

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t x, uint32_t* total) {
    __shared__ uint32_t warp_sums[32];
    const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t inc = x;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, inc, d);
        if (lane >= d) inc += y;
    }
    if (lane == 31) warp_sums[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        uint32_t ws = (lane < (blockDim.x >> 5)) ? warp_sums[lane] : 0;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, ws, d);
            if (lane >= d) ws += y;
        }
        warp_sums[lane] = ws;
    }
    __syncthreads();
    const uint32_t base = wid ? warp_sums[wid - 1] : 0;
    if (total) *total = warp_sums[(blockDim.x >> 5) - 1];
    __syncthreads();
    return base + inc - x;
}

__global__ void __launch_bounds__(kScanThreads) k_scan_tiles(ImportArgs a) {
    pdl_enter();
    const uint32_t base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
    uint32_t sum = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i)
        if (base + i < a.nwords) sum += __popc(a.fbits[base + i]);
    uint32_t total;
    block_exclusive_scan(sum, &total);
    if (threadIdx.x == 0) a.block_sums[blockIdx.x] = total;
}

// Second (last) scan launch, one block per tile.  Every block re-derives its own offset and the grand total
// from the few thousand tile sums (cheaper than a third launch), writes its tile's exclusive prefixes, zeroes
// its share of lastface[0, total), and the first blocks also rank the shape boundaries straight from the
// bitmap (so no block has to wait for another block's prefixes).
__global__ void __launch_bounds__(kScanThreads) k_scan_finish(ImportArgs a, uint32_t ntiles, bool reset_lastface) {
    pdl_enter();
    __shared__ uint32_t red[2][32];
    uint32_t before = 0, all = 0;
    for (uint32_t i = threadIdx.x; i < ntiles; i += kScanThreads) {
        const uint32_t x = a.block_sums[i];
        all += x;
        if (i < blockIdx.x) before += x;
    }
#pragma unroll
    for (int d = 16; d; d >>= 1) {
        before += __shfl_xor_sync(0xFFFFFFFFu, before, d);
        all += __shfl_xor_sync(0xFFFFFFFFu, all, d);
    }
    if ((threadIdx.x & 31) == 0) {
        red[0][threadIdx.x >> 5] = before;
        red[1][threadIdx.x >> 5] = all;
    }
    __syncthreads();
    before = all = 0;
#pragma unroll
    for (int w = 0; w < kScanThreads / 32; ++w) {
        before += red[0][w];
        all += red[1][w];
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        a.fprefix[a.nwords] = all;  // rank(ncorners) when ncorners % 32 == 0
        a.status->total_vertices = all;
        if (all > a.vertex_capacity) a.status->capacity_exceeded = 1;
    }
    // this tile's prefixes
    const uint32_t base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
    uint32_t pc[kScanItems];
    uint32_t sum = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
        pc[i] = (base + i < a.nwords) ? __popc(a.fbits[base + i]) : 0;
        sum += pc[i];
    }
    uint32_t run = before + block_exclusive_scan(sum, nullptr);
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
        if (base + i < a.nwords) a.fprefix[base + i] = run;
        run += pc[i];
    }
    // lastface[0, total) = 0
    const uint32_t nv = reset_lastface ? min(all, a.vertex_capacity) : 0u;  // (the record schedule keeps no lastface)
    for (uint32_t i = blockIdx.x * kScanThreads + threadIdx.x; i < nv; i += ntiles * kScanThreads) a.lastface[i] = 0;
    // shape boundaries: one warp per shape, ranks computed from tile sums + the bitmap itself
    const uint32_t gwarp = (blockIdx.x * kScanThreads + threadIdx.x) >> 5, nwarps = (ntiles * kScanThreads) >> 5;
    const uint32_t lane = threadIdx.x & 31;
    for (uint32_t sidx = gwarp; sidx <= a.nshapes; sidx += nwarps) {
        const uint32_t c = a.shape_off[sidx];
        const uint32_t w = c >> 5, tile = w / kScanTile;
        uint32_t r = 0;
        for (uint32_t i = lane; i < tile; i += 32) r += a.block_sums[i];
        for (uint32_t i = tile * kScanTile + lane; i < w; i += 32) r += __popc(a.fbits[i]);
#pragma unroll
        for (int d = 16; d; d >>= 1) r += __shfl_xor_sync(0xFFFFFFFFu, r, d);
        if (lane == 0) a.vbase[sidx] = r + __popc(a.fbits[w] & ((1u << (c & 31)) - 1u));
    }
}

// ---------------------------------------------------------------------------------------------
// Final indices
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) k_final_index(ImportArgs a) {
    pdl_enter();
    if (import_is_clean(a)) return;  // clean import: k_write_indices ranks the slot's first corner itself
    for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < a.m1; s += gridDim.x * blockDim.x) {
        const uint32_t first = a.table1[s];
        if (first == kEmpty32) continue;
        const uint32_t x = a.aux1[s];
        if (x == kUnmatchable) continue;
        if (a.status->has_value_dup == 0) {  // every distinct triplet is its own vertex: coherent reads only
            if (first >= a.ncorners) continue;
            a.aux1[s] = rank_of(a, first) - a.vbase[find_shape(a.shape_off, a.nshapes, first)];
            continue;
        }
        if (x >= a.m2) continue;                      // only after an overflow; the host re-runs then
        const uint32_t word = a.table2[x];
        const uint32_t rep = (a.cbits >= 32) ? word : (word & ((1u << a.cbits) - 1u));  // first corner with this VALUE key
        if (rep >= a.ncorners) continue;
        const uint32_t sh = find_shape(a.shape_off, a.nshapes, rep);
        a.aux1[s] = rank_of(a, rep) - a.vbase[sh];
    }
}

__global__ void __launch_bounds__(kThreads) k_write_indices(ImportArgs a) {
    pdl_enter();
    const bool clean = import_is_clean(a);
    const uint32_t ngroups = (a.ncorners + 3) >> 2;
    for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < ngroups; g += gridDim.x * blockDim.x) {
        const uint32_t c0 = g << 2;
        uint32_t s[4] = {0, 0, 0, 0};
        const bool full = (c0 + 4 <= a.ncorners);
        if (full) {
            const uint4 q = __ldcs(reinterpret_cast<const uint4*>(a.slot1 + c0));
            s[0] = q.x; s[1] = q.y; s[2] = q.z; s[3] = q.w;
        } else {
            for (int i = 0; i < 4 && c0 + i < a.ncorners; ++i) s[i] = a.slot1[c0 + i];
        }
        uint32_t li[4];
        if (clean) {
            // every distinct triplet is its own vertex: global rank of the slot's first corner (coherent gathers)
#pragma unroll
            for (int i = 0; i < 4; ++i) li[i] = (c0 + i < a.ncorners) ? rank_of(a, min(a.table1[s[i]], a.ncorners - 1)) : 0;
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) li[i] = (c0 + i < a.ncorners) ? a.aux1[s[i]] : 0;
        }
        uint32_t sh = find_shape(a.shape_off, a.nshapes, c0);
        uint32_t sh_end = a.shape_off[sh + 1], base = a.vbase[sh];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t c = c0 + i;
            if (c >= a.ncorners) break;
            if (c >= sh_end) {
                sh = find_shape(a.shape_off, a.nshapes, c);
                sh_end = a.shape_off[sh + 1];
                base = a.vbase[sh];
            }
            if (clean) li[i] -= base;                   // global rank -> index local to the shape
            else if (li[i] == kUnmatchable) li[i] = 0;  // mesh_loader.cpp:285 default-inserts value 0 for a NaN key
            const uint32_t gv = base + li[i];
            if (gv < a.vertex_capacity) atomicMax(a.lastface + gv, c / 3 + 1);
        }
        if (full) {
            __stcs(reinterpret_cast<uint4*>(a.out_idx + c0), make_uint4(li[0], li[1], li[2], li[3]));
        } else {
            for (int i = 0; i < 4 && c0 + i < a.ncorners; ++i) a.out_idx[c0 + i] = li[i];
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Tangent / normal arithmetic.  Mirrors geometry.cpp:26-47 / :72-75 operation by operation with
// round-to-nearest intrinsics, which the compiler never contracts into FMAs (the reference CPU build has none).
// ---------------------------------------------------------------------------------------------
struct Tri {
    float px[3], py[3], pz[3], u[3], v[3];
};

__device__ __forceinline__ void load_tri(const float4* __restrict__ vtx, const uint32_t i[3], Tri& t) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float4 a = vtx[4ull * i[k]];                                         // position.xyz, normal.x
        const float2 b = *reinterpret_cast<const float2*>(reinterpret_cast<const float*>(vtx + 4ull * i[k]) + 6);
        t.px[k] = a.x; t.py[k] = a.y; t.pz[k] = a.z;
        t.u[k] = b.x; t.v[k] = b.y;
    }
}

__device__ __forceinline__ float3 normalize_rn(float x, float y, float z) {
    // glm::normalize = v * inversesqrt(dot(v, v)); dot3 = (x*x + y*y) + z*z; inversesqrt = 1 / sqrt
    const float d = __fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), __fmul_rn(z, z));
    const float inv = __fdiv_rn(1.0f, __fsqrt_rn(d));
    return make_float3(__fmul_rn(x, inv), __fmul_rn(y, inv), __fmul_rn(z, inv));
}

__device__ __forceinline__ float4 tangent_of(const Tri& t) {
    const float e1x = __fsub_rn(t.px[1], t.px[0]), e1y = __fsub_rn(t.py[1], t.py[0]), e1z = __fsub_rn(t.pz[1], t.pz[0]);
    const float e2x = __fsub_rn(t.px[2], t.px[0]), e2y = __fsub_rn(t.py[2], t.py[0]), e2z = __fsub_rn(t.pz[2], t.pz[0]);
    const float du1 = __fsub_rn(t.u[1], t.u[0]), dv1 = __fsub_rn(t.v[1], t.v[0]);
    const float du2 = __fsub_rn(t.u[2], t.u[0]), dv2 = __fsub_rn(t.v[2], t.v[0]);
    const float dividend = __fsub_rn(__fmul_rn(du1, dv2), __fmul_rn(du2, dv1));
    const float fc = __fdiv_rn(1.0f, dividend);
    const float tx = __fmul_rn(fc, __fsub_rn(__fmul_rn(dv2, e1x), __fmul_rn(dv1, e2x)));
    const float ty = __fmul_rn(fc, __fsub_rn(__fmul_rn(dv2, e1y), __fmul_rn(dv1, e2y)));
    const float tz = __fmul_rn(fc, __fsub_rn(__fmul_rn(dv2, e1z), __fmul_rn(dv1, e2z)));
    const float3 n = normalize_rn(tx, ty, tz);
    return make_float4(n.x, n.y, n.z, (dividend < 0.0f) ? -1.0f : 1.0f);
}

__device__ __forceinline__ float3 normal_of(const Tri& t) {
    const float ax = __fsub_rn(t.px[1], t.px[0]), ay = __fsub_rn(t.py[1], t.py[0]), az = __fsub_rn(t.pz[1], t.pz[0]);
    const float bx = __fsub_rn(t.px[2], t.px[0]), by = __fsub_rn(t.py[2], t.py[0]), bz = __fsub_rn(t.pz[2], t.pz[0]);
    // glm::cross, glm/detail/func_geometric.inl:68-79
    const float nx = __fsub_rn(__fmul_rn(ay, bz), __fmul_rn(by, az));
    const float ny = __fsub_rn(__fmul_rn(az, bx), __fmul_rn(bz, ax));
    const float nz = __fsub_rn(__fmul_rn(ax, by), __fmul_rn(bx, ay));
    return normalize_rn(nx, ny, nz);
}

// ---------------------------------------------------------------------------------------------
// Vertex records (A2/A8).  Slot domain: the slot whose first corner carries a flag bit writes the 64-byte
// Vertex3D at the corner's rank.  Slot order ~ corner order for coherent meshes, so the 64-byte records of
// neighbouring threads land next to each other.  NaN/Inf corners (every occurrence is a vertex) are
// written by the corner-domain k_write_unmatchable_vertices, which only runs when any exist.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void store_vertex(float4* dst, const uint32_t k[8], float4 tangent) {
    // Vertex3D: position @0, normal @12, texcoord @24, color @32 (zero), tangent @48
    __stcs(dst + 0, make_float4(__uint_as_float(k[0]), __uint_as_float(k[1]), __uint_as_float(k[2]), __uint_as_float(k[3])));
    __stcs(dst + 1, make_float4(__uint_as_float(k[4]), __uint_as_float(k[5]), __uint_as_float(k[6]), __uint_as_float(k[7])));
    __stcs(dst + 2, make_float4(0.f, 0.f, 0.f, 0.f));
    __stcs(dst + 3, tangent);
}

// Short schedule only: flag the first corner of every distinct triplet (speculation: the import is clean).
__global__ void __launch_bounds__(kThreads) k_flag_firsts(ImportArgs a) {
    pdl_enter();
    const uint32_t npairs = a.m1 >> 1;
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < npairs; p += gridDim.x * blockDim.x) {
        const uint2 cc = __ldcs(reinterpret_cast<const uint2*>(a.table1) + p);
        if (cc.x != kEmpty32 && cc.x < a.ncorners) atomicOr(a.fbits + (cc.x >> 5), 1u << (cc.x & 31));
        if (cc.y != kEmpty32 && cc.y < a.ncorners) atomicOr(a.fbits + (cc.y >> 5), 1u << (cc.y & 31));
    }
}

// DETECT (short schedule): this pass is also where the import is PROVEN clean -- it gathers every distinct
// triplet's 8 floats anyway, so NaN/Inf screening, the bounding box and the value-key insert (which raises
// has_value_dup when two triplets carry one value) ride along here instead of costing a pass of their own.  The
// passes before it ran on the assumption "clean"; if the status block says otherwise, vkm_plan_finish re-runs
// the import with the full schedule (k_resolve_slots first).
template <bool DETECT>
__global__ void __launch_bounds__(kThreads, 6) k_write_vertices(ImportArgs a) {
    pdl_enter();
    const uint32_t npairs = a.m1 >> 1;  // one thread per pair of slots, as in k_resolve_slots
    const bool fuse = import_is_clean(a);
    const bool fast = trip_is_fast(a);
    const bool spatial = DETECT && a.cell_vpc == 1 && __ldg(a.hbox + 8) != 0;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t neutral = 0x80000000u;
    const uint32_t padded = DETECT ? ((npairs + 31) & ~31u) : npairs;  // whole warps for the box reduction
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < padded; p += gridDim.x * blockDim.x) {
        uint2 cc = make_uint2(kEmpty32, kEmpty32);
        if (p < npairs) cc = __ldcs(reinterpret_cast<const uint2*>(a.table1) + p);
        uint32_t mn[3] = {neutral, neutral, neutral}, mx[3] = {neutral, neutral, neutral};
        uint32_t bsh = 0xFFFFFFFFu;
        bool mixed = false;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const uint32_t c = j ? cc.y : cc.x;
            if (c == kEmpty32 || c >= a.ncorners) continue;
            const uint32_t w = c >> 5, bits = a.fbits[w];
            if (!((bits >> (c & 31)) & 1u)) continue;  // a value duplicate of an earlier triplet
            const uint32_t id = a.fprefix[w] + __popc(bits & ((1u << (c & 31)) - 1u));
            if (id >= a.vertex_capacity) continue;
            const uint4 tr = slot_triplet(a, 2 * p + j, c, fast);
            uint32_t k[8];
            gather_key(a, tr.x, tr.y, tr.z, k);
            if (DETECT) {
                bool bad = false;
#pragma unroll
                for (int i = 0; i < 8; ++i) bad |= nonfinite_bits(k[i]);
                uint32_t pmn[3] = {neutral, neutral, neutral}, pmx[3] = {neutral, neutral, neutral};
#pragma unroll
                for (int i = 0; i < 3; ++i)
                    if ((k[i] & 0x7FFFFFFFu) <= 0x7F800000u) pmn[i] = pmx[i] = enc_float(__uint_as_float(k[i]));
                if (bsh == 0xFFFFFFFFu || bsh == tr.w) {
                    bsh = tr.w;
#pragma unroll
                    for (int i = 0; i < 3; ++i) {
                        mn[i] = min(mn[i], pmn[i]);
                        mx[i] = max(mx[i], pmx[i]);
                    }
                } else {
                    mixed = true;
                    box_atomics(a.ext_enc + 6u * tr.w, pmn, pmx);
                }
                if (bad) {
                    a.status->has_unmatchable = 1;
                } else {
                    uint32_t kc[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) kc[i] = canon_bits(k[i]);
                    insert_value(a, c, kc, tr.w, a.shape_off[tr.w], a.shape_off[tr.w + 1], spatial);
                }
            }
            float4 tangent = make_float4(0.f, 0.f, 0.f, 0.f);  // a vertex no triangle references keeps zeros
            const uint32_t lf = fuse ? a.lastface[id] : 0;
            if (lf) {
                // the winning (highest-numbered) triangle: its corners' slots give their triplets
                const uint32_t* sl = a.slot1 + 3ull * (lf - 1);
                Tri t;
#pragma unroll
                for (int q = 0; q < 3; ++q) {
                    // the corner's own triplet, via its slot's record when that is a single aligned read
                    const uint4 ct = slot_triplet(a, sl[q], 3 * (lf - 1) + q, fast);
                    const float* pp = a.pos + 3ull * min(ct.x, a.np - 1);
                    const float2 uu = __ldg(reinterpret_cast<const float2*>(a.uv) + min(ct.z, a.nt - 1));
                    t.px[q] = __ldg(pp); t.py[q] = __ldg(pp + 1); t.pz[q] = __ldg(pp + 2);
                    t.u[q] = uu.x; t.v[q] = uu.y;
                }
                tangent = tangent_of(t);
            }
            store_vertex(a.out_vtx + 4ull * id, k, tangent);
        }
        if (DETECT) {  // bounding box: one set of (improvement-checked) atomics per warp when its slots share a shape
            const bool have = (bsh != 0xFFFFFFFFu);
            const uint32_t any = __ballot_sync(0xFFFFFFFFu, have);
            if (any) {
                const uint32_t sh0 = __shfl_sync(0xFFFFFFFFu, bsh, __ffs(any) - 1);
                if (__all_sync(0xFFFFFFFFu, !have || (bsh == sh0 && !mixed))) {
#pragma unroll
                    for (int i = 0; i < 3; ++i) {
                        mn[i] = __reduce_min_sync(0xFFFFFFFFu, mn[i]);
                        mx[i] = __reduce_max_sync(0xFFFFFFFFu, mx[i]);
                    }
                    if (lane == 0) box_atomics(a.ext_enc + 6u * sh0, mn, mx);
                } else if (have) {
                    box_atomics(a.ext_enc + 6u * bsh, mn, mx);
                }
            }
        }
    }
}

__global__ void __launch_bounds__(kThreads) k_write_unmatchable_vertices(ImportArgs a) {
    pdl_enter();
    if (a.status->has_unmatchable == 0) return;
    for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < a.ncorners; c += gridDim.x * blockDim.x) {
        if (a.aux1[a.slot1[c]] != kUnmatchable) continue;
        const uint32_t id = rank_of(a, c);
        if (id >= a.vertex_capacity) continue;
        const int32_t* o = a.idx + 3ull * c;
        uint32_t k[8];
        gather_key(a, (uint32_t)__ldg(o), (uint32_t)__ldg(o + 1), (uint32_t)__ldg(o + 2), k);
        store_vertex(a.out_vtx + 4ull * id, k, make_float4(0.f, 0.f, 0.f, 0.f));
    }
}

// ---------------------------------------------------------------------------------------------
// Tangents / normals as a separate pass: one thread per vertex evaluates the winning triangle from the
// finished vertex records (the stand-alone utilities, and imports with value duplicates or NaN/Inf keys).
// ---------------------------------------------------------------------------------------------
// lastface[v] = 1 + highest triangle id that references vertex v (0 = unreferenced).
__global__ void __launch_bounds__(kThreads) k_last_face(const uint32_t* __restrict__ indices, uint32_t index_count,
                                                        uint32_t vertex_count, uint32_t* lastface,
                                                        uint32_t* bad_flag) {
    const uint32_t ntri = index_count / 3;
    for (uint32_t f = blockIdx.x * blockDim.x + threadIdx.x; f < ntri; f += gridDim.x * blockDim.x) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const uint32_t v = indices[3ull * f + k];
            if (v < vertex_count) atomicMax(lastface + v, f + 1);
            else *bad_flag = 1;
        }
    }
}

// mode 0: tangents (writes float4 @48), mode 1: normals (writes 3 floats @12).
// shape_off / vbase may be null for a single index space (the stand-alone utilities).
template <int MODE>
__global__ void __launch_bounds__(kThreads) k_per_vertex_frames(float4* vtx, const uint32_t* __restrict__ indices,
                                                                const uint32_t* __restrict__ lastface,
                                                                const uint32_t* __restrict__ total_ptr, uint32_t total_max,
                                                                const uint32_t* __restrict__ shape_off,
                                                                const uint32_t* __restrict__ vbase, uint32_t nshapes,
                                                                uint32_t vertex_limit, const Status* fused_status) {
    pdl_enter();
    // import pipeline: nothing to do when k_write_vertices already produced the tangents
    if (fused_status && fused_status->has_value_dup == 0 && fused_status->has_unmatchable == 0) return;
    const uint32_t total = total_ptr ? min(*total_ptr, total_max) : total_max;
    for (uint32_t v = blockIdx.x * blockDim.x + threadIdx.x; v < total; v += gridDim.x * blockDim.x) {
        const uint32_t lf = lastface[v];
        if (lf == 0) continue;  // referenced by no triangle: the reference never touches it
        const uint32_t f = lf - 1;
        uint32_t base = 0;
        if (shape_off) base = vbase[find_shape(shape_off, nshapes, 3 * f)];
        uint32_t i[3];
        bool ok = true;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            i[k] = indices[3ull * f + k] + base;
            ok &= (i[k] < vertex_limit);
        }
        if (!ok) continue;
        Tri t;
        load_tri(vtx, i, t);
        if (MODE == 0) {
            vtx[4ull * v + 3] = tangent_of(t);
        } else {
            const float3 n = normal_of(t);
            float* dst = reinterpret_cast<float*>(vtx + 4ull * v) + 3;
            dst[0] = n.x; dst[1] = n.y; dst[2] = n.z;
        }
    }
}

// The same result from the face side (stand-alone utilities): one thread per triangle reads its three indices
// (coalesced, every sector used -- the per-vertex variant above gathers 12 bytes out of one or two 32-byte sectors
// per vertex), asks for each corner whether this triangle is that vertex's last one, and only then evaluates the
// frame once and writes it to the vertices it wins.  A triangle that repeats a vertex writes the same value twice.
template <int MODE>
__global__ void __launch_bounds__(kThreads) k_face_frames(float4* vtx, const uint32_t* __restrict__ indices,
                                                          uint32_t index_count, const uint32_t* __restrict__ lastface,
                                                          uint32_t vertex_limit) {
    const uint32_t ntri = index_count / 3;
    for (uint32_t f = blockIdx.x * blockDim.x + threadIdx.x; f < ntri; f += gridDim.x * blockDim.x) {
        uint32_t i[3];
        bool ok = true, wins[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            i[k] = __ldg(indices + 3ull * f + k);
            ok &= (i[k] < vertex_limit);
        }
        if (!ok) continue;  // k_last_face has raised the flag; the reference would read out of bounds here
        bool any = false;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            wins[k] = (__ldg(lastface + i[k]) == f + 1);
            any |= wins[k];
        }
        if (!any) continue;
        Tri t;
        load_tri(vtx, i, t);
        if (MODE == 0) {
            const float4 tg = tangent_of(t);
#pragma unroll
            for (int k = 0; k < 3; ++k)
                if (wins[k]) vtx[4ull * i[k] + 3] = tg;
        } else {
            const float3 n = normal_of(t);
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                if (!wins[k]) continue;
                float* dst = reinterpret_cast<float*>(vtx + 4ull * i[k]) + 3;
                dst[0] = n.x; dst[1] = n.y; dst[2] = n.z;
            }
        }
    }
}

// Packs one column of the 64-byte vertex records (word offset `woff`, `nw` words: tangent = 12/4, normal = 3/3) into a
// dense array: the host utilities bring back only what they changed.
__global__ void __launch_bounds__(kThreads) k_pack_column(const float* __restrict__ vtx, uint32_t vertex_count, uint32_t woff,
                                                          uint32_t nw, float* __restrict__ out) {
    const uint64_t total = (uint64_t)vertex_count * nw;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t v = i / nw;
        out[i] = vtx[v * 16 + woff + (uint32_t)(i - v * nw)];
    }
}

// ---------------------------------------------------------------------------------------------
// Batches of small meshes (vkm_import_batch): several meshes, each with attribute arrays of its own, are laid
// end to end in one set of device arrays and imported as ONE multi-shape stream.  This kernel turns every
// mesh's file-local indices into indices of the concatenated arrays (in the library's own device copy of the
// stream) and checks them against that mesh's own record counts, so that a bad index cannot alias a neighbour.
// ---------------------------------------------------------------------------------------------
struct MeshSpan {
    uint32_t c_begin, c_end;               // corner range of the mesh in the concatenated stream
    uint32_t pos_base, nrm_base, uv_base;  // first record of the mesh in the concatenated attribute arrays
    uint32_t np, nn, nt;                   // the mesh's own record counts
};

__global__ void __launch_bounds__(kThreads) k_rebase_indices(int32_t* idx, const MeshSpan* __restrict__ spans,
                                                             uint32_t nmesh, uint32_t ncorners, uint32_t* bad) {
    for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < ncorners; c += gridDim.x * blockDim.x) {
        uint32_t lo = 0, hi = nmesh;  // largest m with c_begin <= c (meshes without corners are skipped)
        while (hi - lo > 1) {
            const uint32_t mid = (lo + hi) >> 1;
            if (__ldg(&spans[mid].c_begin) <= c) lo = mid; else hi = mid;
        }
        const MeshSpan sp = spans[lo];
        int32_t* o = idx + 3ull * c;
        const uint32_t v = (uint32_t)o[0], n = (uint32_t)o[1], t = (uint32_t)o[2];
        if (v >= sp.np || n >= sp.nn || t >= sp.nt) {  // negative ints are huge as unsigned
            bad[lo] = 1;
            continue;
        }
        o[0] = (int32_t)(v + sp.pos_base);
        o[1] = (int32_t)(n + sp.nrm_base);
        o[2] = (int32_t)(t + sp.uv_base);
    }
}

}  // namespace vkm
