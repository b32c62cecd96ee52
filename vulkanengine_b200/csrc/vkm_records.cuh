// Record schedule of the import path (sm_100a): the schedule an import runs with while the plan expects a
// "clean" mesh (no two distinct triplets with equal values, no NaN/Inf key).  SURVEY.md section 8(a) rows A2-A5.
//
// The general schedule (vkm_kernels.cuh) keeps a hashed u32 table and decides "same triplet" by a second, dependent
// read of the occupant's triplet from the index stream.  Here every record k of the KEY attribute (the largest of
// the v / vn / vt arrays) owns slot k directly -- no hashing, no probing:
//     rfirst[k]  u32    lowest corner seen so far of the record's first variant   (dense: 8 slots per sector)
//     ident[k]   uint2  (a, b) = that variant's other two indices, published by the thread that claimed rfirst[k]
// so a corner decides from two loads whose addresses it knows from its own indices, and the four corners of a thread
// have all eight loads in flight together.  Until the claimer's identity store is visible (a window of one round
// trip) a reader falls back to the general schedule's exact compare through the index stream.  Triplets that are not
// the first variant of their key record (same v, other vn/vt: seams) go to a small hashed side table.  `first` is
// lowered with atomicMin: integer, so the result does not depend on scheduling.  A slot (key record or side-table
// entry) is one distinct (shape, v, n, t).  Measured on B200: 16-byte {first,a,b,last} records claimed with a
// 128-bit CAS cost 61 M L2 atomics per 50 M corners and ran 1.6x slower than this layout.
//
//   k_clear_records      slots / side table / fingerprint table -> empty, bitmap -> 0, boxes, status
//   k_insert_records     corner domain: 3 x 128-bit index loads per 4 corners, slot per corner
//   k_flag_records       slot domain: flag bit of every slot's first corner
//   (k_scan_tiles / k_scan_finish of vkm_kernels.cuh: stable first-occurrence ranks)
//   k_rank_records       slot domain: vid_of[slot] = local vertex index
//   k_write_indices_rec  corner domain: out_idx[c] = vid_of[slot[c]] (a gather) + atomicMax(last face of the vertex)
//   k_write_vertices_rec slot domain: gather 8 floats, box, NaN/Inf screen, fingerprint of the value key (the proof
//                        that the import IS clean), tangent of the vertex's last triangle, 64-byte record
//
// If the status block says the mesh was not clean (or a table overflowed), vkm_plan_finish re-runs the import
// with the general schedule of vkm_kernels.cuh; results are only observable after finish.
#pragma once

#include "vkm_kernels.cuh"

namespace vkm {

// KEY = which index addresses the slot (0: v, 1: vn, 2: vt); (a, b) are the other two in v, vn, vt order.
template <int KEY>
__device__ __forceinline__ void split_triplet(uint32_t v, uint32_t n, uint32_t t, uint32_t& k, uint32_t& x, uint32_t& y) {
    if (KEY == 0) { k = v; x = n; y = t; }
    else if (KEY == 1) { k = n; x = v; y = t; }
    else { k = t; x = v; y = n; }
}
template <int KEY>
__device__ __forceinline__ void join_triplet(uint32_t k, uint32_t x, uint32_t y, uint32_t& v, uint32_t& n, uint32_t& t) {
    if (KEY == 0) { v = k; n = x; t = y; }
    else if (KEY == 1) { n = k; v = x; t = y; }
    else { t = k; v = x; n = y; }
}

__global__ void __launch_bounds__(kThreads) k_clear_records(ImportArgs a) {
    pdl_enter();
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    const uint4 ff = make_uint4(kEmpty32, kEmpty32, kEmpty32, kEmpty32);
    // (the plan's arrays are padded, so whole uint4 stores past nkey stay in bounds)
    uint4* p1 = reinterpret_cast<uint4*>(a.rfirst);
    for (uint32_t i = tid; i < (a.nkey + 3) / 4; i += nth) p1[i] = ff;
    uint4* p2 = reinterpret_cast<uint4*>(a.ident);
    for (uint32_t i = tid; i < (a.nkey + 1) / 2; i += nth) p2[i] = ff;
    for (uint32_t i = tid; i < a.msec; i += nth) a.sec_first[i] = kEmpty32;
    uint4* t2 = reinterpret_cast<uint4*>(a.table2);
    for (uint32_t i = tid; i < (a.m2 >> 2); i += nth) t2[i] = ff;  // m2 is a multiple of 8 here
    for (uint32_t i = tid; i <= a.nwords; i += nth) a.fbits[i] = 0;
    for (uint32_t i = tid; i < a.nshapes * 6u; i += nth) a.ext_enc[i] = 0x80000000u;
    if (tid < sizeof(Status) / 4) reinterpret_cast<uint32_t*>(a.status)[tid] = 0;
    if (tid < 6) a.hbox[tid] = tid < 3 ? 0xFFFFFFFFu : 0u;  // min words / max words of the position box (k_flag_records)
}

// Side table: triplets whose key record is owned by another variant (or by another shape).  Same protocol as
// the general schedule's triplet table: CAS claim, exact compare against the occupant's own triplet, atomicMin.
// (scalar arguments: a by-reference ImportArgs would force a local-memory copy of the kernel parameters)
__device__ __noinline__ uint32_t insert_side(uint32_t* sec_first, uint32_t msec, uint32_t seed, const int32_t* idx,
                                             uint32_t c, uint32_t v, uint32_t n, uint32_t t, uint32_t sh,
                                             uint32_t sh_begin, uint32_t sh_end) {
    uint32_t s = reduce32((uint32_t)hash_triplet(v, n, t, sh, seed), msec);
    for (int probe = 0; probe < kMaxProbe; ++probe) {
        uint32_t cur = __ldcg(sec_first + s);
        if (cur == kEmpty32) {
            const uint32_t prev = atomicCAS(sec_first + s, kEmpty32, c);
            cur = (prev == kEmpty32) ? c : prev;
        }
        if (cur == c || same_triplet(idx, cur, v, n, t, sh_begin, sh_end)) {
            if (cur > c) atomicMin(sec_first + s, c);
            return s;
        }
        s = (s + 1 == msec) ? 0 : s + 1;
    }
    return kEmpty32;  // table full
}

#ifndef VKM_REC_MINB
#define VKM_REC_MINB 4
#endif
// Per thread and iteration: 4 consecutive corners = 48 bytes of the index stream = three 128-bit loads.  An
// iteration costs three memory round trips, not one per corner: (1) the index words, (2) the slots' first-corner and
// identity words, all eight loads in flight together, (3) the claims of whichever slots were empty, all issued before
// the first result is used.  Measured on B200 and dropped (16.8 M-triangle grid, this kernel 0.277 ms): index words
// staged through a per-thread cp.async ring in shared memory 0.38 ms; fully coalesced index loads through a
// warp-private shared-memory transpose 0.317 ms; L1-cached slot loads 0.273 ms (no gain); prefetch.global.L2 of the
// index words four iterations ahead 0.277 ms (no gain); 16-byte AoS records with a 128-bit CAS 0.43-0.56 ms.
template <int KEY, bool MULTI>
__global__ void __launch_bounds__(kThreads, VKM_REC_MINB) k_insert_records(ImportArgs a) {
    pdl_enter();
    const uint32_t ngroups = (a.ncorners + 3) >> 2;
    // Blocked, not grid-strided: every CTA walks ONE contiguous piece of the stream front to back.  With a grid stride
    // the whole grid sweeps a window of ~1 M consecutive corners at once, so all the corners that share a vertex (a
    // few thousand corners apart in a coherent mesh) meet their slot unclaimed at the same moment: measured 3.1 claim
    // attempts and 2 atomicMin per vertex.  Walking in order, the later ones arrive after the claim is visible and
    // need no atomic at all.
    const uint32_t per_cta = (uint32_t)(((uint64_t)ngroups + gridDim.x - 1) / gridDim.x);
    const uint32_t g_end = min(ngroups, (blockIdx.x + 1) * per_cta);
    const uint32_t stride = blockDim.x;
    for (uint32_t g = blockIdx.x * per_cta + threadIdx.x; g < g_end; g += stride) {
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
            for (int i = 0; i < 12; ++i) w[i] = (c0 + i / 3 < a.ncorners) ? __ldg(a.idx + 3ull * c0 + i) : -1;
        }
        uint32_t key[4], oa[4], ob[4], f[4];
        uint2 id[4];
        bool valid[4], tried[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t v = (uint32_t)w[3 * i], n = (uint32_t)w[3 * i + 1], t = (uint32_t)w[3 * i + 2];
            valid[i] = (c0 + i < a.ncorners) && v < a.np && n < a.nn && t < a.nt;  // negative ints are huge as unsigned
            split_triplet<KEY>(v, n, t, key[i], oa[i], ob[i]);
            if (!valid[i]) key[i] = 0;
            f[i] = __ldcg(a.rfirst + key[i]);
            id[i] = __ldcg(a.ident + key[i]);
        }
        // claims: a corner that saw an unclaimed slot tries to become its first corner; f[i] becomes what the slot held
        // before (still all-ones: claimed).  Two corners of this thread with one key: the second claim fails and
        // returns the first corner.
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            tried[i] = valid[i] && f[i] == kEmpty32;
            if (tried[i]) f[i] = atomicCAS(a.rfirst + key[i], kEmpty32, c0 + i);
        }
        uint32_t sh = 0, sh_begin = 0, sh_end = a.ncorners;
        if (MULTI) {
            sh = find_shape_near(a.shape_off, a.nshapes, min(c0, a.ncorners - 1));
            sh_begin = a.shape_off[sh];
            sh_end = a.shape_off[sh + 1];
        }
        uint32_t slots[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t c = c0 + i;
            slots[i] = 0;
            if (c >= a.ncorners) continue;
            if (!valid[i]) {
                a.status->bad_index = 1;
                continue;
            }
            if (MULTI && c >= sh_end) {
                sh = find_shape(a.shape_off, a.nshapes, c);
                sh_begin = a.shape_off[sh];
                sh_end = a.shape_off[sh + 1];
            }
            if (tried[i] && f[i] == kEmpty32) {  // claimed: publish the variant's identity
                __stcg(a.ident + key[i], make_uint2(oa[i], ob[i]));
                slots[i] = key[i];
                continue;
            }
            // identity not published when it was read (the slot was claimed in this very moment, typically by a
            // neighbouring corner of this warp): compare with the claimer's own triplet in the index stream, which is
            // exact and, for a neighbour, an L1 hit
            const uint2 q = id[i];
            uint32_t v, n, t;
            join_triplet<KEY>(key[i], oa[i], ob[i], v, n, t);
            bool match;
            if (q.x != kEmpty32) match = (q.x == oa[i] && q.y == ob[i] && (!MULTI || (f[i] >= sh_begin && f[i] < sh_end)));
            else match = same_triplet(a.idx, f[i], v, n, t, sh_begin, sh_end);
            if (match) {
                if (c < f[i]) atomicMin(a.rfirst + key[i], c);
                slots[i] = key[i];
            } else {
                const uint32_t s2 = insert_side(a.sec_first, a.msec, a.seed, a.idx, c, v, n, t, sh, sh_begin, sh_end);
                if (s2 == kEmpty32) a.status->overflow1 = 1;
                else slots[i] = a.nkey + s2;
            }
        }
        if (c0 + 4 <= a.ncorners) {
            *reinterpret_cast<uint4*>(a.slot1 + c0) = make_uint4(slots[0], slots[1], slots[2], slots[3]);
        } else {
            for (int i = 0; i < 4 && c0 + i < a.ncorners; ++i) a.slot1[c0 + i] = slots[i];
        }
    }
}

// First corner of slot s (kEmpty32 = unused).
__device__ __forceinline__ uint32_t slot_first(const ImportArgs& a, uint32_t s) {
    return (s < a.nkey) ? __ldg(a.rfirst + s) : __ldg(a.sec_first + (s - a.nkey));
}

// The slot-domain helper passes handle FOUR key slots per thread (one 128-bit load of rfirst): with one 4-byte load per
// thread they spend their time waiting for that single load (ncu: 49 % of the samples on it).
// Flag the first corner of every slot (speculation: the import is clean, so every distinct triplet is a vertex).
__global__ void __launch_bounds__(kThreads) k_flag_records(ImportArgs a) {
    pdl_enter();
    const uint32_t nq = (a.nkey + 3) >> 2, tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    for (uint32_t q = tid; q < nq; q += nth) {
        const uint4 c4 = __ldcs(reinterpret_cast<const uint4*>(a.rfirst) + q);  // padding past nkey is all-ones
        const uint32_t c[4] = {c4.x, c4.y, c4.z, c4.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (4 * q + i < a.nkey && c[i] < a.ncorners) atomicOr(a.fbits + (c[i] >> 5), 1u << (c[i] & 31));
    }
    for (uint32_t s = tid; s < a.msec; s += nth) {
        const uint32_t c = __ldcs(a.sec_first + s);
        if (c < a.ncorners) atomicOr(a.fbits + (c >> 5), 1u << (c & 31));  // kEmpty32 >= ncorners
    }
    if (a.cell_vpc) sample_position_box(a, tid, nth);  // (steers the fingerprint table's bucket choice: vkm_device.cuh)
}

// vid_of[slot] = the slot's GLOBAL vertex id (its rank among all vertex-creating corners; kEmpty32 for unused slots).
// The index pass subtracts the shape's base, which it tracks per thread anyway.
__global__ void __launch_bounds__(kThreads) k_rank_records(ImportArgs a) {
    pdl_enter();
    if (blockIdx.x == 0 && threadIdx.x == 0) store_cell_map(a);
    const uint32_t nq = (a.nkey + 3) >> 2, tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    for (uint32_t q = tid; q < nq; q += nth) {
        const uint4 c4 = __ldg(reinterpret_cast<const uint4*>(a.rfirst) + q);
        const uint32_t c[4] = {c4.x, c4.y, c4.z, c4.w};
        uint32_t bits[4], pre[4], vid[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {  // all eight gathers in flight
            const uint32_t w = min(c[i], a.ncorners - 1) >> 5;
            bits[i] = a.fbits[w];
            pre[i] = a.fprefix[w];
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            vid[i] = kEmpty32;
            if (c[i] < a.ncorners) {
                vid[i] = pre[i] + __popc(bits[i] & ((1u << (c[i] & 31)) - 1u));
            }
        }
        if (4 * q + 4 <= a.nkey) {
            *reinterpret_cast<uint4*>(a.vid_of + 4 * q) = make_uint4(vid[0], vid[1], vid[2], vid[3]);
        } else {
            for (int i = 0; i < 4 && 4 * q + i < a.nkey; ++i) a.vid_of[4 * q + i] = vid[i];
        }
    }
    for (uint32_t s = tid; s < a.msec; s += nth) {
        const uint32_t c = __ldg(a.sec_first + s);
        uint32_t vid = kEmpty32;
        if (c < a.ncorners) vid = rank_of(a, c);
        a.vid_of[a.nkey + s] = vid;
    }
}

// Corner domain: the index buffer is a gather of the slots' vertex indices; the vertex's last triangle is an atomicMax.
// (Measured and dropped: two groups of four corners per thread, 0.134 -> 0.162 ms.)
template <bool MULTI>
__global__ void __launch_bounds__(kThreads) k_write_indices_rec(ImportArgs a) {
    pdl_enter();
    const uint32_t nslots = a.nkey + a.msec;
    const uint32_t ngroups = (a.ncorners + 3) >> 2;
    for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < ngroups; g += gridDim.x * blockDim.x) {
        const uint32_t c0 = g << 2;
        uint32_t s[4] = {0, 0, 0, 0}, li[4];
        const bool full = (c0 + 4 <= a.ncorners);
        if (full) {
            const uint4 q = __ldcs(reinterpret_cast<const uint4*>(a.slot1 + c0));
            s[0] = q.x; s[1] = q.y; s[2] = q.z; s[3] = q.w;
        } else {
            for (int i = 0; i < 4 && c0 + i < a.ncorners; ++i) s[i] = a.slot1[c0 + i];
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) li[i] = __ldg(a.vid_of + min(s[i], nslots - 1));
        uint32_t sh = 0, sh_end = a.ncorners, base = 0;
        if (MULTI) {
            sh = find_shape_near(a.shape_off, a.nshapes, min(c0, a.ncorners - 1));
            sh_end = a.shape_off[sh + 1];
            base = a.vbase[sh];
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t c = c0 + i;
            if (c >= a.ncorners) break;
            if (MULTI && c >= sh_end) {
                sh = find_shape(a.shape_off, a.nshapes, c);
                sh_end = a.shape_off[sh + 1];
                base = a.vbase[sh];
            }
            // a later corner of this thread that names the same vertex raises the same word higher: leave it to that one
            bool later = false;
#pragma unroll
            for (int j = i + 1; j < 4; ++j) later |= (s[j] == s[i] && c0 + j < sh_end);
            const uint32_t gv = li[i];
            if (!later && gv < a.vertex_capacity) atomicMax(a.lastface + gv, c / 3 + 1);
            li[i] = gv - base;  // index local to the shape
        }
        if (full) {
            __stcs(reinterpret_cast<uint4*>(a.out_idx + c0), make_uint4(li[0], li[1], li[2], li[3]));
        } else {
            for (int i = 0; i < 4 && c0 + i < a.ncorners; ++i) a.out_idx[c0 + i] = li[i];
        }
    }
}

// The clean-import proof of the record schedule: every distinct triplet's canonical 8-float key leaves a 32-bit
// fingerprint in a bucketed table (bucket = one 32-byte sector = 8 words, read with one 256-bit load).  Two EQUAL keys
// hash to the same bucket sequence and the same fingerprint, and the later of them is guaranteed to see the earlier one
// (every key walks the words of a bucket in an order of its own that equal keys share; a claim at a word has looked at
// every word before it in that order, either in its snapshot or as the value its failed CAS returned, and words never
// change once set): it raises has_value_dup, and vkm_plan_finish re-runs the import with
// the general schedule, which decides exactly.  Two DIFFERENT keys with one fingerprint in one bucket (about
// U * 4 / 2^32 per import) only cost that re-run, never a wrong result.  Against linear probing over single words
// this needs one round trip for the whole bucket and, at load 1/2, almost never a second bucket, so the lanes of
// a warp stay together.
// Multi-shape imports (MULTI): a shape is its own dedup domain, so it also gets its own stretch of the table, sized
// by its share of the corners.  The slots of one shape are neighbours in the slot domain, hence a warp's probes stay
// inside a few hundred kilobytes instead of scattering over the whole table (10 000-mesh batch: an 850 MB table, every
// probe a DRAM and TLB miss otherwise).  A shape that needs more than its stretch holds overflows, and the re-run uses a
// larger table / the general schedule.
// Which bucket a value key starts in.  A scrambled hash sends neighbouring vertices to opposite ends of the table:
// every probe is a random DRAM access (measured: 0.17 of the vertex pass's 0.38 ms on the 16.8 M grid).  For a single
// shape in hash generation 0 the bucket comes from the key's POSITION instead -- its cell on a 2-D lattice over the two
// longest axes of the position box, row-major -- so that vertices that are neighbours in space, which for a
// coherently stored mesh are neighbours in the slot domain, start in neighbouring buckets: the table is swept, not
// hammered.  Equal keys have equal positions, hence equal cells: the proof is untouched.  Meshes whose vertices pile up
// in few cells (thin volumes, far outliers in the box) overflow the probe limit; the re-run's generation is >= 1 and
// scrambles.
// The map is worked out once per import by one thread of k_rank_records (box and vertex count are final by then) and
// parked behind the box words: hbox[8..13] = {on, axes, cells per row, min0, min1, 1 / cell size}.
__device__ __forceinline__ uint32_t cell_bucket(const uint32_t* __restrict__ cm, const uint32_t kc[8], uint32_t nb) {
    const uint32_t axes = __ldg(cm + 1), g0 = __ldg(cm + 2);
    const float min0 = __uint_as_float(__ldg(cm + 3)), min1 = __uint_as_float(__ldg(cm + 4)), inv_h = __uint_as_float(__ldg(cm + 5));
    const float p0 = pick3(axes & 3u, __uint_as_float(kc[0]), __uint_as_float(kc[1]), __uint_as_float(kc[2]));
    const float p1 = pick3(axes >> 2, __uint_as_float(kc[0]), __uint_as_float(kc[1]), __uint_as_float(kc[2]));
    const float c0 = fminf(fmaxf(__fmul_rn(__fsub_rn(p0, min0), inv_h), 0.0f), 4.0e6f);
    const float c1 = fminf(fmaxf(__fmul_rn(__fsub_rn(p1, min1), inv_h), 0.0f), 4.0e6f);
    // Four CONSECUTIVE cells of a lattice row share a bucket, each with its own pair of words (returned in bits 30-31):
    // the lanes of a warp -- consecutive slots, for a coherently stored mesh consecutive cells of a row -- then touch 8
    // sectors of the table instead of 32 and do not meet at the same word, and the table is swept once.  (Four lattice
    // ROWS per bucket, one cell of each: every lane its own sector, vertex pass 0.286 -> 0.264 ms this way round.  A
    // scrambled start per row group: runs overlap like a Poisson process, a quarter of the table ends up over-full and
    // the probes crawl, 0.31 -> 0.42 ms.)
    const uint64_t id = (uint64_t)((uint32_t)c0 >> 2) + (uint64_t)(uint32_t)c1 * ((g0 + 3u) >> 2);
    return (uint32_t)(id % nb) | (((uint32_t)c0 & 3u) << 30);
}

// Where the probe of a value key starts: fingerprint, first bucket, and (MULTI) the shape's stretch of the table.
template <bool MULTI>
__device__ __forceinline__ void fingerprint_start(const ImportArgs& a, bool spatial, const uint32_t kc[8], uint32_t sh,
                                                  uint32_t& fp, uint32_t& b, uint32_t& b_lo, uint32_t& nb) {
    const uint64_t h = hash_key8(kc, sh, a.seed);
    fp = (uint32_t)(h >> 32);
    if (fp == kEmpty32) fp = 0xFFFFFFFEu;
    b_lo = 0;
    nb = a.m2 >> 3;
    if (MULTI) {
        const uint64_t all = nb;
        b_lo = (uint32_t)(all * __ldg(a.shape_off + sh) / a.ncorners);
        const uint32_t b_hi = (uint32_t)(all * __ldg(a.shape_off + sh + 1) / a.ncorners);
        nb = max(b_hi, b_lo + 1) - b_lo;  // (b_lo < all because the shape has corners)
    }
    b = (!MULTI && spatial) ? cell_bucket(a.hbox + 8, kc, nb) : reduce32((uint32_t)h, nb);
}

__device__ __forceinline__ void probe_fingerprint(const ImportArgs& a, bool spatial, uint32_t fp, uint32_t b, uint32_t b_lo, uint32_t nb) {
    const int limit = (int)min(nb, 256u);
    const uint32_t rot = (b >> 30) * 2u;  // the key's own first word of every bucket (0 for hashed buckets)
    b &= 0x3FFFFFFFu;
    // A full bucket sends the key on.  Hashed buckets: to the next one.  Position-ordered buckets: to the same column
    // of the NEXT LATTICE ROW, not along the row -- the lattice has a few more rows than the table (margins, rounding up
    // to whole buckets), so the last rows wrap onto the first and fill them exactly; along such a row every bucket is
    // full and one extra key would crawl to the probe limit (67 M adversarial grid), while the row below has room.
    uint32_t step = 1;
    if (spatial) {
        const uint32_t per_row = (__ldg(a.hbox + 10) + 3u) >> 2;
        if (per_row < nb) step = per_row;
    }
    for (int probe = 0; probe < limit; ++probe) {
        uint32_t* base = a.table2 + 8ull * (b_lo + b);
        uint32_t w[8];
        asm volatile("ld.global.cg.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7])
                     : "l"(base));
        bool seen = false;
        uint32_t empties = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            seen |= (w[j] == fp);
            empties |= (w[j] == kEmpty32 ? 1u : 0u) << j;
        }
        // ONE claim per lane at the lane's own first empty word (a per-word loop would make the warp walk all eight
        // words one atomic round trip at a time); a second turn only when another thread took that word meanwhile
        bool done = false;
        // word order of this key: its own pair first, then round the bucket (equal keys have equal orders: the proof's
        // argument runs along that order)
        empties = ((empties >> rot) | (empties << (8u - rot))) & 0xFFu;
        while (!seen && !done && empties) {
            const int j = ((__ffs(empties) - 1) + (int)rot) & 7;
            const uint32_t prev = atomicCAS(base + j, kEmpty32, fp);
            if (prev == kEmpty32) done = true;
            else if (prev == fp) seen = true;
            empties &= empties - 1;
        }
        if (seen) {
            a.status->has_value_dup = 1;
            return;
        }
        if (done) return;
        b += step;
        if (b >= nb) b -= nb;
    }
    a.status->overflow2 = 1;
}

// (v, vt) of corner `c`, whose slot is `s`: from the slot's identity when it is a key record, else from the index stream.
template <int KEY>
__device__ __forceinline__ void corner_v_t(const ImportArgs& a, uint32_t s, uint32_t c, uint32_t& v, uint32_t& t) {
    if (s < a.nkey) {
        const uint2 ab = __ldg(a.ident + s);
        if (KEY == 0) { v = s; t = ab.y; }
        else if (KEY == 1) { v = ab.x; t = ab.y; }
        else { t = s; v = ab.x; }
    } else {
        const int32_t* o = a.idx + 3ull * min(c, a.ncorners - 1);
        v = (uint32_t)__ldg(o);
        t = (uint32_t)__ldg(o + 2);
    }
}

#ifndef VKM_VTX_MINB
#define VKM_VTX_MINB 5
#endif
// Slot domain.  Also the pass that PROVES the import clean (NaN/Inf screen + fingerprint of every value key).
template <int KEY, bool MULTI>
__global__ void __launch_bounds__(kThreads, VKM_VTX_MINB) k_write_vertices_rec(ImportArgs a) {
    pdl_enter();
    const uint32_t nslots = a.nkey + a.msec;
    const uint32_t padded = (nslots + 31) & ~31u;  // whole warps for the box reduction
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t neutral = 0x80000000u;
    const bool spatial = !MULTI && __ldg(a.hbox + 8) != 0;
    for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < padded; s += gridDim.x * blockDim.x) {
        uint32_t first = kEmpty32, lid = kEmpty32, v = 0, n = 0, t = 0;
        if (s < a.nkey) {
            // (a used slot has a vertex id; the first corner itself is only needed to find the shape: 34 MB less to read
            // on the 16.8 M grid, vertex pass 0.264 -> 0.259 ms)
            lid = __ldg(a.vid_of + s);
            first = MULTI ? __ldg(a.rfirst + s) : (lid != kEmpty32 ? 0u : kEmpty32);
            const uint2 ab = __ldg(a.ident + s);
            join_triplet<KEY>(s, ab.x, ab.y, v, n, t);
        } else if (s < nslots) {
            first = __ldg(a.sec_first + (s - a.nkey));
            if (first < a.ncorners) {
                lid = __ldg(a.vid_of + s);
                const int32_t* o = a.idx + 3ull * first;
                v = (uint32_t)__ldg(o);
                n = (uint32_t)__ldg(o + 1);
                t = (uint32_t)__ldg(o + 2);
            }
        }
        uint32_t mn[3] = {neutral, neutral, neutral}, mx[3] = {neutral, neutral, neutral};
        uint32_t bsh = 0xFFFFFFFFu;
        if (first < a.ncorners) {
            const uint32_t sh = MULTI ? find_shape_near(a.shape_off, a.nshapes, first) : 0u;
            const uint32_t id = lid;
            if (id < a.vertex_capacity) {
                const uint32_t lf = a.lastface[id];  // >= 1: the slot's own corners reference the vertex
                uint32_t k[8];
                gather_key(a, v, n, t, k);
                // single shape: the fingerprint bucket is known as soon as the attributes are; ask L2 for it now, so that the
                // tangent chain's round trips below cover its way in from DRAM (vertex pass 0.305 -> 0.286 ms on the
                // 16.8 M grid, 1.21 -> 1.12 on the 67 M grid, together with 51 registers instead of 40)
                uint32_t pf_fp = 0, pf_b = 0;
                if (!MULTI) {
                    uint32_t kc0[8], b_lo0, nb0;
#pragma unroll
                    for (int i = 0; i < 8; ++i) kc0[i] = canon_bits(k[i]);
                    fingerprint_start<false>(a, spatial, kc0, 0u, pf_fp, pf_b, b_lo0, nb0);
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(a.table2 + 8ull * (pf_b & 0x3FFFFFFFu)));
                }
                // the winning (highest-numbered) triangle: its gathers are issued before the fingerprint probe so that
                // both chains are in flight together
                const uint32_t f = min(max(lf, 1u) - 1u, (a.ncorners - 1) / 3);
                const uint32_t* sl = a.slot1 + 3ull * f;
                Tri tr;
#pragma unroll
                for (int q = 0; q < 3; ++q) {
                    const uint32_t sq = min(__ldg(sl + q), nslots - 1);
                    if (sq == s) {
                        // this corner IS the vertex being written (a vertex is always a corner of its own last
                        // triangle): its position and texcoord are already in registers
                        tr.px[q] = __uint_as_float(k[0]); tr.py[q] = __uint_as_float(k[1]); tr.pz[q] = __uint_as_float(k[2]);
                        tr.u[q] = __uint_as_float(k[6]); tr.v[q] = __uint_as_float(k[7]);
                        continue;
                    }
                    uint32_t cv, ct;
                    // (reading (v, vt) straight from the index stream instead: one hop less, 0.40 -> 0.44 ms)
                    corner_v_t<KEY>(a, sq, 3 * f + q, cv, ct);
                    const float* pp = a.pos + 3ull * min(cv, a.np - 1);
                    const float2 uu = __ldg(reinterpret_cast<const float2*>(a.uv) + min(ct, a.nt - 1));
                    tr.px[q] = __ldg(pp); tr.py[q] = __ldg(pp + 1); tr.pz[q] = __ldg(pp + 2);
                    tr.u[q] = uu.x; tr.v[q] = uu.y;
                }
                bool bad = false;
#pragma unroll
                for (int i = 0; i < 8; ++i) bad |= nonfinite_bits(k[i]);
#pragma unroll
                for (int i = 0; i < 3; ++i)
                    if ((k[i] & 0x7FFFFFFFu) <= 0x7F800000u) mn[i] = mx[i] = enc_float(__uint_as_float(k[i]));
                bsh = sh;
                if (bad) {
                    a.status->has_unmatchable = 1;
                } else if (MULTI) {
                    // (starting the probe early as above was measured on the 10 000-mesh batch: 5.56 -> 5.61 ms, no gain --
                    // a shape's stretch of the table is small enough to sit in L2 while the shape's slots go by)
                    uint32_t kc[8], fp, b, b_lo, nb;
#pragma unroll
                    for (int i = 0; i < 8; ++i) kc[i] = canon_bits(k[i]);
                    fingerprint_start<MULTI>(a, spatial, kc, sh, fp, b, b_lo, nb);
                    probe_fingerprint(a, false, fp, b, b_lo, nb);
                } else {
                    probe_fingerprint(a, spatial, pf_fp, pf_b, 0u, a.m2 >> 3);
                }
#ifndef VKM_VTX_ST128  // two 256-bit streaming stores per record (four 128-bit ones: 0.399 -> 0.387 ms the other way round)
                {
                    const float4 tg = tangent_of(tr);
                    float* dst = reinterpret_cast<float*>(a.out_vtx + 4ull * id);
                    asm volatile("st.global.cs.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(dst), "r"(k[0]), "r"(k[1]), "r"(k[2]),
                                 "r"(k[3]), "r"(k[4]), "r"(k[5]), "r"(k[6]), "r"(k[7]) : "memory");
                    asm volatile("st.global.cs.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(dst + 8), "f"(0.f), "f"(0.f), "f"(0.f),
                                 "f"(0.f), "f"(tg.x), "f"(tg.y), "f"(tg.z), "f"(tg.w) : "memory");
                }
#else
                store_vertex(a.out_vtx + 4ull * id, k, tangent_of(tr));
#endif
            }
        }
        // bounding box: one set of (improvement-checked) atomics per warp when its slots share a shape
        const bool have = (bsh != 0xFFFFFFFFu);
        const uint32_t any = __ballot_sync(0xFFFFFFFFu, have);
        if (any) {
            const uint32_t sh0 = __shfl_sync(0xFFFFFFFFu, bsh, __ffs(any) - 1);
            if (__all_sync(0xFFFFFFFFu, !have || bsh == sh0)) {
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

}  // namespace vkm
