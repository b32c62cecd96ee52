// Host side of libvkmesh_b200.so: the import plan (device buffers + kernel schedule) and the C ABI
// declared in include/vkmesh_b200.h.  CUDA runtime only -- no PyTorch, no CPU fallback.
#include "../../include/vkmesh_b200.h"

#include <cuda_runtime.h>

#include <sys/mman.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <functional>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "vkm_kernels.cuh"
#include "vkm_records.cuh"
#include "vkm_dedup_util.cuh"

using namespace vkm;

namespace {

thread_local std::string g_err;
thread_local vkm_plan_stats g_last_stats{};  // of this thread's last host-entry import

int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CU(call)                                                                                        \
    do {                                                                                                \
        cudaError_t e_ = (call);                                                                        \
        if (e_ != cudaSuccess) {                                                                        \
            int code_ = (e_ == cudaErrorMemoryAllocation) ? VKM_ERR_OOM                                 \
                        : (e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver) ? VKM_ERR_NO_DEVICE \
                                                                                         : VKM_ERR_CUDA; \
            return fail(code_, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
        }                                                                                               \
    } while (0)

constexpr uint64_t kMaxCorners = 4000000000ull;

// VKM_TRACE=1 prints a host-side time line of the host entry points to stderr.
struct Trace {
    bool on;
    std::chrono::steady_clock::time_point t0;
    Trace() : on(getenv("VKM_TRACE") != nullptr), t0(std::chrono::steady_clock::now()) {}
    void mark(const char* what) {
        if (!on) return;
        const auto t = std::chrono::steady_clock::now();
        fprintf(stderr, "[vkm] %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(t - t0).count());
        t0 = t;
    }
};

// Large result buffers are malloc()ed (the engine free()s them) but first-touch page faults on 4 KiB pages cost
// more than the PCIe transfer; asking for transparent huge pages on the interior of the block removes most of them.
void* malloc_result(uint64_t bytes) {
    void* p = malloc(std::max<uint64_t>(bytes, 1));
#ifdef MADV_HUGEPAGE
    if (p && bytes >= (8ull << 20)) {
        const uintptr_t a = ((uintptr_t)p + (2u << 20) - 1) & ~(uintptr_t)((2u << 20) - 1);
        const uintptr_t e = ((uintptr_t)p + bytes) & ~(uintptr_t)((2u << 20) - 1);
        if (e > a) madvise((void*)a, e - a, MADV_HUGEPAGE);
    }
#endif
    return p;
}

// Optional home for the results of the host entry points (vkm_set_result_memory(1)): page-locked blocks that are
// recycled through vkm_free_shape().  One block holds every buffer of one import call (vertices of all shapes, then
// indices of all shapes), so the device-to-host leg is two DMA transfers straight into the caller-visible memory:
// no bounce buffer, no host copy, no first-touch page faults.  Blocks are cudaHostAllocPortable (any device of the
// process may write them) and are kept for reuse up to `max_idle` bytes.
struct ResultBlock {
    char* base = nullptr;
    uint64_t cap = 0;
    uint32_t refs = 0;  // caller-visible pointers into the block that have not been released yet
};
struct ResultPool {
    std::mutex mu;
    std::map<uintptr_t, ResultBlock> live;
    std::vector<ResultBlock> idle;
    uint64_t idle_bytes = 0;
    uint64_t max_idle = 8ull << 30;

    ResultPool() {
        // default: a quarter of the machine's memory, at least 8 GiB
        const long pages = sysconf(_SC_PHYS_PAGES), psz = sysconf(_SC_PAGESIZE);
        if (pages > 0 && psz > 0) max_idle = std::max<uint64_t>(max_idle, (uint64_t)pages * (uint64_t)psz / 4);
        if (const char* e = getenv("VKM_RESULT_POOL_MAX_MB")) max_idle = strtoull(e, nullptr, 10) << 20;
    }
    char* acquire(uint64_t bytes, uint32_t refs) {
        std::lock_guard<std::mutex> lk(mu);
        size_t best = idle.size();
        for (size_t i = 0; i < idle.size(); ++i)
            if (idle[i].cap >= bytes && (best == idle.size() || idle[i].cap < idle[best].cap)) best = i;
        ResultBlock b;
        if (best != idle.size() && idle[best].cap <= 2 * bytes + (4ull << 20)) {
            b = idle[best];
            idle[best] = idle.back();
            idle.pop_back();
            idle_bytes -= b.cap;
        } else {
            b.cap = (bytes + (2ull << 20) - 1) & ~((2ull << 20) - 1);
            void* q = nullptr;
            if (cudaHostAlloc(&q, b.cap, cudaHostAllocPortable) != cudaSuccess) {
                cudaGetLastError();
                // make room: drop every idle block and try once more
                for (auto& i : idle) cudaFreeHost(i.base);
                idle.clear();
                idle_bytes = 0;
                if (cudaHostAlloc(&q, b.cap, cudaHostAllocPortable) != cudaSuccess) {
                    cudaGetLastError();
                    return nullptr;
                }
            }
            b.base = (char*)q;
        }
        b.refs = refs;
        live[(uintptr_t)b.base] = b;
        return b.base;
    }
    // true when p pointed into a pool block (one reference dropped), false when it is not ours
    bool release(const void* p, uint32_t n = 1) {
        if (!p) return false;
        std::lock_guard<std::mutex> lk(mu);
        auto it = live.upper_bound((uintptr_t)p);
        if (it == live.begin()) return false;
        --it;
        ResultBlock& b = it->second;
        if ((uintptr_t)p >= (uintptr_t)b.base + b.cap) return false;
        b.refs = b.refs > n ? b.refs - n : 0;
        if (b.refs == 0) {
            ResultBlock done = b;
            live.erase(it);
            if (idle_bytes + done.cap <= max_idle) {
                idle.push_back(done);
                idle_bytes += done.cap;
            } else {
                cudaFreeHost(done.base);
            }
        }
        return true;
    }
    void trim() {
        std::lock_guard<std::mutex> lk(mu);
        for (auto& i : idle) cudaFreeHost(i.base);
        idle.clear();
        idle_bytes = 0;
    }
};
ResultPool g_results;
std::atomic<int> g_result_mode{0};  // 0: malloc()ed results, 1: pooled page-locked results

// Persistent helper threads for the host-side staging copies.  Starting and joining std::threads costs tens of
// microseconds apiece, more than the copy itself for a mesh of a few megabytes; these helpers sleep on a condition
// variable between jobs.  run(n, fn) executes fn(0) .. fn(n-1), index 0 on the caller; the caller also takes over any
// index no helper has picked up yet, so a job completes even when every helper is busy with another lane's job.
// The pool is created on first use and never destroyed (its threads are detached and idle at process exit).
class HelperPool {
public:
    static HelperPool& get() {
        static HelperPool* p = new HelperPool();
        return *p;
    }
    void run(unsigned n, const std::function<void(unsigned)>& fn) {
        if (n <= 1) {
            fn(0);
            return;
        }
        Job job{&fn, 1, n, {n - 1}};
        {
            std::lock_guard<std::mutex> lk(mu_);
            const unsigned want = std::min(kMaxHelpers, n - 1);
            while (workers_ < want) {
                std::thread(&HelperPool::worker, this).detach();
                ++workers_;
            }
            jobs_.push_back(&job);
        }
        cv_.notify_all();
        fn(0);
        for (;;) {  // help with what is left, then wait for the indices that helpers are running
            unsigned idx;
            {
                std::lock_guard<std::mutex> lk(mu_);
                if (!claim(&job, idx)) break;
            }
            fn(idx);
            job.left.fetch_sub(1, std::memory_order_release);
        }
        while (job.left.load(std::memory_order_acquire) != 0) std::this_thread::yield();
    }

private:
    struct Job {
        const std::function<void(unsigned)>* fn;
        unsigned next, n;           // next unclaimed index (under mu_)
        std::atomic<unsigned> left;  // indices 1 .. n-1 not finished yet
    };
    static constexpr unsigned kMaxHelpers = 15;
    std::mutex mu_;
    std::condition_variable cv_;
    std::deque<Job*> jobs_;  // jobs with unclaimed indices
    unsigned workers_ = 0;

    // under mu_: take the next index of `job` if it is still queued; drops the job from the queue with its last index
    bool claim(Job* job, unsigned& idx) {
        for (auto it = jobs_.begin(); it != jobs_.end(); ++it) {
            if (*it != job) continue;
            idx = job->next++;
            if (job->next >= job->n) jobs_.erase(it);
            return true;
        }
        return false;
    }
    void worker() {
        for (;;) {
            Job* job;
            unsigned idx;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return !jobs_.empty(); });
                job = jobs_.front();
                if (!claim(job, idx)) continue;
            }
            (*job->fn)(idx);
            job->left.fetch_sub(1, std::memory_order_release);  // the job may be gone right after this
        }
    }
};

int sm_count(int device) {
    static int cached[64] = {0};
    if (device >= 0 && device < 64 && cached[device]) return cached[device];
    int n = 148;
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device);
    if (device >= 0 && device < 64) cached[device] = n;
    return n;
}

// Grid for a grid-stride kernel: enough CTAs to cover `items`, capped at a whole number of waves.
unsigned grid_for(uint64_t items, int threads, int device, int ctas_per_sm = 8) {
    const uint64_t need = (items + threads - 1) / threads;
    const uint64_t cap = (uint64_t)sm_count(device) * ctas_per_sm;
    return (unsigned)std::max<uint64_t>(1, std::min(need, cap));
}

template <typename T>
struct DevBuf {
    T* p = nullptr;
    uint64_t n = 0;
    int ensure(uint64_t count) {
        if (count <= n && p) return VKM_OK;
        if (p) cudaFree(p);
        p = nullptr;
        n = 0;
        cudaError_t e = cudaMalloc(&p, std::max<uint64_t>(count, 1) * sizeof(T));
        if (e != cudaSuccess) {
            p = nullptr;
            return fail(VKM_ERR_OOM, "cudaMalloc(%llu bytes) failed: %s", (unsigned long long)(count * sizeof(T)),
                        cudaGetErrorString(e));
        }
        n = count;
        return VKM_OK;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        n = 0;
    }
};

enum Phase { PH_CLEAR = 0, PH_INSERT, PH_VALUES, PH_SCAN, PH_INDICES, PH_VERTICES, PH_TANGENTS, PH_COUNT };
const char* kPhaseNames[PH_COUNT] = {"clear", "insert_triplets", "resolve_values", "mark_scan", "indices",
                                     "vertices", "tangents"};

}  // namespace

struct vkm_plan {
    vkm_plan_desc desc{};
    cudaStream_t own_stream = nullptr;
    cudaStream_t last_stream = nullptr;
    uint32_t m1 = 0, m2 = 0;       // allocated slots
    uint32_t t1_mode = 0;          // sticky: 0 = locality-preserving triplet table, 1 = scrambled (after clustering)
    bool assume_clean = true;      // speculation for the next import: no value duplicates / NaN-Inf keys (true for nearly
                                   // every mesh; after a finished import: what that import turned out to be)
    bool force_full = false;       // re-run in progress: launch the complete schedule
    bool ran_full = false;         // what the import in flight was launched with
    bool ran_records = false;      // the import in flight used the record schedule (vkm_records.cuh)
    bool records_off = false;      // sticky: the side table would outgrow the slot arrays -> general schedule only
    bool records_off_env = false;  // VKM_RECORDS=0
    uint32_t msec = 0;             // side-table slots of the record schedule
    uint32_t m2_pct_env = 200;     // the configured value (VKM_M2_PCT)
    uint32_t m2_pct = 200;         // record schedule: value-table slots per 100 key records
    DevBuf<uint32_t> sec_first, hbox;
    bool cell_on = true;           // record schedule: position-ordered bucket choice in the fingerprint table (VKM_CELL=0: off)
    bool cell_off = false;         // sticky: a mesh piled its vertices up in few cells and overflowed the probe limit
    uint32_t seed = 0;             // hash generation (bumped when a table overflows: see hash_triplet)
    vkm_plan_stats stats{};        // of the import in flight / last finished
    DevBuf<uint32_t> table1, table2;
    DevBuf<uint32_t> slot1, aux1, fbits, fprefix, block_sums, ext_enc, vbase, lastface, out_idx, shape_off;
    DevBuf<float4> out_vtx;
    DevBuf<uint4> trip;
    DevBuf<Status> status;
    // pinned host mirrors
    Status* h_status = nullptr;
    uint32_t* h_vbase = nullptr;
    uint32_t* h_ext = nullptr;
    uint32_t* h_shape_off = nullptr;  // pinned, kRing x (max_shapes + 1): one staging row per in-flight import
    // ring of per-import phase events, so that imports can be queued back to back and still be timed
    static constexpr int kRing = 32;
    cudaEvent_t ev[kRing][PH_COUNT + 1]{};
    bool ev_used[kRing]{};     // phase events of the slot hold an unharvested measurement
    int ev_phase[kRing][PH_COUNT]{};  // phase that the interval between event i and i + 1 of the slot belongs to
    bool slot_busy[kRing]{};   // the slot's import may still be running (its end event was recorded)
    uint64_t seq = 0;
    double phase_sum[8]{};
    uint32_t phase_n = 0;
    cudaEvent_t done = nullptr;
    bool timing = false;
    bool pending = false;
    bool grow_vertices = false;  // on "vertex capacity exceeded": reallocate to the reported count and re-run
    // last import (kept for the re-run after a table overflow)
    ImportArgs args{};
    std::vector<uint32_t> vertex_offsets;
    std::vector<vkm_extent> extents;
    float phase_ms[8]{};
};

namespace {

int plan_size_tables(vkm_plan* p, uint64_t slots) {
    slots = std::min<uint64_t>(std::max<uint64_t>(slots, 1024), 0xFFFFFF00ull);
    p->m1 = p->m2 = (uint32_t)slots;
    int rc;
    if ((rc = p->table1.ensure(slots))) return rc;
    if ((rc = p->table2.ensure(slots))) return rc;
    if ((rc = p->aux1.ensure(slots))) return rc;
    if ((rc = p->trip.ensure(slots / 2 + 1))) return rc;
    return VKM_OK;
}

uint64_t initial_slots(const vkm_plan_desc& d) {
    // Distinct (v,n,t) triplets T are at least the largest attribute array (if all records are referenced)
    // and at most the corner count.  Start optimistic, grow on overflow.
    uint64_t est = d.max_corners;
    if (d.max_attrib) est = std::min<uint64_t>(d.max_corners, d.max_attrib + d.max_attrib / 4 + 1024);
    return 2 * est + 64;
}

// Folds the phase times of ring slot `slot` into the running sums (waits for that import to finish).
void plan_harvest(vkm_plan* p, int slot) {
    if (!p->ev_used[slot]) return;
    if (cudaEventSynchronize(p->ev[slot][PH_COUNT]) == cudaSuccess) {
        for (int i = 0; i < PH_COUNT; ++i) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, p->ev[slot][i], p->ev[slot][i + 1]) == cudaSuccess) p->phase_sum[p->ev_phase[slot][i]] += ms;
        }
        p->phase_n++;
    }
    p->ev_used[slot] = false;
}

// Kernel launch of the import schedule.  With `pdl` the launch carries the programmatic-stream-serialisation
// attribute: the grid may become resident while its predecessor in the stream drains (the predecessor's first
// statement allows it) and blocks in pdl_enter() until that predecessor has completed -- launch latency and the
// predecessor's tail are overlapped, the data dependency is unchanged.
template <typename... KArgs, typename... Args>
cudaError_t launch_k(bool pdl, void (*kernel)(KArgs...), unsigned grid, unsigned block, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    int n = 0;
    if (pdl) {
        attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[n].val.programmaticStreamSerializationAllowed = 1;
        ++n;
    }
    cfg.attrs = attr;
    cfg.numAttrs = n;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

int plan_enqueue(vkm_plan* p) {
    ImportArgs& a = p->args;
    cudaStream_t st = p->last_stream;
    const int dev = p->desc.device;
    const int slot = (int)(p->seq % vkm_plan::kRing);
    cudaEvent_t* ev = p->ev[slot];
    if (p->timing) p->ev_used[slot] = true;
    a.table1 = p->table1.p;
    a.table2 = p->table2.p;
    a.aux1 = p->aux1.p;
    a.trip = p->trip.p;
    a.m2 = p->m2;
    a.seed = p->seed;
    p->stats.runs++;
    a.cbits = 1;
    while (a.cbits < 32 && (1ull << a.cbits) <= (uint64_t)a.ncorners) a.cbits++;  // 2^cbits > ncorners
    {
        // Triplet table addressing.  Mode 0: home slot = (index into the largest attribute array) * scale, which
        // keeps the table access pattern as coherent as the mesh's own index order.  Needs scale slots per record.
        const uint32_t na = std::max(a.np, std::max(a.nn, a.nt));
        const uint64_t need = 2ull * na + 1024;
        if (p->t1_mode == 0 && need <= p->m1) {
            a.t1_mode = 0;
            a.t1_sel = (a.np >= a.nn && a.np >= a.nt) ? 0u : (a.nn >= a.nt ? 1u : 2u);
            a.t1_scale = 2;
            a.m1 = (uint32_t)need;
        } else {
            a.t1_mode = 1;
            a.t1_sel = 0;
            a.t1_scale = 1;
            a.m1 = p->m1 & ~1u;  // slot-domain kernels walk pairs of slots
        }
    }
    const bool T = p->timing;
    // Dirty-only kernels (value duplicates, NaN/Inf keys) are not even launched while the plan expects a clean mesh;
    // vkm_plan_finish() checks the status block and re-runs with the full schedule if that was wrong.
    const bool full = p->force_full || !p->assume_clean;
    p->ran_full = full;
    // Record schedule (vkm_records.cuh): the clean-mesh speculation with 16-byte key records instead of the hashed table.
    const uint32_t na_key = std::max(a.np, std::max(a.nn, a.nt));
    const bool use_rec = !full && !p->records_off && p->t1_mode == 0 && a.ncorners > 0 && (uint64_t)na_key + 8 <= p->table1.n && (uint64_t)na_key / 2 + 2 <= p->trip.n &&
                         (uint64_t)na_key + 64 <= p->aux1.n && p->msec > 0;
    p->ran_records = use_rec;
    for (int i = 0; i < PH_COUNT; ++i) p->ev_phase[slot][i] = i;
    cudaError_t le = cudaSuccess;
    auto L = [&](cudaError_t e) { if (le == cudaSuccess) le = e; };
    // Programmatic dependent launch between the kernels of one import (not across the event records of timing mode, and
    // not for the clear kernel, whose predecessor in the stream is a copy): 1 M triangles 0.112 -> 0.103 ms per import,
    // 16.8 M neutral (0.822 -> 0.821).  VKM_PDL=0 switches it off.
    static const bool pdl_on = !(getenv("VKM_PDL") && atoi(getenv("VKM_PDL")) == 0);
    const bool P = !T && pdl_on;
    if (use_rec) {
        a.t1_mode = 0;
        a.t1_sel = (a.np >= a.nn && a.np >= a.nt) ? 0u : (a.nn >= a.nt ? 1u : 2u);
        a.rfirst = p->table1.p;                           // (the general schedule's buffers, idle in this one)
        a.ident = reinterpret_cast<uint2*>(p->trip.p);
        a.nkey = na_key;
        a.sec_first = p->sec_first.p;
        a.msec = (uint32_t)std::min<uint64_t>(p->msec, p->aux1.n - na_key);  // vid_of (= aux1) holds nkey + msec slots
        a.hbox = p->hbox.p;
        a.vid_of = p->aux1.p;
        a.m2 = (uint32_t)std::min<uint64_t>(p->m2, (uint64_t)na_key * p->m2_pct / 100 + 1024) & ~7u;  // whole 8-word buckets
        // position-ordered buckets pay when the table is too large to live in L2 (random probes then go to DRAM); a
        // small table is cheaper scrambled (no position box to compute)
        a.cell_vpc = (p->cell_on && !p->cell_off && a.nshapes == 1 && p->seed == 0 && (uint64_t)a.m2 * 4 >= (32ull << 20)) ? 4u : 0u;
        const uint32_t nslots = a.nkey + a.msec;
        const bool multi = a.nshapes > 1;
        const int key2 = (int)a.t1_sel * 2 + (multi ? 1 : 0);
        // in time order: clear, insert, flags, scan + ranks, indices, vertices
        const int order[PH_COUNT] = {PH_CLEAR, PH_INSERT, PH_VALUES, PH_SCAN, PH_INDICES, PH_VERTICES, PH_TANGENTS};
        for (int i = 0; i < PH_COUNT; ++i) p->ev_phase[slot][i] = order[i];
        int e = 0;
        if (T) CU(cudaEventRecord(ev[e++], st));
        L(launch_k(false, k_clear_records, grid_for(std::max<uint64_t>(a.nkey / 2, a.m2 / 4), kThreads, dev), kThreads, st, a));
        if (T) CU(cudaEventRecord(ev[e++], st));
        {
            const unsigned g1 = grid_for(((uint64_t)a.ncorners + 3) / 4, kThreads, dev, VKM_REC_MINB);  // one resident wave
#define VKM_K1(K, M) L(launch_k(P, k_insert_records<K, M>, g1, kThreads, st, a))
            switch (key2) {
                case 0: VKM_K1(0, false); break;
                case 1: VKM_K1(0, true); break;
                case 2: VKM_K1(1, false); break;
                case 3: VKM_K1(1, true); break;
                case 4: VKM_K1(2, false); break;
                default: VKM_K1(2, true); break;
            }
#undef VKM_K1
        }
        if (T) CU(cudaEventRecord(ev[e++], st));
        L(launch_k(P, k_flag_records, grid_for(nslots / 4 + 1, kThreads, dev), kThreads, st, a));
        if (T) CU(cudaEventRecord(ev[e++], st));
        const uint32_t ntiles = std::max<uint32_t>(1, (a.nwords + kScanTile - 1) / kScanTile);
        L(launch_k(P, k_scan_tiles, ntiles, kScanThreads, st, a));
        L(launch_k(P, k_scan_finish, ntiles, kScanThreads, st, a, ntiles, true));
        L(launch_k(P, k_rank_records, grid_for(nslots / 4 + 1, kThreads, dev), kThreads, st, a));
        if (T) CU(cudaEventRecord(ev[e++], st));
        {
            const unsigned g5 = grid_for(((uint64_t)a.ncorners + 3) / 4, kThreads, dev);
            if (multi) L(launch_k(P, k_write_indices_rec<true>, g5, kThreads, st, a));
            else L(launch_k(P, k_write_indices_rec<false>, g5, kThreads, st, a));
        }
        if (T) CU(cudaEventRecord(ev[e++], st));
        {
            const unsigned g4 = grid_for(nslots, kThreads, dev, VKM_VTX_MINB);
#define VKM_K4(K, M) L(launch_k(P, k_write_vertices_rec<K, M>, g4, kThreads, st, a))
            switch (key2) {
                case 0: VKM_K4(0, false); break;
                case 1: VKM_K4(0, true); break;
                case 2: VKM_K4(1, false); break;
                case 3: VKM_K4(1, true); break;
                case 4: VKM_K4(2, false); break;
                default: VKM_K4(2, true); break;
            }
#undef VKM_K4
        }
        if (T) CU(cudaEventRecord(ev[e++], st));
    } else {
    const bool Pg = P && a.ncorners < (6u << 20);  // general schedule: only for short kernels (measured in round 1, below)
    if (T) CU(cudaEventRecord(ev[PH_CLEAR], st));
    // Programmatic dependent launch between the kernels of one import (not across the event records of timing mode,
    // and not for k_clear, whose predecessor in the stream is a copy).  Only for imports whose kernels are short:
    // there the launch gaps are a tenth of the import (1 M triangles: 0.121 -> 0.109 ms).  On long kernels it LOSES
    // (16 M triangles: 1.02 -> 1.07 ms): the dependent grid's CTAs move in wherever a slot frees up first, so its
    // 8-per-SM grid-stride split is no longer even and the SMs that took the extra CTAs finish late.
    // position-ordered homes in the value table (insert_value): same conditions as the record schedule's buckets
    a.hbox = p->hbox.p;
    a.cell_vpc = (p->cell_on && !p->cell_off && a.nshapes == 1 && p->seed == 0 && (uint64_t)a.m2 * 4 >= (32ull << 20)) ? 1u : 0u;
    L(launch_k(false, k_clear, grid_for(std::max(a.m1, a.m2) / 4, kThreads, dev), kThreads, st, a));
    if (a.cell_vpc) {
        L(launch_k(Pg, k_sample_box, grid_for(a.np / 8 + 1, kThreads, dev), kThreads, st, a));
        L(launch_k(Pg, k_cell_map, 1, 32, st, a));
    }
    if (T) CU(cudaEventRecord(ev[PH_INSERT], st));
    {
        const unsigned g1 = grid_for(((uint64_t)a.ncorners + 3) / 4, kThreads, dev);
        const int addr = a.t1_mode == 0 ? (int)a.t1_sel : 3;
        const bool multi = a.nshapes > 1;
#define VKM_K1(A, M) L(launch_k(Pg, k_insert_triplets<A, M>, g1, kThreads, st, a))
        switch (addr * 2 + (multi ? 1 : 0)) {
            case 0: VKM_K1(0, false); break;
            case 1: VKM_K1(0, true); break;
            case 2: VKM_K1(1, false); break;
            case 3: VKM_K1(1, true); break;
            case 4: VKM_K1(2, false); break;
            case 5: VKM_K1(2, true); break;
            case 6: VKM_K1(3, false); break;
            default: VKM_K1(3, true); break;
        }
#undef VKM_K1
    }
    if (T) CU(cudaEventRecord(ev[PH_VALUES], st));
    // short schedule: the value-level proof rides along in k_write_vertices<true>; only the flag bits are set here
    if (full) L(launch_k(Pg, k_resolve_slots, grid_for(a.m1 / 2, kThreads, dev), kThreads, st, a, true));
    else L(launch_k(Pg, k_flag_firsts, grid_for(a.m1 / 2, kThreads, dev), kThreads, st, a));
    if (T) CU(cudaEventRecord(ev[PH_SCAN], st));
    if (full) {
        L(launch_k(Pg, k_redo_flags_clear, grid_for(a.nwords, kThreads, dev), kThreads, st, a));
        L(launch_k(Pg, k_redo_flags_mark, grid_for(a.m2, kThreads, dev), kThreads, st, a));
        L(launch_k(Pg, k_mark_unmatchable, grid_for(a.ncorners, kThreads, dev), kThreads, st, a));
    }
    const uint32_t ntiles = std::max<uint32_t>(1, (a.nwords + kScanTile - 1) / kScanTile);
    L(launch_k(Pg, k_scan_tiles, ntiles, kScanThreads, st, a));
    L(launch_k(Pg, k_scan_finish, ntiles, kScanThreads, st, a, ntiles, true));
    if (T) CU(cudaEventRecord(ev[PH_INDICES], st));
    if (full) L(launch_k(Pg, k_final_index, grid_for(a.m1, kThreads, dev), kThreads, st, a));
    L(launch_k(Pg, k_write_indices, grid_for(((uint64_t)a.ncorners + 3) / 4, kThreads, dev), kThreads, st, a));
    if (T) CU(cudaEventRecord(ev[PH_VERTICES], st));
    if (full) L(launch_k(Pg, k_write_vertices<false>, grid_for(a.m1 / 2, kThreads, dev, 6), kThreads, st, a));
    else L(launch_k(Pg, k_write_vertices<true>, grid_for(a.m1 / 2, kThreads, dev, 6), kThreads, st, a));
    if (full) L(launch_k(Pg, k_write_unmatchable_vertices, grid_for(a.ncorners, kThreads, dev), kThreads, st, a));
    if (T) CU(cudaEventRecord(ev[PH_TANGENTS], st));
    if (full)
        L(launch_k(Pg, k_per_vertex_frames<0>, grid_for(a.vertex_capacity, kThreads, dev), kThreads, st, a.out_vtx,
                   (const uint32_t*)a.out_idx, (const uint32_t*)a.lastface, (const uint32_t*)&p->status.p->total_vertices,
                   a.vertex_capacity, (const uint32_t*)a.shape_off, (const uint32_t*)a.vbase, a.nshapes, a.vertex_capacity,
                   (const Status*)p->status.p));
    }
    if (le != cudaSuccess) return fail(VKM_ERR_CUDA, "kernel launch failed: %s", cudaGetErrorString(le));
    CU(cudaEventRecord(ev[PH_COUNT], st));
    p->slot_busy[slot] = true;
    p->seq++;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(p->h_status, p->status.p, sizeof(Status), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(p->h_vbase, a.vbase, ((uint64_t)a.nshapes + 1) * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(p->h_ext, a.ext_enc, (uint64_t)a.nshapes * 6 * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaEventRecord(p->done, st));
    p->pending = true;
    return VKM_OK;
}

}  // namespace

// =================================================================================================
// C ABI
// =================================================================================================
extern "C" {

const char* vkm_last_error(void) { return g_err.c_str(); }
void vkm_internal_set_error(const char* msg) { g_err = msg ? msg : ""; }  // for the library's other translation units
const char* vkm_version(void) { return "vkmesh_b200 0.1 (sm_100a)"; }

int vkm_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

const char* vkm_plan_phase_name(int phase) { return (phase >= 0 && phase < PH_COUNT) ? kPhaseNames[phase] : nullptr; }

int vkm_plan_create(const vkm_plan_desc* desc, vkm_plan** out) {
    if (!desc || !out) return fail(VKM_ERR_INVALID_ARG, "vkm_plan_create: null argument");
    if (desc->max_corners > kMaxCorners) return fail(VKM_ERR_INVALID_ARG, "vkm_plan_create: more than %llu corners", (unsigned long long)kMaxCorners);
    if (vkm_device_count() <= desc->device) return fail(VKM_ERR_NO_DEVICE, "vkm_plan_create: CUDA device %d not present (there is no CPU path)", desc->device);
    CU(cudaSetDevice(desc->device));
    vkm_plan* p = new vkm_plan();
    p->desc = *desc;
    if (p->desc.max_shapes == 0) p->desc.max_shapes = 1;
    if (p->desc.max_vertices == 0 || p->desc.max_vertices > p->desc.max_corners) p->desc.max_vertices = p->desc.max_corners;
    const uint64_t C = std::max<uint64_t>(p->desc.max_corners, 1), V = std::max<uint64_t>(p->desc.max_vertices, 1);
    const uint64_t W = (C + 31) / 32;
    const uint32_t S = p->desc.max_shapes;
    int rc = VKM_OK;
    auto chk = [&](int r) { if (rc == VKM_OK) rc = r; };
    chk(plan_size_tables(p, initial_slots(p->desc)));
    chk(p->slot1.ensure(C + 4));
    chk(p->fbits.ensure(W + 1));
    chk(p->fprefix.ensure(W + 1));
    chk(p->block_sums.ensure((W + kScanTile - 1) / kScanTile + 1));
    chk(p->ext_enc.ensure((uint64_t)S * 6));
    chk(p->vbase.ensure((uint64_t)S + 1));
    chk(p->shape_off.ensure((uint64_t)S + 1));
    chk(p->lastface.ensure(V));
    chk(p->out_idx.ensure(C + 4));
    chk(p->out_vtx.ensure(V * 4));
    chk(p->status.ensure(1));
    // record schedule: side table for triplets that are not their key record's first variant (grown on overflow)
    p->msec = (uint32_t)std::min<uint64_t>(std::max<uint64_t>(4096, p->desc.max_attrib / 8), 1u << 30);
    if (const char* e = getenv("VKM_RECORDS")) p->records_off = p->records_off_env = atoi(e) == 0;
    if (const char* e = getenv("VKM_M2_PCT")) p->m2_pct = p->m2_pct_env = (uint32_t)std::max(100, atoi(e));
    chk(p->sec_first.ensure(p->msec));
    chk(p->hbox.ensure(16));
    if (const char* e = getenv("VKM_CELL")) p->cell_on = atoi(e) != 0;
    if (rc != VKM_OK) {
        vkm_plan_destroy(p);
        return rc;
    }
    cudaError_t e = cudaSuccess;
    auto cu = [&](cudaError_t r) { if (e == cudaSuccess) e = r; };
    cu(cudaStreamCreateWithFlags(&p->own_stream, cudaStreamNonBlocking));
    cu(cudaMallocHost(&p->h_status, sizeof(Status)));
    cu(cudaMallocHost(&p->h_vbase, ((uint64_t)S + 1) * 4));
    cu(cudaMallocHost(&p->h_ext, (uint64_t)S * 6 * 4));
    cu(cudaMallocHost(&p->h_shape_off, (uint64_t)vkm_plan::kRing * ((uint64_t)S + 1) * 4));
    for (auto& ring : p->ev) for (auto& ev : ring) cu(cudaEventCreate(&ev));
    cu(cudaEventCreateWithFlags(&p->done, cudaEventDisableTiming));
    if (e != cudaSuccess) {
        vkm_plan_destroy(p);
        return fail(VKM_ERR_CUDA, "vkm_plan_create: %s", cudaGetErrorString(e));
    }
    *out = p;
    return VKM_OK;
}

void vkm_plan_destroy(vkm_plan* p) {
    if (!p) return;
    cudaSetDevice(p->desc.device);
    if (p->pending && p->done) cudaEventSynchronize(p->done);
    p->table1.release(); p->table2.release(); p->slot1.release(); p->aux1.release(); p->fbits.release();
    p->fprefix.release(); p->block_sums.release(); p->ext_enc.release(); p->vbase.release(); p->lastface.release();
    p->out_idx.release(); p->shape_off.release(); p->out_vtx.release(); p->trip.release(); p->status.release();
    p->sec_first.release(); p->hbox.release();
    if (p->h_status) cudaFreeHost(p->h_status);
    if (p->h_vbase) cudaFreeHost(p->h_vbase);
    if (p->h_ext) cudaFreeHost(p->h_ext);
    if (p->h_shape_off) cudaFreeHost(p->h_shape_off);
    for (auto& ring : p->ev) for (auto& ev : ring) if (ev) cudaEventDestroy(ev);
    if (p->done) cudaEventDestroy(p->done);
    if (p->own_stream) cudaStreamDestroy(p->own_stream);
    delete p;
}

void vkm_plan_set_timing(vkm_plan* p, int enabled) { if (p) p->timing = enabled != 0; }

int vkm_plan_get_stats(const vkm_plan* p, vkm_plan_stats* out) {
    if (!p || !out) return fail(VKM_ERR_INVALID_ARG, "vkm_plan_get_stats: null argument");
    *out = p->stats;
    return VKM_OK;
}

int vkm_plan_import(vkm_plan* p, const float* d_pos, uint64_t np, const float* d_nrm, uint64_t nn, const float* d_uv,
                    uint64_t nt, const int32_t* d_idx, const uint64_t* shape_corner_offsets, uint32_t nshapes,
                    void* stream) {
    if (!p || !shape_corner_offsets) return fail(VKM_ERR_INVALID_ARG, "vkm_plan_import: null argument");
    if (nshapes == 0 || nshapes > p->desc.max_shapes) return fail(VKM_ERR_CAPACITY, "vkm_plan_import: %u shapes, plan holds %u", nshapes, p->desc.max_shapes);
    if (shape_corner_offsets[0] != 0) return fail(VKM_ERR_INVALID_ARG, "vkm_plan_import: shape_corner_offsets[0] must be 0");
    const uint64_t C = shape_corner_offsets[nshapes];
    if (C > p->desc.max_corners) return fail(VKM_ERR_CAPACITY, "vkm_plan_import: %llu corners, plan holds %llu", (unsigned long long)C, (unsigned long long)p->desc.max_corners);
    if (np >= 0x7FFFFFFFull || nn >= 0x7FFFFFFFull || nt >= 0x7FFFFFFFull) return fail(VKM_ERR_INVALID_ARG, "vkm_plan_import: attribute arrays beyond the int32 index range");
    if (((uintptr_t)d_idx & 15) || ((uintptr_t)d_uv & 7) || ((uintptr_t)d_pos & 3) || ((uintptr_t)d_nrm & 3))
        return fail(VKM_ERR_INVALID_ARG, "vkm_plan_import: d_idx_vnt must be 16-byte, d_uv 8-byte, d_pos/d_nrm 4-byte aligned (vector loads)");
    cudaStream_t st = stream ? (cudaStream_t)stream : p->own_stream;
    if (p->pending && st != p->last_stream) {  // same-stream imports are ordered by the stream itself
        CU(cudaEventSynchronize(p->done));
        p->pending = false;
    }
    for (uint32_t s = 0; s < nshapes; ++s) {
        const uint64_t a = shape_corner_offsets[s], b = shape_corner_offsets[s + 1];
        if (b < a || (b - a) % 3 != 0) return fail(VKM_ERR_INVALID_ARG, "vkm_plan_import: shape %u corner range is not a whole number of triangles", s);
    }
    // claim this import's ring slot: its staging row and events are free once the import that last used it is done
    const int slot = (int)(p->seq % vkm_plan::kRing);
    if (p->ev_used[slot]) plan_harvest(p, slot);
    else if (p->slot_busy[slot]) CU(cudaEventSynchronize(p->ev[slot][PH_COUNT]));
    p->slot_busy[slot] = false;
    uint32_t* stage = p->h_shape_off + (uint64_t)slot * ((uint64_t)p->desc.max_shapes + 1);
    for (uint32_t s = 0; s <= nshapes; ++s) stage[s] = (uint32_t)shape_corner_offsets[s];
    if (C > 0 && (np == 0 || nn == 0 || nt == 0 || !d_pos || !d_nrm || !d_uv || !d_idx))
        return fail(VKM_ERR_BAD_INDEX, "vkm_plan_import: corners reference an empty attribute array");
    CU(cudaSetDevice(p->desc.device));
    p->last_stream = st;
    ImportArgs& a = p->args;
    a = ImportArgs{};
    a.pos = d_pos; a.nrm = d_nrm; a.uv = d_uv; a.idx = d_idx;
    a.np = (uint32_t)np; a.nn = (uint32_t)nn; a.nt = (uint32_t)nt;
    a.ncorners = (uint32_t)C;
    a.shape_off = p->shape_off.p;
    a.nshapes = nshapes;
    a.slot1 = p->slot1.p; a.fbits = p->fbits.p; a.fprefix = p->fprefix.p; a.block_sums = p->block_sums.p;
    a.nwords = (uint32_t)((C + 31) / 32);
    a.ext_enc = p->ext_enc.p; a.vbase = p->vbase.p; a.lastface = p->lastface.p;
    a.out_idx = p->out_idx.p; a.out_vtx = p->out_vtx.p;
    a.vertex_capacity = (uint32_t)std::min<uint64_t>(p->desc.max_vertices, 0xFFFFFFFEull);
    a.status = p->status.p;
    CU(cudaMemcpyAsync(p->shape_off.p, stage, ((uint64_t)nshapes + 1) * 4, cudaMemcpyHostToDevice, st));
    p->stats = vkm_plan_stats{};
    return plan_enqueue(p);
}

int vkm_plan_finish(vkm_plan* p, vkm_plan_result* r) {
    if (!p) return fail(VKM_ERR_INVALID_ARG, "vkm_plan_finish: null plan");
    if (!p->pending) return fail(VKM_ERR_INVALID_ARG, "vkm_plan_finish: no import in flight");
    CU(cudaSetDevice(p->desc.device));
    for (int attempt = 0;; ++attempt) {
        CU(cudaEventSynchronize(p->done));
        p->pending = false;
        const Status s = *p->h_status;
        if (s.bad_index) return fail(VKM_ERR_BAD_INDEX, "import: a vertex/normal/texcoord index is negative or out of range");
        if (p->ran_records && (s.overflow1 || s.overflow2)) {
            // record schedule: the side table (seam variants) or the compact fingerprint table was too small -- or
            // the mesh's keys cluster in this hash generation: grow AND move to the next generation
            CU(cudaStreamSynchronize(p->last_stream));
            p->stats.table_overflows++;
            bool again = attempt <= 16;
            if (s.overflow1) {
                const uint64_t cap = p->aux1.n > p->args.nkey ? p->aux1.n - p->args.nkey : 0;  // vid_of holds nkey + msec
                if (p->args.msec < cap) {
                    const uint64_t want = std::min<uint64_t>(cap, 4ull * std::max(p->msec, p->args.msec));
                    int rc;
                    if ((rc = p->sec_first.ensure(want))) return rc;
                    p->msec = (uint32_t)want;
                } else {
                    // more seam variants than key records: the general schedule handles it -- with scrambled addressing
                    // straight away (so many triplets per key record are what overflows its coherent addressing too)
                    p->records_off = true;
                    p->t1_mode = 1;
                }
            }
            if (s.overflow2) {
                if (p->args.cell_vpc) p->cell_off = true;  // vertices piled up in few cells: scrambled buckets from now on
                if (p->args.m2 < (p->m2 & ~7u)) p->m2_pct = 1u << 20;  // use the whole value table from now on
                else again = false;                                    // fall through to the general growth below
            }
            if (again) {
                p->seed++;
                p->stats.reseeds++;
                const int rc = plan_enqueue(p);
                if (rc) return rc;
                continue;
            }
        }
        if (s.overflow1 || s.overflow2) {
            // optimistic table sizing failed: grow (bounded by 2 slots per corner) and run the import again.  At the
            // worst-case size a table is at most half full, so an overflow there means that this mesh's keys cluster in
            // the current hash generation: every re-run also moves to the next one.
            const uint64_t worst = 2ull * p->args.ncorners + 64;
            CU(cudaStreamSynchronize(p->last_stream));
            p->stats.table_overflows++;
            int rc = VKM_OK;
            if (s.overflow1 && p->args.t1_mode == 0 && !s.overflow2 && !p->ran_records) {
                p->t1_mode = 1;  // the coherent addressing clustered (many triplets per key record): scramble instead
            } else {
                if (attempt > 16) return fail(VKM_ERR_CAPACITY, "import: hash table overflow persists at worst-case size");
                if ((uint64_t)p->m1 < worst) rc = plan_size_tables(p, std::min<uint64_t>(worst, 2ull * p->m1));
                p->seed++;
                p->stats.reseeds++;
            }
            if (rc) return rc;
            if ((rc = plan_enqueue(p))) return rc;
            continue;
        }
        const bool dirty = s.has_value_dup || s.has_unmatchable;
        if (dirty && !p->ran_full) {  // the short schedule was launched for a mesh that needs the general path
            p->stats.mis_speculated = 1;
            p->force_full = true;
            const int rc = plan_enqueue(p);
            p->force_full = false;
            if (rc) return rc;
            continue;
        }
        p->assume_clean = !dirty;
        if (s.capacity_exceeded && p->grow_vertices && attempt <= 12) {
            // the host entry points size the vertex buffer from the attribute arrays (an estimate): the status block
            // carries the true count, so allocate that and run the import once more
            CU(cudaStreamSynchronize(p->last_stream));
            const uint64_t want = std::min<uint64_t>(p->args.ncorners, (uint64_t)s.total_vertices + s.total_vertices / 16 + 64);
            int rc;
            if ((rc = p->lastface.ensure(want)) || (rc = p->out_vtx.ensure(want * 4))) return rc;
            p->desc.max_vertices = want;
            p->args.lastface = p->lastface.p;
            p->args.out_vtx = p->out_vtx.p;
            p->args.vertex_capacity = (uint32_t)std::min<uint64_t>(want, 0xFFFFFFFEull);
            if ((rc = plan_enqueue(p))) return rc;
            continue;
        }
        if (s.capacity_exceeded) return fail(VKM_ERR_CAPACITY, "import: %u unique vertices exceed the plan capacity %llu", s.total_vertices, (unsigned long long)p->desc.max_vertices);
        break;
    }
    p->stats.schedule = p->ran_records ? 0u : (p->ran_full ? 2u : 1u);
    p->stats.hash_seed = p->seed;
    const uint32_t S = p->args.nshapes;
    p->vertex_offsets.assign(p->h_vbase, p->h_vbase + S + 1);
    p->extents.resize(S);
    for (uint32_t s = 0; s < S; ++s) {
        vkm_extent& e = p->extents[s];
        for (int k = 0; k < 3; ++k) {
            float mn = dec_float(p->h_ext[6 * s + k]), mx = dec_float(p->h_ext[6 * s + 3 + k]);
            if (mn == 0.0f) mn = 0.0f;  // -0 never replaces the initial +0 in the reference (strict <)
            if (mx == 0.0f) mx = 0.0f;
            e.min_extents[k] = mn;
            e.max_extents[k] = mx;
            e.center[k] = (mn + mx) / 2.0f;  // mesh_loader.cpp:290-292
        }
    }
    // average phase times over the imports queued since the previous finish
    for (int slot = 0; slot < vkm_plan::kRing; ++slot) plan_harvest(p, slot);
    for (int i = 0; i < 8; ++i) {
        p->phase_ms[i] = p->phase_n ? (float)(p->phase_sum[i] / p->phase_n) : 0.f;
        p->phase_sum[i] = 0.0;
    }
    p->phase_n = 0;
    if (r) {
        r->total_vertices = p->h_status->total_vertices;
        r->total_indices = p->args.ncorners;
        r->d_vertices = p->out_vtx.p;
        r->d_indices = p->out_idx.p;
        r->shape_vertex_offsets = p->vertex_offsets.data();
        r->shape_extents = p->extents.data();
        memcpy(r->phase_ms, p->phase_ms, sizeof p->phase_ms);
    }
    return VKM_OK;
}

// ------------------------------------------------------------------------------------------------
// Stand-alone geometry utilities
// ------------------------------------------------------------------------------------------------
static int frames_dev(int mode, uint32_t vertex_count, void* d_vertices, uint32_t index_count, const uint32_t* d_indices,
                      uint32_t* d_scratch, cudaStream_t st, uint32_t* d_bad) {
    if (vertex_count == 0 || index_count < 3) return VKM_OK;
    int dev = 0;
    CU(cudaGetDevice(&dev));
    CU(cudaMemsetAsync(d_scratch, 0, (uint64_t)vertex_count * 4, st));
    k_last_face<<<grid_for(index_count / 3, kThreads, dev), kThreads, 0, st>>>(d_indices, index_count, vertex_count,
                                                                             d_scratch, d_bad);
    if (mode == 0)
        k_face_frames<0><<<grid_for(index_count / 3, kThreads, dev), kThreads, 0, st>>>((float4*)d_vertices, d_indices,
                                                                                      index_count, d_scratch, vertex_count);
    else
        k_face_frames<1><<<grid_for(index_count / 3, kThreads, dev), kThreads, 0, st>>>((float4*)d_vertices, d_indices,
                                                                                      index_count, d_scratch, vertex_count);
    CU(cudaGetLastError());
    return VKM_OK;
}

// Device variants report out-of-range indices by skipping them (a device-side caller owns its buffers);
// the host variants below check the flag and fail.
__device__ uint32_t g_frames_bad_sink;

int vkm_generate_tangents_dev(uint32_t vc, void* d_v, uint32_t ic, const uint32_t* d_i, uint32_t* d_scratch, void* stream) {
    if (vkm_device_count() == 0) return fail(VKM_ERR_NO_DEVICE, "no CUDA device (there is no CPU path)");
    uint32_t* sink = nullptr;
    CU(cudaGetSymbolAddress((void**)&sink, g_frames_bad_sink));
    return frames_dev(0, vc, d_v, ic, d_i, d_scratch, (cudaStream_t)stream, sink);
}
int vkm_generate_normals_dev(uint32_t vc, void* d_v, uint32_t ic, const uint32_t* d_i, uint32_t* d_scratch, void* stream) {
    if (vkm_device_count() == 0) return fail(VKM_ERR_NO_DEVICE, "no CUDA device (there is no CPU path)");
    uint32_t* sink = nullptr;
    CU(cudaGetSymbolAddress((void**)&sink, g_frames_bad_sink));
    return frames_dev(1, vc, d_v, ic, d_i, d_scratch, (cudaStream_t)stream, sink);
}

// ------------------------------------------------------------------------------------------------
// Tangents without any device allocation (see the header for who needs this and why).
//
// The per-vertex "last triangle" words live in a file-scope device array instead of caller scratch.  In host
// code the name `g_noalloc_scratch` is only a handle for the symbol, not something a kernel can dereference:
// the real device address differs per GPU and has to be asked for with cudaGetSymbolAddress while that GPU is
// current.  That lookup happens once per device (first call) and the address is cached below.
// Layout: [0, VKM_NOALLOC_MAX_VERTICES) last-triangle words, then one word that collects the
// "index out of range" flag (ignored, like the sink of vkm_generate_tangents_dev).
// ------------------------------------------------------------------------------------------------
static __device__ __align__(256) uint32_t g_noalloc_scratch[VKM_NOALLOC_MAX_VERTICES + 64];

namespace {
std::mutex g_noalloc_mu;
uint32_t* g_noalloc_addr[64] = {nullptr};  // device address of g_noalloc_scratch, per CUDA ordinal
}  // namespace

int vkm_generate_tangents_dev_noalloc(uint32_t vertex_count, void* d_vertices, uint32_t index_count,
                                      const uint32_t* d_indices, void* stream) {
    if (vkm_device_count() == 0) return fail(VKM_ERR_NO_DEVICE, "no CUDA device (there is no CPU path)");
    if (vertex_count > VKM_NOALLOC_MAX_VERTICES)
        return fail(VKM_ERR_CAPACITY,
                    "vkm_generate_tangents_dev_noalloc: %u vertices exceed the built-in scratch (%u); use "
                    "vkm_generate_tangents_dev with caller scratch",
                    vertex_count, VKM_NOALLOC_MAX_VERTICES);
    if (vertex_count == 0 || index_count < 3) return VKM_OK;
    if (!d_vertices || !d_indices) return fail(VKM_ERR_INVALID_ARG, "vkm_generate_tangents_dev_noalloc: null buffer");
    int dev = 0;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(VKM_ERR_INVALID_ARG, "vkm_generate_tangents_dev_noalloc: device ordinal %d", dev);
    uint32_t* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        if (!g_noalloc_addr[dev]) {  // first call on this device: resolve the symbol for the current device
            uint32_t* p = nullptr;
            CU(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
            g_noalloc_addr[dev] = p;
        }
        scratch = g_noalloc_addr[dev];
    }
    cudaStream_t st = (cudaStream_t)stream;
    uint32_t* flag = scratch + VKM_NOALLOC_MAX_VERTICES;
    // The array still holds the previous call's words: zero exactly what this call reads.
    CU(cudaMemsetAsync(scratch, 0, (uint64_t)vertex_count * 4, st));
    CU(cudaMemsetAsync(flag, 0, 4, st));
    k_last_face<<<grid_for(index_count / 3, kThreads, dev), kThreads, 0, st>>>(d_indices, index_count, vertex_count,
                                                                             scratch, flag);
    k_per_vertex_frames<0><<<grid_for(vertex_count, kThreads, dev), kThreads, 0, st>>>(
        (float4*)d_vertices, d_indices, scratch, nullptr, vertex_count, nullptr, nullptr, 1, vertex_count, nullptr);
    CU(cudaGetLastError());
    return VKM_OK;
}

// ------------------------------------------------------------------------------------------------
// Host entry points
// ------------------------------------------------------------------------------------------------
namespace {

// Host threads for the staging copies (VKM_COPY_THREADS overrides; the helper pool holds 15 + the caller): 16 for the
// single-call context (one import at a time: 16.8 M mesh, pageable in / malloc out 49.9 -> 45.7 ms against 8), 8 per
// lane of vkm_import_batch (two lanes per GPU run at once).
inline unsigned default_copy_threads(bool lane = false) {
    if (const char* e = getenv("VKM_COPY_THREADS")) return (unsigned)std::max(1, std::min(16, atoi(e)));
    return lane ? 8 : 16;
}

struct HostCtx {
    std::mutex mu;
    std::vector<int> devices{0};
    // cached per-process staging (grown on demand) for the host entry points, device = devices[0]
    vkm_plan* plan = nullptr;
    DevBuf<float> pos, nrm, uv;
    DevBuf<int32_t> idx;
    DevBuf<float4> vtx;
    DevBuf<uint32_t> ind, scratch, flag, col;
    DevBuf<uint32_t> dd_head, dd_next, dd_m, dd_rf, dd_cnt;  // vkm_deduplicate_vertices scratch, kept between calls
    DevBuf<float4> dd_out;
    cudaStream_t st = nullptr;
    int bound_device = -1;
    // pinned bounce buffer + per-chunk events for the device-to-host leg of vkm_import_shapes
    char* h_stage = nullptr;
    uint64_t h_stage_cap = 0;
    std::vector<cudaEvent_t> chunk_ev;
    unsigned copy_threads = default_copy_threads();  // host threads that stage uploads / move finished chunks into the result buffers
    // vkm_import_batch: the two lanes of one device take turns on each PCIe direction, so that one lane's upload
    // runs against the other lane's download (full duplex) instead of both sharing one direction in lock step
    std::mutex* h2d_gate = nullptr;
    std::mutex* d2h_gate = nullptr;
    // page-locked ring for uploads from pageable caller memory (the engine's std::vector storage): kUpSlots pieces
    static constexpr uint64_t kUpPiece = 4ull << 20;
    static constexpr unsigned kUpSlots = 32;
    char* h_up = nullptr;
    std::vector<cudaEvent_t> up_ev;
    // vkm_import_batch groups: per-mesh spans / bad-index flags of a coalesced group
    DevBuf<MeshSpan> spans;
    DevBuf<uint32_t> span_bad;
    uint64_t up_issued = 0;                    // bins queued so far, over all calls (bin g uses slot g % kUpSlots)
    std::atomic<uint64_t> up_done[kUpSlots]{};  // per slot: 1 + the last piece whose DMA and event have been queued
};
constexpr uint64_t kUpPiece = HostCtx::kUpPiece;
constexpr unsigned kUpSlots = HostCtx::kUpSlots;
struct Gate {  // lock_guard over an optional mutex
    std::mutex* m;
    explicit Gate(std::mutex* mu) : m(mu) { if (m) m->lock(); }
    ~Gate() { if (m) m->unlock(); }
    Gate(const Gate&) = delete;
    Gate& operator=(const Gate&) = delete;
};
HostCtx g_ctx;                  // the lane behind the single-call host entry points
std::mutex g_lanes_mu;
std::vector<HostCtx*> g_lanes;  // vkm_import_batch: 2 per device, created on first use, released by vkm_shutdown

// Device-to-host leg of the host entry point.  The engine wants malloc()ed buffers; a direct copy into fresh
// pageable memory runs at a few GB/s (driver bounce buffers + first-touch page faults on one thread).  Instead
// the results are DMA'd in 16 MiB chunks into a cached pinned buffer at PCIe speed, and a few host threads copy
// finished chunks into the malloc()ed destinations while later chunks are still in flight.
struct Seg {
    uint64_t off;  // offset in the staging image
    char* dst;
    uint64_t bytes;
};

void copy_stage_range(const char* stage, const std::vector<Seg>& segs, uint64_t a, uint64_t b) {
    size_t lo = 0, hi = segs.size();
    while (lo < hi) {  // first segment that ends after a
        const size_t mid = (lo + hi) / 2;
        if (segs[mid].off + segs[mid].bytes <= a) lo = mid + 1; else hi = mid;
    }
    for (size_t i = lo; i < segs.size() && segs[i].off < b; ++i) {
        const uint64_t s0 = std::max(a, segs[i].off), s1 = std::min(b, segs[i].off + segs[i].bytes);
        if (s1 > s0) memcpy(segs[i].dst + (s0 - segs[i].off), stage + s0, s1 - s0);
    }
}

int download_staged(HostCtx& c, const void* d_vertices, uint64_t vbytes, const void* d_indices, uint64_t ibytes,
                    const std::vector<Seg>& segs) {
    const uint64_t total = vbytes + ibytes;
    if (total == 0) return VKM_OK;
    constexpr uint64_t kChunk = 16ull << 20;
    if (total > c.h_stage_cap) {
        if (c.h_stage) cudaFreeHost(c.h_stage);
        c.h_stage = nullptr;
        c.h_stage_cap = 0;
        const uint64_t want = total + total / 8;
        if (cudaMallocHost((void**)&c.h_stage, want) != cudaSuccess) {
            cudaGetLastError();
            return fail(VKM_ERR_OOM, "pinned staging buffer (%llu bytes) could not be allocated", (unsigned long long)want);
        }
        c.h_stage_cap = want;
    }
    const uint64_t nchunks = (total + kChunk - 1) / kChunk;
    while (c.chunk_ev.size() < nchunks) {
        cudaEvent_t e;
        CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        c.chunk_ev.push_back(e);
    }
    for (uint64_t k = 0; k < nchunks; ++k) {
        const uint64_t a = k * kChunk, b = std::min(total, a + kChunk);
        if (a < vbytes) CU(cudaMemcpyAsync(c.h_stage + a, (const char*)d_vertices + a, std::min(b, vbytes) - a, cudaMemcpyDeviceToHost, c.st));
        if (b > vbytes) {
            const uint64_t ia = std::max(a, vbytes);
            CU(cudaMemcpyAsync(c.h_stage + ia, (const char*)d_indices + (ia - vbytes), b - ia, cudaMemcpyDeviceToHost, c.st));
        }
        CU(cudaEventRecord(c.chunk_ev[k], c.st));
    }
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const unsigned nthreads = (unsigned)std::min<uint64_t>(std::min(std::max(1u, c.copy_threads), hw), nchunks);
    const int dev = c.bound_device;
    std::vector<int> errs(nthreads, 0);
    auto work = [&](unsigned t) {
        cudaSetDevice(dev);
        for (uint64_t k = t; k < nchunks; k += nthreads) {
            if (cudaEventSynchronize(c.chunk_ev[k]) != cudaSuccess) {
                errs[t] = 1;
                return;
            }
            copy_stage_range(c.h_stage, segs, k * kChunk, std::min(total, (k + 1) * kChunk));
        }
    };
    HelperPool::get().run(nthreads, work);
    for (int e : errs)
        if (e) return fail(VKM_ERR_CUDA, "device-to-host copy failed: %s", cudaGetErrorString(cudaGetLastError()));
    return VKM_OK;
}

void release_upload_ring(HostCtx& c) {
    if (c.h_up) cudaFreeHost(c.h_up);
    c.h_up = nullptr;
    for (auto e : c.up_ev) cudaEventDestroy(e);
    c.up_ev.clear();
    c.up_issued = 0;
    for (auto& d : c.up_done) d.store(0);
}

bool is_page_locked(const void* p) {
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeManaged;
}

// Host-to-device copies of a list of input arrays.  Page-locked sources go straight to the copy engine.  Pageable
// sources (what a drop-in caller has: tinyobj's vectors) would be bounced by the driver on the calling thread at a
// few GB/s; here they are cut into pieces of at most 4 MiB, the pieces are packed into 4 MiB bins, and a few host
// threads copy one bin each into a slot of a page-locked ring and queue its DMAs as soon as it is staged (so many
// small arrays cost one staging pass, not one blocking driver copy each).  A ring slot is reused once the DMAs that
// last read it have completed (per-slot event + sequence number, carried across calls).
struct UpSeg {
    char* dst;
    const char* src;
    uint64_t bytes;
};

int upload_segments(HostCtx& c, const std::vector<UpSeg>& segs) {
    std::vector<UpSeg> pieces;  // pageable, <= kUpPiece each
    uint64_t staged_bytes = 0;
    for (const UpSeg& sg : segs) {
        if (sg.bytes == 0) continue;
        // (the driver query costs microseconds: small arrays are staged whatever memory they are in)
        if (sg.bytes >= (1ull << 20) && is_page_locked(sg.src)) {
            CU(cudaMemcpyAsync(sg.dst, sg.src, sg.bytes, cudaMemcpyHostToDevice, c.st));
            continue;
        }
        for (uint64_t a = 0; a < sg.bytes; a += kUpPiece)
            pieces.push_back(UpSeg{sg.dst + a, sg.src + a, std::min(kUpPiece, sg.bytes - a)});
        staged_bytes += sg.bytes;
    }
    if (pieces.empty()) return VKM_OK;
    if (!c.h_up) {
        if (cudaMallocHost((void**)&c.h_up, kUpPiece * kUpSlots) != cudaSuccess) {
            cudaGetLastError();
            c.h_up = nullptr;  // no page-locked memory to be had: plain copies
            for (const UpSeg& pc : pieces) CU(cudaMemcpyAsync(pc.dst, pc.src, pc.bytes, cudaMemcpyHostToDevice, c.st));
            return VKM_OK;
        }
        for (unsigned i = 0; i < kUpSlots; ++i) {
            cudaEvent_t e;
            CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            c.up_ev.push_back(e);
        }
    }
    // bins: runs of consecutive pieces that fit one slot.  A piece whose destination continues the previous one's
    // (the arrays of a group's meshes are laid end to end on the device) is staged right behind it and shares its
    // DMA; any other piece starts on a 256-byte boundary of the slot.
    auto continues = [&](uint32_t i) { return i > 0 && pieces[i].dst == pieces[i - 1].dst + pieces[i - 1].bytes; };
    const uint64_t bin_cap = kUpPiece;
    std::vector<std::pair<uint32_t, uint32_t>> bins;
    {
        uint64_t fill = 0;
        uint32_t first = 0;
        for (uint32_t i = 0; i < pieces.size(); ++i) {
            uint64_t start = (i > first && continues(i)) ? fill : ((fill + 255) & ~255ull);
            if (start + pieces[i].bytes > bin_cap && i > first) {
                bins.push_back({first, i});
                first = i;
                start = 0;
            }
            fill = start + pieces[i].bytes;
        }
        bins.push_back({first, (uint32_t)pieces.size()});
    }
    const uint64_t nbins = bins.size();
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    // small uploads are not worth waking threads for
    const uint64_t by_size = std::max<uint64_t>(1, staged_bytes / (2 * kUpPiece));
    const unsigned nthreads = (unsigned)std::min<uint64_t>(std::min<uint64_t>(std::min(std::max(1u, c.copy_threads), hw), nbins), by_size);
    std::atomic<uint64_t> next{0};
    std::atomic<int> err{0};
    const uint64_t base = c.up_issued;
    const int dev = c.bound_device;
    auto work = [&](unsigned) {
        cudaSetDevice(dev);
        for (;;) {
            const uint64_t k = next.fetch_add(1);
            if (k >= nbins) return;
            const uint64_t g = base + k;
            const unsigned slot = (unsigned)(g % kUpSlots);
            if (g >= kUpSlots) {
                // the slot's previous user (bin g - kUpSlots, possibly of an earlier call) has queued its DMAs ...
                while (c.up_done[slot].load(std::memory_order_acquire) < g - kUpSlots + 1) std::this_thread::yield();
                // ... and they have finished reading the slot
                if (cudaEventSynchronize(c.up_ev[slot]) != cudaSuccess) err.store(1);
            }
            char* stage = c.h_up + (uint64_t)slot * kUpPiece;
            uint64_t off = 0;
            for (uint32_t i = bins[k].first; i < bins[k].second && !err.load();) {
                off = (off + 255) & ~255ull;
                const uint64_t run_at = off;
                char* run_dst = pieces[i].dst;
                do {  // a run of pieces with one contiguous destination: one DMA
                    memcpy(stage + off, pieces[i].src, pieces[i].bytes);
                    off += pieces[i].bytes;
                    ++i;
                } while (i < bins[k].second && continues(i));
                if (cudaMemcpyAsync(run_dst, stage + run_at, off - run_at, cudaMemcpyHostToDevice, c.st) != cudaSuccess) err.store(1);
            }
            if (cudaEventRecord(c.up_ev[slot], c.st) != cudaSuccess) err.store(1);
            c.up_done[slot].store(g + 1, std::memory_order_release);  // always, so that no later bin waits forever
        }
    };
    HelperPool::get().run(nthreads, work);
    c.up_issued = base + nbins;
    if (err.load()) return fail(VKM_ERR_CUDA, "staged host-to-device copy failed: %s", cudaGetErrorString(cudaGetLastError()));
    return VKM_OK;
}

int ctx_bind(HostCtx& c) {
    if (vkm_device_count() == 0) return fail(VKM_ERR_NO_DEVICE, "no CUDA device: vkmesh_b200 has no CPU path");
    const int dev = c.devices[0];
    CU(cudaSetDevice(dev));
    if (c.bound_device != dev) {
        if (c.plan) { vkm_plan_destroy(c.plan); c.plan = nullptr; }
        c.pos.release(); c.nrm.release(); c.uv.release(); c.idx.release(); c.vtx.release(); c.ind.release();
        c.scratch.release(); c.flag.release(); c.spans.release(); c.span_bad.release(); c.col.release();
        c.dd_head.release(); c.dd_next.release(); c.dd_m.release(); c.dd_rf.release(); c.dd_cnt.release(); c.dd_out.release();
        for (auto e : c.chunk_ev) cudaEventDestroy(e);  // events belong to the device they were created on
        c.chunk_ev.clear();
        release_upload_ring(c);
        if (c.st) cudaStreamDestroy(c.st);
        c.st = nullptr;
        CU(cudaStreamCreateWithFlags(&c.st, cudaStreamNonBlocking));
        c.bound_device = dev;
    }
    return VKM_OK;
}

int frames_host(int mode, uint32_t vc, void* vertices, uint32_t ic, const uint32_t* indices) {
    if (ic >= 3 && (!vertices || !indices)) return fail(VKM_ERR_INVALID_ARG, "null vertices/indices");
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    int rc = ctx_bind(g_ctx);
    if (rc) return rc;
    if (vc == 0 || ic < 3) return VKM_OK;
    HostCtx& c = g_ctx;
    if ((rc = c.vtx.ensure((uint64_t)vc * 4))) return rc;
    if ((rc = c.ind.ensure(ic))) return rc;
    if ((rc = c.scratch.ensure(vc))) return rc;
    if ((rc = c.flag.ensure(1))) return rc;
    // Only the column the utility changes comes back: tangent (16 of the 64 bytes) or normal (12).  It is packed on
    // the device, copied in one piece into page-locked staging together with the error flag (one synchronisation for
    // the whole call), and scattered into the caller's interleaved array by a few host threads.
    const uint32_t col_bytes = mode == 0 ? 16u : 12u, col_off = mode == 0 ? 48u : 12u;
    const uint64_t packed_bytes = (uint64_t)vc * col_bytes;
    if ((rc = c.col.ensure((packed_bytes + 3) / 4 + 4))) return rc;
    if (packed_bytes + 64 > c.h_stage_cap) {
        if (c.h_stage) cudaFreeHost(c.h_stage);
        c.h_stage = nullptr;
        c.h_stage_cap = 0;
        const uint64_t want = packed_bytes + packed_bytes / 8 + 4096;
        if (cudaMallocHost((void**)&c.h_stage, want) != cudaSuccess) {
            cudaGetLastError();
            return fail(VKM_ERR_OOM, "pinned staging buffer (%llu bytes) could not be allocated", (unsigned long long)want);
        }
        c.h_stage_cap = want;
    }
    CU(cudaMemcpyAsync(c.vtx.p, vertices, (uint64_t)vc * 64, cudaMemcpyHostToDevice, c.st));
    CU(cudaMemcpyAsync(c.ind.p, indices, (uint64_t)ic * 4, cudaMemcpyHostToDevice, c.st));
    CU(cudaMemsetAsync(c.flag.p, 0, 4, c.st));
    if ((rc = frames_dev(mode, vc, c.vtx.p, ic, c.ind.p, c.scratch.p, c.st, c.flag.p))) return rc;
    k_pack_column<<<grid_for(vc, kThreads, c.bound_device), kThreads, 0, c.st>>>((const float*)c.vtx.p, vc, col_off / 4,
                                                                                col_bytes / 4, (float*)c.col.p);
    CU(cudaGetLastError());
    uint32_t* h_flag = reinterpret_cast<uint32_t*>(c.h_stage + ((packed_bytes + 15) & ~15ull));
    CU(cudaMemcpyAsync(c.h_stage, c.col.p, packed_bytes, cudaMemcpyDeviceToHost, c.st));
    CU(cudaMemcpyAsync(h_flag, c.flag.p, 4, cudaMemcpyDeviceToHost, c.st));
    CU(cudaStreamSynchronize(c.st));
    if (*h_flag) return fail(VKM_ERR_BAD_INDEX, "an index is >= vertex_count (undefined behaviour in the reference)");
    {
        const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
        const unsigned nthreads = (unsigned)std::min<uint64_t>(std::min(8u, hw), std::max<uint64_t>(1, vc / 65536));
        char* dst = (char*)vertices;
        const char* src = c.h_stage;
        auto work = [&](unsigned t) {
            const uint64_t a = (uint64_t)vc * t / nthreads, b = (uint64_t)vc * (t + 1) / nthreads;
            for (uint64_t v = a; v < b; ++v) memcpy(dst + v * 64 + col_off, src + v * col_bytes, col_bytes);
        };
        HelperPool::get().run(nthreads, work);
    }
    return VKM_OK;
}

}  // namespace

int vkm_init(const int* device_ordinals, int count) {
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    const int n = vkm_device_count();
    if (n == 0) return fail(VKM_ERR_NO_DEVICE, "no CUDA device: vkmesh_b200 has no CPU path");
    std::vector<int> d;
    if (!device_ordinals || count <= 0) d.push_back(0);
    for (int i = 0; device_ordinals && i < count; ++i) {
        if (device_ordinals[i] < 0 || device_ordinals[i] >= n) return fail(VKM_ERR_INVALID_ARG, "device ordinal %d out of range (%d devices)", device_ordinals[i], n);
        d.push_back(device_ordinals[i]);
    }
    g_ctx.devices = d;
    return ctx_bind(g_ctx);
}

static void release_ctx(HostCtx& c);
void vkm_internal_objparse_release(void);  // vkm_objparse.cu

void vkm_shutdown(void) {
    {
        std::lock_guard<std::mutex> lk(g_lanes_mu);
        for (HostCtx* l : g_lanes) {
            std::lock_guard<std::mutex> lk2(l->mu);
            release_ctx(*l);
        }
    }
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    release_ctx(g_ctx);
    vkm_internal_objparse_release();
    g_results.trim();  // idle blocks only; blocks the caller still holds stay valid until vkm_free_shape()
}

static void release_ctx(HostCtx& c) {
    if (c.bound_device < 0) return;
    cudaSetDevice(c.bound_device);
    if (c.plan) { vkm_plan_destroy(c.plan); c.plan = nullptr; }
    c.pos.release(); c.nrm.release(); c.uv.release(); c.idx.release(); c.vtx.release(); c.ind.release();
    c.scratch.release(); c.flag.release(); c.spans.release(); c.span_bad.release(); c.col.release();
    c.dd_head.release(); c.dd_next.release(); c.dd_m.release(); c.dd_rf.release(); c.dd_cnt.release(); c.dd_out.release();
    if (c.h_stage) cudaFreeHost(c.h_stage);
    c.h_stage = nullptr;
    c.h_stage_cap = 0;
    for (auto e : c.chunk_ev) cudaEventDestroy(e);
    c.chunk_ev.clear();
    release_upload_ring(c);
    if (c.st) cudaStreamDestroy(c.st);
    c.st = nullptr;
    c.bound_device = -1;
}

int vkm_generate_tangents(uint32_t vc, void* v, uint32_t ic, const uint32_t* idx) { return frames_host(0, vc, v, ic, idx); }
int vkm_generate_normals(uint32_t vc, void* v, uint32_t ic, const uint32_t* idx) { return frames_host(1, vc, v, ic, idx); }

static int dedup_host(uint32_t vc, const void* vertices, uint32_t ic, uint32_t* indices, void** out_vertices,
                      uint32_t* out_count) {
    if (!out_vertices || (vc && !vertices) || (ic && !indices)) return fail(VKM_ERR_INVALID_ARG, "vkm_deduplicate_vertices: null argument");
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    HostCtx& c = g_ctx;
    int rc = ctx_bind(c);
    if (rc) return rc;
    *out_count = 0;
    if (vc == 0) {
        *out_vertices = malloc(1);
        return *out_vertices ? VKM_OK : fail(VKM_ERR_OOM, "host allocation failed");
    }
    uint32_t buckets = 1024;
    while (buckets < 2ull * vc && buckets < (1u << 30)) buckets <<= 1;
    // scratch lives in the host context and is reused (six cudaMalloc / cudaFree per call cost more than the kernels)
    DevBuf<float4>& out = c.dd_out;
    DevBuf<uint32_t>&head = c.dd_head, &next = c.dd_next, &m = c.dd_m, &rf = c.dd_rf, &cnt = c.dd_cnt;
    auto cleanup = [&]() {};
    if ((rc = c.vtx.ensure((uint64_t)vc * 4)) || (rc = c.ind.ensure(std::max<uint32_t>(ic, 1))) ||
        (rc = out.ensure((uint64_t)vc * 4)) || (rc = head.ensure(buckets)) || (rc = next.ensure(vc)) ||
        (rc = m.ensure(vc)) || (rc = rf.ensure(vc)) || (rc = cnt.ensure(1))) {
        cleanup();
        return rc;
    }
    const int dev = c.devices[0];
    cudaError_t e = cudaSuccess;
    auto cu = [&](cudaError_t r) { if (e == cudaSuccess) e = r; };
    cu(cudaMemcpyAsync(c.vtx.p, vertices, (uint64_t)vc * 64, cudaMemcpyHostToDevice, c.st));
    if (ic) cu(cudaMemcpyAsync(c.ind.p, indices, (uint64_t)ic * 4, cudaMemcpyHostToDevice, c.st));
    cu(cudaMemsetAsync(head.p, 0xFF, (uint64_t)buckets * 4, c.st));
    k_dd_build<<<grid_for(vc, kThreads, dev), kThreads, 0, c.st>>>(c.vtx.p, vc, head.p, next.p, buckets - 1);
    k_dd_match<<<grid_for(vc, kThreads, dev), kThreads, 0, c.st>>>(c.vtx.p, vc, head.p, next.p, buckets - 1, m.p);
    k_dd_sequence<<<1, 32, 0, c.st>>>(m.p, vc, rf.p, cnt.p);
    k_dd_emit<<<grid_for(vc, kThreads, dev), kThreads, 0, c.st>>>(c.vtx.p, vc, rf.p, out.p);
    if (ic) k_dd_remap<<<grid_for(ic, kThreads, dev), kThreads, 0, c.st>>>(c.ind.p, ic, vc, rf.p, m.p, cnt.p);
    cu(cudaGetLastError());
    uint32_t n = 0;
    cu(cudaMemcpyAsync(&n, cnt.p, 4, cudaMemcpyDeviceToHost, c.st));
    cu(cudaStreamSynchronize(c.st));
    if (e == cudaSuccess) {
        // Platform::AllocMemory == malloc (platform_Win32.cpp:367-369): the caller frees with free()
        *out_vertices = malloc(std::max<uint64_t>((uint64_t)n * 64, 1));
        if (!*out_vertices) {
            cleanup();
            return fail(VKM_ERR_OOM, "host allocation failed");
        }
        if (n) cu(cudaMemcpyAsync(*out_vertices, out.p, (uint64_t)n * 64, cudaMemcpyDeviceToHost, c.st));
        if (ic) cu(cudaMemcpyAsync(indices, c.ind.p, (uint64_t)ic * 4, cudaMemcpyDeviceToHost, c.st));
        cu(cudaStreamSynchronize(c.st));
    }
    cleanup();
    if (e != cudaSuccess) return fail(VKM_ERR_CUDA, "vkm_deduplicate_vertices: %s", cudaGetErrorString(e));
    *out_count = n;
    return VKM_OK;
}

uint32_t vkm_deduplicate_vertices(uint32_t vc, const void* vertices, uint32_t ic, uint32_t* indices, void** out_vertices) {
    uint32_t n = 0;
    if (dedup_host(vc, vertices, ic, indices, out_vertices, &n) != VKM_OK) return 0xFFFFFFFFu;
    return n;
}

void vkm_free_shape(vkm_shape_out* s) {
    if (!s) return;
    // pooled page-locked results go back to the pool, everything else came from malloc()
    if (!g_results.release(s->vertices)) free(s->vertices);
    if (!g_results.release(s->indices)) free(s->indices);
    s->vertices = nullptr;
    s->indices = nullptr;
}

// The part of the host entry point that runs once the streams are in device memory: plan, kernels, results into
// malloc()ed buffers.  Caller holds g_ctx.mu and has bound the context.
static int import_resident(HostCtx& c, const float* d_pos, uint64_t np, const float* d_nrm, uint64_t nn, const float* d_uv,
                           uint64_t nt, const int32_t* d_idx, const uint64_t* shape_off, uint32_t nshapes,
                           vkm_shape_out* out, Trace& tr) {
    int rc;
    const uint64_t C = shape_off[nshapes];
    const uint64_t max_attrib = std::max(np, std::max(nn, nt));
    if (!c.plan || c.plan->desc.max_corners < C || c.plan->desc.max_shapes < nshapes || c.plan->desc.max_attrib < max_attrib) {
        vkm_plan_desc d{};
        d.device = c.devices[0];
        d.max_corners = C;
        d.max_attrib = max_attrib;
        d.max_shapes = nshapes;
        if (c.plan) {
            // a caller with inputs of varying size (a batch): never shrink, and leave headroom, so that the plan is
            // rebuilt a few times at most instead of once per call
            const vkm_plan_desc& o = c.plan->desc;
            d.max_corners = std::min<uint64_t>(kMaxCorners, std::max<uint64_t>(C + C / 4, o.max_corners));
            d.max_attrib = std::max<uint64_t>(max_attrib + max_attrib / 4, o.max_attrib);
            d.max_shapes = std::max<uint32_t>(nshapes < 0x7FFFFFFFu ? 2 * nshapes : nshapes, o.max_shapes);
            vkm_plan_destroy(c.plan);
        }
        c.plan = nullptr;
        // unique vertices are usually close to the largest attribute array; the plan grows when they are not
        d.max_vertices = std::min<uint64_t>(d.max_corners, d.max_attrib + d.max_attrib / 8 + 1024);
        if ((rc = vkm_plan_create(&d, &c.plan))) return rc;
        c.plan->grow_vertices = true;
    }
    // every host call is a new mesh: whatever an earlier mesh forced on the plan (scrambled addressing, a later hash
    // generation, no position-ordered buckets, no record schedule) is not held against this one
    c.plan->t1_mode = 0;
    c.plan->seed = 0;
    c.plan->cell_off = false;
    c.plan->records_off = c.plan->records_off_env;
    c.plan->m2_pct = c.plan->m2_pct_env;
    if ((rc = vkm_plan_import(c.plan, d_pos, np, d_nrm, nn, d_uv, nt, d_idx, shape_off, nshapes, c.st))) return rc;
    vkm_plan_result r{};
    rc = vkm_plan_finish(c.plan, &r);
    g_last_stats = c.plan->stats;
    if (rc) return rc;
    tr.mark("H2D + kernels (finish)");
    const uint64_t vbytes = r.total_vertices * 64, ibytes = C * 4;
    if (g_result_mode.load() == 1) {
        // one page-locked block: [vertices of every shape][indices of every shape], DMA'd in place
        char* base = g_results.acquire(vbytes + ibytes + 64, 2 * nshapes);
        if (!base) return fail(VKM_ERR_OOM, "vkm_import_shapes: page-locked result block (%llu bytes) could not be allocated",
                               (unsigned long long)(vbytes + ibytes + 64));
        for (uint32_t s = 0; s < nshapes; ++s) {
            vkm_shape_out& o = out[s];
            const uint32_t v0 = r.shape_vertex_offsets[s], v1 = r.shape_vertex_offsets[s + 1];
            o.vertex_count = v1 - v0;
            o.index_count = (uint32_t)(shape_off[s + 1] - shape_off[s]);
            o.extent = r.shape_extents[s];
            o.vertices = base + (uint64_t)v0 * 64;
            o.indices = base + vbytes + shape_off[s] * 4;
        }
        tr.mark("result block");
        cudaError_t e = cudaSuccess;
        {
            Gate g(c.d2h_gate);
            if (vbytes) e = cudaMemcpyAsync(base, r.d_vertices, vbytes, cudaMemcpyDeviceToHost, c.st);
            if (e == cudaSuccess && ibytes) e = cudaMemcpyAsync(base + vbytes, r.d_indices, ibytes, cudaMemcpyDeviceToHost, c.st);
            if (e == cudaSuccess) e = cudaStreamSynchronize(c.st);
        }
        if (e != cudaSuccess) {
            g_results.release(base, 2 * nshapes);
            for (uint32_t k = 0; k < nshapes; ++k) out[k].vertices = out[k].indices = nullptr;
            return fail(VKM_ERR_CUDA, "device-to-host copy failed: %s", cudaGetErrorString(e));
        }
        tr.mark("D2H (in place)");
        return VKM_OK;
    }
    std::vector<Seg> segs;
    segs.reserve(2ull * nshapes);
    for (uint32_t s = 0; s < nshapes; ++s) {
        vkm_shape_out& o = out[s];
        const uint32_t v0 = r.shape_vertex_offsets[s], v1 = r.shape_vertex_offsets[s + 1];
        o.vertex_count = v1 - v0;
        o.index_count = (uint32_t)(shape_off[s + 1] - shape_off[s]);
        o.extent = r.shape_extents[s];
        // malloc(): the engine frees these with free() (Platform::FrMemory, mesh_resource.cpp:13-18)
        o.vertices = malloc_result((uint64_t)o.vertex_count * 64);
        o.indices = malloc_result((uint64_t)o.index_count * 4);
        if (!o.vertices || !o.indices) {
            for (uint32_t k = 0; k <= s; ++k) vkm_free_shape(&out[k]);
            return fail(VKM_ERR_OOM, "vkm_import_shapes: host allocation failed");
        }
        if (o.vertex_count) segs.push_back(Seg{(uint64_t)v0 * 64, (char*)o.vertices, (uint64_t)o.vertex_count * 64});
    }
    for (uint32_t s = 0; s < nshapes; ++s)
        if (out[s].index_count) segs.push_back(Seg{vbytes + shape_off[s] * 4, (char*)out[s].indices, (uint64_t)out[s].index_count * 4});
    tr.mark("malloc results");
    {
        Gate g(c.d2h_gate);
        rc = download_staged(c, r.d_vertices, vbytes, r.d_indices, ibytes, segs);
    }
    if (rc) {
        for (uint32_t k = 0; k < nshapes; ++k) vkm_free_shape(&out[k]);
        return rc;
    }
    tr.mark("D2H + copy-out");
    return VKM_OK;
}

static void init_shape_outs(vkm_shape_out* out, uint32_t nshapes) {
    for (uint32_t s = 0; s < nshapes; ++s) {
        memset(&out[s], 0, sizeof(vkm_shape_out));
        out[s].vertex_size = 64;
        out[s].index_size = 4;
    }
}

// vkm_import_shapes on one lane (a HostCtx: device, stream, plan, staging).  Takes the lane's lock.
static int lane_import(HostCtx& c, const float* pos, uint64_t np, const float* nrm, uint64_t nn, const float* uv,
                       uint64_t nt, const int32_t* idx, const uint64_t* shape_off, uint32_t nshapes, vkm_shape_out* out) {
    if (nshapes == 0) return VKM_OK;
    if (!shape_off || !out) return fail(VKM_ERR_INVALID_ARG, "vkm_import_shapes: null argument");
    init_shape_outs(out, nshapes);
    const uint64_t C = shape_off[nshapes];
    if (C > kMaxCorners) return fail(VKM_ERR_INVALID_ARG, "vkm_import_shapes: too many corners");
    if (C > 0 && (!pos || !nrm || !uv || !idx)) return fail(VKM_ERR_INVALID_ARG, "vkm_import_shapes: null input stream");
    std::lock_guard<std::mutex> lk(c.mu);
    Trace tr;
    int rc = ctx_bind(c);
    if (rc) return rc;
    if ((rc = c.pos.ensure(3 * np + 4))) return rc;
    if ((rc = c.nrm.ensure(3 * nn + 4))) return rc;
    if ((rc = c.uv.ensure(2 * nt + 4))) return rc;
    if ((rc = c.idx.ensure(3 * C + 16))) return rc;
    {
        Gate g(c.h2d_gate);
        const std::vector<UpSeg> segs{{(char*)c.pos.p, (const char*)pos, np * 12}, {(char*)c.nrm.p, (const char*)nrm, nn * 12},
                                      {(char*)c.uv.p, (const char*)uv, nt * 8}, {(char*)c.idx.p, (const char*)idx, C * 12}};
        if ((rc = upload_segments(c, segs))) return rc;
        if (c.h2d_gate) CU(cudaStreamSynchronize(c.st));  // hold the direction until the upload has landed
    }
    tr.mark("enqueue H2D");
    return import_resident(c, c.pos.p, np, c.nrm.p, nn, c.uv.p, nt, c.idx.p, shape_off, nshapes, out, tr);
}

// Several small meshes as ONE import (vkm_import_batch): their arrays are laid end to end in the lane's device
// buffers, k_rebase_indices makes the indices global (and checks them per mesh), and the import runs once over
// all their shapes -- per-mesh launch, synchronisation and allocation costs are paid once per group.  Returns
// VKM_OK with every member's shapes filled, or a non-zero code WITHOUT touching any output: the caller then imports
// the members one by one, which also yields the per-mesh statuses and messages.
static int lane_import_group(HostCtx& c, const vkm_mesh_in* meshes, const uint32_t* ids, uint32_t n,
                             vkm_shape_out* const* outs) {
    uint64_t NP = 0, NN = 0, NT = 0, C = 0, S = 0;
    std::vector<MeshSpan> spans(n);
    for (uint32_t k = 0; k < n; ++k) {
        const vkm_mesh_in& m = meshes[ids[k]];
        if (m.nshapes == 0 || !m.shape_corner_offsets || !outs[ids[k]] || m.shape_corner_offsets[0] != 0)
            return VKM_ERR_INVALID_ARG;
        const uint64_t mc = m.shape_corner_offsets[m.nshapes];
        if (mc > 0 && (!m.pos || !m.nrm || !m.uv || !m.idx_vnt)) return VKM_ERR_INVALID_ARG;
        for (uint32_t s = 0; s < m.nshapes; ++s) {
            const uint64_t a = m.shape_corner_offsets[s], b = m.shape_corner_offsets[s + 1];
            if (b < a || (b - a) % 3 != 0) return VKM_ERR_INVALID_ARG;
        }
        spans[k] = MeshSpan{(uint32_t)C, (uint32_t)(C + mc), (uint32_t)NP, (uint32_t)NN, (uint32_t)NT,
                            (uint32_t)m.np, (uint32_t)m.nn, (uint32_t)m.nt};
        NP += m.np; NN += m.nn; NT += m.nt; C += mc; S += m.nshapes;
    }
    if (C == 0 || NP == 0 || NN == 0 || NT == 0 || C > 0x7FFFFFF0ull || NP >= 0x7FFFFFFFull || NN >= 0x7FFFFFFFull ||
        NT >= 0x7FFFFFFFull)
        return VKM_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(c.mu);
    Trace tr;
    int rc = ctx_bind(c);
    if (rc) return rc;
    if ((rc = c.pos.ensure(3 * NP + 4)) || (rc = c.nrm.ensure(3 * NN + 4)) || (rc = c.uv.ensure(2 * NT + 4)) ||
        (rc = c.idx.ensure(3 * C + 16)) || (rc = c.spans.ensure(n)) || (rc = c.span_bad.ensure(n)))
        return rc;
    std::vector<uint64_t> shape_off;
    shape_off.reserve(S + 1);
    shape_off.push_back(0);
    std::vector<uint32_t> bad(n, 0);
    {
        Gate g(c.h2d_gate);
        std::vector<UpSeg> segs;
        segs.reserve(4ull * n);
        // array kind by array kind: consecutive meshes then have consecutive destinations and share DMAs
        for (int kind = 0; kind < 4; ++kind)
            for (uint32_t k = 0; k < n; ++k) {
                const vkm_mesh_in& m = meshes[ids[k]];
                const MeshSpan& sp = spans[k];
                if (kind == 0) segs.push_back(UpSeg{(char*)(c.pos.p + 3ull * sp.pos_base), (const char*)m.pos, m.np * 12});
                if (kind == 1) segs.push_back(UpSeg{(char*)(c.nrm.p + 3ull * sp.nrm_base), (const char*)m.nrm, m.nn * 12});
                if (kind == 2) segs.push_back(UpSeg{(char*)(c.uv.p + 2ull * sp.uv_base), (const char*)m.uv, m.nt * 8});
                if (kind == 3) segs.push_back(UpSeg{(char*)(c.idx.p + 3ull * sp.c_begin), (const char*)m.idx_vnt, (uint64_t)(sp.c_end - sp.c_begin) * 12});
            }
        for (uint32_t k = 0; k < n; ++k) {
            const vkm_mesh_in& m = meshes[ids[k]];
            for (uint32_t s = 1; s <= m.nshapes; ++s) shape_off.push_back(spans[k].c_begin + m.shape_corner_offsets[s]);
        }
        if ((rc = upload_segments(c, segs))) return rc;
        CU(cudaMemcpyAsync(c.spans.p, spans.data(), n * sizeof(MeshSpan), cudaMemcpyHostToDevice, c.st));
        CU(cudaMemsetAsync(c.span_bad.p, 0, (uint64_t)n * 4, c.st));
        k_rebase_indices<<<grid_for(C, kThreads, c.bound_device), kThreads, 0, c.st>>>(c.idx.p, c.spans.p, n, (uint32_t)C,
                                                                                        c.span_bad.p);
        CU(cudaGetLastError());
        CU(cudaMemcpyAsync(bad.data(), c.span_bad.p, (uint64_t)n * 4, cudaMemcpyDeviceToHost, c.st));
        CU(cudaStreamSynchronize(c.st));  // spans / bad live on this stack frame; the uploads have landed
    }
    tr.mark("group upload + rebase");
    for (uint32_t k = 0; k < n; ++k)
        if (bad[k]) return VKM_ERR_BAD_INDEX;  // the one-by-one path names the mesh
    std::vector<vkm_shape_out> flat(S);
    init_shape_outs(flat.data(), (uint32_t)S);
    if ((rc = import_resident(c, c.pos.p, NP, c.nrm.p, NN, c.uv.p, NT, c.idx.p, shape_off.data(), (uint32_t)S, flat.data(), tr)))
        return rc;
    uint64_t at = 0;
    for (uint32_t k = 0; k < n; ++k) {
        const vkm_mesh_in& m = meshes[ids[k]];
        for (uint32_t s = 0; s < m.nshapes; ++s) outs[ids[k]][s] = flat[at++];
    }
    return VKM_OK;
}

int vkm_import_shapes(const float* pos, uint64_t np, const float* nrm, uint64_t nn, const float* uv, uint64_t nt,
                      const int32_t* idx, const uint64_t* shape_off, uint32_t nshapes, vkm_shape_out* out) {
    return lane_import(g_ctx, pos, np, nrm, nn, uv, nt, idx, shape_off, nshapes, out);
}

// ------------------------------------------------------------------------------------------------
// Batches of independent meshes over every selected GPU: the natural shard.
// Meshes are partitioned over the devices of vkm_init() by longest-processing-time (corner count); every device
// runs TWO lanes (own stream, plan, device inputs, pinned staging, host thread), so the host->device copy of one
// mesh overlaps the kernels and the device->host copy of another (the copy engines are full duplex).  Small
// meshes are coalesced into groups that run as one segmented (multi-shape) import each, see lane_import_group.  No
// inter-GPU traffic: results land in the caller's per-mesh output arrays.
// ------------------------------------------------------------------------------------------------
int vkm_import_batch(const vkm_mesh_in* meshes, uint32_t count, vkm_shape_out* const* outs, int* statuses) {
    if (count == 0) return VKM_OK;
    if (!meshes || !outs) return fail(VKM_ERR_INVALID_ARG, "vkm_import_batch: null argument");
    if (vkm_device_count() == 0) return fail(VKM_ERR_NO_DEVICE, "no CUDA device: vkmesh_b200 has no CPU path");
    std::vector<int> devices;
    {
        std::lock_guard<std::mutex> lk(g_ctx.mu);
        devices = g_ctx.devices;
    }
    const size_t nd = devices.size();
    std::lock_guard<std::mutex> batch_lock(g_lanes_mu);  // one batch at a time
    while (g_lanes.size() < 2 * nd) g_lanes.push_back(new HostCtx());
    std::vector<std::mutex> gates(2 * nd);
    for (size_t k = 0; k < 2 * nd; ++k) {
        g_lanes[k]->devices[0] = devices[k / 2];
        g_lanes[k]->copy_threads = default_copy_threads(true);
        // (with malloc()ed results the download leg is mostly host copying, which overlaps better ungated)
        const bool gated = g_result_mode.load() == 1;
        g_lanes[k]->h2d_gate = gated ? &gates[2 * (k / 2)] : nullptr;
        g_lanes[k]->d2h_gate = gated ? &gates[2 * (k / 2) + 1] : nullptr;
    }
    // LPT partition by corner count (deterministic: ties by mesh id)
    std::vector<uint32_t> order(count);
    for (uint32_t i = 0; i < count; ++i) order[i] = i;
    auto size_of = [&](uint32_t i) { return meshes[i].nshapes ? meshes[i].shape_corner_offsets[meshes[i].nshapes] : 0; };
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return size_of(a) > size_of(b); });
    std::vector<std::vector<uint32_t>> queue(nd);
    std::vector<uint64_t> load(nd, 0);
    for (uint32_t i : order) {
        size_t best = 0;
        for (size_t d = 1; d < nd; ++d)
            if (load[d] < load[best]) best = d;
        queue[best].push_back(i);
        load[best] += size_of(i);
    }
    // Work units per device: a large mesh is a unit of its own; small meshes (below kSoloCorners) are coalesced, in
    // queue order, into groups of up to kGroupCorners corners / kGroupMeshes meshes that run as one import each.
    constexpr uint64_t kSoloCorners = 1ull << 20, kGroupCorners = 6ull << 20;
    constexpr size_t kGroupMeshes = 1024;
    std::vector<std::vector<std::vector<uint32_t>>> units(nd);
    for (size_t d = 0; d < nd; ++d) {
        std::vector<uint32_t> group;
        uint64_t group_corners = 0;
        auto flush = [&]() {
            if (!group.empty()) units[d].push_back(group);
            group.clear();
            group_corners = 0;
        };
        for (uint32_t i : queue[d]) {
            const uint64_t sz = size_of(i);
            if (sz >= kSoloCorners || sz == 0) {  // (empty meshes keep the one-by-one path's early outs)
                units[d].push_back({i});
                continue;
            }
            if (group_corners + sz > kGroupCorners || group.size() >= kGroupMeshes) flush();
            group.push_back(i);
            group_corners += sz;
        }
        flush();
    }
    std::vector<int> st(count, VKM_OK);
    std::vector<std::string> msgs(count);
    std::vector<std::atomic<uint32_t>> next(nd);
    for (auto& n : next) n.store(0);
    auto work = [&](size_t lane) {
        const size_t d = lane / 2;
        for (;;) {
            const uint32_t k = next[d].fetch_add(1);
            if (k >= units[d].size()) return;
            const std::vector<uint32_t>& unit = units[d][k];
            if (unit.size() > 1 && lane_import_group(*g_lanes[lane], meshes, unit.data(), (uint32_t)unit.size(), outs) == VKM_OK)
                continue;
            for (uint32_t i : unit) {  // a single mesh, or a group that has to be taken apart
                const vkm_mesh_in& m = meshes[i];
                st[i] = lane_import(*g_lanes[lane], m.pos, m.np, m.nrm, m.nn, m.uv, m.nt, m.idx_vnt, m.shape_corner_offsets,
                                    m.nshapes, outs[i]);
                if (st[i] != VKM_OK) msgs[i] = g_err;  // thread-local message of this worker
            }
        }
    };
    std::vector<std::thread> pool;
    for (size_t lane = 1; lane < 2 * nd; ++lane) pool.emplace_back(work, lane);
    work(0);
    for (auto& t : pool) t.join();
    for (size_t k = 0; k < 2 * nd; ++k) g_lanes[k]->h2d_gate = g_lanes[k]->d2h_gate = nullptr;
    int rc = VKM_OK;
    for (uint32_t i = 0; i < count; ++i) {
        if (statuses) statuses[i] = st[i];
        if (st[i] != VKM_OK && rc == VKM_OK) {
            rc = st[i];
            g_err = "mesh " + std::to_string(i) + ": " + msgs[i];
        }
    }
    return rc;
}

// For the library's OBJ-text entry point: the same import, with streams that already live on the device
// (device pointers on g_ctx's device; the default stream must have finished writing them).
int vkm_internal_import_device(int device, const float* d_pos, uint64_t np, const float* d_nrm, uint64_t nn,
                               const float* d_uv, uint64_t nt, const int32_t* d_idx, const uint64_t* shape_off,
                               uint32_t nshapes, vkm_shape_out* out) {
    if (nshapes == 0) return VKM_OK;
    init_shape_outs(out, nshapes);
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    HostCtx& c = g_ctx;
    if (c.devices[0] != device) c.devices[0] = device;
    Trace tr;
    int rc = ctx_bind(c);
    if (rc) return rc;
    return import_resident(c, d_pos, np, d_nrm, nn, d_uv, nt, d_idx, shape_off, nshapes, out, tr);
}

int vkm_last_import_stats(vkm_plan_stats* out) {
    if (!out) return fail(VKM_ERR_INVALID_ARG, "vkm_last_import_stats: null argument");
    *out = g_last_stats;
    return VKM_OK;
}

int vkm_set_result_memory(int mode) {
    if (mode != 0 && mode != 1) return fail(VKM_ERR_INVALID_ARG, "vkm_set_result_memory: mode must be 0 (malloc) or 1 (pooled page-locked)");
    g_result_mode.store(mode);
    if (mode == 0) g_results.trim();
    return VKM_OK;
}

// for the library's other translation units: the device chosen with vkm_init (first ordinal) and its SM count
int vkm_internal_device(int* sm) {
    int dev;
    {
        std::lock_guard<std::mutex> lk(g_ctx.mu);
        dev = g_ctx.devices[0];
    }
    if (sm) *sm = sm_count(dev);
    return dev;
}

// for the library's other translation units: fn(ctx, 0) .. fn(ctx, n-1) on the helper pool
void vkm_internal_parallel_run(unsigned n, void (*fn)(void*, unsigned), void* ctx) {
    HelperPool::get().run(n, [&](unsigned i) { fn(ctx, i); });
}

void* vkm_host_alloc(uint64_t bytes) {
    void* p = nullptr;
    if (cudaMallocHost(&p, std::max<uint64_t>(bytes, 1)) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return p;
}
void vkm_host_free(void* p) { if (p) cudaFreeHost(p); }

int vkm_dev_alloc(int device, uint64_t bytes, void** d_ptr) {
    if (!d_ptr) return fail(VKM_ERR_INVALID_ARG, "null");
    if (vkm_device_count() <= device) return fail(VKM_ERR_NO_DEVICE, "CUDA device %d not present", device);
    CU(cudaSetDevice(device));
    CU(cudaMalloc(d_ptr, std::max<uint64_t>(bytes, 1)));
    return VKM_OK;
}
int vkm_dev_free(int device, void* d_ptr) {
    CU(cudaSetDevice(device));
    CU(cudaFree(d_ptr));
    return VKM_OK;
}
int vkm_dev_upload(int device, void* d_dst, const void* h_src, uint64_t bytes) {
    CU(cudaSetDevice(device));
    CU(cudaMemcpy(d_dst, h_src, bytes, cudaMemcpyHostToDevice));
    return VKM_OK;
}
int vkm_dev_download(int device, void* h_dst, const void* d_src, uint64_t bytes) {
    CU(cudaSetDevice(device));
    CU(cudaMemcpy(h_dst, d_src, bytes, cudaMemcpyDeviceToHost));
    return VKM_OK;
}
int vkm_dev_mem_info(int device, uint64_t* free_bytes, uint64_t* total_bytes) {
    CU(cudaSetDevice(device));
    size_t f = 0, t = 0;
    CU(cudaMemGetInfo(&f, &t));
    if (free_bytes) *free_bytes = f;
    if (total_bytes) *total_bytes = t;
    return VKM_OK;
}
int vkm_dev_sync(int device) {
    CU(cudaSetDevice(device));
    CU(cudaDeviceSynchronize());
    return VKM_OK;
}

}  // extern "C"
