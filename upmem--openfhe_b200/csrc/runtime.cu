// runtime.cu — the C ABI (include/fhe_b200.h): device context, stream pool, arena allocator,
// ring/table management and the host-operand pipelines.  Replaces the reference's UPMEM host
// runtime (src/core/include/pim/PimManager.h, src/core/pim/host/PimManager.cpp, pim/DpuMemory.h).
// No CPU fallback lives here: every compute entry point launches a CUDA kernel or fails.
#include "fhe_b200.h"
#include "pim/common.h"

#include "hostmath.h"
#include "kernels.h"
#include "modarith.cuh"
#include "serial.h"

#include <algorithm>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

using namespace fhe;

// ---------------------------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

static int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
#define CUDA_TRY(expr)                                                                        \
    do {                                                                                      \
        cudaError_t e__ = (expr);                                                             \
        if (e__ != cudaSuccess)                                                               \
            return fail(FHE_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__),   \
                        __FILE__, __LINE__);                                                  \
    } while (0)

extern "C" const char* fhe_last_error(void) { return g_err; }
extern "C" const char* fhe_version(void) { return "fhe_b200 0.1 (sm_100a)"; }
extern "C" long fhe_launch_count(void) { return kernel_launch_count(); }

// ---------------------------------------------------------------------------------------------
// capture gate.  While one thread records a CUDA graph on its stream (keyswitch_graphed), nothing else in the process may
// synchronise the device or that stream, or record an event on it (the arena's markers do, on every stream that carried
// work): such calls fail and invalidate the capture.  Every API call therefore runs "inside" the gate (GateHold, also held
// by StreamRef); a capture steps out, waits until it is alone, records, and lets the others back in.  Captures are rare
// (once per plan and key), so the others spin for some tens of microseconds at most.  Re-entrant per thread.
// ---------------------------------------------------------------------------------------------
namespace {
struct CaptureGate {
    std::atomic<int> inside{0};
    std::atomic<int> capturing{0};
    std::mutex one;
    static int& depth() { static thread_local int d = 0; return d; }
    void enter() {
        if (depth()++) return;
        for (;;) {
            while (capturing.load(std::memory_order_acquire)) std::this_thread::yield();
            inside.fetch_add(1, std::memory_order_acq_rel);
            if (!capturing.load(std::memory_order_acquire)) return;
            inside.fetch_sub(1, std::memory_order_acq_rel);
        }
    }
    void leave() {
        if (--depth() == 0) inside.fetch_sub(1, std::memory_order_release);
    }
    void begin_capture() {                         // the caller is inside (it holds a StreamRef)
        inside.fetch_sub(1, std::memory_order_acq_rel);
        one.lock();
        capturing.store(1, std::memory_order_release);
        while (inside.load(std::memory_order_acquire) != 0) std::this_thread::yield();
    }
    void end_capture() {
        inside.fetch_add(1, std::memory_order_acq_rel);
        capturing.store(0, std::memory_order_release);
        one.unlock();
    }
};
CaptureGate g_gate;
struct GateHold {
    GateHold() { g_gate.enter(); }
    ~GateHold() { g_gate.leave(); }
    GateHold(const GateHold&) = delete;
    GateHold& operator=(const GateHold&) = delete;
};
}  // namespace

// ---------------------------------------------------------------------------------------------
// device arena: slabs obtained with cudaMalloc, first-fit free list with coalescing inside a slab.
// Takes the place of DpuMemory's 64 MB / 4 KB block map (pim/DpuMemory.h:11-82), sized for HBM.
// ---------------------------------------------------------------------------------------------
namespace {

constexpr size_t ARENA_ALIGN = 256;

struct Slab {
    char* base = nullptr;
    size_t size = 0;
    std::map<size_t, size_t> free_;            // offset -> length
};

struct Arena {
    std::mutex mu;
    std::vector<Slab> slabs;
    std::map<uintptr_t, std::pair<int, size_t>> live;   // ptr -> (slab, length)
    size_t slab_bytes = 256ull << 20;
    size_t grow_limit = 0;                         // bytes the arena may reserve before it prefers waiting to growing
    size_t reserved = 0, in_use = 0, stalls = 0;

    bool take_locked(size_t bytes, void** out) {
        for (size_t s = 0; s < slabs.size(); s++) {
            auto& fl = slabs[s].free_;
            for (auto it = fl.begin(); it != fl.end(); ++it) {
                if (it->second < bytes) continue;
                size_t off = it->first, len = it->second;
                fl.erase(it);
                if (len > bytes) fl[off + bytes] = len - bytes;
                *out = slabs[s].base + off;
                live[(uintptr_t)*out] = {(int)s, bytes};
                in_use += bytes;
                return true;
            }
        }
        return false;
    }

    int alloc(size_t bytes, void** out) {
        if (bytes == 0) bytes = ARENA_ALIGN;
        bytes = (bytes + ARENA_ALIGN - 1) / ARENA_ALIGN * ARENA_ALIGN;
        std::lock_guard<std::mutex> lk(mu);
        if (!parked.empty()) reap_locked();
        bool sealed = false, grown = false, waited = false;
        for (;;) {
            if (take_locked(bytes, out)) return FHE_OK;
            // No free block.  First the blocks freed since the last markers get theirs (see `pending`).  Then, below
            // grow_limit, a new slab is taken: waiting for the parked blocks instead would stall the host behind
            // everything the device has queued, on every allocation of a loop whose working set is a little larger than
            // the arena (seen as a 5x cliff between N = 2^15 and 2^16 in the scheme-level runs).  Above the limit the
            // blocks parked behind unfinished work are waited for first.
            if (!sealed && !pending.empty()) {
                sealed = true;
                if (seal) seal();
                reap_locked();
                continue;
            }
            const bool may_grow = reserved + std::max(slab_bytes, bytes) <= grow_limit;
            const bool can_wait = !waited && (!parked.empty() || !pending.empty());
            if (!grown && (may_grow || !can_wait)) {
                grown = true;
                Slab sl;
                sl.size = std::max(std::max(slab_bytes, bytes), std::min(reserved / 2, (size_t)2 << 30) / ARENA_ALIGN * ARENA_ALIGN);
                cudaError_t e = cudaMalloc((void**)&sl.base, sl.size);
                if (e == cudaSuccess) {
                    sl.free_[0] = sl.size;
                    reserved += sl.size;
                    slabs.push_back(std::move(sl));
                    continue;
                }
                cudaGetLastError();
                if (!can_wait)
                    return fail(FHE_ENOMEM, "device arena: cudaMalloc(%zu) failed: %s", sl.size, cudaGetErrorString(e));
            }
            if (can_wait) {
                waited = true;
                cudaDeviceSynchronize();
                flush_locked();
                stalls++;
                continue;
            }
            break;
        }
        return fail(FHE_ENOMEM, "device arena: no block of %zu bytes", bytes);
    }

    // give the block back to the free list NOW (the caller knows no device work can still touch it)
    int release_locked(void* p) {
        auto it = live.find((uintptr_t)p);
        if (it == live.end()) return fail(FHE_EINVAL, "fhe_free: %p is not a live arena block", p);
        int s = it->second.first;
        size_t len = it->second.second, off = (char*)p - slabs[s].base;
        live.erase(it);
        in_use -= len;
        auto& fl = slabs[s].free_;
        auto nx = fl.lower_bound(off);
        if (nx != fl.end() && off + len == nx->first) { len += nx->second; nx = fl.erase(nx); }
        if (nx != fl.begin()) {
            auto pv = std::prev(nx);
            if (pv->first + pv->second == off) { off = pv->first; len += pv->second; fl.erase(pv); }
        }
        fl[off] = len;
        return FHE_OK;
    }
    // ---- stream-ordered release ----------------------------------------------------------------
    // A released block is parked behind one CUDA event per stream of the context that has carried work since that
    // stream's last marker (markers are shared between blocks: a stream with no new work costs nothing), and goes
    // back to the free list only when all of them have completed.  So a temporary may be freed right after the
    // launches that use it were enqueued, on any stream of the pool or the legacy stream, and a block handed out by
    // alloc() is never touched by work that was enqueued before it was freed.
    struct Marker {
        cudaEvent_t ev = nullptr;
        bool done = false;
        uint64_t asked = 0;                        // the reap pass that last found it unfinished: blocks share markers
        bool finished(uint64_t pass = 0) {
            if (!done && !(pass && asked == pass)) {
                cudaError_t e = cudaEventQuery(ev);
                if (e == cudaSuccess) done = true;
                else if (e != cudaErrorNotReady) { cudaGetLastError(); done = true; }   // a dead context frees nothing later either
                else asked = pass;
            }
            return done;
        }
    };
    struct Parked { void* p; std::vector<std::shared_ptr<Marker>> waits; };
    std::deque<Parked> parked;
    // Freed blocks that have no markers yet.  Recording one event per busy stream on every free costs tens of microseconds
    // under this mutex when sixteen host threads free small blocks; a marker recorded later covers at least as much, so the
    // frees are collected and sealed together (arena_seal_locked: every 32 blocks or 256 MB, or when alloc finds nothing).
    std::vector<void*> pending;
    size_t pending_bytes = 0;
    std::function<void()> seal;                    // set by fhe_init: needs the context's streams
    std::vector<cudaEvent_t> spare_events;
    struct StreamMark { std::shared_ptr<Marker> last; uint64_t epoch = 0; };
    std::vector<StreamMark> marks;                 // one per stream slot (pool streams, then the legacy stream)

    std::shared_ptr<Marker> new_marker(cudaStream_t st) {
        cudaEvent_t ev = nullptr;
        if (!spare_events.empty()) { ev = spare_events.back(); spare_events.pop_back(); }
        else if (cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        if (cudaEventRecord(ev, st) != cudaSuccess) { cudaGetLastError(); spare_events.push_back(ev); return nullptr; }
        auto* sp = &spare_events;
        return std::shared_ptr<Marker>(new Marker{ev, false}, [sp](Marker* m) { sp->push_back(m->ev); delete m; });
    }

    // blocks whose markers have all completed return to the free list (called with mu held)
    uint64_t reap_pass = 0;
    void reap_locked() {
        const uint64_t pass = ++reap_pass;
        for (auto it = parked.begin(); it != parked.end();) {
            bool ready = true;
            for (auto& w : it->waits)
                if (!w->finished(pass)) { ready = false; break; }
            if (ready) { release_locked(it->p); it = parked.erase(it); }
            else ++it;
        }
    }
    void flush_locked() {                          // after a device-wide synchronisation: nothing is pending any more
        for (auto& pk : parked) release_locked(pk.p);
        parked.clear();
        for (void* p : pending) release_locked(p);
        pending.clear();
        pending_bytes = 0;
    }

    void destroy() {
        parked.clear();
        pending.clear();
        pending_bytes = 0;
        marks.clear();
        for (cudaEvent_t e : spare_events) cudaEventDestroy(e);
        spare_events.clear();
        for (auto& s : slabs) cudaFree(s.base);
        slabs.clear();
        live.clear();
        reserved = in_use = 0;
    }
};

struct Timer { cudaEvent_t a = nullptr, b = nullptr; };

}  // namespace

struct fhe_ctx {
    int device = 0;
    std::vector<cudaStream_t> streams;
    // epoch[slot] counts the API calls that have enqueued work on a stream (slot = pool index; the legacy stream
    // is the last slot).  It is bumped when such a call returns, so "epoch unchanged since a marker was recorded"
    // means the marker covers everything the stream carries (Arena::park).
    std::unique_ptr<std::atomic<uint64_t>[]> epoch;
    Arena arena;
    std::mutex mu;
    std::map<int, Timer> timers;
    std::vector<void*> pinned;
    size_t chunk_bytes = 32ull << 20;   // host pipelines: bytes per operand per chunk (opts: chunk_mb)
};

struct fhe_ring {
    fhe_ctx* ctx = nullptr;
    RingDev dev{};
    std::vector<uint64_t> q, psi;
    bool has_ntt = false;      // psi given and primitive for every tower
    bool ntt_identity = false; // every psi is 0 or 1: SwitchFormat leaves data untouched (transformnat-impl.h:580-582)
    void* d_mods = nullptr;
    void* d_twF = nullptr;
    void* d_twI = nullptr;
    void* d_invc = nullptr;
    std::mutex layout_mu;      // wire-format layout of one poly over this ring, built on first use (serial.h)
    std::unique_ptr<serial::Layout> layout;
};

// A stream of the context as an API call uses it: converts to cudaStream_t, and when the call returns (after its
// launches were enqueued) marks the stream as carrying new work, which is what fhe_free orders itself against.
struct StreamRef {
    cudaStream_t st = nullptr;
    fhe_ctx* ctx = nullptr;
    int slot = -1;
    GateHold gate;                                 // see CaptureGate
    StreamRef() = default;
    StreamRef(fhe_ctx* c, int pool_index) : st(c->streams[pool_index]), ctx(c), slot(pool_index) {}
    StreamRef(const StreamRef&) = delete;
    StreamRef& operator=(const StreamRef&) = delete;
    operator cudaStream_t() const { return st; }
    ~StreamRef() { if (ctx && slot >= 0) ctx->epoch[slot].fetch_add(1, std::memory_order_release); }
};

static int get_stream(fhe_ctx* ctx, int stream, StreamRef* out) {
    if (!ctx) return fail(FHE_EINVAL, "null context");
    if (stream == FHE_STREAM_LEGACY) {
        out->st = cudaStreamLegacy; out->ctx = ctx; out->slot = (int)ctx->streams.size();
        return FHE_OK;
    }
    if (stream < 0 || stream >= (int)ctx->streams.size())
        return fail(FHE_EINVAL, "stream index %d out of range [0,%zu)", stream, ctx->streams.size());
    out->st = ctx->streams[stream]; out->ctx = ctx; out->slot = stream;
    return FHE_OK;
}

// markers for everything freed since the last call (Arena::pending); A.mu is held
static void arena_seal_locked(fhe_ctx* ctx) {
    Arena& A = ctx->arena;
    if (A.pending.empty()) return;
    const size_t nslots = ctx->streams.size() + 1;
    if (A.marks.size() != nslots) A.marks.resize(nslots);
    std::vector<std::shared_ptr<Arena::Marker>> waits;
    for (size_t i = 0; i < nslots; i++) {
        uint64_t e = ctx->epoch[i].load(std::memory_order_acquire);
        if (e == 0) continue;                                   // this stream has never carried work
        Arena::StreamMark& m = A.marks[i];
        if (!m.last || m.epoch != e) {
            auto mk = A.new_marker(i + 1 == nslots ? cudaStreamLegacy : ctx->streams[i]);
            if (!mk) {                                          // cannot record: fall back to waiting for the device
                cudaDeviceSynchronize();
                A.flush_locked();
                return;
            }
            m.last = mk; m.epoch = e;
        }
        if (!m.last->finished()) waits.push_back(m.last);
    }
    for (void* p : A.pending) {
        if (waits.empty()) A.release_locked(p);
        else A.parked.push_back(Arena::Parked{p, waits});
    }
    A.pending.clear();
    A.pending_bytes = 0;
    if (A.parked.size() > 64) A.reap_locked();
}

// fhe_free and every internal release of a block that launches may still be using: see Arena::parked.
static int arena_park(fhe_ctx* ctx, void* p) {
    GateHold gate_;
    if (!p) return FHE_OK;
    Arena& A = ctx->arena;
    std::lock_guard<std::mutex> lk(A.mu);
    auto it = A.live.find((uintptr_t)p);
    if (it == A.live.end()) return fail(FHE_EINVAL, "fhe_free: %p is not a live arena block", p);
    for (auto& pk : A.parked)
        if (pk.p == p) return fail(FHE_EINVAL, "fhe_free: %p was already freed", p);
    for (void* q : A.pending)
        if (q == p) return fail(FHE_EINVAL, "fhe_free: %p was already freed", p);
    A.pending.push_back(p);
    A.pending_bytes += it->second.second;
    // how many frees share one set of markers: 32 when many streams are busy (a set costs one event per busy stream), every
    // free when at most two are (one caller thread: the block is back in the free list as soon as its stream has passed it)
    static const long fixed = [] { const char* e = getenv("FHE_B200_ARENA_SEAL"); return e ? atol(e) : 0l; }();
    size_t group = 32;
    if (fixed > 0) group = (size_t)fixed;
    else {
        const size_t nslots = ctx->streams.size() + 1;
        size_t busy = 0;
        for (size_t i = 0; i < nslots && busy <= 2; i++) {
            const uint64_t e = ctx->epoch[i].load(std::memory_order_acquire);
            if (e != 0 && (i >= A.marks.size() || !A.marks[i].last || A.marks[i].epoch != e)) busy++;
        }
        if (busy <= 2) group = 1;
    }
    if (A.pending.size() >= group || A.pending_bytes >= ((size_t)256 << 20)) arena_seal_locked(ctx);
    return FHE_OK;
}

static long opt_value(const char* opts, const char* key, long dflt) {
    if (!opts) return dflt;
    std::string s(opts), k = std::string(key) + "=";
    size_t p = 0;
    while ((p = s.find(k, p)) != std::string::npos) {
        if (p == 0 || s[p - 1] == ',' || s[p - 1] == ' ') return atol(s.c_str() + p + k.size());
        p += k.size();
    }
    return dflt;
}

// ---------------------------------------------------------------------------------------------
// runtime
// ---------------------------------------------------------------------------------------------
extern "C" int fhe_init(int device, const char* opts, fhe_ctx** out) {
    GateHold gate_;
    if (!out) return fail(FHE_EINVAL, "fhe_init: out is null");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail(FHE_ENODEV, "fhe_init: no CUDA device (%s); this backend has no CPU fallback",
                    e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    }
    if (device < 0 || device >= n) return fail(FHE_EINVAL, "fhe_init: device %d not in [0,%d)", device, n);
    CUDA_TRY(cudaSetDevice(device));
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        return fail(FHE_ENODEV, "fhe_init: device %d is sm_%d%d; this library holds sm_100a code only",
                    device, prop.major, prop.minor);
    auto* ctx = new fhe_ctx;
    ctx->device = device;
    long ns = std::max(1l, std::min(64l, opt_value(opts, "streams", 4)));
    ctx->epoch.reset(new std::atomic<uint64_t>[ns + 1]);
    for (long i = 0; i <= ns; i++) ctx->epoch[i].store(0);
    ctx->arena.slab_bytes = (size_t)std::max(1l, opt_value(opts, "arena_mb", 256)) << 20;
    ctx->arena.seal = [ctx] { arena_seal_locked(ctx); };
    long grow_mb = (long)(prop.totalGlobalMem >> 22);        // default: a quarter of the device
    if (const char* e = getenv("FHE_B200_ARENA_GROW_MB")) grow_mb = atol(e);
    ctx->arena.grow_limit = (size_t)std::max(0l, opt_value(opts, "arena_grow_mb", grow_mb)) << 20;
    ctx->chunk_bytes = (size_t)std::max(1l, std::min(1024l, opt_value(opts, "chunk_mb", 32))) << 20;
    for (long i = 0; i < ns; i++) {
        cudaStream_t s;
        e = cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
        if (e != cudaSuccess) {
            delete ctx;
            return fail(FHE_ECUDA, "cudaStreamCreate failed: %s", cudaGetErrorString(e));
        }
        ctx->streams.push_back(s);
    }
    *out = ctx;
    return FHE_OK;
}

extern "C" int fhe_shutdown(fhe_ctx* ctx) {
    GateHold gate_;
    if (!ctx) return FHE_OK;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    for (auto& t : ctx->timers) { cudaEventDestroy(t.second.a); cudaEventDestroy(t.second.b); }
    for (auto s : ctx->streams) cudaStreamDestroy(s);
    for (void* p : ctx->pinned) cudaFreeHost(p);
    ctx->arena.destroy();
    delete ctx;
    return FHE_OK;
}

extern "C" int fhe_device(const fhe_ctx* ctx) { return ctx ? ctx->device : -1; }
extern "C" int fhe_num_streams(const fhe_ctx* ctx) { return ctx ? (int)ctx->streams.size() : 0; }

extern "C" int fhe_sync(fhe_ctx* ctx, int stream) {
    GateHold gate_;
    if (!ctx) return fail(FHE_EINVAL, "null context");
    CUDA_TRY(cudaSetDevice(ctx->device));
    if (stream == -2) { CUDA_TRY(cudaDeviceSynchronize()); return FHE_OK; }
    StreamRef s;
    int rc = get_stream(ctx, stream, &s);
    if (rc) return rc;
    CUDA_TRY(cudaStreamSynchronize(s));
    return FHE_OK;
}

// ---------------------------------------------------------------------------------------------
// memory
// ---------------------------------------------------------------------------------------------
extern "C" int fhe_alloc(fhe_ctx* ctx, size_t bytes, fhe_dptr* out) {
    GateHold gate_;
    if (!ctx || !out) return fail(FHE_EINVAL, "fhe_alloc: null argument");
    CUDA_TRY(cudaSetDevice(ctx->device));
    void* p = nullptr;
    int rc = ctx->arena.alloc(bytes, &p);
    if (rc) return rc;
    *out = (fhe_dptr)(uintptr_t)p;
    return FHE_OK;
}

extern "C" int fhe_free(fhe_ctx* ctx, fhe_dptr p) {
    GateHold gate_;
    if (!ctx) return fail(FHE_EINVAL, "null context");
    if (!p) return FHE_OK;
    cudaSetDevice(ctx->device);
    return arena_park(ctx, (void*)(uintptr_t)p);
}

extern "C" int fhe_arena_stats(fhe_ctx* ctx, size_t* reserved_bytes, size_t* in_use_bytes) {
    if (!ctx) return fail(FHE_EINVAL, "null context");
    std::lock_guard<std::mutex> lk(ctx->arena.mu);
    if (reserved_bytes) *reserved_bytes = ctx->arena.reserved;
    if (in_use_bytes) *in_use_bytes = ctx->arena.in_use;
    return FHE_OK;
}

extern "C" int fhe_host_alloc(fhe_ctx* ctx, size_t bytes, void** out) {
    GateHold gate_;
    if (!ctx || !out) return fail(FHE_EINVAL, "fhe_host_alloc: null argument");
    CUDA_TRY(cudaSetDevice(ctx->device));
    CUDA_TRY(cudaMallocHost(out, bytes ? bytes : 1));
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->pinned.push_back(*out);
    return FHE_OK;
}

extern "C" int fhe_host_free(fhe_ctx* ctx, void* p) {
    GateHold gate_;
    if (!ctx) return fail(FHE_EINVAL, "null context");
    if (!p) return FHE_OK;
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        auto it = std::find(ctx->pinned.begin(), ctx->pinned.end(), p);
        if (it == ctx->pinned.end()) return fail(FHE_EINVAL, "fhe_host_free: unknown pointer");
        ctx->pinned.erase(it);
    }
    CUDA_TRY(cudaFreeHost(p));
    return FHE_OK;
}

static int copy_async(fhe_ctx* ctx, void* dst, const void* src, size_t bytes, cudaMemcpyKind kind, int stream) {
    StreamRef s;
    int rc = get_stream(ctx, stream, &s);
    if (rc) return rc;
    if (bytes == 0) return FHE_OK;
    if (!dst || !src) return fail(FHE_EINVAL, "copy: null pointer");
    CUDA_TRY(cudaSetDevice(ctx->device));
    CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, kind, s));
    return FHE_OK;
}
extern "C" int fhe_h2d(fhe_ctx* ctx, fhe_dptr dst, const void* src, size_t bytes, int stream) {
    return copy_async(ctx, (void*)(uintptr_t)dst, src, bytes, cudaMemcpyHostToDevice, stream);
}
extern "C" int fhe_d2h(fhe_ctx* ctx, void* dst, fhe_dptr src, size_t bytes, int stream) {
    return copy_async(ctx, dst, (const void*)(uintptr_t)src, bytes, cudaMemcpyDeviceToHost, stream);
}
extern "C" int fhe_d2d(fhe_ctx* ctx, fhe_dptr dst, fhe_dptr src, size_t bytes, int stream) {
    return copy_async(ctx, (void*)(uintptr_t)dst, (const void*)(uintptr_t)src, bytes, cudaMemcpyDeviceToDevice, stream);
}

// ---------------------------------------------------------------------------------------------
// timing
// ---------------------------------------------------------------------------------------------
extern "C" int fhe_timer_start(fhe_ctx* ctx, int stream) {
    GateHold gate_;
    StreamRef s;
    int rc = get_stream(ctx, stream, &s);
    if (rc) return rc;
    CUDA_TRY(cudaSetDevice(ctx->device));
    Timer* t;
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        t = &ctx->timers[stream];
    }
    if (!t->a) { CUDA_TRY(cudaEventCreate(&t->a)); CUDA_TRY(cudaEventCreate(&t->b)); }
    CUDA_TRY(cudaEventRecord(t->a, s));
    return FHE_OK;
}

extern "C" int fhe_timer_stop(fhe_ctx* ctx, int stream, float* elapsed_ms) {
    GateHold gate_;
    StreamRef s;
    int rc = get_stream(ctx, stream, &s);
    if (rc) return rc;
    Timer* t;
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        auto it = ctx->timers.find(stream);
        if (it == ctx->timers.end() || !it->second.a) return fail(FHE_EINVAL, "fhe_timer_stop without fhe_timer_start");
        t = &it->second;
    }
    CUDA_TRY(cudaEventRecord(t->b, s));
    CUDA_TRY(cudaEventSynchronize(t->b));
    float ms = 0;
    CUDA_TRY(cudaEventElapsedTime(&ms, t->a, t->b));
    if (elapsed_ms) *elapsed_ms = ms;
    return FHE_OK;
}

// ---------------------------------------------------------------------------------------------
// parameters
// ---------------------------------------------------------------------------------------------
extern "C" int fhe_gen_params(uint32_t nbits, uint32_t N, uint32_t towers, uint64_t* q, uint64_t* psi) {
    if (!q || !psi) return fail(FHE_EINVAL, "fhe_gen_params: null output");
    if (nbits < 4 || nbits > 60) return fail(FHE_EINVAL, "fhe_gen_params: nbits %u not in [4,60]", nbits);
    if (N == 0 || (N & (N - 1))) return fail(FHE_EINVAL, "fhe_gen_params: N must be a power of two");
    uint64_t m = 2ull * N, prev = host::first_prime(nbits, m);
    for (uint32_t i = 0; i < towers; i++) {
        prev = host::previous_prime(prev, m);
        if (prev <= m) return fail(FHE_EINVAL, "fhe_gen_params: ran out of %u-bit primes = 1 mod %llu", nbits, (unsigned long long)m);
        q[i] = prev;
        psi[i] = host::root_of_unity(m, prev);
    }
    return FHE_OK;
}

extern "C" uint64_t fhe_root_of_unity(uint64_t m, uint64_t q) { return host::root_of_unity(m, q); }

// ---------------------------------------------------------------------------------------------
// ring
// ---------------------------------------------------------------------------------------------
static TowerMod make_mod(uint64_t q) {
    TowerMod m;
    unsigned b = host::msb(q);
    m.q = q;
    m.mu = (uint64_t)(((host::u128)1 << (2 * b + 3)) / q);
    m.n = b - 2;
    m.s = b + 5;
    towermod_set_reduce(m);
    return m;
}

extern "C" int fhe_ring_destroy(fhe_ring* r) {
    GateHold gate_;
    if (!r) return FHE_OK;
    fhe_ctx* c = r->ctx;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    if (r->d_mods) arena_park(c, r->d_mods);
    if (r->d_twF) arena_park(c, r->d_twF);
    if (r->d_twI) arena_park(c, r->d_twI);
    if (r->d_invc) arena_park(c, r->d_invc);
    delete r;
    return FHE_OK;
}

extern "C" int fhe_ring_create(fhe_ctx* ctx, uint32_t N, uint32_t T, const uint64_t* q, const uint64_t* psi,
                               fhe_ring** out) {
    GateHold gate_;
    if (!ctx || !q || !out) return fail(FHE_EINVAL, "fhe_ring_create: null argument");
    *out = nullptr;
    if (N == 0 || (N & (N - 1)) || N > 65536) return fail(FHE_EINVAL, "fhe_ring_create: N=%u must be a power of two <= 65536", N);
    if (T == 0 || T > 4096) return fail(FHE_EINVAL, "fhe_ring_create: towers=%u not in [1,4096]", T);
    for (uint32_t t = 0; t < T; t++)
        if (q[t] < 4 || q[t] >= (1ull << 61))
            return fail(FHE_EINVAL, "fhe_ring_create: modulus %u = %llu not in [4, 2^61)", t, (unsigned long long)q[t]);
    CUDA_TRY(cudaSetDevice(ctx->device));
    auto* r = new fhe_ring;
    r->ctx = ctx;
    r->q.assign(q, q + T);
    r->dev.N = N; r->dev.T = T; r->dev.logN = host::msb(N) - 1;

    std::vector<TowerMod> mods(T);
    for (uint32_t t = 0; t < T; t++) mods[t] = make_mod(q[t]);
    int rc = ctx->arena.alloc(sizeof(TowerMod) * T, &r->d_mods);
    if (rc) { fhe_ring_destroy(r); return rc; }
    cudaError_t e = cudaMemcpy(r->d_mods, mods.data(), sizeof(TowerMod) * T, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaStreamSynchronize(cudaStreamLegacy);      // pageable source: see upload()
    if (e != cudaSuccess) { fhe_ring_destroy(r); return fail(FHE_ECUDA, "upload moduli: %s", cudaGetErrorString(e)); }
    r->dev.mods = (const TowerMod*)r->d_mods;

    if (psi) {
        r->psi.assign(psi, psi + T);
        uint32_t trivial = 0;
        for (uint32_t t = 0; t < T; t++) trivial += (psi[t] == 0 || psi[t] == 1);
        if (trivial == T) {
            r->ntt_identity = true;
        } else if (trivial != 0) {
            fhe_ring_destroy(r);
            return fail(FHE_EUNSUPPORTED, "fhe_ring_create: mixing trivial (0/1) and real roots of unity is not supported");
        } else {
            for (uint32_t t = 0; t < T; t++) {
                if (q[t] >= (1ull << 60)) {      // lazy butterflies keep values below 16q in a 64-bit word
                    fhe_ring_destroy(r);
                    return fail(FHE_EUNSUPPORTED, "fhe_ring_create: NTT needs moduli below 2^60 (the reference's MAX_MODULUS_SIZE); modulus %u = %llu",
                                t, (unsigned long long)q[t]);
                }
                bool ok = N == 1 ? (psi[t] % q[t] == q[t] - 1) : (host::powmod(psi[t], N, q[t]) == q[t] - 1);
                if (!ok || psi[t] >= q[t]) {
                    fhe_ring_destroy(r);
                    return fail(FHE_EINVAL, "fhe_ring_create: psi[%u]=%llu is not a primitive %u-th root of unity mod %llu",
                                t, (unsigned long long)psi[t], 2 * N, (unsigned long long)q[t]);
                }
            }
            size_t tabBytes = (size_t)T * N * 16;
            std::vector<uint64_t> hF((size_t)T * N * 2), hI((size_t)T * N * 2), hC((size_t)T * 4);
            unsigned nth = std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), T);
            std::vector<std::thread> pool;
            for (unsigned w = 0; w < nth; w++)
                pool.emplace_back([&, w]() {
                    host::TowerTables tt;
                    for (uint32_t t = w; t < T; t += nth) {
                        host::precompute(psi[t], N, q[t], tt);
                        memcpy(&hF[(size_t)t * N * 2], tt.twF.data(), (size_t)N * 16);
                        memcpy(&hI[(size_t)t * N * 2], tt.twI.data(), (size_t)N * 16);
                        memcpy(&hC[(size_t)t * 4], tt.invc, 32);
                    }
                });
            for (auto& th : pool) th.join();
            if ((rc = ctx->arena.alloc(tabBytes, &r->d_twF)) || (rc = ctx->arena.alloc(tabBytes, &r->d_twI)) ||
                (rc = ctx->arena.alloc((size_t)T * 32, &r->d_invc))) {
                fhe_ring_destroy(r);
                return rc;
            }
            if ((e = cudaMemcpy(r->d_twF, hF.data(), tabBytes, cudaMemcpyHostToDevice)) != cudaSuccess ||
                (e = cudaMemcpy(r->d_twI, hI.data(), tabBytes, cudaMemcpyHostToDevice)) != cudaSuccess ||
                (e = cudaMemcpy(r->d_invc, hC.data(), (size_t)T * 32, cudaMemcpyHostToDevice)) != cudaSuccess ||
                (e = cudaStreamSynchronize(cudaStreamLegacy)) != cudaSuccess) {
                fhe_ring_destroy(r);
                return fail(FHE_ECUDA, "upload twiddle tables: %s", cudaGetErrorString(e));
            }
            r->dev.twF = (const ulonglong2*)r->d_twF;
            r->dev.twI = (const ulonglong2*)r->d_twI;
            r->dev.invc = (const ulonglong2*)r->d_invc;
            r->has_ntt = true;
        }
    }
    *out = r;
    return FHE_OK;
}

extern "C" uint32_t fhe_ring_n(const fhe_ring* r) { return r ? r->dev.N : 0; }
extern "C" uint32_t fhe_ring_towers(const fhe_ring* r) { return r ? r->dev.T : 0; }

extern "C" int fhe_ring_table(const fhe_ring* r, uint32_t tower, int which, uint64_t* out) {
    GateHold gate_;
    if (!r || !out) return fail(FHE_EINVAL, "fhe_ring_table: null argument");
    if (!r->has_ntt) return fail(FHE_EINVAL, "fhe_ring_table: ring has no NTT tables");
    if (tower >= r->dev.T || which < 0 || which > 3) return fail(FHE_EINVAL, "fhe_ring_table: bad tower/table");
    size_t N = r->dev.N;
    std::vector<uint64_t> tmp(2 * N);
    const void* src = (const char*)(which < 2 ? r->d_twF : r->d_twI) + (size_t)tower * N * 16;
    CUDA_TRY(cudaSetDevice(r->ctx->device));
    CUDA_TRY(cudaMemcpy(tmp.data(), src, N * 16, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < N; i++) out[i] = tmp[2 * i + (which & 1)];
    return FHE_OK;
}

// ---------------------------------------------------------------------------------------------
// hot path, device-resident operands
// ---------------------------------------------------------------------------------------------
static int ring_stream(fhe_ring* r, int stream, StreamRef* s) {
    if (!r) return fail(FHE_EINVAL, "null ring");
    int rc = get_stream(r->ctx, stream, s);
    if (rc) return rc;
    CUDA_TRY(cudaSetDevice(r->ctx->device));
    return FHE_OK;
}

extern "C" int fhe_dcrt_op(fhe_ring* r, int op, fhe_dptr dst, fhe_dptr a, fhe_dptr b, size_t batch, int stream) {
    StreamRef s;
    int rc = ring_stream(r, stream, &s);
    if (rc) return rc;
    if (op < FHE_OP_ADD || op > FHE_OP_MUL) return fail(FHE_EINVAL, "fhe_dcrt_op: unknown op %d", op);
    if (batch == 0) return FHE_OK;
    if (!dst || !a || !b) return fail(FHE_EINVAL, "fhe_dcrt_op: null device pointer");
    if (r->dev.N == 1 && ((batch * r->dev.T) & 1)) return fail(FHE_EUNSUPPORTED, "fhe_dcrt_op: N=1 needs an even number of towers*batch");
    CUDA_TRY(launch_elementwise(r->dev, op, (uint64_t*)(uintptr_t)dst, (const uint64_t*)(uintptr_t)a,
                                (const uint64_t*)(uintptr_t)b, batch, s));
    return FHE_OK;
}

extern "C" int fhe_dcrt_op_scalars(fhe_ring* r, int op, fhe_dptr dst, fhe_dptr a, const uint64_t* scalars, size_t batch, int stream) {
    StreamRef s;
    int rc = ring_stream(r, stream, &s);
    if (rc) return rc;
    if (op != FHE_OP_ADD && op != FHE_OP_SUB) return fail(FHE_EINVAL, "fhe_dcrt_op_scalars: op must be FHE_OP_ADD or FHE_OP_SUB");
    if (!scalars) return fail(FHE_EINVAL, "fhe_dcrt_op_scalars: null scalars");
    if (batch == 0) return FHE_OK;
    if (!dst || !a) return fail(FHE_EINVAL, "fhe_dcrt_op_scalars: null device pointer");
    if ((dst | a) & 15) return fail(FHE_EINVAL, "fhe_dcrt_op_scalars: device pointers must be 16-byte aligned");
    if (r->dev.T > EW_SCALARS_MAX) return fail(FHE_EUNSUPPORTED, "fhe_dcrt_op_scalars: more than %d towers", EW_SCALARS_MAX);
    if (r->dev.N < 2) return fail(FHE_EUNSUPPORTED, "fhe_dcrt_op_scalars: N must be at least 2");
    uint64_t red[EW_SCALARS_MAX];
    for (uint32_t i = 0; i < r->dev.T; i++) red[i] = scalars[i] % r->q[i];          // NativeVector::ModAdd / ModSub reduce the scalar first
    CUDA_TRY(launch_tower_scalars(r->dev, op, (uint64_t*)(uintptr_t)dst, (const uint64_t*)(uintptr_t)a, red, batch, s));
    return FHE_OK;
}

extern "C" int fhe_ntt_algo(fhe_ring* r, fhe_dptr x, size_t batch, int inverse, int algo, int stream) {
    StreamRef s;
    int rc = ring_stream(r, stream, &s);
    if (rc) return rc;
    if (r->ntt_identity) return FHE_OK;       // psi in {0,1}: the reference returns untouched
    if (!r->has_ntt) return fail(FHE_EINVAL, "fhe_ntt: ring was created without roots of unity");
    if (batch == 0) return FHE_OK;
    if (!x) return fail(FHE_EINVAL, "fhe_ntt: null device pointer");
    if (r->dev.N > 1 && (x & 15)) return fail(FHE_EINVAL, "fhe_ntt: device pointers must be 16-byte aligned");
    if (r->dev.N == 1) return FHE_OK;
    CUDA_TRY(launch_ntt(r->dev, (uint64_t*)(uintptr_t)x, batch, inverse != 0, s, algo));
    return FHE_OK;
}
extern "C" int fhe_ntt_forward(fhe_ring* r, fhe_dptr x, size_t batch, int stream) { return fhe_ntt_algo(r, x, batch, 0, 0, stream); }
extern "C" int fhe_ntt_inverse(fhe_ring* r, fhe_dptr x, size_t batch, int stream) { return fhe_ntt_algo(r, x, batch, 1, 0, stream); }

extern "C" int fhe_poly_mul(fhe_ring* r, fhe_dptr c, fhe_dptr a, fhe_dptr b, size_t batch, int stream) {
    if (c == a || c == b) return fail(FHE_EINVAL, "fhe_poly_mul: c must not alias an input");
    int rc;
    if ((rc = fhe_ntt_forward(r, a, batch, stream))) return rc;
    if (a != b && (rc = fhe_ntt_forward(r, b, batch, stream))) return rc;
    if ((rc = fhe_dcrt_op(r, FHE_OP_MUL, c, a, b, batch, stream))) return rc;
    return fhe_ntt_inverse(r, c, batch, stream);
}

extern "C" int fhe_automorphism(fhe_ring* r, fhe_dptr dst, fhe_dptr src, size_t batch, uint32_t k, int format, int stream) {
    StreamRef s;
    int rc = ring_stream(r, stream, &s);
    if (rc) return rc;
    if (format != 0 && format != 1) return fail(FHE_EINVAL, "fhe_automorphism: format must be 0 (EVALUATION) or 1 (COEFFICIENT)");
    if (k % 2 == 0) return fail(FHE_EINVAL, "Automorphism index not odd");          // poly-impl.h:346-347
    if (batch == 0) return FHE_OK;
    if (!dst || !src) return fail(FHE_EINVAL, "fhe_automorphism: null device pointer");
    if (r->dev.N > 1 && ((dst | src) & 15)) return fail(FHE_EINVAL, "fhe_automorphism: device pointers must be 16-byte aligned");
    if (dst == src) return fail(FHE_EINVAL, "fhe_automorphism: dst must not alias src (it is a permutation)");
    CUDA_TRY(launch_automorphism(r->dev, (uint64_t*)(uintptr_t)dst, (const uint64_t*)(uintptr_t)src, batch, k, format, s));
    return FHE_OK;
}

extern "C" int fhe_precompute_automap(uint32_t N, uint32_t k, uint32_t* map) {
    GateHold gate_;
    if (!map) return fail(FHE_EINVAL, "fhe_precompute_automap: null output");
    if (N == 0 || (N & (N - 1))) return fail(FHE_EINVAL, "fhe_precompute_automap: N must be a power of two");
    uint32_t logn = 0;
    while ((1u << logn) < N) logn++;
    auto rev = [logn](uint32_t x) { uint32_t r = 0; for (uint32_t i = 0; i < logn; i++) { r = (r << 1) | (x & 1); x >>= 1; } return r; };
    for (uint32_t j = 0; j < N; j++) {
        uint32_t idx = (((2 * j + 1) * k) & (2 * N - 1)) >> 1;
        map[rev(j)] = rev(idx);
    }
    return FHE_OK;
}

extern "C" int fhe_automorphism_map(fhe_ring* r, fhe_dptr dst, fhe_dptr src, size_t batch, fhe_dptr map, int stream) {
    StreamRef s;
    int rc = ring_stream(r, stream, &s);
    if (rc) return rc;
    if (batch == 0) return FHE_OK;
    if (!dst || !src || !map) return fail(FHE_EINVAL, "fhe_automorphism_map: null device pointer");
    if (r->dev.N > 1 && ((dst | src) & 15)) return fail(FHE_EINVAL, "fhe_automorphism_map: device pointers must be 16-byte aligned");
    if (dst == src) return fail(FHE_EINVAL, "fhe_automorphism_map: dst must not alias src");
    CUDA_TRY(launch_automorphism_map(r->dev, (uint64_t*)(uintptr_t)dst, (const uint64_t*)(uintptr_t)src, batch,
                                     (const uint32_t*)(uintptr_t)map, s));
    return FHE_OK;
}

extern "C" int fhe_ct_tensor_product(fhe_ring* r, const fhe_dptr* c, const fhe_dptr* a, const fhe_dptr* b,
                                     size_t batch, int stream) {
    StreamRef s;
    int rc = ring_stream(r, stream, &s);
    if (rc) return rc;
    if (!c || !a || !b) return fail(FHE_EINVAL, "fhe_ct_tensor_product: null argument");
    if (batch == 0) return FHE_OK;
    if (!c[0] || !c[1] || !c[2] || !a[0] || !a[1] || !b[0] || !b[1])
        return fail(FHE_EINVAL, "fhe_ct_tensor_product: null device pointer");
    if (c[0] == c[1] || c[0] == c[2] || c[1] == c[2]) return fail(FHE_EINVAL, "fhe_ct_tensor_product: outputs must be distinct");
    if (r->dev.N == 1 && ((batch * r->dev.T) & 1)) return fail(FHE_EUNSUPPORTED, "fhe_ct_tensor_product: N=1 needs an even number of towers*batch");
    auto P = [](fhe_dptr p) { return (uint64_t*)(uintptr_t)p; };
    CUDA_TRY(launch_tensor_product(r->dev, P(c[0]), P(c[1]), P(c[2]), P(a[0]), P(a[1]), P(b[0]), P(b[1]), batch, s));
    return FHE_OK;
}

extern "C" int fhe_ct_multiply(fhe_ring* r, const fhe_dptr* c, const fhe_dptr* a, const fhe_dptr* b, size_t batch,
                               int transform, int stream) {
    if (transform < 0 || transform > 2) return fail(FHE_EINVAL, "fhe_ct_multiply: transform must be 0, 1 or 2");
    if (!c || !a || !b) return fail(FHE_EINVAL, "fhe_ct_multiply: null argument");
    int rc;
    if (transform) {
        const int inv_in = transform == 2;       // 1: NTT in, INTT out;  2: INTT in, NTT out
        const fhe_dptr in[4] = {a[0], a[1], b[0], b[1]};
        for (int i = 0; i < 4; i++) {
            bool seen = false;                   // squaring passes the same buffers twice: transform once
            for (int j = 0; j < i; j++) seen |= in[j] == in[i];
            if (!seen && (rc = fhe_ntt_algo(r, in[i], batch, inv_in, 0, stream))) return rc;
        }
    }
    if ((rc = fhe_ct_tensor_product(r, c, a, b, batch, stream))) return rc;
    if (transform)
        for (int i = 0; i < 3; i++)
            if ((rc = fhe_ntt_algo(r, c[i], batch, transform == 1, 0, stream))) return rc;
    return FHE_OK;
}

extern "C" int fhe_fill_uniform(fhe_ring* r, fhe_dptr dst, size_t batch, uint64_t seed, int stream) {
    StreamRef s;
    int rc = ring_stream(r, stream, &s);
    if (rc) return rc;
    if (batch && !dst) return fail(FHE_EINVAL, "fhe_fill_uniform: null device pointer");
    CUDA_TRY(launch_fill_uniform(r->dev, (uint64_t*)(uintptr_t)dst, batch, seed, s));
    return FHE_OK;
}

// ---------------------------------------------------------------------------------------------
// flat vectors with explicit modulus
// ---------------------------------------------------------------------------------------------
static int check_vec_mod(int op, uint64_t q) {
    if (op < FHE_OP_ADD || op > FHE_OP_MUL) return fail(FHE_EINVAL, "unknown op %d", op);
    if (q == 0) return fail(FHE_EINVAL, "modulus is zero");
    if (op == FHE_OP_MUL && (q < 4 || q >= (1ull << 61)))
        return fail(FHE_EUNSUPPORTED, "modular multiply needs 4 <= modulus < 2^61 (got %llu)", (unsigned long long)q);
    return FHE_OK;
}

extern "C" int fhe_vec_op(fhe_ctx* ctx, int op, fhe_dptr dst, fhe_dptr a, fhe_dptr b, size_t n, uint64_t q, int stream) {
    StreamRef s;
    int rc = get_stream(ctx, stream, &s);
    if (rc || (rc = check_vec_mod(op, q))) return rc;
    if (n == 0) return FHE_OK;
    if (!dst || !a || !b) return fail(FHE_EINVAL, "fhe_vec_op: null device pointer");
    CUDA_TRY(cudaSetDevice(ctx->device));
    CUDA_TRY(launch_vec(op, (uint64_t*)(uintptr_t)dst, (const uint64_t*)(uintptr_t)a, (const uint64_t*)(uintptr_t)b, n, q, s));
    return FHE_OK;
}

// kind = enum mod_kernel (pim/common.h:32-46).  Host-side argument handling as the reference's vector
// methods do it: ModExp reduces the exponent mod the modulus first (mubintvecnat.cpp:391-392).
extern "C" int fhe_vec_modop(fhe_ctx* ctx, int kind, fhe_dptr dst, fhe_dptr src, size_t n, uint64_t mod,
                             uint64_t p, uint64_t q, int stream) {
    StreamRef s;
    int rc = get_stream(ctx, stream, &s);
    if (rc) return rc;
    if (kind < SWITCHMODULUS || kind > DIVIDEANDROUNDEQ) return fail(FHE_EINVAL, "fhe_vec_modop: unknown mod_kernel %d", kind);
    if (mod == 0) return fail(FHE_EINVAL, "fhe_vec_modop: modulus is zero");
    if (kind <= MODEQ && p == 0) return fail(FHE_EINVAL, "fhe_vec_modop: new modulus is zero");
    if ((kind == MULTIPLYANDROUND || kind == MULTIPLYANDROUNDEQ) && q == 0)
        return fail(FHE_EINVAL, "NativeIntegerT MultiplyAndRound: Divide by zero");
    if ((kind == DIVIDEANDROUND || kind == DIVIDEANDROUNDEQ) && p == 0)
        return fail(FHE_EINVAL, "NativeIntegerT DivideAndRound: zero");
    if (n == 0) return FHE_OK;
    if (!dst || !src) return fail(FHE_EINVAL, "fhe_vec_modop: null device pointer");
    if ((kind == MODEXP || kind == MODEXPEQ) && p >= mod) p %= mod;
    CUDA_TRY(cudaSetDevice(ctx->device));
    bool inv = kind == MODINVERSE || kind == MODINVERSEEQ;
    void* err = nullptr;                         // the error word is this call's own: two threads may invert at once
    if (inv && (rc = ctx->arena.alloc(256, &err))) return rc;
    cudaError_t e = inv ? cudaMemsetAsync(err, 0, sizeof(int), s) : cudaSuccess;
    if (e == cudaSuccess)
        e = launch_modop(kind, (uint64_t*)(uintptr_t)dst, (const uint64_t*)(uintptr_t)src, n, mod, p, q, (int*)err, s);
    int flag = 0;
    if (inv) {                                   // the reference throws math_error: report it synchronously
        if (e == cudaSuccess) e = cudaMemcpyAsync(&flag, err, sizeof(int), cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        arena_park(ctx, err);
    }
    if (e != cudaSuccess) return fail(FHE_ECUDA, "fhe_vec_modop: %s", cudaGetErrorString(e));
    if (flag) return fail(FHE_EINVAL, "fhe_vec_modop: an element does not have a ModInverse using %llu", (unsigned long long)mod);
    return FHE_OK;
}

extern "C" int fhe_vec_op_scalar(fhe_ctx* ctx, int op, fhe_dptr dst, fhe_dptr a, uint64_t scalar, size_t n,
                                 uint64_t q, int stream) {
    StreamRef s;
    int rc = get_stream(ctx, stream, &s);
    if (rc || (rc = check_vec_mod(op, q))) return rc;
    if (n == 0) return FHE_OK;
    if (!dst || !a) return fail(FHE_EINVAL, "fhe_vec_op_scalar: null device pointer");
    CUDA_TRY(cudaSetDevice(ctx->device));
    CUDA_TRY(launch_vec_scalar(op, (uint64_t*)(uintptr_t)dst, (const uint64_t*)(uintptr_t)a, scalar, n, q, s));
    return FHE_OK;
}

extern "C" int fhe_checksum(fhe_ctx* ctx, fhe_dptr src, size_t nwords, uint64_t* sum_xor, int stream) {
    StreamRef s;
    int rc = get_stream(ctx, stream, &s);
    if (rc) return rc;
    if (!sum_xor) return fail(FHE_EINVAL, "fhe_checksum: null output");
    if (nwords && !src) return fail(FHE_EINVAL, "fhe_checksum: null device pointer");
    CUDA_TRY(cudaSetDevice(ctx->device));
    void* d = nullptr;
    if ((rc = ctx->arena.alloc(256, &d))) return rc;
    cudaError_t e = launch_checksum((const uint64_t*)(uintptr_t)src, nwords, (uint64_t*)d, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(sum_xor, d, 16, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    arena_park(ctx, d);
    if (e != cudaSuccess) return fail(FHE_ECUDA, "fhe_checksum: %s", cudaGetErrorString(e));
    return FHE_OK;
}

// ---------------------------------------------------------------------------------------------
// legacy descriptor path (PimData::common_do_ops, pim/PimData.h:275-305)
// ---------------------------------------------------------------------------------------------
extern "C" int fhe_launch(fhe_ctx* ctx, int kernel_id, const void* desc, fhe_dptr heap_base) {
    GateHold gate_;
    if (!ctx || !desc) return fail(FHE_EINVAL, "fhe_launch: null argument");
    if (kernel_id == FHE_KERNEL_BOOT) return FHE_OK;          // boot only reset the DPU heap
    int op;
    switch (kernel_id) {
        case FHE_KERNEL_MOD_ADD: op = FHE_OP_ADD; break;
        case FHE_KERNEL_MOD_SUB: op = FHE_OP_SUB; break;
        case FHE_KERNEL_MOD_MULT: op = FHE_OP_MUL; break;
        case FHE_KERNEL_MOD_OPS: {
            const mod_ops* o = (const mod_ops*)desc;
            bool eq = o->kernel == SWITCHMODULUS || o->kernel == MODEQ || o->kernel == MODBYTWOEQ || o->kernel == MODEXPEQ ||
                      o->kernel == MODINVERSEEQ || o->kernel == MULTIPLYANDROUNDEQ || o->kernel == DIVIDEANDROUNDEQ;
            fhe_dptr src = heap_base + o->op.start, dst = eq ? src : heap_base + o->res.start;
            int rc = fhe_vec_modop(ctx, (int)o->kernel, dst, src, o->op.size / sizeof(uint64_t), o->mod, o->p, o->q, FHE_STREAM_LEGACY);
            if (rc) return rc;
            CUDA_TRY(cudaStreamSynchronize(cudaStreamLegacy));
            return FHE_OK;
        }
        default: return fail(FHE_EINVAL, "fhe_launch: unknown kernel id %d", kernel_id);
    }
    const mod_add_sub_mult* m = (const mod_add_sub_mult*)desc;
    size_t n = m->op1.size / sizeof(uint64_t);
    fhe_dptr a = heap_base + m->op1.start, dst;
    int rc;
    switch (m->kernel) {
        case VECTOR:    dst = heap_base + m->res.start; rc = fhe_vec_op(ctx, op, dst, a, heap_base + m->op2.start, n, m->mod, FHE_STREAM_LEGACY); break;
        case VECTOR_EQ: rc = fhe_vec_op(ctx, op, a, a, heap_base + m->op2.start, n, m->mod, FHE_STREAM_LEGACY); break;
        case SCALAR:    dst = heap_base + m->res.start; rc = fhe_vec_op_scalar(ctx, op, dst, a, m->op2.start, n, m->mod, FHE_STREAM_LEGACY); break;
        case SCALAR_EQ: rc = fhe_vec_op_scalar(ctx, op, a, a, m->op2.start, n, m->mod, FHE_STREAM_LEGACY); break;
        default: return fail(FHE_EINVAL, "fhe_launch: unknown add_sub_mul_kernel %d", (int)m->kernel);
    }
    if (rc) return rc;
    CUDA_TRY(cudaStreamSynchronize(cudaStreamLegacy));        // start_kernel() is DPU_SYNCHRONOUS
    return FHE_OK;
}

// ---------------------------------------------------------------------------------------------
// RNS basis conversion and scaling
// ---------------------------------------------------------------------------------------------
extern "C" int fhe_ring_view(fhe_ring* parent, uint32_t first, uint32_t count, fhe_ring** out) {
    GateHold gate_;
    if (!parent || !out) return fail(FHE_EINVAL, "fhe_ring_view: null argument");
    *out = nullptr;
    if (count == 0 || first + count > parent->dev.T || first + count < first)
        return fail(FHE_EINVAL, "fhe_ring_view: towers [%u, %u) not inside the parent's %u", first, first + count, parent->dev.T);
    auto* r = new fhe_ring;
    r->ctx = parent->ctx;
    r->dev = parent->dev;
    r->dev.T = count;
    r->dev.mods = parent->dev.mods + first;
    if (parent->dev.twF) r->dev.twF = parent->dev.twF + (size_t)first * parent->dev.N;
    if (parent->dev.twI) r->dev.twI = parent->dev.twI + (size_t)first * parent->dev.N;
    if (parent->dev.invc) r->dev.invc = parent->dev.invc + 2 * (size_t)first;
    r->q.assign(parent->q.begin() + first, parent->q.begin() + first + count);
    if (!parent->psi.empty()) r->psi.assign(parent->psi.begin() + first, parent->psi.begin() + first + count);
    r->has_ntt = parent->has_ntt;
    r->ntt_identity = parent->ntt_identity;
    *out = r;                                  // owns no device memory: fhe_ring_destroy only deletes the handle
    return FHE_OK;
}

namespace {
// grow-only device scratch owned by a plan; released when the plan is destroyed
struct Scratch {
    void* p = nullptr;
    size_t bytes = 0;
    int need(fhe_ctx* ctx, size_t n) {
        if (n <= bytes) return FHE_OK;
        GateHold gate_;
        if (p) { cudaDeviceSynchronize(); arena_park(ctx, p); p = nullptr; bytes = 0; }
        int rc = ctx->arena.alloc(n, &p);
        if (rc) return rc;
        bytes = n;
        return FHE_OK;
    }
    void drop(fhe_ctx* ctx) { if (p) arena_park(ctx, p); p = nullptr; bytes = 0; }
};

template <class T>
int upload(fhe_ctx* ctx, const std::vector<T>& h, void** d) {
    GateHold gate_;
    if (!*d) {
        int rc = ctx->arena.alloc(std::max<size_t>(h.size() * sizeof(T), 16), d);
        if (rc) return rc;
    } else {
        cudaDeviceSynchronize();      // a table being replaced may still be read by launches on the (non-blocking) pool streams
    }
    cudaError_t e = cudaMemcpy(*d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice);
    // A copy from pageable memory returns once the data sits in the driver's staging buffer; the DMA may still be in
    // flight, and the (non-blocking) pool streams do not wait for the legacy stream.  Wait for it here: the caller
    // launches on a pool stream next.
    if (e == cudaSuccess) e = cudaStreamSynchronize(cudaStreamLegacy);
    return e == cudaSuccess ? FHE_OK : fail(FHE_ECUDA, "upload constants: %s", cudaGetErrorString(e));
}

uint64_t prod_mod(const std::vector<uint64_t>& f, size_t skip, uint64_t m) {   // prod_{k != skip} f[k] mod m
    uint64_t r = 1 % m;
    for (size_t k = 0; k < f.size(); k++)
        if (k != skip) r = host::mulmod(r, f[k] % m, m);
    return r;
}
}  // namespace

// The carry-free 30-bit split accumulation of basis.cu needs every factor below 2^60 (true of every modulus the reference
// can produce, basicint.h:44-45); tables that are not take the 128-bit multiply-accumulate kernels instead.
// FHE_B200_BASIS_WIDE=1 forces that form (tests run both).
static bool all_below_2_60(const std::vector<uint64_t>& v) {
    for (uint64_t x : v)
        if (x >> 60) return false;
    return true;
}
static bool basis_force_wide() {
    const char* e = getenv("FHE_B200_BASIS_WIDE");
    return e && *e && *e != '0';
}

struct fhe_basis {
    fhe_ctx* ctx = nullptr;
    fhe_ring *from = nullptr, *to = nullptr;
    std::vector<uint64_t> hat_inv, hat_modp;      // [sq], [sq][sp]
    BasisDev dev{};                               // plain tables (t = 0)
    void *d_q = nullptr, *d_pre = nullptr, *d_hat = nullptr, *d_tgt = nullptr;
    // ApproxModDown extras: [P^-1]_{q_i} records and the t-folded variant of pre / tgt (cached for the last t)
    void *d_pinv = nullptr, *d_pre_t = nullptr, *d_tgt_t = nullptr;
    void *d_qinv = nullptr, *d_alphaq = nullptr;   // exact form (SwitchCRTBasis): 1/q_i and [alpha*Q]_{p_j}
    void* d_tc = nullptr;                          // the table as 8-bit limbs for the tensor-core kernel (basis.cu)
    uint64_t t_cached = 0;
    bool pinv_ready = false;
    Scratch s0, s1;
};

static void basis_fill(const fhe_basis* b, uint64_t t, std::vector<ulonglong2>& pre, std::vector<BasisTarget>& tgt) {
    const auto& q = b->from->q;
    const auto& p = b->to->q;
    pre.resize(q.size());
    tgt.resize(p.size());
    for (size_t i = 0; i < q.size(); i++) {
        uint64_t c = b->hat_inv[i] % q[i];
        if (t) c = host::mulmod(c, host::invmod(t % q[i], q[i]), q[i]);     // partP *= [t^-1]_{p_j} folded in
        pre[i] = make_ulonglong2(c, host::shoup_companion(c, q[i]));
    }
    for (size_t j = 0; j < p.size(); j++) {
        BasisTarget g{};
        basis_target_set(g, p[j]);
        if (t) { g.post = t % p[j]; g.postp = host::shoup_companion(g.post, p[j]); }
        tgt[j] = g;
    }
}

// The [Q/q_i]_{p_j} table as the B operand of basis_switch_tc_kernel: 8-bit limbs, Toeplitz per target
// (row 16j + d, column 8i + a holds limb d - a of the entry), stored in the kernel's shared-memory layout
// (16-byte K chunks, chunk-major), one block per group of at most 16 targets.
bool fhe::basis_tc_build(const uint64_t* hat, uint32_t sq, uint32_t sp, uint32_t& groups, uint32_t& jg, uint32_t& kpad,
                         std::vector<uint8_t>& out) {
    if (sq == 0 || sq > 32 || sp == 0) return false;
    const uint32_t ng = (sp + 15) / 16;
    jg = (sp + ng - 1) / ng;
    groups = (sp + jg - 1) / jg;
    kpad = (8 * sq + 31) / 32 * 32;
    const uint32_t ngemm = 16 * jg;
    out.assign((size_t)groups * kpad * ngemm, 0);
    for (uint32_t g = 0; g < groups; g++)
        for (uint32_t jj = 0; jj < jg && g * jg + jj < sp; jj++)
            for (uint32_t i = 0; i < sq; i++) {
                const uint64_t h = hat[(size_t)i * sp + g * jg + jj];
                for (uint32_t a = 0; a < 8; a++)
                    for (uint32_t bb = 0; bb < 8; bb++) {
                        const uint32_t d = a + bb, k = 8 * i + a, nrow = 16 * jj + d;
                        out[(size_t)g * kpad * ngemm + ((size_t)(k / 16) * ngemm + nrow) * 16 + (k % 16)] = (uint8_t)(h >> (8 * bb));
                    }
            }
    return true;
}

extern "C" int fhe_basis_destroy(fhe_basis* b) {
    GateHold gate_;
    if (!b) return FHE_OK;
    cudaSetDevice(b->ctx->device);
    cudaDeviceSynchronize();
    for (void* p : {b->d_q, b->d_pre, b->d_hat, b->d_tgt, b->d_pinv, b->d_pre_t, b->d_tgt_t, b->d_qinv, b->d_alphaq, b->d_tc})
        if (p) arena_park(b->ctx, p);
    b->s0.drop(b->ctx);
    b->s1.drop(b->ctx);
    delete b;
    return FHE_OK;
}

extern "C" int fhe_basis_create(fhe_ring* from, fhe_ring* to, const uint64_t* qhat_inv_modq, const uint64_t* qhat_modp,
                                fhe_basis** out) {
    GateHold gate_;
    if (!from || !to || !out) return fail(FHE_EINVAL, "fhe_basis_create: null argument");
    *out = nullptr;
    if (from->ctx != to->ctx) return fail(FHE_EINVAL, "fhe_basis_create: rings live on different contexts");
    if (from->dev.N != to->dev.N) return fail(FHE_EINVAL, "fhe_basis_create: ring dimensions differ");
    const size_t sq = from->q.size(), sp = to->q.size();
    for (uint64_t p : to->q)
        if ((p & (p - 1)) == 0) return fail(FHE_EINVAL, "fhe_basis_create: target modulus %llu is a power of two", (unsigned long long)p);
    CUDA_TRY(cudaSetDevice(from->ctx->device));
    auto* b = new fhe_basis;
    b->ctx = from->ctx; b->from = from; b->to = to;
    b->hat_inv.resize(sq);
    b->hat_modp.resize(sq * sp);
    for (size_t i = 0; i < sq; i++) {
        if (qhat_inv_modq) {
            b->hat_inv[i] = qhat_inv_modq[i];
        } else {
            uint64_t h = prod_mod(from->q, i, from->q[i]);
            if (h == 0 || std::__gcd(h, from->q[i]) != 1) {
                delete b;
                return fail(FHE_EINVAL, "fhe_basis_create: source moduli are not pairwise coprime (tower %zu)", i);
            }
            b->hat_inv[i] = host::invmod(h, from->q[i]);
        }
        for (size_t j = 0; j < sp; j++)
            b->hat_modp[i * sp + j] = qhat_modp ? qhat_modp[i * sp + j] : prod_mod(from->q, i, to->q[j]);
    }
    std::vector<ulonglong2> pre;
    std::vector<BasisTarget> tgt;
    basis_fill(b, 0, pre, tgt);
    int rc;
    if ((rc = upload(b->ctx, from->q, &b->d_q)) || (rc = upload(b->ctx, pre, &b->d_pre)) ||
        (rc = upload(b->ctx, b->hat_modp, &b->d_hat)) || (rc = upload(b->ctx, tgt, &b->d_tgt))) {
        fhe_basis_destroy(b);
        return rc;
    }
    b->dev.sq = (uint32_t)sq; b->dev.sp = (uint32_t)sp;
    b->dev.q = (const uint64_t*)b->d_q;
    b->dev.pre = (const ulonglong2*)b->d_pre;
    b->dev.hat = (const uint64_t*)b->d_hat;
    b->dev.tgt = (const BasisTarget*)b->d_tgt;
    b->dev.narrow = !basis_force_wide() && all_below_2_60(from->q) && all_below_2_60(b->hat_modp);
    {   // tensor-core form of the table (sq <= 32; ring dimension checked at launch)
        std::vector<uint8_t> tc;
        uint32_t groups = 0, jg = 0, kpad = 0;
        if (basis_tc_build(b->hat_modp.data(), (uint32_t)sq, (uint32_t)sp, groups, jg, kpad, tc)) {
            if ((rc = upload(b->ctx, tc, &b->d_tc))) { fhe_basis_destroy(b); return rc; }
            b->dev.tc_b = (const uint8_t*)b->d_tc;
            b->dev.tc_groups = groups; b->dev.tc_jg = jg; b->dev.tc_kpad = kpad;
        }
    }
    *out = b;
    return FHE_OK;
}

extern "C" int fhe_basis_tables(const fhe_basis* b, uint64_t* qhat_inv_modq, uint64_t* qhat_modp) {
    if (!b) return fail(FHE_EINVAL, "null basis");
    if (qhat_inv_modq) memcpy(qhat_inv_modq, b->hat_inv.data(), b->hat_inv.size() * 8);
    if (qhat_modp) memcpy(qhat_modp, b->hat_modp.data(), b->hat_modp.size() * 8);
    return FHE_OK;
}

extern "C" int fhe_approx_switch_crt_basis(fhe_basis* b, fhe_dptr dst, fhe_dptr src, size_t batch, int stream) {
    if (!b) return fail(FHE_EINVAL, "null basis");
    StreamRef s;
    int rc = ring_stream(b->from, stream, &s);
    if (rc) return rc;
    if (batch == 0) return FHE_OK;
    if (!dst || !src) return fail(FHE_EINVAL, "fhe_approx_switch_crt_basis: null device pointer");
    if ((dst | src) & 15) return fail(FHE_EINVAL, "fhe_approx_switch_crt_basis: device pointers must be 16-byte aligned");
    CUDA_TRY(launch_basis_switch(b->dev, (uint64_t*)(uintptr_t)dst, b->dev.sp, 0, (const uint64_t*)(uintptr_t)src, b->dev.sq, 0,
                                 batch, b->from->dev.logN, false, s));
    return FHE_OK;
}

// strided tower copy between [batch][dT][N] (towers d0..) and [batch][sT][N] (towers s0..), cnt towers per item
static cudaError_t copy_towers(uint64_t* dst, uint32_t dT, uint32_t d0, const uint64_t* src, uint32_t sT, uint32_t s0,
                               uint32_t cnt, size_t batch, uint32_t N, cudaStream_t s) {
    return cudaMemcpy2DAsync(dst + (size_t)d0 * N, (size_t)dT * N * 8, src + (size_t)s0 * N, (size_t)sT * N * 8,
                             (size_t)cnt * N * 8, batch, cudaMemcpyDeviceToDevice, s);
}

// NTT of the r->T towers that start at x (outT towers apart per batch item; 0 = packed).  With src the transform reads
// there (inT towers apart) and leaves src untouched: no staging copy, and a result can be produced straight into its
// towers of a wider poly.
static int ring_ntt(fhe_ring* r, uint64_t* x, size_t batch, bool inverse, cudaStream_t s, const char* who,
                    const uint64_t* src = nullptr, uint32_t outT = 0, uint32_t inT = 0) {
    const uint32_t T = r->dev.T, N = r->dev.N;
    if (!outT) outT = T;
    if (!inT) inT = src ? T : outT;
    if (r->ntt_identity || N == 1) {
        if (src && src != x)
            CUDA_TRY(cudaMemcpy2DAsync(x, (size_t)outT * N * 8, src, (size_t)inT * N * 8, (size_t)T * N * 8, batch,
                                       cudaMemcpyDeviceToDevice, s));
        return FHE_OK;
    }
    if (!r->has_ntt) return fail(FHE_EINVAL, "%s: a ring was created without roots of unity", who);
    CUDA_TRY(launch_ntt(r->dev, x, batch, inverse, s, 0, src, outT, inT));
    return FHE_OK;
}

extern "C" int fhe_approx_mod_up(fhe_basis* b, fhe_dptr dst_qp, fhe_dptr src_q, size_t batch, int format, int stream) {
    if (!b) return fail(FHE_EINVAL, "null basis");
    StreamRef s;
    int rc = ring_stream(b->from, stream, &s);
    if (rc) return rc;
    if (format != 0 && format != 1) return fail(FHE_EINVAL, "fhe_approx_mod_up: format must be 0 or 1");
    if (batch == 0) return FHE_OK;
    if (!dst_qp || !src_q) return fail(FHE_EINVAL, "fhe_approx_mod_up: null device pointer");
    if ((dst_qp | src_q) & 15) return fail(FHE_EINVAL, "fhe_approx_mod_up: device pointers must be 16-byte aligned");
    const uint32_t N = b->from->dev.N, logN = b->from->dev.logN, sq = b->dev.sq, sp = b->dev.sp;
    uint64_t* dst = (uint64_t*)(uintptr_t)dst_qp;
    const uint64_t* src = (const uint64_t*)(uintptr_t)src_q;
    const size_t wq = batch * sq * (size_t)N;
    const char* who = "fhe_approx_mod_up";
    const uint64_t* coef = src;                              // COEFFICIENT form of the Q towers (dcrtpoly-impl.h:1096-1099)
    if (format == 0) {
        if ((rc = b->s0.need(b->ctx, wq * 8))) return rc;
        if ((rc = ring_ntt(b->from, (uint64_t*)b->s0.p, batch, true, s, who, src))) return rc;
        coef = (const uint64_t*)b->s0.p;
    }
    // the P towers are produced where they belong in dst and transformed there (:1105-1106, :1113-1116)
    CUDA_TRY(launch_basis_switch(b->dev, dst, sq + sp, sq, coef, sq, 0, batch, logN, false, s));
    if ((rc = ring_ntt(b->to, dst + (size_t)sq * N, batch, false, s, who, nullptr, sq + sp))) return rc;
    if (format == 0) {
        CUDA_TRY(copy_towers(dst, sq + sp, 0, src, sq, 0, sq, batch, N, s));                   // :1119-1123
    } else {
        if ((rc = ring_ntt(b->from, dst, batch, false, s, who, src, sq + sp, sq))) return rc;  // :1124-1130
    }
    return FHE_OK;
}

// ApproxModDown in three steps, so that a composition can run the middle one over several polys at once and from a captured
// graph (fhe_keyswitch_hybrid): the tables, P part -> the correction term over Q in EVALUATION form, the final subtraction.
static int mod_down_tables(fhe_basis* b, uint64_t t, cudaStream_t s, BasisDev* out) {
    fhe_ring *P = b->from, *Q = b->to;                      // the basis switches P -> Q
    const uint32_t sq = b->dev.sp;
    int rc;
    if (!b->pinv_ready) {                                    // [P^-1]_{q_i}, rns-cryptoparameters.cpp:190-197
        std::vector<ulonglong4> k(sq);
        for (uint32_t i = 0; i < sq; i++) {
            uint64_t pm = prod_mod(P->q, (size_t)-1, Q->q[i]);
            if (pm == 0 || std::__gcd(pm, Q->q[i]) != 1) return fail(FHE_EINVAL, "fhe_approx_mod_down: P is not invertible modulo q_%u", i);
            uint64_t c = host::invmod(pm, Q->q[i]);
            k[i] = make_ulonglong4(c, host::shoup_companion(c, Q->q[i]), 0, 0);
        }
        if ((rc = upload(b->ctx, k, &b->d_pinv))) return rc;
        b->pinv_ready = true;
    }
    BasisDev bd = b->dev;
    if (t) {
        if (b->t_cached != t) {
            for (uint64_t pj : P->q)
                if (std::__gcd(t % pj, pj) != 1) return fail(FHE_EINVAL, "fhe_approx_mod_down: t has no inverse modulo %llu", (unsigned long long)pj);
            std::vector<ulonglong2> pre;
            std::vector<BasisTarget> tgt;
            basis_fill(b, t, pre, tgt);
            CUDA_TRY(cudaStreamSynchronize(s));              // an earlier launch may still read the cached tables
            if ((rc = upload(b->ctx, pre, &b->d_pre_t)) || (rc = upload(b->ctx, tgt, &b->d_tgt_t))) return rc;
            b->t_cached = t;
        }
        bd.pre = (const ulonglong2*)b->d_pre_t;
        bd.tgt = (const BasisTarget*)b->d_tgt_t;
    }
    *out = bd;
    return FHE_OK;
}

// sw[batch][sq][N] <- the P towers of src[batch][sq+sp][N] taken to COEFFICIENT form (:1150-1151), switched to Q
// (:1152-1159, :1170-1171) and back to EVALUATION (:1172); partP[batch][sp][N] is scratch
static int mod_down_inner(fhe_basis* b, const BasisDev& bd, uint64_t* sw, uint64_t* partP, const uint64_t* src, size_t batch, cudaStream_t s) {
    fhe_ring *P = b->from, *Q = b->to;
    const uint32_t N = P->dev.N, logN = P->dev.logN, sp = b->dev.sq, sq = b->dev.sp;
    int rc;
    if ((rc = ring_ntt(P, partP, batch, true, s, "fhe_approx_mod_down", src + (size_t)sq * N, sp, sq + sp))) return rc;
    CUDA_TRY(launch_basis_switch(bd, sw, sq, 0, partP, sp, 0, batch, logN, false, s));
    return ring_ntt(Q, sw, batch, false, s, "fhe_approx_mod_down");
}

// dst = (src's Q towers - sw) * [P^-1]_q (:1173)
static int mod_down_final(fhe_basis* b, uint64_t* dst, const uint64_t* src, const uint64_t* sw, size_t batch, cudaStream_t s,
                          bool accumulate = false) {
    const uint32_t sp = b->dev.sq, sq = b->dev.sp;
    CUDA_TRY(launch_tower_affine(accumulate ? 3 : 0, dst, src, sw, batch, b->from->dev.logN, sq, sq + sp, b->to->dev.mods,
                                 (const ulonglong4*)b->d_pinv, s));
    return FHE_OK;
}

// both halves of a key switch: one launch when the caller's two results are adjacent ([2][batch][sq][N]), else two
static int mod_down_final_pair(fhe_basis* b, uint64_t* ct0, uint64_t* ct1, const uint64_t* c01, const uint64_t* sw, size_t batch,
                               size_t item, cudaStream_t s, bool accumulate = false) {
    const size_t half = batch * b->dev.sp * (size_t)b->from->dev.N;        // words of one half over Q
    if (ct1 == ct0 + half) return mod_down_final(b, ct0, c01, sw, 2 * batch, s, accumulate);
    int rc = mod_down_final(b, ct0, c01, sw, batch, s, accumulate);
    return rc ? rc : mod_down_final(b, ct1, c01 + batch * item, sw + half, batch, s, accumulate);
}

extern "C" int fhe_approx_mod_down(fhe_basis* b, fhe_dptr dst_q, fhe_dptr src_qp, size_t batch, uint64_t t, int stream) {
    if (!b) return fail(FHE_EINVAL, "null basis");
    StreamRef s;
    int rc = ring_stream(b->from, stream, &s);
    if (rc) return rc;
    if (batch == 0) return FHE_OK;
    if (!dst_q || !src_qp) return fail(FHE_EINVAL, "fhe_approx_mod_down: null device pointer");
    if ((dst_q | src_qp) & 15) return fail(FHE_EINVAL, "fhe_approx_mod_down: device pointers must be 16-byte aligned");
    const uint32_t N = b->from->dev.N, sp = b->dev.sq, sq = b->dev.sp;
    BasisDev bd;
    if ((rc = mod_down_tables(b, t, s, &bd))) return rc;
    uint64_t* dst = (uint64_t*)(uintptr_t)dst_q;
    const uint64_t* src = (const uint64_t*)(uintptr_t)src_qp;
    const size_t wq = batch * sq * (size_t)N, wp = batch * sp * (size_t)N;
    if ((rc = b->s0.need(b->ctx, wp * 8)) || (rc = b->s1.need(b->ctx, wq * 8))) return rc;
    if ((rc = mod_down_inner(b, bd, (uint64_t*)b->s1.p, (uint64_t*)b->s0.p, src, batch, s))) return rc;
    return mod_down_final(b, dst, src, (const uint64_t*)b->s1.p, batch, s);
}

struct fhe_rescale {
    fhe_ctx* ctx = nullptr;
    fhe_ring* ring = nullptr;
    fhe_ring *rest = nullptr, *last = nullptr;     // views: towers [0, T-1) and [T-1, T)
    std::vector<uint64_t> qlql, qlinv;
    void *d_sw = nullptr, *d_k = nullptr, *d_kt = nullptr;
    uint64_t t_cached = 0;
    ulonglong2 pre_t{0, 0};
    Scratch s0, s1;
};

extern "C" int fhe_rescale_destroy(fhe_rescale* p) {
    GateHold gate_;
    if (!p) return FHE_OK;
    cudaSetDevice(p->ctx->device);
    cudaDeviceSynchronize();
    for (void* d : {p->d_sw, p->d_k, p->d_kt})
        if (d) arena_park(p->ctx, d);
    p->s0.drop(p->ctx);
    p->s1.drop(p->ctx);
    delete p->rest;
    delete p->last;
    delete p;
    return FHE_OK;
}

extern "C" int fhe_rescale_create(fhe_ring* ring, const uint64_t* qlql, const uint64_t* qlinv, fhe_rescale** out) {
    GateHold gate_;
    if (!ring || !out) return fail(FHE_EINVAL, "fhe_rescale_create: null argument");
    *out = nullptr;
    const uint32_t T = ring->dev.T;
    if (T < 2) return fail(FHE_EINVAL, "Removing last element of DCRTPoly object renders it invalid!");   // dcrtpoly-impl.h:722-723
    CUDA_TRY(cudaSetDevice(ring->ctx->device));
    auto* p = new fhe_rescale;
    p->ctx = ring->ctx; p->ring = ring;
    int rc;
    if ((rc = fhe_ring_view(ring, 0, T - 1, &p->rest)) || (rc = fhe_ring_view(ring, T - 1, 1, &p->last))) {
        fhe_rescale_destroy(p);
        return rc;
    }
    const uint32_t l = T - 1;
    const uint64_t ql = ring->q[l];
    p->qlql.resize(l); p->qlinv.resize(l);
    std::vector<ulonglong4> sw(l), k(l);
    for (uint32_t i = 0; i < l; i++) {
        const uint64_t qi = ring->q[i];
        if (std::__gcd(ql % qi, qi) != 1) { fhe_rescale_destroy(p); return fail(FHE_EINVAL, "fhe_rescale_create: q_%u and the last modulus are not coprime", i); }
        // Ql*[Ql^-1]_{ql} = 1 + m*ql, so floor(./ql) = m = -ql^-1 (mod q_i): q_i divides Ql
        p->qlinv[i] = qlinv ? qlinv[i] : host::invmod(ql % qi, qi);
        p->qlql[i] = qlql ? qlql[i] : (qi - host::invmod(ql % qi, qi)) % qi;
        const bool up = qi > ql;                     // NativeVectorT::SwitchModulus, mubintvecnat.cpp:112-135
        sw[i] = make_ulonglong4(ql >> 1, qi, up ? qi - ql : qi - (ql % qi), up ? 1 : 0);
        k[i] = make_ulonglong4(p->qlql[i] % qi, host::shoup_companion(p->qlql[i] % qi, qi), 0, 0);
    }
    if ((rc = upload(p->ctx, sw, &p->d_sw))) { fhe_rescale_destroy(p); return rc; }
    // k: {qlql, companion} for the spread; the affine tail wants {qlinv, companion, t, companion}
    std::vector<ulonglong4> kk(2 * l);
    for (uint32_t i = 0; i < l; i++) {
        kk[i] = k[i];
        kk[l + i] = make_ulonglong4(p->qlinv[i] % ring->q[i], host::shoup_companion(p->qlinv[i] % ring->q[i], ring->q[i]), 0, 0);
    }
    if ((rc = upload(p->ctx, kk, &p->d_k))) { fhe_rescale_destroy(p); return rc; }
    *out = p;
    return FHE_OK;
}

extern "C" int fhe_rescale_tables(const fhe_rescale* p, uint64_t* qlql, uint64_t* qlinv) {
    if (!p) return fail(FHE_EINVAL, "null rescale plan");
    if (qlql) memcpy(qlql, p->qlql.data(), p->qlql.size() * 8);
    if (qlinv) memcpy(qlinv, p->qlinv.data(), p->qlinv.size() * 8);
    return FHE_OK;
}

static int rescale_run(fhe_rescale* p, fhe_dptr dst_, fhe_dptr src_, size_t batch, uint64_t t, int format, int stream,
                       const char* who) {
    if (!p) return fail(FHE_EINVAL, "null rescale plan");
    StreamRef s;
    int rc = ring_stream(p->ring, stream, &s);
    if (rc) return rc;
    if (format != 0 && format != 1) return fail(FHE_EINVAL, "%s: format must be 0 or 1", who);
    if (batch == 0) return FHE_OK;
    if (!dst_ || !src_) return fail(FHE_EINVAL, "%s: null device pointer", who);
    if ((dst_ | src_) & 15) return fail(FHE_EINVAL, "%s: device pointers must be 16-byte aligned", who);
    const uint32_t N = p->ring->dev.N, logN = p->ring->dev.logN, T = p->ring->dev.T, l = T - 1;
    const uint64_t ql = p->ring->q[l];
    uint64_t* dst = (uint64_t*)(uintptr_t)dst_;
    const uint64_t* src = (const uint64_t*)(uintptr_t)src_;
    const ulonglong4* kspread = (const ulonglong4*)p->d_k;
    const ulonglong4* ktail = kspread + l;
    if (t) {                                        // bgvrns-cryptoparameters.cpp:91-99
        if (p->t_cached != t) {
            if (std::__gcd(t % ql, ql) != 1) return fail(FHE_EINVAL, "%s: t has no inverse modulo the last modulus", who);
            uint64_t negtinv = ql - host::invmod(t % ql, ql);
            std::vector<ulonglong4> kt(l);
            for (uint32_t i = 0; i < l; i++) {
                const uint64_t qi = p->ring->q[i], c = p->qlinv[i] % qi, tt = t % qi;
                kt[i] = make_ulonglong4(c, host::shoup_companion(c, qi), tt, host::shoup_companion(tt, qi));
            }
            CUDA_TRY(cudaStreamSynchronize(s));
            if ((rc = upload(p->ctx, kt, &p->d_kt))) return rc;
            p->pre_t = make_ulonglong2(negtinv, host::shoup_companion(negtinv, ql));
            p->t_cached = t;
        }
        ktail = (const ulonglong4*)p->d_kt;
    }
    if ((rc = p->s0.need(p->ctx, batch * (size_t)N * 8)) || (rc = p->s1.need(p->ctx, batch * l * (size_t)N * 8))) return rc;
    uint64_t* lastp = (uint64_t*)p->s0.p;
    uint64_t* tmp = (uint64_t*)p->s1.p;
    if (format == 0) {                                                                         // SetFormat(COEFFICIENT)
        if ((rc = ring_ntt(p->last, lastp, batch, true, s, who, src + (size_t)l * N, 1, T))) return rc;
    } else {
        CUDA_TRY(copy_towers(lastp, 1, 0, src, T, l, 1, batch, N, s));
    }
    CUDA_TRY(launch_rescale_spread(tmp, lastp, batch, logN, l, ql, (const ulonglong4*)p->d_sw, kspread, p->pre_t, t ? 1 : 0, s));
    if (format == 0 && (rc = ring_ntt(p->rest, tmp, batch, false, s, who))) return rc;
    CUDA_TRY(launch_tower_affine(t ? 2 : 1, dst, src, tmp, batch, logN, l, T, p->rest->dev.mods, ktail, s));
    if (format == 1 && !t && (rc = ring_ntt(p->rest, dst, batch, false, s, who))) return rc;   // dcrtpoly-impl.h:765-766
    return FHE_OK;
}

extern "C" int fhe_drop_last_element_and_scale(fhe_rescale* p, fhe_dptr dst, fhe_dptr src, size_t batch, int format, int stream) {
    return rescale_run(p, dst, src, batch, 0, format, stream, "fhe_drop_last_element_and_scale");
}
extern "C" int fhe_mod_reduce(fhe_rescale* p, fhe_dptr dst, fhe_dptr src, size_t batch, uint64_t t, int format, int stream) {
    if (t == 0) return fail(FHE_EINVAL, "fhe_mod_reduce: t must be positive");
    return rescale_run(p, dst, src, batch, t, format, stream, "fhe_mod_reduce");
}

// ---- BFV (HPS) set: exact basis switch, ExpandCRTBasis, ScaleAndRound ----------------------------------------
static int basis_exact_ready(fhe_basis* b) {
    if (b->dev.qinv) return FHE_OK;
    const auto& q = b->from->q;
    const auto& p = b->to->q;
    std::vector<double> qinv(q.size());
    for (size_t i = 0; i < q.size(); i++) qinv[i] = 1. / static_cast<double>(q[i]);       // bfvrns-cryptoparameters.cpp:270-273
    std::vector<uint64_t> aq((q.size() + 1) * p.size());
    for (size_t j = 0; j < p.size(); j++) {                                                 // :343-350
        uint64_t Qm = prod_mod(q, (size_t)-1, p[j]);
        for (size_t a = 0; a <= q.size(); a++) aq[a * p.size() + j] = host::mulmod(Qm, a % p[j], p[j]);
    }
    int rc;
    if ((rc = upload(b->ctx, qinv, &b->d_qinv)) || (rc = upload(b->ctx, aq, &b->d_alphaq))) return rc;
    b->dev.qinv = (const double*)b->d_qinv;
    b->dev.alphaq = (const uint64_t*)b->d_alphaq;
    return FHE_OK;
}

extern "C" int fhe_switch_crt_basis(fhe_basis* b, fhe_dptr dst, fhe_dptr src, size_t batch, int stream) {
    if (!b) return fail(FHE_EINVAL, "null basis");
    StreamRef s;
    int rc = ring_stream(b->from, stream, &s);
    if (rc) return rc;
    if (batch == 0) return FHE_OK;
    if (!dst || !src) return fail(FHE_EINVAL, "fhe_switch_crt_basis: null device pointer");
    if ((dst | src) & 15) return fail(FHE_EINVAL, "fhe_switch_crt_basis: device pointers must be 16-byte aligned");
    if ((rc = basis_exact_ready(b))) return rc;
    CUDA_TRY(launch_basis_switch(b->dev, (uint64_t*)(uintptr_t)dst, b->dev.sp, 0, (const uint64_t*)(uintptr_t)src, b->dev.sq, 0,
                                 batch, b->from->dev.logN, true, s));
    return FHE_OK;
}

// ExpandCRTBasis (reverse = false: towers [Q | P]) and ExpandCRTBasisReverseOrder (reverse = true: [P | Q])
static int expand_run(fhe_basis* b, fhe_dptr dst_, fhe_dptr src_q, size_t batch, int format, int result_format, int stream,
                      bool reverse, const char* who) {
    if (!b) return fail(FHE_EINVAL, "null basis");
    StreamRef s;
    int rc = ring_stream(b->from, stream, &s);
    if (rc) return rc;
    if ((format != 0 && format != 1) || (result_format != 0 && result_format != 1))
        return fail(FHE_EINVAL, "%s: formats must be 0 or 1", who);
    if (batch == 0) return FHE_OK;
    if (!dst_ || !src_q) return fail(FHE_EINVAL, "%s: null device pointer", who);
    if ((dst_ | src_q) & 15) return fail(FHE_EINVAL, "%s: device pointers must be 16-byte aligned", who);
    if ((rc = basis_exact_ready(b))) return rc;
    const uint32_t N = b->from->dev.N, logN = b->from->dev.logN, sq = b->dev.sq, sp = b->dev.sp;
    const uint32_t q0 = reverse ? sp : 0, p0 = reverse ? 0 : sq;                               // first tower of each part
    uint64_t* dst = (uint64_t*)(uintptr_t)dst_;
    const uint64_t* src = (const uint64_t*)(uintptr_t)src_q;
    const size_t wq = batch * sq * (size_t)N;
    const uint64_t* coef = src;
    if (format == 0) {                                                                         // dcrtpoly-impl.h:1321-1324
        if ((rc = b->s0.need(b->ctx, wq * 8))) return rc;
        if ((rc = ring_ntt(b->from, (uint64_t*)b->s0.p, batch, true, s, who, src))) return rc;
        coef = (const uint64_t*)b->s0.p;
    }
    CUDA_TRY(launch_basis_switch(b->dev, dst, sq + sp, p0, coef, sq, 0, batch, logN, true, s));          // :1326-1327
    if (result_format == 0 && (rc = ring_ntt(b->to, dst + (size_t)p0 * N, batch, false, s, who, nullptr, sq + sp))) return rc;  // :1338-1341
    if (result_format == 0 && format == 0) {
        CUDA_TRY(copy_towers(dst, sq + sp, q0, src, sq, 0, sq, batch, N, s));                  // :1346-1349
    } else if (result_format == 0) {
        if ((rc = ring_ntt(b->from, dst + (size_t)q0 * N, batch, false, s, who, src, sq + sp, sq))) return rc;  // :1350-1355
    } else {
        CUDA_TRY(copy_towers(dst, sq + sp, q0, coef, sq, 0, sq, batch, N, s));                 // COEFFICIENT result keeps coef
    }
    return FHE_OK;
}

extern "C" int fhe_expand_crt_basis(fhe_basis* b, fhe_dptr dst_qp, fhe_dptr src_q, size_t batch, int format, int result_format,
                                    int stream) {
    return expand_run(b, dst_qp, src_q, batch, format, result_format, stream, false, "fhe_expand_crt_basis");
}
extern "C" int fhe_expand_crt_basis_reverse_order(fhe_basis* b, fhe_dptr dst_pq, fhe_dptr src_q, size_t batch, int format,
                                                  int result_format, int stream) {
    return expand_run(b, dst_pq, src_q, batch, format, result_format, stream, true, "fhe_expand_crt_basis_reverse_order");
}

// DCRTPolyImpl::FastExpandCRTBasisPloverQ (dcrtpoly-impl.h:1413-1468): the Pl towers are produced where they belong in
// dst (:1420-1444, the approximate switch with the scheme's [-Pl*(Q/q_i)^-1]_{q_i} and [q_i^-1]_{p_j}), and the exact
// switch Pl -> Ql (:1448-1450) reads them there and fills the first towers.
extern "C" int fhe_fast_expand_crt_basis_pl_over_q(fhe_basis* q2pl, fhe_basis* pl2ql, fhe_dptr dst_qlpl, fhe_dptr src_q,
                                                   size_t batch, int stream) {
    const char* who = "fhe_fast_expand_crt_basis_pl_over_q";
    if (!q2pl || !pl2ql) return fail(FHE_EINVAL, "null basis");
    if (q2pl->ctx != pl2ql->ctx || q2pl->from->dev.N != pl2ql->from->dev.N)
        return fail(FHE_EINVAL, "%s: the two bases must share context and ring dimension", who);
    if (q2pl->to->q != pl2ql->from->q) return fail(FHE_EINVAL, "%s: the second basis must start at the first one's target moduli", who);
    StreamRef s;
    int rc = ring_stream(q2pl->from, stream, &s);
    if (rc) return rc;
    if (batch == 0) return FHE_OK;
    if (!dst_qlpl || !src_q) return fail(FHE_EINVAL, "%s: null device pointer", who);
    if ((dst_qlpl | src_q) & 15) return fail(FHE_EINVAL, "%s: device pointers must be 16-byte aligned", who);
    if (dst_qlpl == src_q) return fail(FHE_EINVAL, "%s: dst must not be src (different tower counts)", who);
    if (pl2ql->to->dev.T != q2pl->dev.sp) return fail(FHE_EINVAL, "%s: Ql and Pl must have the same number of towers", who);
    if ((rc = basis_exact_ready(pl2ql))) return rc;
    const uint32_t logN = q2pl->from->dev.logN, sq = q2pl->dev.sq, spl = q2pl->dev.sp, sql = pl2ql->dev.sp;
    uint64_t* dst = (uint64_t*)(uintptr_t)dst_qlpl;
    CUDA_TRY(launch_basis_switch(q2pl->dev, dst, sql + spl, sql, (const uint64_t*)(uintptr_t)src_q, sq, 0, batch, logN, false, s));
    CUDA_TRY(launch_basis_switch(pl2ql->dev, dst, sql + spl, 0, dst, sql + spl, sql, batch, logN, true, s));
    return FHE_OK;
}

struct fhe_tower_scale {
    fhe_ctx* ctx = nullptr;
    fhe_ring* ring = nullptr;
    std::vector<uint64_t> c;
    void* d_k = nullptr;
};

extern "C" int fhe_tower_scale_destroy(fhe_tower_scale* p) {
    GateHold gate_;
    if (!p) return FHE_OK;
    cudaSetDevice(p->ctx->device);
    cudaDeviceSynchronize();
    if (p->d_k) arena_park(p->ctx, p->d_k);
    delete p;
    return FHE_OK;
}

extern "C" int fhe_tower_scale_create(fhe_ring* ring, const uint64_t* c, fhe_tower_scale** out) {
    GateHold gate_;
    if (!ring || !c || !out) return fail(FHE_EINVAL, "fhe_tower_scale_create: null argument");
    *out = nullptr;
    CUDA_TRY(cudaSetDevice(ring->ctx->device));
    const uint32_t T = ring->dev.T;
    std::vector<ulonglong4> k(T);
    for (uint32_t i = 0; i < T; i++) {
        const uint64_t qi = ring->q[i];
        if (c[i] >= qi) return fail(FHE_EINVAL, "fhe_tower_scale_create: constant %u is not reduced modulo its tower", i);
        k[i] = make_ulonglong4(c[i], host::shoup_companion(c[i], qi), qi * ((1ull << 63) / qi), 0);   // {c, c', M_i}
    }
    auto* p = new fhe_tower_scale;
    p->ctx = ring->ctx; p->ring = ring;
    p->c.assign(c, c + T);
    int rc = upload(p->ctx, k, &p->d_k);
    if (rc) { fhe_tower_scale_destroy(p); return rc; }
    *out = p;
    return FHE_OK;
}

static int tower_scale_run(fhe_tower_scale* p, int mode, uint32_t sl, fhe_dptr dst, fhe_dptr src, size_t batch, int stream,
                           const char* who) {
    if (!p) return fail(FHE_EINVAL, "null tower-scale plan");
    StreamRef s;
    int rc = ring_stream(p->ring, stream, &s);
    if (rc) return rc;
    const uint32_t T = p->ring->dev.T;
    if (mode == 0 && (sl == 0 || sl > T)) return fail(FHE_EINVAL, "%s: size_ql must be in [1, %u]", who, T);
    if (batch == 0) return FHE_OK;
    if (!dst || !src) return fail(FHE_EINVAL, "%s: null device pointer", who);
    if ((dst | src) & 15) return fail(FHE_EINVAL, "%s: device pointers must be 16-byte aligned", who);
    if (dst == src && !(mode == 0 && sl == T))
        return fail(FHE_EINVAL, "%s: dst must not be src (different tower counts)", who);
    CUDA_TRY(launch_tower_scale(mode, (uint64_t*)(uintptr_t)dst, (const uint64_t*)(uintptr_t)src, batch, p->ring->dev.logN, sl, T,
                                p->ring->dev.mods, (const ulonglong4*)p->d_k, s));
    return FHE_OK;
}

extern "C" int fhe_expand_crt_basis_ql_hat(fhe_tower_scale* p, uint32_t size_ql, fhe_dptr dst_q, fhe_dptr src_ql, size_t batch,
                                           int stream) {
    return tower_scale_run(p, 0, size_ql, dst_q, src_ql, batch, stream, "fhe_expand_crt_basis_ql_hat");
}
extern "C" int fhe_dcrt_times_towers(fhe_tower_scale* p, fhe_dptr dst, fhe_dptr src, size_t batch, int stream) {
    if (!p) return fail(FHE_EINVAL, "null tower-scale plan");
    return tower_scale_run(p, 0, p->ring->dev.T, dst, src, batch, stream, "fhe_dcrt_times_towers");
}
extern "C" int fhe_scale_and_round_p_over_q(fhe_tower_scale* p, fhe_dptr dst_q, fhe_dptr src_qp, size_t batch, int stream) {
    return tower_scale_run(p, 1, 0, dst_q, src_qp, batch, stream, "fhe_scale_and_round_p_over_q");
}

struct fhe_scale_round {
    fhe_ctx* ctx = nullptr;
    fhe_ring* ring = nullptr;
    ScaleDev dev{};
    void *d_frac = nullptr, *d_modo = nullptr, *d_tgt = nullptr;
};

extern "C" int fhe_scale_round_destroy(fhe_scale_round* p) {
    GateHold gate_;
    if (!p) return FHE_OK;
    cudaSetDevice(p->ctx->device);
    cudaDeviceSynchronize();
    for (void* d : {p->d_frac, p->d_modo, p->d_tgt})
        if (d) arena_park(p->ctx, d);
    delete p;
    return FHE_OK;
}

extern "C" int fhe_scale_round_create(fhe_ring* ring, uint32_t size_out, int out_last, const double* frac, const uint64_t* modo,
                                      fhe_scale_round** out) {
    GateHold gate_;
    if (!ring || !frac || !modo || !out) return fail(FHE_EINVAL, "fhe_scale_round_create: null argument");
    *out = nullptr;
    const uint32_t T = ring->dev.T;
    if (size_out == 0 || size_out >= T) return fail(FHE_EINVAL, "fhe_scale_round_create: size_out must be in [1, towers)");
    const uint32_t sI = T - size_out, sO = size_out;
    CUDA_TRY(cudaSetDevice(ring->ctx->device));
    auto* p = new fhe_scale_round;
    p->ctx = ring->ctx; p->ring = ring;
    std::vector<double> f(frac, frac + sI);
    std::vector<uint64_t> m(modo, modo + (size_t)sO * (sI + 1));
    std::vector<BasisTarget> tgt(sO);
    const uint32_t out0 = out_last ? sI : 0;
    for (uint32_t j = 0; j < sO; j++) {
        BasisTarget g{};
        g.p = ring->q[out0 + j];
        if ((g.p & (g.p - 1)) == 0) { delete p; return fail(FHE_EINVAL, "fhe_scale_round_create: output modulus is a power of two"); }
        basis_target_set(g, g.p);
        tgt[j] = g;
    }
    int rc;
    if ((rc = upload(p->ctx, f, &p->d_frac)) || (rc = upload(p->ctx, m, &p->d_modo)) || (rc = upload(p->ctx, tgt, &p->d_tgt))) {
        fhe_scale_round_destroy(p);
        return rc;
    }
    p->dev.sI = sI; p->dev.sO = sO; p->dev.in0 = out_last ? 0 : sO; p->dev.out0 = out0;
    p->dev.frac = (const double*)p->d_frac;
    p->dev.modo = (const uint64_t*)p->d_modo;
    p->dev.tgt = (const BasisTarget*)p->d_tgt;
    p->dev.narrow = !basis_force_wide() && all_below_2_60(ring->q) && all_below_2_60(m);
    *out = p;
    return FHE_OK;
}

extern "C" int fhe_scale_and_round(fhe_scale_round* p, fhe_dptr dst, fhe_dptr src, size_t batch, int stream) {
    if (!p) return fail(FHE_EINVAL, "null scale-and-round plan");
    StreamRef s;
    int rc = ring_stream(p->ring, stream, &s);
    if (rc) return rc;
    if (batch == 0) return FHE_OK;
    if (!dst || !src) return fail(FHE_EINVAL, "fhe_scale_and_round: null device pointer");
    if ((dst | src) & 15) return fail(FHE_EINVAL, "fhe_scale_and_round: device pointers must be 16-byte aligned");
    CUDA_TRY(launch_scale_round(p->dev, (uint64_t*)(uintptr_t)dst, (const uint64_t*)(uintptr_t)src, batch, p->ring->dev.logN, s));
    return FHE_OK;
}

// ---------------------------------------------------------------------------------------------
// host-operand pipelines: chunk the batch, rotate over the stream pool so that the H2D copy of
// chunk k+1, the kernel of chunk k and the D2H copy of chunk k-1 overlap.
// ---------------------------------------------------------------------------------------------
namespace {
struct HostPipe {
    fhe_ring* r;
    size_t item_words, chunk_items, nslots;
    std::vector<void*> bufA, bufB;
    HostPipe(fhe_ring* ring, bool two_operands) : r(ring) {
        item_words = (size_t)r->dev.T * r->dev.N;
        size_t item_bytes = item_words * 8;
        chunk_items = std::max<size_t>(1, r->ctx->chunk_bytes / item_bytes);
        nslots = std::min<size_t>(r->ctx->streams.size(), 4);
        (void)two_operands;
    }
    int reserve(size_t batch, bool two_operands) {
        chunk_items = std::min(chunk_items, std::max<size_t>(1, batch));
        size_t nchunks = (batch + chunk_items - 1) / chunk_items;
        nslots = std::max<size_t>(1, std::min(nslots, nchunks));
        for (size_t i = 0; i < nslots; i++) {
            void* p = nullptr;
            int rc = r->ctx->arena.alloc(chunk_items * item_words * 8, &p);
            if (rc) return rc;
            bufA.push_back(p);
            if (two_operands) {
                rc = r->ctx->arena.alloc(chunk_items * item_words * 8, &p);
                if (rc) return rc;
                bufB.push_back(p);
            }
        }
        return FHE_OK;
    }
    ~HostPipe() {
        for (void* p : bufA) arena_park(r->ctx, p);
        for (void* p : bufB) arena_park(r->ctx, p);
    }
};
}  // namespace

extern "C" int fhe_dcrt_multi_op_host(fhe_ring* r, const int* ops, int nops, uint64_t* const* dsts,
                                      const uint64_t* a, const uint64_t* b, size_t batch) {
    GateHold gate_;
    if (!r) return fail(FHE_EINVAL, "null ring");
    if (!ops || !dsts || nops < 1 || nops > 4) return fail(FHE_EINVAL, "fhe_dcrt_multi_op_host: need 1..4 ops");
    for (int i = 0; i < nops; i++) {
        if (ops[i] < FHE_OP_ADD || ops[i] > FHE_OP_MUL) return fail(FHE_EINVAL, "fhe_dcrt_multi_op_host: unknown op %d", ops[i]);
        if (!dsts[i]) return fail(FHE_EINVAL, "fhe_dcrt_multi_op_host: null destination");
    }
    if (batch == 0) return FHE_OK;
    if (!a || !b) return fail(FHE_EINVAL, "fhe_dcrt_multi_op_host: null host pointer");
    CUDA_TRY(cudaSetDevice(r->ctx->device));
    // slot = {a, b, result 1..nops-1}; the last op writes over a
    HostPipe hp(r, true);
    int rc = hp.reserve(batch, true);
    if (rc) return rc;
    std::vector<void*> extra;
    struct Guard { fhe_ring* r; std::vector<void*>& v; ~Guard() { for (void* p : v) arena_park(r->ctx, p); } } guard{r, extra};
    for (size_t i = 0; i < hp.nslots * (size_t)(nops - 1); i++) {
        void* p = nullptr;
        if ((rc = r->ctx->arena.alloc(hp.chunk_items * hp.item_words * 8, &p))) return rc;
        extra.push_back(p);
    }
    size_t k = 0;
    for (size_t b0 = 0; b0 < batch; b0 += hp.chunk_items, k++) {
        size_t items = std::min(hp.chunk_items, batch - b0), words = items * hp.item_words;
        size_t slot = k % hp.nslots;
        StreamRef s(r->ctx, (int)slot);
        uint64_t* da = (uint64_t*)hp.bufA[slot];
        uint64_t* db = (uint64_t*)hp.bufB[slot];
        CUDA_TRY(cudaMemcpyAsync(da, a + b0 * hp.item_words, words * 8, cudaMemcpyHostToDevice, s));
        CUDA_TRY(cudaMemcpyAsync(db, b + b0 * hp.item_words, words * 8, cudaMemcpyHostToDevice, s));
        for (int i = 0; i < nops; i++) {
            uint64_t* dd = (i == nops - 1) ? da : (uint64_t*)extra[slot * (nops - 1) + i];
            CUDA_TRY(launch_elementwise(r->dev, ops[i], dd, da, db, items, s));
            CUDA_TRY(cudaMemcpyAsync(dsts[i] + b0 * hp.item_words, dd, words * 8, cudaMemcpyDeviceToHost, s));
        }
    }
    for (size_t i = 0; i < hp.nslots; i++) CUDA_TRY(cudaStreamSynchronize(r->ctx->streams[i]));
    return FHE_OK;
}

extern "C" int fhe_dcrt_chain_host(fhe_ring* r, const int* ops, int nops, uint64_t* a, const uint64_t* b, size_t batch) {
    GateHold gate_;
    if (!r) return fail(FHE_EINVAL, "null ring");
    if (!ops || nops < 1 || nops > 8) return fail(FHE_EINVAL, "fhe_dcrt_chain_host: need 1..8 ops");
    bool any_ntt = false, any_ew = false;
    for (int i = 0; i < nops; i++) {
        if (ops[i] < FHE_OP_ADD || ops[i] > FHE_OP_INTT) return fail(FHE_EINVAL, "fhe_dcrt_chain_host: unknown op %d", ops[i]);
        any_ntt |= ops[i] >= FHE_OP_NTT;
        any_ew |= ops[i] <= FHE_OP_MUL;
    }
    if (batch == 0) return FHE_OK;
    if (!a || (any_ew && !b)) return fail(FHE_EINVAL, "fhe_dcrt_chain_host: null host pointer");
    const bool transform = any_ntt && !r->ntt_identity && r->dev.N > 1;   // psi in {0,1}: SwitchFormat leaves the data alone
    if (transform && !r->has_ntt) return fail(FHE_EINVAL, "fhe_dcrt_chain_host: ring was created without roots of unity");
    CUDA_TRY(cudaSetDevice(r->ctx->device));
    HostPipe hp(r, any_ew);
    int rc = hp.reserve(batch, any_ew);
    if (rc) return rc;
    size_t k = 0;
    for (size_t b0 = 0; b0 < batch; b0 += hp.chunk_items, k++) {
        size_t items = std::min(hp.chunk_items, batch - b0), words = items * hp.item_words;
        size_t slot = k % hp.nslots;
        StreamRef s(r->ctx, (int)slot);
        uint64_t* da = (uint64_t*)hp.bufA[slot];
        uint64_t* db = any_ew ? (uint64_t*)hp.bufB[slot] : nullptr;
        CUDA_TRY(cudaMemcpyAsync(da, a + b0 * hp.item_words, words * 8, cudaMemcpyHostToDevice, s));
        if (any_ew) CUDA_TRY(cudaMemcpyAsync(db, b + b0 * hp.item_words, words * 8, cudaMemcpyHostToDevice, s));
        for (int i = 0; i < nops; i++) {
            if (ops[i] <= FHE_OP_MUL) CUDA_TRY(launch_elementwise(r->dev, ops[i], da, da, db, items, s));
            else if (transform) CUDA_TRY(launch_ntt(r->dev, da, items, ops[i] == FHE_OP_INTT, s, 0));
        }
        CUDA_TRY(cudaMemcpyAsync(a + b0 * hp.item_words, da, words * 8, cudaMemcpyDeviceToHost, s));
    }
    for (size_t i = 0; i < hp.nslots; i++) CUDA_TRY(cudaStreamSynchronize(r->ctx->streams[i]));
    return FHE_OK;
}

extern "C" int fhe_dcrt_op_host(fhe_ring* r, int op, uint64_t* dst, const uint64_t* a, const uint64_t* b, size_t batch) {
    uint64_t* d[1] = {dst};
    return fhe_dcrt_multi_op_host(r, &op, 1, d, a, b, batch);
}

extern "C" int fhe_ntt_host(fhe_ring* r, uint64_t* x, size_t batch, int inverse) {
    GateHold gate_;
    if (!r) return fail(FHE_EINVAL, "null ring");
    if (r->ntt_identity || batch == 0 || r->dev.N == 1) return FHE_OK;
    if (!r->has_ntt) return fail(FHE_EINVAL, "fhe_ntt_host: ring was created without roots of unity");
    if (!x) return fail(FHE_EINVAL, "fhe_ntt_host: null host pointer");
    CUDA_TRY(cudaSetDevice(r->ctx->device));
    HostPipe hp(r, false);
    int rc = hp.reserve(batch, false);
    if (rc) return rc;
    size_t k = 0;
    for (size_t b0 = 0; b0 < batch; b0 += hp.chunk_items, k++) {
        size_t items = std::min(hp.chunk_items, batch - b0), words = items * hp.item_words;
        size_t slot = k % hp.nslots;
        StreamRef s(r->ctx, (int)slot);
        uint64_t* d = (uint64_t*)hp.bufA[slot];
        CUDA_TRY(cudaMemcpyAsync(d, x + b0 * hp.item_words, words * 8, cudaMemcpyHostToDevice, s));
        CUDA_TRY(launch_ntt(r->dev, d, items, inverse != 0, s, 0));
        CUDA_TRY(cudaMemcpyAsync(x + b0 * hp.item_words, d, words * 8, cudaMemcpyDeviceToHost, s));
    }
    for (size_t i = 0; i < hp.nslots; i++) CUDA_TRY(cudaStreamSynchronize(r->ctx->streams[i]));
    return FHE_OK;
}

// ---------------------------------------------------------------------------------------------
// wire format: Serial::Serialize / Deserialize (SerType::BINARY) of device-resident polys (serial.h)
// ---------------------------------------------------------------------------------------------
static const serial::Layout& ring_layout(fhe_ring* r) {
    std::lock_guard<std::mutex> g(r->layout_mu);
    if (!r->layout) r->layout.reset(new serial::Layout(serial::make_layout(r->dev.N, r->q, r->psi)));
    return *r->layout;
}

// towers 1.. of the canonical stream are equally spaced, so one byte-granular 2-D copy moves them all
static cudaError_t copy_payload(uint8_t* blob, const std::vector<size_t>& off, uint64_t* dev, uint32_t T, uint32_t N, bool to_host,
                                cudaStream_t s) {
    const size_t row = (size_t)N * 8;
    auto one = [&](uint32_t t) {
        return to_host ? cudaMemcpyAsync(blob + off[t], dev + (size_t)t * N, row, cudaMemcpyDeviceToHost, s)
                       : cudaMemcpyAsync(dev + (size_t)t * N, blob + off[t], row, cudaMemcpyHostToDevice, s);
    };
    cudaError_t e = one(0);
    if (e != cudaSuccess || T == 1) return e;
    bool even = true;
    for (uint32_t t = 2; t < T; t++) even = even && (off[t] - off[t - 1] == off[2] - off[1]);
    if (T == 2 || !even) {
        for (uint32_t t = 1; t < T && e == cudaSuccess; t++) e = one(t);
        return e;
    }
    const size_t pitch = off[2] - off[1];
    return to_host ? cudaMemcpy2DAsync(blob + off[1], pitch, dev + N, row, row, T - 1, cudaMemcpyDeviceToHost, s)
                   : cudaMemcpy2DAsync(dev + N, row, blob + off[1], pitch, row, T - 1, cudaMemcpyHostToDevice, s);
}

// host-only views of the format (no device needed): the header bytes around the payload holes, and the checker
extern "C" size_t fhe_serial_template(uint32_t N, uint32_t T, const uint64_t* q, const uint64_t* psi, int format, void* blob,
                                      size_t cap, size_t* data_off) {
    if (!q || T == 0 || N == 0 || (format != 0 && format != 1)) { fail(FHE_EINVAL, "fhe_serial_template: bad argument"); return 0; }
    std::vector<uint64_t> vq(q, q + T), vpsi;
    if (psi) vpsi.assign(psi, psi + T);
    serial::Layout L = serial::make_layout(N, vq, vpsi);
    if (data_off) std::copy(L.data_off.begin(), L.data_off.end(), data_off);
    if (blob && cap >= L.bytes) serial::write_headers(L, (uint8_t*)blob, (uint32_t)format);
    return L.bytes;
}
extern "C" int fhe_serial_parse(uint32_t N, uint32_t T, const uint64_t* q, const uint64_t* psi, const void* blob, size_t len,
                                size_t* data_off, int* format, size_t* consumed) {
    if (!q || !blob || T == 0) return fail(FHE_EINVAL, "fhe_serial_parse: bad argument");
    std::vector<uint64_t> vq(q, q + T), vpsi;
    if (psi) vpsi.assign(psi, psi + T);
    serial::Parsed ps;
    std::string err;
    if (!serial::parse((const uint8_t*)blob, len, N, vq, vpsi, ps, err)) return fail(FHE_EINVAL, "fhe_serial_parse: %s", err.c_str());
    if (data_off) std::copy(ps.data_off.begin(), ps.data_off.end(), data_off);
    if (format) *format = (int)ps.format;
    if (consumed) *consumed = ps.consumed;
    return FHE_OK;
}

extern "C" size_t fhe_serial_bytes(fhe_ring* r) { return r ? ring_layout(r).bytes : 0; }

extern "C" int fhe_serialize(fhe_ring* r, void* dst_host, fhe_dptr src, size_t batch, int format, int stream) {
    StreamRef s;
    int rc = ring_stream(r, stream, &s);
    if (rc) return rc;
    if (format != 0 && format != 1) return fail(FHE_EINVAL, "fhe_serialize: format must be 0 or 1");
    if (batch == 0) return FHE_OK;
    if (!dst_host || !src) return fail(FHE_EINVAL, "fhe_serialize: null pointer");
    const serial::Layout& L = ring_layout(r);
    const uint32_t N = r->dev.N, T = r->dev.T;
    uint8_t* out = (uint8_t*)dst_host;
    uint64_t* dev = (uint64_t*)(uintptr_t)src;
    for (size_t b = 0; b < batch; b++) {
        serial::write_headers(L, out + b * L.bytes, (uint32_t)format);
        CUDA_TRY(copy_payload(out + b * L.bytes, L.data_off, dev + b * (size_t)T * N, T, N, true, s));
    }
    return FHE_OK;
}

extern "C" int fhe_deserialize(fhe_ring* r, fhe_dptr dst, const void* src_host, size_t len, size_t batch, int* format,
                               size_t* consumed, int stream) {
    StreamRef s;
    int rc = ring_stream(r, stream, &s);
    if (rc) return rc;
    if (consumed) *consumed = 0;
    if (batch == 0) return FHE_OK;
    if (!dst || !src_host) return fail(FHE_EINVAL, "fhe_deserialize: null pointer");
    const uint32_t N = r->dev.N, T = r->dev.T;
    const uint8_t* in = (const uint8_t*)src_host;
    uint64_t* dev = (uint64_t*)(uintptr_t)dst;
    // every stream is checked before the first byte moves, so a bad batch leaves dst untouched
    std::vector<serial::Parsed> ps(batch);
    size_t pos = 0;
    for (size_t b = 0; b < batch; b++) {
        std::string err;
        if (!serial::parse(in + pos, len - pos, N, r->q, r->psi, ps[b], err))
            return fail(FHE_EINVAL, "fhe_deserialize: poly %zu: %s", b, err.c_str());
        if (ps[b].format != ps[0].format) return fail(FHE_EINVAL, "fhe_deserialize: poly %zu: format differs from poly 0's", b);
        for (auto& o : ps[b].data_off) o += pos;
        pos += ps[b].consumed;
    }
    for (size_t b = 0; b < batch; b++)
        CUDA_TRY(copy_payload(const_cast<uint8_t*>(in), ps[b].data_off, dev + b * (size_t)T * N, T, N, false, s));
    if (format) *format = (int)ps[0].format;
    if (consumed) *consumed = pos;
    return FHE_OK;
}

// ---------------------------------------------------------------------------------------------
// diagnostics
// ---------------------------------------------------------------------------------------------
extern "C" int fhe_imad_probe(fhe_ctx* ctx, int blocks, int threads, int iters, float* elapsed_ms) {
    GateHold gate_;
    if (!ctx) return fail(FHE_EINVAL, "null context");
    if (blocks < 1 || threads < 1 || threads > 1024 || iters < 1) return fail(FHE_EINVAL, "fhe_imad_probe: bad launch shape");
    CUDA_TRY(cudaSetDevice(ctx->device));
    void* out = nullptr;
    int rc = ctx->arena.alloc(256, &out);
    if (rc) return rc;
    StreamRef s(ctx, 0);
    cudaError_t e = launch_imad_probe((uint64_t*)out, blocks, threads, 16, s);   // warm-up
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) { arena_park(ctx, out); return fail(FHE_ECUDA, "imad probe: %s", cudaGetErrorString(e)); }
    rc = fhe_timer_start(ctx, 0);
    if (!rc) {
        e = launch_imad_probe((uint64_t*)out, blocks, threads, iters, s);
        if (e != cudaSuccess) rc = fail(FHE_ECUDA, "imad probe: %s", cudaGetErrorString(e));
    }
    if (!rc) rc = fhe_timer_stop(ctx, 0, elapsed_ms);
    arena_park(ctx, out);
    return rc;
}

// ---------------------------------------------------------------------------------------------
// BFV decryption: DCRTPolyImpl::ScaleAndRound(t, ...) -> NativePoly (dcrtpoly-impl.h:1537-1816)
// ---------------------------------------------------------------------------------------------
struct fhe_decrypt_scale {
    fhe_ctx* ctx = nullptr;
    fhe_ring* ring = nullptr;
    DecryptDev dev{};
    std::vector<uint64_t> modt, bmodt;
    std::vector<double> frac, bfrac;
    void *d_k = nullptr, *d_f = nullptr;
};

extern "C" int fhe_decrypt_scale_destroy(fhe_decrypt_scale* p) {
    GateHold gate_;
    if (!p) return FHE_OK;
    cudaSetDevice(p->ctx->device);
    if (p->d_k) arena_park(p->ctx, p->d_k);
    if (p->d_f) arena_park(p->ctx, p->d_f);
    delete p;
    return FHE_OK;
}

extern "C" int fhe_decrypt_scale_create(fhe_ring* ring, uint64_t t, const uint64_t* modt, const uint64_t* b_modt,
                                        const double* frac, const double* b_frac, fhe_decrypt_scale** out) {
    GateHold gate_;
    if (!ring || !out) return fail(FHE_EINVAL, "fhe_decrypt_scale_create: null argument");
    *out = nullptr;
    if (t < 2 || t >= (1ull << 60)) return fail(FHE_EINVAL, "fhe_decrypt_scale_create: plaintext modulus %llu not in [2, 2^60)", (unsigned long long)t);
    const auto& q = ring->q;
    const uint32_t T = (uint32_t)q.size();
    if ((modt == nullptr) != (frac == nullptr) || (b_modt == nullptr) != (b_frac == nullptr))
        return fail(FHE_EINVAL, "fhe_decrypt_scale_create: pass the integer and the fractional table of a pair together");
    // the reference's branch (dcrtpoly-impl.h:1546-1556, 1628-1631): bit lengths of q_0, t and the tower count
    const unsigned qMSB = host::msb(q[0]), tMSB = host::msb(t), sMSB = host::msb(T);
    auto* p = new fhe_decrypt_scale;
    p->ctx = ring->ctx; p->ring = ring;
    DecryptDev& d = p->dev;
    d.T = T; d.t = t;
    d.pow2 = (t & (t - 1)) == 0;
    d.split = !(qMSB + sMSB < 52);
    d.hbits = qMSB >> 1;
    const unsigned width = (d.split ? d.hbits : qMSB) + tMSB + sMSB;
    d.reduce = d.pow2 ? !(width < (d.split ? 62u : 63u)) : !(width < 52u);
    d.td = (double)t;
    d.tinv = 1. / d.td;
    if (d.split && modt && !b_modt) {
        delete p;
        return fail(FHE_EINVAL, "fhe_decrypt_scale_create: these moduli take the split branch (qMSB + sizeQMSB >= 52): the B tables are needed");
    }
    p->modt.assign(T, 0); p->bmodt.assign(T, 0); p->frac.assign(T, 0.0); p->bfrac.assign(T, 0.0);
    for (uint32_t i = 0; i < T; i++) {
        if (modt) {
            p->modt[i] = modt[i]; p->frac[i] = frac[i];
            if (b_modt) { p->bmodt[i] = b_modt[i]; p->bfrac[i] = b_frac[i]; }
            continue;
        }
        // bfvrns-cryptoparameters.cpp:447-494 with 128-bit integers: X = [(Q/q_i)^-1]_{q_i} * t = D * q_i + R
        const uint64_t qi = q[i], h = prod_mod(q, i, qi);
        if (h == 0 || std::__gcd(h, qi) != 1) { delete p; return fail(FHE_EINVAL, "fhe_decrypt_scale_create: moduli are not pairwise coprime (tower %u)", i); }
        const host::u128 X = (host::u128)host::invmod(h, qi) * t;
        const host::u128 D = X / qi;
        const uint64_t R = (uint64_t)(X % qi);
        p->modt[i] = (uint64_t)(D % t);
        p->frac[i] = static_cast<double>((int64_t)R) / static_cast<double>((int64_t)qi);
        if (d.split) {                               // (X << hbits) = (D << hbits) * q_i + (R << hbits)
            const host::u128 RB = (host::u128)R << d.hbits;
            const uint64_t two_h = (uint64_t)(((host::u128)1 << d.hbits) % t);
            p->bmodt[i] = (uint64_t)(((host::u128)(uint64_t)(D % t) * two_h + (RB / qi) % t) % t);
            p->bfrac[i] = static_cast<double>((int64_t)(uint64_t)(RB % qi)) / static_cast<double>((int64_t)qi);
        }
    }
    for (uint32_t i = 0; i < T; i++)
        if (p->modt[i] >= t || p->bmodt[i] >= t) { delete p; return fail(FHE_EINVAL, "fhe_decrypt_scale_create: a table entry is not reduced modulo t"); }
    std::vector<ulonglong4> k(T);
    std::vector<double2> f(T);
    for (uint32_t i = 0; i < T; i++) {
        k[i] = make_ulonglong4(p->modt[i], host::shoup_companion(p->modt[i], t), p->bmodt[i], host::shoup_companion(p->bmodt[i], t));
        f[i] = make_double2(p->frac[i], p->bfrac[i]);
    }
    CUDA_TRY(cudaSetDevice(p->ctx->device));
    int rc;
    if ((rc = upload(p->ctx, k, &p->d_k)) || (rc = upload(p->ctx, f, &p->d_f))) { fhe_decrypt_scale_destroy(p); return rc; }
    d.k = (const ulonglong4*)p->d_k;
    d.f = (const double2*)p->d_f;
    *out = p;
    return FHE_OK;
}

extern "C" int fhe_decrypt_scale_tables(const fhe_decrypt_scale* p, uint64_t* modt, uint64_t* b_modt, double* frac, double* b_frac,
                                        int* branch) {
    if (!p) return fail(FHE_EINVAL, "null plan");
    const size_t T = p->modt.size();
    if (modt) memcpy(modt, p->modt.data(), T * 8);
    if (b_modt) memcpy(b_modt, p->bmodt.data(), T * 8);
    if (frac) memcpy(frac, p->frac.data(), T * 8);
    if (b_frac) memcpy(b_frac, p->bfrac.data(), T * 8);
    if (branch) *branch = (p->dev.pow2 ? 4 : 0) | (p->dev.split ? 2 : 0) | (p->dev.reduce ? 1 : 0);
    return FHE_OK;
}

extern "C" int fhe_scale_and_round_native(fhe_decrypt_scale* p, fhe_dptr dst, fhe_dptr src, size_t batch, int stream) {
    if (!p) return fail(FHE_EINVAL, "null plan");
    StreamRef s;
    int rc = ring_stream(p->ring, stream, &s);
    if (rc) return rc;
    if (batch == 0) return FHE_OK;
    if (!dst || !src) return fail(FHE_EINVAL, "fhe_scale_and_round_native: null device pointer");
    if ((dst | src) & 15) return fail(FHE_EINVAL, "fhe_scale_and_round_native: device pointers must be 16-byte aligned");
    CUDA_TRY(launch_decrypt_scale(p->dev, (uint64_t*)(uintptr_t)dst, (const uint64_t*)(uintptr_t)src, batch, p->ring->dev.logN, s));
    return FHE_OK;
}

// ---------------------------------------------------------------------------------------------
// hybrid key switching as one device composition (KeySwitchHYBRID, src/pke/lib/keyswitch/keyswitch-hybrid.cpp:324-480)
// ---------------------------------------------------------------------------------------------
struct fhe_keyswitch {
    fhe_ctx* ctx = nullptr;
    fhe_ring* qlp = nullptr;                      // caller's ring over the towers of Ql followed by P
    uint32_t size_ql = 0, size_p = 0, alpha = 0;
    fhe_ring *ql = nullptr, *p = nullptr;         // views
    struct Part {
        uint32_t s = 0, e = 0;                    // the digit's towers [s, e)
        fhe_ring *part = nullptr, *lo = nullptr, *hi = nullptr;       // views: the digit, the towers below it, the towers above it (incl. P)
        fhe_basis *to_lo = nullptr, *to_hi = nullptr;                 // ApproxSwitchCRTBasis digit -> complement, in two contiguous pieces
    };
    std::vector<Part> parts;
    fhe_basis* p_to_ql = nullptr;                 // ApproxModDown
    Scratch coef, digits, c01;                    // c01: both halves of the inner product, [2][batch][T][N]
    // fhe_keyswitch_hybrid's middle (everything between the first read of c and the last write of ct0/ct1 touches only the
    // plan's own buffers and the key) as captured graphs, one per (batch, key, t, buffer addresses)
    struct Graph {
        std::vector<uint64_t> key;
        cudaGraphExec_t exec = nullptr;
        int kernels = 0;
        uint64_t used = 0;
    };
    std::vector<Graph> graphs;
    uint64_t tick = 0;
    bool warmed = false, graph_off = false;
};

extern "C" int fhe_keyswitch_destroy(fhe_keyswitch* k) {
    GateHold gate_;
    if (!k) return FHE_OK;
    cudaSetDevice(k->ctx->device);
    cudaDeviceSynchronize();
    for (auto& pt : k->parts) {
        fhe_basis_destroy(pt.to_lo); fhe_basis_destroy(pt.to_hi);
        fhe_ring_destroy(pt.part); fhe_ring_destroy(pt.lo); fhe_ring_destroy(pt.hi);
    }
    fhe_basis_destroy(k->p_to_ql);
    fhe_ring_destroy(k->ql); fhe_ring_destroy(k->p);
    for (auto& g : k->graphs)
        if (g.exec) cudaGraphExecDestroy(g.exec);
    k->coef.drop(k->ctx); k->digits.drop(k->ctx); k->c01.drop(k->ctx);
    delete k;
    return FHE_OK;
}

extern "C" int fhe_keyswitch_create(fhe_ring* ring_qlp, uint32_t size_ql, uint32_t alpha, fhe_keyswitch** out) {
    GateHold gate_;
    if (!ring_qlp || !out) return fail(FHE_EINVAL, "fhe_keyswitch_create: null argument");
    *out = nullptr;
    const uint32_t T = ring_qlp->dev.T;
    if (size_ql == 0 || size_ql >= T) return fail(FHE_EINVAL, "fhe_keyswitch_create: size_ql=%u must leave at least one tower of P in the ring's %u", size_ql, T);
    if (alpha == 0) return fail(FHE_EINVAL, "fhe_keyswitch_create: alpha (towers per digit) is zero");
    if (ring_qlp->dev.N < 2) return fail(FHE_EUNSUPPORTED, "fhe_keyswitch_create: N must be at least 2");
    auto* k = new fhe_keyswitch;
    k->ctx = ring_qlp->ctx; k->qlp = ring_qlp; k->size_ql = size_ql; k->size_p = T - size_ql; k->alpha = alpha;
    int rc;
    if ((rc = fhe_ring_view(ring_qlp, 0, size_ql, &k->ql)) || (rc = fhe_ring_view(ring_qlp, size_ql, T - size_ql, &k->p)) ||
        (rc = fhe_basis_create(k->p, k->ql, nullptr, nullptr, &k->p_to_ql))) {
        fhe_keyswitch_destroy(k);
        return rc;
    }
    for (uint32_t s = 0; s < size_ql; s += alpha) {
        fhe_keyswitch::Part pt;
        pt.s = s; pt.e = std::min(s + alpha, size_ql);
        k->parts.push_back(pt);
        auto& P = k->parts.back();
        // the digit's own tables, products over its (possibly truncated) basis: rns-cryptoparameters.cpp:289-340
        if ((rc = fhe_ring_view(ring_qlp, P.s, P.e - P.s, &P.part)) ||
            (P.s > 0 && ((rc = fhe_ring_view(ring_qlp, 0, P.s, &P.lo)) || (rc = fhe_basis_create(P.part, P.lo, nullptr, nullptr, &P.to_lo)))) ||
            (rc = fhe_ring_view(ring_qlp, P.e, T - P.e, &P.hi)) || (rc = fhe_basis_create(P.part, P.hi, nullptr, nullptr, &P.to_hi))) {
            fhe_keyswitch_destroy(k);
            return rc;
        }
    }
    *out = k;
    return FHE_OK;
}

extern "C" uint32_t fhe_keyswitch_digits(const fhe_keyswitch* k) { return k ? (uint32_t)k->parts.size() : 0; }

// EvalKeySwitchPrecomputeCore (keyswitch-hybrid.cpp:330-420) on stream s: digits [np][batch][T][N], in two steps.
// keyswitch_read: the only launches that read c - its towers to COEFFICIENT form, one transform for all digits
// (:392-393), and each digit's own stored towers copied where they belong (:410-412).
static int keyswitch_read(fhe_keyswitch* k, uint64_t* digits, const uint64_t* c, size_t batch, cudaStream_t s) {
    const uint32_t T = k->qlp->dev.T;
    int rc;
    if ((rc = ring_ntt(k->ql, (uint64_t*)k->coef.p, batch, true, s, "fhe_keyswitch_precompute", c, k->size_ql, k->size_ql))) return rc;
    CUDA_TRY(launch_ks_own_towers(digits, c, k->size_ql, T, k->alpha, k->qlp->dev.logN, batch, s));
    return FHE_OK;
}
// keyswitch_extend: ApproxSwitchCRTBasis of every digit to its complement basis (:396-402), written where the towers belong
// in the digit, then EVALUATION (:404) in place there.  Reads and writes the plan's buffers only.
static int keyswitch_extend(fhe_keyswitch* k, uint64_t* digits, size_t batch, cudaStream_t s) {
    const uint32_t N = k->qlp->dev.N, logN = k->qlp->dev.logN, T = k->qlp->dev.T;
    const char* who = "fhe_keyswitch_precompute";
    const uint64_t* coef = (const uint64_t*)k->coef.p;
    int rc;
    for (size_t j = 0; j < k->parts.size(); j++) {
        auto& P = k->parts[j];
        uint64_t* d = digits + j * batch * (size_t)T * N;
        if (P.lo) {
            CUDA_TRY(launch_basis_switch(P.to_lo->dev, d, T, 0, coef, k->size_ql, P.s, batch, logN, false, s));
            if ((rc = ring_ntt(P.lo, d, batch, false, s, who, nullptr, T))) return rc;
        }
        CUDA_TRY(launch_basis_switch(P.to_hi->dev, d, T, P.e, coef, k->size_ql, P.s, batch, logN, false, s));
        if ((rc = ring_ntt(P.hi, d + (size_t)P.e * N, batch, false, s, who, nullptr, T))) return rc;
    }
    return FHE_OK;
}
static int keyswitch_precompute(fhe_keyswitch* k, uint64_t* digits, const uint64_t* c, size_t batch, cudaStream_t s) {
    int rc;
    if ((rc = k->coef.need(k->ctx, batch * k->size_ql * (size_t)k->qlp->dev.N * 8))) return rc;
    if ((rc = keyswitch_read(k, digits, c, batch, s))) return rc;
    return keyswitch_extend(k, digits, batch, s);
}

extern "C" int fhe_keyswitch_precompute(fhe_keyswitch* k, fhe_dptr digits, fhe_dptr c, size_t batch, int stream) {
    if (!k) return fail(FHE_EINVAL, "null key-switching plan");
    StreamRef s;
    int rc = ring_stream(k->qlp, stream, &s);
    if (rc) return rc;
    if (batch == 0) return FHE_OK;
    if (!digits || !c) return fail(FHE_EINVAL, "fhe_keyswitch_precompute: null device pointer");
    if ((digits | c) & 15) return fail(FHE_EINVAL, "fhe_keyswitch_precompute: device pointers must be 16-byte aligned");
    return keyswitch_precompute(k, (uint64_t*)(uintptr_t)digits, (const uint64_t*)(uintptr_t)c, batch, s);
}

extern "C" int fhe_keyswitch_inner(fhe_keyswitch* k, fhe_dptr ctilda0, fhe_dptr ctilda1, fhe_dptr digits, fhe_dptr evk_b,
                                   fhe_dptr evk_a, size_t batch, int stream) {
    if (!k) return fail(FHE_EINVAL, "null key-switching plan");
    StreamRef s;
    int rc = ring_stream(k->qlp, stream, &s);
    if (rc) return rc;
    if (batch == 0) return FHE_OK;
    if (!ctilda0 || !ctilda1 || !digits || !evk_b || !evk_a) return fail(FHE_EINVAL, "fhe_keyswitch_inner: null device pointer");
    if (ctilda0 == ctilda1) return fail(FHE_EINVAL, "fhe_keyswitch_inner: the two outputs must be distinct");
    if ((ctilda0 | ctilda1 | digits | evk_b | evk_a) & 15) return fail(FHE_EINVAL, "fhe_keyswitch_inner: device pointers must be 16-byte aligned");
    CUDA_TRY(launch_ks_inner(k->qlp->dev, (uint64_t*)(uintptr_t)ctilda0, (uint64_t*)(uintptr_t)ctilda1, (const uint64_t*)(uintptr_t)digits,
                             (const uint64_t*)(uintptr_t)evk_b, (const uint64_t*)(uintptr_t)evk_a, (uint32_t)k->parts.size(), batch, s));
    return FHE_OK;
}

// Runs `body` (launches on s that touch nothing but plan-owned buffers and the addresses in `key`) from a captured graph:
// the first call of a plan runs it directly (function attributes, lazily built tables), the first call with a new key
// captures it, later ones replay it with one driver call.  FHE_B200_KS_GRAPH=0 turns this off.
template <class Body>
static int keyswitch_graphed(fhe_keyswitch* k, const std::vector<uint64_t>& key, cudaStream_t s, Body body) {
    // 0: never, 1 (default): where the launches cost more than the kernels (measured: up to 2^20 words per poly over Ql u P;
    // beyond that the replay is a few percent slower than direct launches from 16 host threads), 2: always
    static const int mode = [] { const char* e = getenv("FHE_B200_KS_GRAPH"); return e && e[0] >= '0' && e[0] <= '2' ? e[0] - '0' : 1; }();
    const bool launch_bound = key[0] * k->qlp->dev.T * (uint64_t)k->qlp->dev.N <= (1ull << 20);
    if (mode == 0 || (mode == 1 && !launch_bound) || k->graph_off || !k->warmed || s == cudaStreamLegacy || s == nullptr) {
        k->warmed = true;
        return body();
    }
    for (auto& g : k->graphs)
        if (g.key == key) {
            g.used = ++k->tick;
            CUDA_TRY(cudaGraphLaunch(g.exec, s));
            kernel_launch_count_add(g.kernels);
            return FHE_OK;
        }
    g_gate.begin_capture();                                  // alone in the library from here (CaptureGate)
    if (cudaStreamBeginCapture(s, cudaStreamCaptureModeRelaxed) != cudaSuccess) {
        g_gate.end_capture();
        cudaGetLastError();
        k->graph_off = true;
        return body();
    }
    const int rc = body();                                   // recorded, not run (the launchers count their kernels here)
    cudaGraph_t graph = nullptr;
    cudaError_t e = cudaStreamEndCapture(s, &graph);
    g_gate.end_capture();
    cudaGraphExec_t exec = nullptr;
    size_t nodes = 0;
    if (!rc && e == cudaSuccess && graph) {
        e = cudaGraphGetNodes(graph, nullptr, &nodes);
        if (e == cudaSuccess) e = cudaGraphInstantiate(&exec, graph, 0);
    }
    if (graph) cudaGraphDestroy(graph);
    if (rc) return rc;
    if (e != cudaSuccess || !exec) {
        cudaGetLastError();
        k->graph_off = true;
        return body();
    }
    if (k->graphs.size() >= 64) {                            // least recently used out
        size_t v = 0;
        for (size_t i = 1; i < k->graphs.size(); i++)
            if (k->graphs[i].used < k->graphs[v].used) v = i;
        CUDA_TRY(cudaStreamSynchronize(s));
        cudaGraphExecDestroy(k->graphs[v].exec);
        k->graphs.erase(k->graphs.begin() + v);
    }
    fhe_keyswitch::Graph g;
    g.key = key; g.exec = exec; g.kernels = (int)nodes; g.used = ++k->tick;
    k->graphs.push_back(g);
    CUDA_TRY(cudaGraphLaunch(exec, s));
    return FHE_OK;
}

static int keyswitch_hybrid_run(fhe_keyswitch* k, fhe_dptr ct0, fhe_dptr ct1, fhe_dptr c, fhe_dptr evk_b, fhe_dptr evk_a,
                                size_t batch, uint64_t t, int stream, bool accumulate) {
    if (!k) return fail(FHE_EINVAL, "null key-switching plan");
    if (batch == 0) return FHE_OK;
    if (!ct0 || !ct1 || !c || !evk_b || !evk_a) return fail(FHE_EINVAL, "fhe_keyswitch_hybrid: null device pointer");
    if (ct0 == ct1) return fail(FHE_EINVAL, "fhe_keyswitch_hybrid: the two outputs must be distinct");
    if ((ct0 | ct1 | c | evk_b | evk_a) & 15) return fail(FHE_EINVAL, "fhe_keyswitch_hybrid: device pointers must be 16-byte aligned");
    const uint32_t N = k->qlp->dev.N, T = k->qlp->dev.T, sq = k->size_ql, sp = k->size_p;
    const size_t item = (size_t)T * N;                       // words of one poly over Ql u P
    fhe_basis* md = k->p_to_ql;
    StreamRef s;
    int rc;
    if ((rc = ring_stream(k->qlp, stream, &s))) return rc;
    if ((rc = k->coef.need(k->ctx, batch * sq * (size_t)N * 8)) || (rc = k->digits.need(k->ctx, k->parts.size() * batch * item * 8)) ||
        (rc = k->c01.need(k->ctx, 2 * batch * item * 8)) || (rc = md->s0.need(k->ctx, 2 * batch * sp * (size_t)N * 8)) ||
        (rc = md->s1.need(k->ctx, 2 * batch * sq * (size_t)N * 8)))
        return rc;
    BasisDev bd;
    if ((rc = mod_down_tables(md, t, s, &bd))) return rc;
    uint64_t *digits = (uint64_t*)k->digits.p, *c01 = (uint64_t*)k->c01.p, *sw = (uint64_t*)md->s1.p, *partP = (uint64_t*)md->s0.p;
    // 1. the launches that read c
    if ((rc = keyswitch_read(k, digits, (const uint64_t*)(uintptr_t)c, batch, s))) return rc;
    // 2. digits over Ql u P, the inner product with the key (EvalFastKeySwitchCoreExt, :442-480), and ApproxModDown's
    //    correction term of both halves at once (EvalFastKeySwitchCore, :430-440)
    const std::vector<uint64_t> key = {batch, t, evk_b, evk_a, (uint64_t)(uintptr_t)k->coef.p, (uint64_t)(uintptr_t)digits,
                                       (uint64_t)(uintptr_t)c01, (uint64_t)(uintptr_t)sw, (uint64_t)(uintptr_t)partP,
                                       (uint64_t)(uintptr_t)bd.pre, (uint64_t)(uintptr_t)bd.tgt};
    rc = keyswitch_graphed(k, key, s, [&]() -> int {
        int r = keyswitch_extend(k, digits, batch, s);
        if (r) return r;
        CUDA_TRY(launch_ks_inner(k->qlp->dev, c01, c01 + batch * item, digits, (const uint64_t*)(uintptr_t)evk_b,
                                 (const uint64_t*)(uintptr_t)evk_a, (uint32_t)k->parts.size(), batch, s));
        return mod_down_inner(md, bd, sw, partP, c01, 2 * batch, s);
    });
    if (rc) return rc;
    // 3. the launches that write the results
    return mod_down_final_pair(md, (uint64_t*)(uintptr_t)ct0, (uint64_t*)(uintptr_t)ct1, c01, sw, batch, item, s, accumulate);
}

extern "C" int fhe_keyswitch_hybrid(fhe_keyswitch* k, fhe_dptr ct0, fhe_dptr ct1, fhe_dptr c, fhe_dptr evk_b, fhe_dptr evk_a,
                                    size_t batch, uint64_t t, int stream) {
    return keyswitch_hybrid_run(k, ct0, ct1, c, evk_b, evk_a, batch, t, stream, false);
}

// acc0 += ct0, acc1 += ct1 of the key switch of c: what every caller of KeySwitchCore in the schemes does next
// (base-leveledshe.cpp:210-213 and the six other places: cv[0] += (*ab)[0]; cv[1] += (*ab)[1]), folded into the last launch
extern "C" int fhe_keyswitch_hybrid_add(fhe_keyswitch* k, fhe_dptr acc0, fhe_dptr acc1, fhe_dptr c, fhe_dptr evk_b, fhe_dptr evk_a,
                                        size_t batch, uint64_t t, int stream) {
    if (k && k->qlp->dev.N < 2) return fail(FHE_EUNSUPPORTED, "fhe_keyswitch_hybrid_add: N must be at least 2");
    if (acc0 == c || acc1 == c) return fail(FHE_EINVAL, "fhe_keyswitch_hybrid_add: an accumulator must not be the input");
    return keyswitch_hybrid_run(k, acc0, acc1, c, evk_b, evk_a, batch, t, stream, true);
}

// EvalFastKeySwitchCore (keyswitch-hybrid.cpp:415-437), the per-rotation half of a hoisted rotation: the inner product of
// given digits with a key, and ApproxModDown of both halves at once.
extern "C" int fhe_keyswitch_fast(fhe_keyswitch* k, fhe_dptr ct0, fhe_dptr ct1, fhe_dptr digits, fhe_dptr evk_b, fhe_dptr evk_a,
                                  size_t batch, uint64_t t, int stream) {
    if (!k) return fail(FHE_EINVAL, "null key-switching plan");
    if (batch == 0) return FHE_OK;
    if (!ct0 || !ct1 || !digits || !evk_b || !evk_a) return fail(FHE_EINVAL, "fhe_keyswitch_fast: null device pointer");
    if (ct0 == ct1) return fail(FHE_EINVAL, "fhe_keyswitch_fast: the two outputs must be distinct");
    if ((ct0 | ct1 | digits | evk_b | evk_a) & 15) return fail(FHE_EINVAL, "fhe_keyswitch_fast: device pointers must be 16-byte aligned");
    const uint32_t N = k->qlp->dev.N, T = k->qlp->dev.T, sq = k->size_ql, sp = k->size_p;
    const size_t item = (size_t)T * N;
    fhe_basis* md = k->p_to_ql;
    StreamRef s;
    int rc;
    if ((rc = ring_stream(k->qlp, stream, &s))) return rc;
    if ((rc = k->c01.need(k->ctx, 2 * batch * item * 8)) || (rc = md->s0.need(k->ctx, 2 * batch * sp * (size_t)N * 8)) ||
        (rc = md->s1.need(k->ctx, 2 * batch * sq * (size_t)N * 8)))
        return rc;
    BasisDev bd;
    if ((rc = mod_down_tables(md, t, s, &bd))) return rc;
    uint64_t *c01 = (uint64_t*)k->c01.p, *sw = (uint64_t*)md->s1.p, *partP = (uint64_t*)md->s0.p;
    CUDA_TRY(launch_ks_inner(k->qlp->dev, c01, c01 + batch * item, (const uint64_t*)(uintptr_t)digits, (const uint64_t*)(uintptr_t)evk_b,
                             (const uint64_t*)(uintptr_t)evk_a, (uint32_t)k->parts.size(), batch, s));
    if ((rc = mod_down_inner(md, bd, sw, partP, c01, 2 * batch, s))) return rc;
    return mod_down_final_pair(md, (uint64_t*)(uintptr_t)ct0, (uint64_t*)(uintptr_t)ct1, c01, sw, batch, item, s);
}

// ---------------------------------------------------------------------------------------------
// DCRTPolyImpl::CRTDecompose(baseBits = 0) (dcrtpoly-impl.h:266-288): the digits of KeySwitchBV (keyswitch-bv.cpp:308-312)
// ---------------------------------------------------------------------------------------------
struct fhe_crt_decomp {
    fhe_ctx* ctx = nullptr;
    fhe_ring* ring = nullptr;
    void* d_sw = nullptr;                         // [T][T] {half = q_i >> 1, q_k, diff, up} (NativeVector::SwitchModulus, q_i -> q_k)
    Scratch coef;
};

extern "C" int fhe_crt_decompose_destroy(fhe_crt_decomp* p) {
    GateHold gate_;
    if (!p) return FHE_OK;
    cudaSetDevice(p->ctx->device);
    cudaDeviceSynchronize();
    if (p->d_sw) arena_park(p->ctx, p->d_sw);
    p->coef.drop(p->ctx);
    delete p;
    return FHE_OK;
}

extern "C" int fhe_crt_decompose_create(fhe_ring* ring, fhe_crt_decomp** out) {
    GateHold gate_;
    if (!ring || !out) return fail(FHE_EINVAL, "fhe_crt_decompose_create: null argument");
    *out = nullptr;
    if (ring->dev.N < 2) return fail(FHE_EUNSUPPORTED, "fhe_crt_decompose_create: N must be at least 2");
    if (!ring->has_ntt && !ring->ntt_identity) return fail(FHE_EINVAL, "fhe_crt_decompose_create: the ring was created without roots of unity");
    const uint32_t T = ring->dev.T;
    if (T > 1024) return fail(FHE_EUNSUPPORTED, "fhe_crt_decompose_create: more than 1024 towers");
    auto* p = new fhe_crt_decomp;
    p->ctx = ring->ctx; p->ring = ring;
    std::vector<ulonglong4> sw((size_t)T * T);
    for (uint32_t i = 0; i < T; i++)
        for (uint32_t k = 0; k < T; k++) {
            const uint64_t om = ring->q[i], nm = ring->q[k];
            const bool up = nm > om;
            sw[(size_t)i * T + k] = make_ulonglong4(om >> 1, nm, up ? nm - om : nm - (om % nm), up ? 1 : 0);
        }
    CUDA_TRY(cudaSetDevice(p->ctx->device));
    int rc = upload(p->ctx, sw, &p->d_sw);
    if (rc) { fhe_crt_decompose_destroy(p); return rc; }
    *out = p;
    return FHE_OK;
}

extern "C" int fhe_crt_decompose(fhe_crt_decomp* p, fhe_dptr dst_digits, fhe_dptr src, size_t batch, int format, int stream) {
    const char* who = "fhe_crt_decompose";
    if (!p) return fail(FHE_EINVAL, "null CRT decomposition plan");
    StreamRef s;
    int rc = ring_stream(p->ring, stream, &s);
    if (rc) return rc;
    if (format != 0 && format != 1) return fail(FHE_EINVAL, "%s: format must be 0 (EVALUATION) or 1 (COEFFICIENT)", who);
    if (batch == 0) return FHE_OK;
    if (!dst_digits || !src) return fail(FHE_EINVAL, "%s: null device pointer", who);
    if ((dst_digits | src) & 15) return fail(FHE_EINVAL, "%s: device pointers must be 16-byte aligned", who);
    if (dst_digits == src) return fail(FHE_EINVAL, "%s: dst must not alias src", who);
    const uint32_t N = p->ring->dev.N, logN = p->ring->dev.logN, T = p->ring->dev.T;
    uint64_t* dst = (uint64_t*)(uintptr_t)dst_digits;
    const uint64_t* in = (const uint64_t*)(uintptr_t)src;
    const uint64_t* coef = in;
    if (format == 0) {                                   // :267-270 the COEFFICIENT form, in scratch
        if ((rc = p->coef.need(p->ctx, batch * T * (size_t)N * 8))) return rc;
        if ((rc = ring_ntt(p->ring, (uint64_t*)p->coef.p, batch, true, s, who, in, T, T))) return rc;
        coef = (const uint64_t*)p->coef.p;
    }
    CUDA_TRY(launch_crt_spread(dst, coef, batch, logN, T, (const ulonglong4*)p->d_sw, s));      // :276-286
    return ring_ntt(p->ring, dst, (size_t)T * batch, false, s, who);   // every tower of every digit; tower i of digit i comes back as stored
}

// ---------------------------------------------------------------------------------------------
// BEHZ (dcrtpoly-impl.h:2005-2050, 2069-2420): the BFV multiplication technique that the HPS plans above do not cover.
// A plan over the ring Q u Bsk holds the scheme's tables (bfvrns-cryptoparameters.cpp:662-885, derived from the
// moduli and t), three conversions built as fhe_basis objects with those tables, and the correction constants of
// behz.cu.
// ---------------------------------------------------------------------------------------------
enum { BEHZ_TABLES = 17 };
struct fhe_behz {
    fhe_ctx* ctx = nullptr;
    fhe_ring* qbsk = nullptr;                     // caller's ring: nq towers of Q, then B, then msk
    uint32_t nq = 0, nb = 0;
    uint64_t t = 0;
    fhe_ring *q = nullptr, *bsk = nullptr, *b = nullptr;        // views
    fhe_ring* qm = nullptr;                       // Q followed by msk (targets of FastBaseConvSK's two conversions), no NTT
    fhe_basis *mont = nullptr, *floor = nullptr, *sk = nullptr;
    std::vector<uint64_t> tab[BEHZ_TABLES];
    BehzDev dev{};
    void *d_qk = nullptr, *d_bk = nullptr, *d_bsk = nullptr, *d_fk = nullptr, *d_sk = nullptr, *d_dk = nullptr;
    Scratch coef, conv;
};

extern "C" int fhe_behz_destroy(fhe_behz* z) {
    GateHold gate_;
    if (!z) return FHE_OK;
    cudaSetDevice(z->ctx->device);
    cudaDeviceSynchronize();
    fhe_basis_destroy(z->mont); fhe_basis_destroy(z->floor); fhe_basis_destroy(z->sk);
    fhe_ring_destroy(z->q); fhe_ring_destroy(z->bsk); fhe_ring_destroy(z->b); fhe_ring_destroy(z->qm);
    for (void* p : {z->d_qk, z->d_bk, z->d_bsk, z->d_fk, z->d_sk, z->d_dk})
        if (p) arena_park(z->ctx, p);
    z->coef.drop(z->ctx); z->conv.drop(z->ctx);
    delete z;
    return FHE_OK;
}

static uint64_t prod_all_mod(const std::vector<uint64_t>& f, uint64_t m) { return prod_mod(f, (size_t)-1, m); }

extern "C" int fhe_behz_create(fhe_ring* ring_qbsk, uint32_t size_q, uint64_t t, fhe_behz** out) {
    GateHold gate_;
    if (!ring_qbsk || !out) return fail(FHE_EINVAL, "fhe_behz_create: null argument");
    *out = nullptr;
    const uint32_t T = ring_qbsk->dev.T;
    if (size_q == 0 || size_q + 2 > T)
        return fail(FHE_EINVAL, "fhe_behz_create: size_q=%u must leave at least one tower of B and msk in the ring's %u", size_q, T);
    if (T > 256) return fail(FHE_EUNSUPPORTED, "fhe_behz_create: more than 256 towers");
    if (ring_qbsk->dev.N < 2) return fail(FHE_EUNSUPPORTED, "fhe_behz_create: N must be at least 2");
    if (t < 2 || t >= (1ull << 32)) return fail(FHE_EINVAL, "fhe_behz_create: plaintext modulus %llu not in [2, 2^32) (t * 2^26 has to stay below 2^58)", (unsigned long long)t);
    const uint32_t nq = size_q, nb = T - nq, nB = nb - 1;
    const std::vector<uint64_t> q(ring_qbsk->q.begin(), ring_qbsk->q.begin() + nq);
    const std::vector<uint64_t> bsk(ring_qbsk->q.begin() + nq, ring_qbsk->q.end());
    const std::vector<uint64_t> B(bsk.begin(), bsk.begin() + nB);
    const uint64_t msk = bsk[nB], mtilde = 1ull << 16, gamma = 1ull << 26, tgamma = t * gamma;
    for (uint64_t m : ring_qbsk->q)
        if (!(m & 1)) return fail(FHE_EINVAL, "fhe_behz_create: modulus %llu is even", (unsigned long long)m);
    for (uint64_t m : bsk)          // "mtilde = 2^16 < all moduli of Bsk", dcrtpoly-impl.h:2178
        if (m <= mtilde) return fail(FHE_EINVAL, "fhe_behz_create: modulus %llu of Bsk is not above mtilde = 2^16", (unsigned long long)m);
    auto* z = new fhe_behz;
    z->ctx = ring_qbsk->ctx; z->qbsk = ring_qbsk; z->nq = nq; z->nb = nb; z->t = t;
    auto bad = [&](const char* what) { fhe_behz_destroy(z); return fail(FHE_EINVAL, "fhe_behz_create: %s", what); };
    // ---- tables, numbered as include/fhe_b200.h lists them (bfvrns-cryptoparameters.cpp:724-891) ----
    std::vector<uint64_t> hatinv(nq);
    for (uint32_t i = 0; i < nq; i++) {
        hatinv[i] = host::invmod(prod_mod(q, i, q[i]), q[i]);
        if (!hatinv[i]) return bad("the moduli of Q are not pairwise coprime");
    }
    auto& tb = z->tab;
    tb[0].resize(nq); tb[1].resize((size_t)nq * nb); tb[2].resize(nq); tb[6].resize(nq); tb[7].resize((size_t)nq * nb);
    tb[13].resize(nq); tb[15].resize(nq); tb[16].resize(nq);
    for (uint32_t i = 0; i < nq; i++) {
        tb[0][i] = host::mulmod(hatinv[i], mtilde % q[i], q[i]);
        tb[6][i] = host::mulmod(hatinv[i], t % q[i], q[i]);
        tb[2][i] = prod_mod(q, i, mtilde);
        tb[13][i] = prod_all_mod(B, q[i]);
        const uint64_t qinv_tg = host::invmod(q[i] % tgamma, tgamma);
        if (!qinv_tg) return bad("a modulus of Q shares a factor with t");
        tb[15][i] = host::mulmod(tgamma - 1, qinv_tg, tgamma);
        tb[16][i] = host::mulmod(host::mulmod(hatinv[i], gamma % q[i], q[i]), t % q[i], q[i]);
        for (uint32_t j = 0; j < nb; j++) {
            tb[1][(size_t)i * nb + j] = prod_mod(q, i, bsk[j]);
            const uint64_t inv = host::invmod(q[i] % bsk[j], bsk[j]);
            if (!inv) return bad("a modulus of Q is not invertible modulo a modulus of Bsk");
            tb[7][(size_t)i * nb + j] = inv;
        }
    }
    tb[3].resize(nb); tb[5].resize(nb); tb[8].resize(nb);
    for (uint32_t j = 0; j < nb; j++) {
        tb[3][j] = prod_all_mod(q, bsk[j]);
        tb[5][j] = host::invmod(mtilde % bsk[j], bsk[j]);
        const uint64_t qinv = host::invmod(tb[3][j], bsk[j]);
        if (!qinv || !tb[5][j]) return bad("Q or mtilde is not invertible modulo a modulus of Bsk");
        tb[8][j] = host::mulmod(qinv, t % bsk[j], bsk[j]);
    }
    tb[4].assign(1, host::mulmod(mtilde - 1, host::invmod(prod_all_mod(q, mtilde), mtilde), mtilde));
    tb[9].resize(nB); tb[10].resize((size_t)nB * nq); tb[11].resize(nB);
    for (uint32_t i = 0; i < nB; i++) {
        tb[9][i] = host::invmod(prod_mod(B, i, B[i]), B[i]);
        if (!tb[9][i]) return bad("the moduli of B are not pairwise coprime");
        tb[11][i] = prod_mod(B, i, msk);
        for (uint32_t j = 0; j < nq; j++) tb[10][(size_t)i * nq + j] = prod_mod(B, i, q[j]);
    }
    tb[12].assign(1, host::invmod(prod_all_mod(B, msk), msk));
    if (!tb[12][0]) return bad("B is not invertible modulo msk");
    tb[14].assign(1, tgamma);
    // ---- views, conversions, constants ----
    std::vector<uint64_t> qm(q);
    qm.push_back(msk);
    std::vector<uint64_t> sk_hat((size_t)nB * (nq + 1));                  // [B/b_i]_{q_j} with [B/b_i]_{msk} as the last column
    for (uint32_t i = 0; i < nB; i++) {
        for (uint32_t j = 0; j < nq; j++) sk_hat[(size_t)i * (nq + 1) + j] = tb[10][(size_t)i * nq + j];
        sk_hat[(size_t)i * (nq + 1) + nq] = tb[11][i];
    }
    int rc;
    if ((rc = fhe_ring_view(ring_qbsk, 0, nq, &z->q)) || (rc = fhe_ring_view(ring_qbsk, nq, nb, &z->bsk)) ||
        (rc = fhe_ring_view(ring_qbsk, nq, nB, &z->b)) ||
        (rc = fhe_ring_create(z->ctx, ring_qbsk->dev.N, nq + 1, qm.data(), nullptr, &z->qm)) ||
        (rc = fhe_basis_create(z->q, z->bsk, tb[0].data(), tb[1].data(), &z->mont)) ||
        (rc = fhe_basis_create(z->q, z->bsk, tb[6].data(), tb[7].data(), &z->floor)) ||
        (rc = fhe_basis_create(z->b, z->qm, tb[9].data(), sk_hat.data(), &z->sk))) {
        fhe_behz_destroy(z);
        return rc;
    }
    std::vector<ulonglong4> qk(nq), bk(nb), fk(nq + nb), sk(nq), dk(nq);
    for (uint32_t i = 0; i < nq; i++) {
        qk[i] = make_ulonglong4(q[i], tb[0][i], host::shoup_companion(tb[0][i], q[i]), tb[2][i]);
        fk[i] = make_ulonglong4(tb[6][i], host::shoup_companion(tb[6][i], q[i]), q[i], 0);
        sk[i] = make_ulonglong4(tb[13][i], host::shoup_companion(tb[13][i], q[i]), q[i], 0);
        dk[i] = make_ulonglong4(tb[16][i], host::shoup_companion(tb[16][i], q[i]), tb[15][i], host::shoup_companion(tb[15][i], tgamma));
    }
    for (uint32_t j = 0; j < nb; j++) {
        bk[j] = make_ulonglong4(tb[3][j], host::shoup_companion(tb[3][j], bsk[j]), tb[5][j], host::shoup_companion(tb[5][j], bsk[j]));
        fk[nq + j] = make_ulonglong4(tb[8][j], host::shoup_companion(tb[8][j], bsk[j]), bsk[j], 0);
    }
    if ((rc = upload(z->ctx, qk, &z->d_qk)) || (rc = upload(z->ctx, bk, &z->d_bk)) || (rc = upload(z->ctx, bsk, &z->d_bsk)) ||
        (rc = upload(z->ctx, fk, &z->d_fk)) || (rc = upload(z->ctx, sk, &z->d_sk)) || (rc = upload(z->ctx, dk, &z->d_dk))) {
        fhe_behz_destroy(z);
        return rc;
    }
    BehzDev& d = z->dev;
    d.nq = nq; d.nb = nb;
    d.qk = (const ulonglong4*)z->d_qk; d.bk = (const ulonglong4*)z->d_bk; d.bsk = (const uint64_t*)z->d_bsk;
    d.fk = (const ulonglong4*)z->d_fk; d.sk = (const ulonglong4*)z->d_sk; d.dk = (const ulonglong4*)z->d_dk;
    d.neg_qinv_mtilde = tb[4][0];
    d.msk = msk; d.binv_msk = tb[12][0]; d.binv_mskp = host::shoup_companion(tb[12][0], msk);
    d.tgamma = tgamma;
    *out = z;
    return FHE_OK;
}

extern "C" int fhe_behz_table(const fhe_behz* z, int which, uint64_t* out, size_t* len) {
    if (!z) return fail(FHE_EINVAL, "null BEHZ plan");
    if (which < 0 || which >= BEHZ_TABLES) return fail(FHE_EINVAL, "fhe_behz_table: table %d not in [0, %d)", which, (int)BEHZ_TABLES);
    if (len) *len = z->tab[which].size();
    if (out) memcpy(out, z->tab[which].data(), z->tab[which].size() * 8);
    return FHE_OK;
}

static int behz_enter(fhe_behz* z, int stream, StreamRef* s, const char* who, size_t batch, std::initializer_list<fhe_dptr> ptrs) {
    if (!z) return fail(FHE_EINVAL, "null BEHZ plan");
    int rc = ring_stream(z->qbsk, stream, s);
    if (rc) return rc;
    if (batch == 0) return FHE_OK;
    fhe_dptr all = 0;
    for (fhe_dptr p : ptrs) {
        if (!p) return fail(FHE_EINVAL, "%s: null device pointer", who);
        all |= p;
    }
    if (all & 15) return fail(FHE_EINVAL, "%s: device pointers must be 16-byte aligned", who);
    return FHE_OK;
}

extern "C" int fhe_fast_base_conv_q_to_bsk_montgomery(fhe_behz* z, fhe_dptr dst_qbsk, fhe_dptr src_q, size_t batch, int format,
                                                      int stream) {
    const char* who = "fhe_fast_base_conv_q_to_bsk_montgomery";
    StreamRef s;
    int rc = behz_enter(z, stream, &s, who, batch, {dst_qbsk, src_q});
    if (rc || batch == 0) return rc;
    if (format != 0 && format != 1) return fail(FHE_EINVAL, "%s: format must be 0 (EVALUATION) or 1 (COEFFICIENT)", who);
    if (dst_qbsk == src_q) return fail(FHE_EINVAL, "%s: dst must not alias src (the towers are laid out differently)", who);
    const uint32_t N = z->qbsk->dev.N, logN = z->qbsk->dev.logN, nq = z->nq, T = z->nq + z->nb;
    uint64_t* dst = (uint64_t*)(uintptr_t)dst_qbsk;
    const uint64_t* src = (const uint64_t*)(uintptr_t)src_q;
    const uint64_t* coef = src;
    if (format == 0) {                                   // :2087-2090 the input's COEFFICIENT form, in scratch
        if ((rc = z->coef.need(z->ctx, batch * nq * (size_t)N * 8))) return rc;
        if ((rc = ring_ntt(z->q, (uint64_t*)z->coef.p, batch, true, s, who, src, nq, nq))) return rc;
        coef = (const uint64_t*)z->coef.p;
    }
    // step 0 modulo Bsk (:2103-2127), written where the towers of Bsk belong; then modulo mtilde and step 1 (:2148-2190)
    CUDA_TRY(launch_basis_switch(z->mont->dev, dst, T, nq, coef, nq, 0, batch, logN, false, s));
    CUDA_TRY(launch_behz_montgomery(z->dev, dst, coef, batch, logN, s));
    if (format == 0) CUDA_TRY(copy_towers(dst, T, 0, src, nq, 0, nq, batch, N, s));          // :2194-2197
    else if ((rc = ring_ntt(z->q, dst, batch, false, s, who, src, T, nq))) return rc;        // :2198-2202
    return ring_ntt(z->bsk, dst + (size_t)nq * N, batch, false, s, who, nullptr, T);         // :2204-2206
}

extern "C" int fhe_fast_rns_floor_q(fhe_behz* z, fhe_dptr x_qbsk, size_t batch, int update_q, int stream) {
    const char* who = "fhe_fast_rns_floor_q";
    StreamRef s;
    int rc = behz_enter(z, stream, &s, who, batch, {x_qbsk});
    if (rc || batch == 0) return rc;
    const uint32_t N = z->qbsk->dev.N, logN = z->qbsk->dev.logN, nq = z->nq, nb = z->nb, T = nq + nb;
    uint64_t* x = (uint64_t*)(uintptr_t)x_qbsk;
    if ((rc = z->conv.need(z->ctx, batch * (size_t)std::max(nb, nq + 1) * N * 8))) return rc;
    // FastBaseConv(|t*x|_q, q, Bsk) (:2233-2256), then the subtraction (:2260-2269)
    CUDA_TRY(launch_basis_switch(z->floor->dev, (uint64_t*)z->conv.p, nb, 0, x, T, 0, batch, logN, false, s));
    CUDA_TRY(launch_behz_floor(z->dev, x, (const uint64_t*)z->conv.p, batch, logN, update_q != 0, s));
    return FHE_OK;
}

// fhe_fast_rns_floor_q without any allocation (no counterpart in the reference; added
// for callers that may not allocate: CUDA graph capture, serving loops).  The conversion's intermediate
// [batch][nb][N] lives in a file-scope device array instead of the plan's arena scratch.  The array's name is only a
// host-side stand-in in host code: its device address is resolved with cudaGetSymbolAddress, once per device, with
// that device current, and that address is what the kernels receive.
#define FHE_NOALLOC_SCRATCH_WORDS (1u << 20)                       /* 8 MiB per GPU that loads the library */
static __device__ __align__(256) unsigned long long g_noalloc_scratch[FHE_NOALLOC_SCRATCH_WORDS];

static int noalloc_scratch(int device, uint64_t** out) {
    static std::mutex mu;
    static uint64_t* addr[64] = {};
    if (device < 0 || device >= 64) return fail(FHE_EUNSUPPORTED, "no-alloc scratch: device %d not in [0, 64)", device);
    std::lock_guard<std::mutex> lk(mu);
    if (!addr[device]) {
        void* p = nullptr;
        CUDA_TRY(cudaSetDevice(device));
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) return fail(FHE_ECUDA, "no-alloc scratch: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        addr[device] = (uint64_t*)p;
    }
    *out = addr[device];
    return FHE_OK;
}

extern "C" int fhe_fast_rns_floor_q_noalloc(fhe_behz* z, fhe_dptr x_qbsk, size_t batch, int update_q, int stream) {
    const char* who = "fhe_fast_rns_floor_q_noalloc";
    StreamRef s;
    int rc = behz_enter(z, stream, &s, who, batch, {x_qbsk});
    if (rc || batch == 0) return rc;
    const uint32_t N = z->qbsk->dev.N, logN = z->qbsk->dev.logN, nq = z->nq, nb = z->nb, T = nq + nb;
    if (batch > FHE_NOALLOC_SCRATCH_WORDS / ((size_t)nb * N))
        return fail(FHE_EINVAL, "%s: batch * towers of Bsk * N = %zu * %u * %u words exceeds the static scratch of %u words (8 MiB); "
                    "use fhe_fast_rns_floor_q", who, batch, nb, N, (unsigned)FHE_NOALLOC_SCRATCH_WORDS);
    uint64_t* conv = nullptr;
    if ((rc = noalloc_scratch(z->ctx->device, &conv))) return rc;
    uint64_t* x = (uint64_t*)(uintptr_t)x_qbsk;
    // every word of conv that the second kernel reads is written by the first: nothing has to start at zero
    CUDA_TRY(launch_basis_switch(z->floor->dev, conv, nb, 0, x, T, 0, batch, logN, false, s));
    CUDA_TRY(launch_behz_floor(z->dev, x, conv, batch, logN, update_q != 0, s));
    return FHE_OK;
}

extern "C" int fhe_fast_base_conv_sk(fhe_behz* z, fhe_dptr dst_q, fhe_dptr src_qbsk, size_t batch, int stream) {
    const char* who = "fhe_fast_base_conv_sk";
    StreamRef s;
    int rc = behz_enter(z, stream, &s, who, batch, {dst_q, src_qbsk});
    if (rc || batch == 0) return rc;
    if (dst_q == src_qbsk) return fail(FHE_EINVAL, "%s: dst must not alias src (the towers are laid out differently)", who);
    const uint32_t N = z->qbsk->dev.N, logN = z->qbsk->dev.logN, nq = z->nq, nb = z->nb, T = nq + nb;
    const uint64_t* src = (const uint64_t*)(uintptr_t)src_qbsk;
    if ((rc = z->conv.need(z->ctx, batch * (size_t)std::max(nb, nq + 1) * N * 8))) return rc;
    // FastBaseconv(x, B, q) and FastBaseConv(x, B, msk) in one conversion B -> Q u {msk} (:2327-2364), then alpha_sk (:2366-2388)
    CUDA_TRY(launch_basis_switch(z->sk->dev, (uint64_t*)z->conv.p, nq + 1, 0, src, T, nq, batch, logN, false, s));
    CUDA_TRY(launch_behz_sk(z->dev, (uint64_t*)(uintptr_t)dst_q, (const uint64_t*)z->conv.p, src + (size_t)(T - 1) * N, T, batch, logN, s));
    return FHE_OK;
}

extern "C" int fhe_scale_and_round_behz(fhe_behz* z, fhe_dptr dst, fhe_dptr src_q, size_t batch, int stream) {
    StreamRef s;
    int rc = behz_enter(z, stream, &s, "fhe_scale_and_round_behz", batch, {dst, src_q});
    if (rc || batch == 0) return rc;
    CUDA_TRY(launch_behz_decrypt(z->dev, (uint64_t*)(uintptr_t)dst, (const uint64_t*)(uintptr_t)src_q, batch, z->qbsk->dev.logN, s));
    return FHE_OK;
}

// ---------------------------------------------------------------------------------------------
// several devices of one node: the batch is split in contiguous blocks, one per device, every
// device holds the ring's tables, no collective runs inside the path.  NCCL (loaded at run time: libnccl.so.2)
// carries only the optional gather of results over NVLink / NVSwitch, on the library's own streams.
// Replaces the scatter of PimManager::copy_to_pim over `nr_dpus` DPUs (src/core/pim/host/PimManager.cpp:15-24).
// ---------------------------------------------------------------------------------------------
#include <dlfcn.h>
#include <nccl.h>

namespace {
struct NcclApi {
    void* h = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string why;
};

NcclApi* nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
            api.h = dlopen(name, RTLD_NOW | RTLD_LOCAL);
            if (api.h) break;
        }
        if (!api.h) { api.why = dlerror() ? dlerror() : "libnccl.so.2 not found"; return; }
        auto sym = [&](const char* n) { void* p = dlsym(api.h, n); if (!p) api.why = std::string("missing symbol ") + n; return p; };
        api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
        api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
        api.CommInitAll = (decltype(api.CommInitAll))sym("ncclCommInitAll");
        api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
        api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
        api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
        api.AllGather = (decltype(api.AllGather))sym("ncclAllGather");
        api.Send = (decltype(api.Send))sym("ncclSend");
        api.Recv = (decltype(api.Recv))sym("ncclRecv");
        api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
        if (!api.why.empty()) { dlclose(api.h); api.h = nullptr; }
    });
    return api.h ? &api : nullptr;
}
#define NCCL_TRY(api, expr)                                                                              \
    do {                                                                                                 \
        ncclResult_t r__ = (expr);                                                                       \
        if (r__ != ncclSuccess) return fail(FHE_ECUDA, "%s failed: %s", #expr, (api)->GetErrorString(r__)); \
    } while (0)

void partition(size_t batch, int n, int rank, size_t* first, size_t* count) {
    size_t base = batch / n, rem = batch % n;
    *count = base + ((size_t)rank < rem ? 1 : 0);
    *first = (size_t)rank * base + std::min<size_t>(rank, rem);
}
}  // namespace

struct fhe_comm {
    fhe_ctx* ctx = nullptr;
    ncclComm_t comm = nullptr;
    int nranks = 0, rank = 0;
};

extern "C" int fhe_comm_unique_id(void* id, size_t cap) {
    NcclApi* api = nccl_api();
    if (!api) return fail(FHE_EUNSUPPORTED, "NCCL is not available: %s", nccl_api() ? "" : "libnccl.so.2 could not be loaded");
    if (!id || cap < sizeof(ncclUniqueId)) return fail(FHE_EINVAL, "fhe_comm_unique_id: need a buffer of %zu bytes", sizeof(ncclUniqueId));
    ncclUniqueId u;
    NCCL_TRY(api, api->GetUniqueId(&u));
    memcpy(id, &u, sizeof(u));
    return FHE_OK;
}

extern "C" int fhe_comm_create(fhe_ctx* ctx, int nranks, int rank, const void* id, fhe_comm** out) {
    GateHold gate_;
    if (!ctx || !id || !out) return fail(FHE_EINVAL, "fhe_comm_create: null argument");
    *out = nullptr;
    if (nranks < 1 || rank < 0 || rank >= nranks) return fail(FHE_EINVAL, "fhe_comm_create: rank %d not in [0,%d)", rank, nranks);
    NcclApi* api = nccl_api();
    if (!api) return fail(FHE_EUNSUPPORTED, "NCCL is not available (libnccl.so.2 could not be loaded)");
    CUDA_TRY(cudaSetDevice(ctx->device));
    ncclUniqueId u;
    memcpy(&u, id, sizeof(u));
    auto* c = new fhe_comm;
    c->ctx = ctx; c->nranks = nranks; c->rank = rank;
    ncclResult_t r = api->CommInitRank(&c->comm, nranks, u, rank);
    if (r != ncclSuccess) { delete c; return fail(FHE_ECUDA, "ncclCommInitRank failed: %s", api->GetErrorString(r)); }
    *out = c;
    return FHE_OK;
}

extern "C" int fhe_comm_destroy(fhe_comm* c) {
    GateHold gate_;
    if (!c) return FHE_OK;
    NcclApi* api = nccl_api();
    if (api && c->comm) { cudaSetDevice(c->ctx->device); api->CommDestroy(c->comm); }
    delete c;
    return FHE_OK;
}

extern "C" int fhe_comm_all_gather(fhe_comm* c, fhe_dptr dst, fhe_dptr src, size_t words, int stream) {
    if (!c) return fail(FHE_EINVAL, "null communicator");
    StreamRef s;
    int rc = get_stream(c->ctx, stream, &s);
    if (rc) return rc;
    if (words == 0) return FHE_OK;
    if (!dst || !src) return fail(FHE_EINVAL, "fhe_comm_all_gather: null device pointer");
    NcclApi* api = nccl_api();
    CUDA_TRY(cudaSetDevice(c->ctx->device));
    NCCL_TRY(api, api->AllGather((const void*)(uintptr_t)src, (void*)(uintptr_t)dst, words, ncclUint64, c->comm, s));
    return FHE_OK;
}

struct fhe_group {
    std::vector<fhe_ctx*> ctx;
    std::vector<ncclComm_t> comms;       // created on first use (fhe_group_all_gather)
    std::mutex mu;
};
struct fhe_group_ring {
    fhe_group* g = nullptr;
    std::vector<fhe_ring*> rings;
};

extern "C" int fhe_group_destroy(fhe_group* g) {
    GateHold gate_;
    if (!g) return FHE_OK;
    NcclApi* api = nccl_api();
    for (size_t i = 0; i < g->comms.size(); i++)
        if (api && g->comms[i]) { cudaSetDevice(g->ctx[i]->device); api->CommDestroy(g->comms[i]); }
    for (fhe_ctx* c : g->ctx) fhe_shutdown(c);
    delete g;
    return FHE_OK;
}

extern "C" int fhe_group_create(int ndev, const int* devices, const char* opts, fhe_group** out) {
    GateHold gate_;
    if (!out) return fail(FHE_EINVAL, "fhe_group_create: out is null");
    *out = nullptr;
    int have = 0;
    cudaError_t e = cudaGetDeviceCount(&have);
    if (e != cudaSuccess || have == 0) {
        cudaGetLastError();
        return fail(FHE_ENODEV, "fhe_group_create: no CUDA device (%s); this backend has no CPU fallback",
                    e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    }
    if (ndev <= 0) ndev = have;                       // 0: every visible device
    if (ndev > have && !devices) return fail(FHE_EINVAL, "fhe_group_create: %d devices requested, %d visible", ndev, have);
    auto* g = new fhe_group;
    for (int i = 0; i < ndev; i++) {
        int dev = devices ? devices[i] : i;
        for (fhe_ctx* c : g->ctx)
            if (c->device == dev) { fhe_group_destroy(g); return fail(FHE_EINVAL, "fhe_group_create: device %d listed twice", dev); }
        fhe_ctx* c = nullptr;
        int rc = fhe_init(dev, opts, &c);
        if (rc) { fhe_group_destroy(g); return rc; }
        g->ctx.push_back(c);
    }
    *out = g;
    return FHE_OK;
}

extern "C" int fhe_group_size(const fhe_group* g) { return g ? (int)g->ctx.size() : 0; }
extern "C" fhe_ctx* fhe_group_ctx(fhe_group* g, int rank) {
    return (g && rank >= 0 && rank < (int)g->ctx.size()) ? g->ctx[rank] : nullptr;
}

extern "C" int fhe_group_sync(fhe_group* g) {
    if (!g) return fail(FHE_EINVAL, "null group");
    for (fhe_ctx* c : g->ctx) {
        int rc = fhe_sync(c, -2);
        if (rc) return rc;
    }
    return FHE_OK;
}

extern "C" int fhe_group_partition(const fhe_group* g, size_t batch, int rank, size_t* first, size_t* count) {
    if (!g || !first || !count) return fail(FHE_EINVAL, "fhe_group_partition: null argument");
    if (rank < 0 || rank >= (int)g->ctx.size()) return fail(FHE_EINVAL, "fhe_group_partition: rank %d not in [0,%zu)", rank, g->ctx.size());
    partition(batch, (int)g->ctx.size(), rank, first, count);
    return FHE_OK;
}

// Legacy descriptor on every device at once: the descriptor's sizes are per device (PimData divides them by
// getNumDpus(), pim/PimData.h:281-290), heap_bases[rank] is that device's heap.  Blocking, like start_kernel().
extern "C" int fhe_group_launch(fhe_group* g, int kernel_id, const void* desc, const fhe_dptr* heap_bases) {
    // no GateHold here: the per-device worker threads below enter the gate themselves, and a parent that waited for them
    // while counted inside would deadlock with a capture that started in between (CaptureGate)
    if (!g || !desc || !heap_bases) return fail(FHE_EINVAL, "fhe_group_launch: null argument");
    const size_t n = g->ctx.size();
    if (n == 1) return fhe_launch(g->ctx[0], kernel_id, desc, heap_bases[0]);
    std::vector<int> rcs(n, FHE_OK);
    std::vector<std::string> msgs(n);
    std::vector<std::thread> th;
    for (size_t i = 0; i < n; i++)
        th.emplace_back([&, i] {
            rcs[i] = fhe_launch(g->ctx[i], kernel_id, desc, heap_bases[i]);
            if (rcs[i]) msgs[i] = fhe_last_error();
        });
    for (auto& t : th) t.join();
    for (size_t i = 0; i < n; i++)
        if (rcs[i]) return fail(rcs[i], "device %d: %s", g->ctx[i]->device, msgs[i].c_str());
    return FHE_OK;
}

extern "C" int fhe_group_ring_destroy(fhe_group_ring* r) {
    GateHold gate_;
    if (!r) return FHE_OK;
    for (fhe_ring* x : r->rings) fhe_ring_destroy(x);
    delete r;
    return FHE_OK;
}

extern "C" int fhe_group_ring_create(fhe_group* g, uint32_t N, uint32_t towers, const uint64_t* q, const uint64_t* psi,
                                     fhe_group_ring** out) {
    GateHold gate_;
    if (!g || !out) return fail(FHE_EINVAL, "fhe_group_ring_create: null argument");
    *out = nullptr;
    auto* r = new fhe_group_ring;
    r->g = g;
    for (fhe_ctx* c : g->ctx) {
        fhe_ring* x = nullptr;
        int rc = fhe_ring_create(c, N, towers, q, psi, &x);
        if (rc) { fhe_group_ring_destroy(r); return rc; }
        r->rings.push_back(x);
    }
    *out = r;
    return FHE_OK;
}

extern "C" fhe_ring* fhe_group_ring_at(fhe_group_ring* r, int rank) {
    return (r && rank >= 0 && rank < (int)r->rings.size()) ? r->rings[rank] : nullptr;
}

// per-device work on each device's block of the batch; launches are asynchronous, so the devices run concurrently
template <class F>
static int group_each(fhe_group_ring* r, size_t batch, F f) {
    if (!r) return fail(FHE_EINVAL, "null group ring");
    const int n = (int)r->rings.size();
    for (int i = 0; i < n; i++) {
        size_t first, count;
        partition(batch, n, i, &first, &count);
        if (count == 0) continue;
        int rc = f(i, r->rings[i], first, count);
        if (rc) return rc;
    }
    return FHE_OK;
}

extern "C" int fhe_group_dcrt_op(fhe_group_ring* r, int op, const fhe_dptr* dst, const fhe_dptr* a, const fhe_dptr* b,
                                 size_t batch, int stream) {
    if (!dst || !a || !b) return fail(FHE_EINVAL, "fhe_group_dcrt_op: null argument");
    return group_each(r, batch, [&](int i, fhe_ring* x, size_t, size_t count) { return fhe_dcrt_op(x, op, dst[i], a[i], b[i], count, stream); });
}

extern "C" int fhe_group_ntt(fhe_group_ring* r, const fhe_dptr* x, size_t batch, int inverse, int stream) {
    if (!x) return fail(FHE_EINVAL, "fhe_group_ntt: null argument");
    return group_each(r, batch, [&](int i, fhe_ring* ring, size_t, size_t count) { return fhe_ntt_algo(ring, x[i], count, inverse, 0, stream); });
}

extern "C" int fhe_group_alloc(fhe_group_ring* r, size_t batch, fhe_dptr* out) {
    if (!r || !out) return fail(FHE_EINVAL, "fhe_group_alloc: null argument");
    const int n = (int)r->rings.size();
    for (int i = 0; i < n; i++) out[i] = 0;
    for (int i = 0; i < n; i++) {
        size_t first, count;
        partition(batch, n, i, &first, &count);
        int rc = fhe_alloc(r->g->ctx[i], std::max<size_t>(count, 1) * r->rings[i]->dev.T * r->rings[i]->dev.N * 8, &out[i]);
        if (rc) {
            for (int k = 0; k < i; k++) { fhe_free(r->g->ctx[k], out[k]); out[k] = 0; }
            return rc;
        }
    }
    return FHE_OK;
}

extern "C" int fhe_group_free(fhe_group_ring* r, const fhe_dptr* p) {
    if (!r || !p) return fail(FHE_EINVAL, "fhe_group_free: null argument");
    int first_rc = FHE_OK;
    for (size_t i = 0; i < r->rings.size(); i++) {
        int rc = fhe_free(r->g->ctx[i], p[i]);
        if (rc && !first_rc) first_rc = rc;
    }
    return first_rc;
}

// host [batch][T][N] -> each device's block, and back.  Asynchronous on `stream` of every device when the host
// buffer is pinned; fhe_group_sync (or a later blocking call) completes them.
extern "C" int fhe_group_scatter(fhe_group_ring* r, const fhe_dptr* dst, const uint64_t* host, size_t batch, int stream) {
    if (!dst || (batch && !host)) return fail(FHE_EINVAL, "fhe_group_scatter: null argument");
    return group_each(r, batch, [&](int i, fhe_ring* x, size_t first, size_t count) {
        size_t item = (size_t)x->dev.T * x->dev.N;
        return fhe_h2d(r->g->ctx[i], dst[i], host + first * item, count * item * 8, stream);
    });
}
extern "C" int fhe_group_gather(fhe_group_ring* r, uint64_t* host, const fhe_dptr* src, size_t batch, int stream) {
    if (!src || (batch && !host)) return fail(FHE_EINVAL, "fhe_group_gather: null argument");
    return group_each(r, batch, [&](int i, fhe_ring* x, size_t first, size_t count) {
        size_t item = (size_t)x->dev.T * x->dev.N;
        return fhe_d2h(r->g->ctx[i], host + first * item, src[i], count * item * 8, stream);
    });
}

// Every device receives every block: dst[rank] = the whole [batch][T][N] result, HBM to HBM over NVLink / NVSwitch
// (north_star: "NCCL over NVLink only to gather results").  ncclAllGather when the blocks are equal, grouped
// ncclSend / ncclRecv otherwise.  Enqueued on `stream` of every device; complete after fhe_group_sync.
extern "C" int fhe_group_all_gather(fhe_group_ring* r, const fhe_dptr* dst, const fhe_dptr* src, size_t batch, int stream) {
    GateHold gate_;
    if (!r || !dst || !src) return fail(FHE_EINVAL, "fhe_group_all_gather: null argument");
    fhe_group* g = r->g;
    const int n = (int)g->ctx.size();
    if (batch == 0) return FHE_OK;
    const size_t item = (size_t)r->rings[0]->dev.T * r->rings[0]->dev.N;
    if (n == 1) return fhe_d2d(g->ctx[0], dst[0], src[0], batch * item * 8, stream);
    NcclApi* api = nccl_api();
    if (!api) return fail(FHE_EUNSUPPORTED, "NCCL is not available (libnccl.so.2 could not be loaded)");
    {
        std::lock_guard<std::mutex> lk(g->mu);
        if (g->comms.empty()) {
            std::vector<int> devs;
            for (fhe_ctx* c : g->ctx) devs.push_back(c->device);
            g->comms.assign(n, nullptr);
            ncclResult_t rr = api->CommInitAll(g->comms.data(), n, devs.data());
            if (rr != ncclSuccess) { g->comms.clear(); return fail(FHE_ECUDA, "ncclCommInitAll failed: %s", api->GetErrorString(rr)); }
        }
    }
    std::vector<cudaStream_t> st(n);
    for (int i = 0; i < n; i++) {
        if (stream == FHE_STREAM_LEGACY) st[i] = cudaStreamLegacy;
        else if (stream < 0 || stream >= (int)g->ctx[i]->streams.size()) return fail(FHE_EINVAL, "fhe_group_all_gather: bad stream index %d", stream);
        else st[i] = g->ctx[i]->streams[stream];
    }
    const bool even = batch % n == 0;
    NCCL_TRY(api, api->GroupStart());
    for (int i = 0; i < n; i++) {
        cudaSetDevice(g->ctx[i]->device);
        if (even) {
            ncclResult_t rr = api->AllGather((const void*)(uintptr_t)src[i], (void*)(uintptr_t)dst[i], batch / n * item, ncclUint64, g->comms[i], st[i]);
            if (rr != ncclSuccess) { api->GroupEnd(); return fail(FHE_ECUDA, "ncclAllGather failed: %s", api->GetErrorString(rr)); }
            continue;
        }
        size_t myfirst, mycount;
        partition(batch, n, i, &myfirst, &mycount);
        for (int p = 0; p < n; p++) {
            size_t pfirst, pcount;
            partition(batch, n, p, &pfirst, &pcount);
            ncclResult_t rr = ncclSuccess;
            if (p == i) {
                if (mycount && cudaMemcpyAsync((void*)(uintptr_t)(dst[i] + myfirst * item * 8), (const void*)(uintptr_t)src[i], mycount * item * 8,
                                               cudaMemcpyDeviceToDevice, st[i]) != cudaSuccess) rr = ncclUnhandledCudaError;
            } else {
                if (mycount) rr = api->Send((const void*)(uintptr_t)src[i], mycount * item, ncclUint64, p, g->comms[i], st[i]);
                if (rr == ncclSuccess && pcount)
                    rr = api->Recv((void*)(uintptr_t)(dst[i] + pfirst * item * 8), pcount * item, ncclUint64, p, g->comms[i], st[i]);
            }
            if (rr != ncclSuccess) { api->GroupEnd(); return fail(FHE_ECUDA, "NCCL send/recv failed: %s", api->GetErrorString(rr)); }
        }
    }
    NCCL_TRY(api, api->GroupEnd());
    for (int i = 0; i < n; i++) {
        int slot = stream == FHE_STREAM_LEGACY ? (int)g->ctx[i]->streams.size() : stream;
        g->ctx[i]->epoch[slot].fetch_add(1, std::memory_order_release);
    }
    return FHE_OK;
}

// The in-place operator sequence on HOST operands over all devices: each device streams its block of the batch through
// its own pipeline (fhe_dcrt_chain_host), one host thread per device.  Blocking.
extern "C" int fhe_group_chain_host(fhe_group_ring* r, const int* ops, int nops, uint64_t* a, const uint64_t* b, size_t batch) {
    // no GateHold here: the per-device worker threads below enter the gate themselves, and a parent that waited for them
    // while counted inside would deadlock with a capture that started in between (CaptureGate)
    if (!r) return fail(FHE_EINVAL, "null group ring");
    const int n = (int)r->rings.size();
    if (n == 1) return fhe_dcrt_chain_host(r->rings[0], ops, nops, a, b, batch);
    std::vector<int> rcs(n, FHE_OK);
    std::vector<std::string> msgs(n);
    std::vector<std::thread> th;
    for (int i = 0; i < n; i++)
        th.emplace_back([&, i] {
            size_t first, count;
            partition(batch, n, i, &first, &count);
            if (!count) return;
            size_t item = (size_t)r->rings[i]->dev.T * r->rings[i]->dev.N;
            rcs[i] = fhe_dcrt_chain_host(r->rings[i], ops, nops, a + first * item, b ? b + first * item : nullptr, count);
            if (rcs[i]) msgs[i] = fhe_last_error();
        });
    for (auto& t : th) t.join();
    for (int i = 0; i < n; i++)
        if (rcs[i]) return fail(rcs[i], "device %d: %s", r->g->ctx[i]->device, msgs[i].c_str());
    return FHE_OK;
}
