// C ABI of the LD-score engine (include/nldsc_ld.h): shard planning, device buffers, the
// chunked  upload -> count -> stats -> unpack -> band GEMM + epilogue -> finalize  pipeline,
// and the multi-GPU driver (contiguous SNP shards with a window halo, no collective).
//
// Replaces calculate() (reference ldscalc.h:8-65) and ChunkwiseReader (stream.h:106-198):
// the reference slides a cache over the .bed one SNP at a time; here the host computes
// every SNP's window up front (compute_windows), cuts the band into 64x64 SNP-pair tiles
// and streams SNP rows through a ring of unpacked genotype planes in HBM.
#include "session.cuh"

#include <algorithm>
#include <atomic>
#include <cerrno>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <fcntl.h>
#include <limits>
#include <memory>
#include <mutex>
#include <string>
#include <sys/stat.h>
#include <thread>
#include <unistd.h>

#include <nvtx3/nvToolsExt.h>      // header-only NVTX v3: ranges cost nothing unless a profiler is attached

using namespace nldsc;

namespace nldsc {

void compute_windows(const double* pos, int32_t n_snp, double wind, std::vector<int32_t>& lo,
                     std::vector<int32_t>& hi) {
    lo.resize(n_snp);
    hi.resize(n_snp);
    if (n_snp <= 0) return;
    // SNPFilter::filter: std::abs(positions[snd] - positions[fst]) <= ld_wind  (tools.h:43-44).  For sorted
    // positions the predicate is monotone in either argument (IEEE subtraction is monotone and
    // antisymmetric), so lo[] comes from one two-pointer sweep and hi[] is its inverse:
    // hi[i] = #{j : lo[j] <= i} - 1.  The sweep is split over a few threads (it sits on the critical
    // path of every call: nothing can be scored before the band plan exists).
    auto inside = [&](int32_t l, int32_t i) { return std::abs(pos[l] - pos[i]) <= wind; };
    auto sweep = [&](int32_t a, int32_t b) {
        // first l <= a inside the window of a, by bisection on the monotone predicate
        int32_t x = 0, y = a;
        while (x < y) {
            const int32_t m = x + (y - x) / 2;
            if (inside(m, a)) y = m; else x = m + 1;
        }
        int32_t l = x;
        for (int32_t i = a; i < b; ++i) {
            while (l < i && !inside(l, i)) ++l;
            lo[i] = l;
        }
    };
    const int hw = static_cast<int>(std::thread::hardware_concurrency());
    const int nt = n_snp < 32768 ? 1 : std::max(1, std::min(8, hw));
    if (nt == 1) {
        sweep(0, n_snp);
    } else {
        std::vector<std::thread> th;
        for (int t = 0; t < nt; ++t) {
            const int32_t a = static_cast<int32_t>(static_cast<int64_t>(n_snp) * t / nt);
            const int32_t b = static_cast<int32_t>(static_cast<int64_t>(n_snp) * (t + 1) / nt);
            th.emplace_back(sweep, a, b);
        }
        for (auto& t : th) t.join();
    }
    std::vector<int32_t> cnt(static_cast<size_t>(n_snp) + 1, 0);
    for (int32_t j = 0; j < n_snp; ++j) ++cnt[lo[j]];
    int32_t run = 0;
    for (int32_t i = 0; i < n_snp; ++i) {
        run += cnt[i];
        hi[i] = run - 1;
    }
}

// ---- device workspace cache ------------------------------------------------------------------
namespace {
constexpr int kMaxDevices = 64;
std::mutex g_arena_mu;
Arena* g_arena[kMaxDevices] = {nullptr};
bool g_arena_busy[kMaxDevices] = {false};
}  // namespace

void* Arena::alloc(size_t bytes) {
    bytes = (bytes + 511) & ~size_t(511);
    Block* best = nullptr;          // best fit: small buffers must not nibble at the block the plane ring needs
    for (Block& b : blocks)
        if (b.size - b.used >= bytes && (!best || b.size - b.used < best->size - best->used)) best = &b;
    if (best) {
        void* p = best->base + best->used;
        best->used += bytes;
        return p;
    }
    // small requests share 64 MB blocks; large ones (packed rows, plane ring) get their own
    const size_t sz = std::max<size_t>(bytes, size_t(64) << 20);
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, sz);
    if (e == cudaErrorMemoryAllocation) {
        // cached blocks of an earlier, differently shaped call that nothing of this session lives in are only
        // holding memory: give them back and try once more
        cudaGetLastError();
        bool freed = false;
        for (size_t i = 0; i < blocks.size();) {
            if (blocks[i].used == 0) {
                cudaFree(blocks[i].base);
                blocks.erase(blocks.begin() + static_cast<long>(i));
                freed = true;
            } else {
                ++i;
            }
        }
        if (freed) e = cudaMalloc(&p, sz);
    }
    CUDA_CHECK(e);
    blocks.push_back(Block{static_cast<char*>(p), sz, bytes});
    return p;
}

size_t Arena::largest_free() const {
    size_t m = 0;
    for (const Block& b : blocks) m = std::max(m, b.size - b.used);
    return m;
}

size_t Arena::capacity() const {
    size_t c = 0;
    for (const Block& b : blocks) c += b.size;
    return c;
}

Arena* arena_acquire(int device) {
    std::lock_guard<std::mutex> g(g_arena_mu);
    if (device >= 0 && device < kMaxDevices && !g_arena_busy[device]) {
        if (!g_arena[device]) {
            g_arena[device] = new Arena;
            g_arena[device]->device = device;
            g_arena[device]->cached = true;
        }
        g_arena_busy[device] = true;
        return g_arena[device];
    }
    Arena* a = new Arena;
    a->device = device;
    return a;
}

static void arena_free_blocks(Arena* a) {
    for (Arena::Block& b : a->blocks) cudaFree(b.base);
    a->blocks.clear();
}

void arena_release(Arena* a) {
    if (!a) return;
    if (!a->cached) {
        arena_free_blocks(a);
        delete a;
        return;
    }
    // A workload of a different shape leaves blocks that no longer fit: keep at most the blocks in use
    // by the last session plus nothing else (blocks the session never touched are dropped).
    std::vector<Arena::Block> keep;
    for (Arena::Block& b : a->blocks) {
        if (b.used) { b.used = 0; keep.push_back(b); }
        else cudaFree(b.base);
    }
    a->blocks.swap(keep);
    std::lock_guard<std::mutex> g(g_arena_mu);
    g_arena_busy[a->device] = false;
}

size_t arena_cached_bytes(int device) {
    std::lock_guard<std::mutex> g(g_arena_mu);
    if (device < 0 || device >= kMaxDevices || !g_arena[device] || g_arena_busy[device]) return 0;
    return g_arena[device]->capacity();
}

void arena_trim(int device) {
    std::lock_guard<std::mutex> g(g_arena_mu);
    for (int d = 0; d < kMaxDevices; ++d) {
        if ((device >= 0 && d != device) || !g_arena[d] || g_arena_busy[d]) continue;
        cudaSetDevice(d);
        arena_free_blocks(g_arena[d]);
    }
}

// ---- pinned staging cache ---------------------------------------------------------------------
namespace {
HostStage* g_stage[kMaxDevices] = {nullptr};
bool g_stage_busy[kMaxDevices] = {false};

HostStage* host_stage_new(int device) {
    HostStage* h = new HostStage;
    h->device = device;
    h->bytes = size_t(HostStage::kSlots) * (size_t(16) << 20);
    if (cudaHostAlloc(&h->base, h->bytes, cudaHostAllocDefault) != cudaSuccess) {
        delete h;
        throw CudaError("cudaHostAlloc of the .bed staging ring failed");
    }
    for (auto& e : h->ev) {
        CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    return h;
}
}  // namespace

HostStage* host_stage_acquire(int device) {
    {
        std::lock_guard<std::mutex> g(g_arena_mu);
        if (device >= 0 && device < kMaxDevices && !g_stage_busy[device]) {
            g_stage_busy[device] = true;
            if (g_stage[device]) return g_stage[device];
        } else {
            device = -1 - device;       // private stage
        }
    }
    if (device >= 0) {
        HostStage* h = nullptr;
        try {
            h = host_stage_new(device);
        } catch (...) {
            std::lock_guard<std::mutex> g(g_arena_mu);
            g_stage_busy[device] = false;
            throw;
        }
        h->cached = true;
        std::lock_guard<std::mutex> g(g_arena_mu);
        g_stage[device] = h;
        return h;
    }
    return host_stage_new(-1 - device);
}

void host_stage_release(HostStage* h) {
    if (!h) return;
    if (h->cached) {
        std::lock_guard<std::mutex> g(g_arena_mu);
        g_stage_busy[h->device] = false;
        return;
    }
    for (auto& e : h->ev) cudaEventDestroy(e);
    cudaFreeHost(h->base);
    delete h;
}

namespace {
PlanStage* g_pstage[kMaxDevices] = {nullptr};
bool g_pstage_busy[kMaxDevices] = {false};
}  // namespace

void PlanStage::reserve(size_t need) {
    if (need <= bytes) return;
    if (base) cudaFreeHost(base);
    base = nullptr;
    bytes = 0;
    const size_t want = need + need / 4 + 4096;
    void* q = nullptr;
    if (cudaHostAlloc(&q, want, cudaHostAllocPortable) != cudaSuccess) throw CudaError("cudaHostAlloc of the plan staging buffer failed");
    base = static_cast<char*>(q);
    bytes = want;
}

PlanStage* plan_stage_acquire(int device) {
    std::lock_guard<std::mutex> g(g_arena_mu);
    if (device >= 0 && device < kMaxDevices && !g_pstage_busy[device]) {
        g_pstage_busy[device] = true;
        if (!g_pstage[device]) {
            g_pstage[device] = new PlanStage;
            g_pstage[device]->device = device;
            g_pstage[device]->cached = true;
        }
        return g_pstage[device];
    }
    PlanStage* h = new PlanStage;       // the device's cached buffer is in use: a private one
    h->device = device;
    return h;
}

void plan_stage_release(PlanStage* h) {
    if (!h) return;
    if (h->cached) {
        std::lock_guard<std::mutex> g(g_arena_mu);
        g_pstage_busy[h->device] = false;
        return;
    }
    if (h->base) cudaFreeHost(h->base);
    delete h;
}

namespace {

void set_err(char* err, size_t errlen, const char* msg) {
    if (err && errlen) {
        strncpy(err, msg, errlen - 1);
        err[errlen - 1] = 0;
    }
}

template <class F>
int guarded(char* err, size_t errlen, F&& f) {
    try {
        f();
        return NLDSC_OK;
    } catch (const std::invalid_argument& e) {
        set_err(err, errlen, e.what());
        return NLDSC_ERR_INVALID;
    } catch (const std::exception& e) {
        set_err(err, errlen, e.what());
        return NLDSC_ERR_RUNTIME;
    } catch (...) {
        set_err(err, errlen, "unknown error");
        return NLDSC_ERR_RUNTIME;
    }
}

int usable_devices() {
    int n = 0;
    const cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        cudaGetLastError();
        throw CudaError(std::string("no usable CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e));
    }
    if (n <= 0) throw CudaError("no usable CUDA device (there is no CPU fallback)");
    return n;
}

// compute capability and SM count per device, queried once (cudaGetDeviceProperties costs milliseconds)
struct DevInfo { int major = -1, minor = 0, sms = 0; };
DevInfo& dev_info(int device) {
    static DevInfo info[64];
    static std::mutex mu;
    std::lock_guard<std::mutex> g(mu);
    DevInfo& d = info[device & 63];
    if (d.major < 0) {
        CUDA_CHECK(cudaDeviceGetAttribute(&d.major, cudaDevAttrComputeCapabilityMajor, device));
        CUDA_CHECK(cudaDeviceGetAttribute(&d.minor, cudaDevAttrComputeCapabilityMinor, device));
        CUDA_CHECK(cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, device));
    }
    return d;
}

void require_sm100(int device) {
    const DevInfo& d = dev_info(device);
    if (d.major != 10) {
        char buf[160];
        snprintf(buf, sizeof(buf), "device %d is sm_%d%d; this library is built for sm_100a only", device, d.major,
                 d.minor);
        throw CudaError(buf);
    }
}

int device_sm_count(int device) { return dev_info(device).sms; }

// CUDA loads kernel images lazily, at a kernel's first launch, and that load can wait for the device to go idle.  The
// fused band launch keeps the device busy until the unpack stream has run -- so every kernel that stream launches must
// be resident BEFORE the first fused launch, or its first launch deadlocks against the kernel that is waiting for it.
void preload_kernels(int device) {
    static std::mutex mu;
    static bool done[64] = {false};
    std::lock_guard<std::mutex> g(mu);
    if (done[device & 63]) return;
    install_band_kernels();
    preload_unpack_kernels();
    preload_plan_kernels();
    preload_finalize_kernels();
    done[device & 63] = true;
}

void validate_params(const nldsc_ld_params* p, bool allow_unused = false) {
    if (!p) throw InvalidArgument("params is null");
    if (p->n_snp <= 0) throw InvalidArgument("n_snp must be positive");
    if (p->n_org <= 0) throw InvalidArgument("n_org must be positive");
    if (!p->positions) throw InvalidArgument("positions is null");
    if (!(p->ld_wind >= 0.0)) throw InvalidArgument("ld_wind must be non-negative");
    double last = -1.0;
    for (int32_t i = 0; i < p->n_snp; ++i) {
        const double x = p->positions[i];
        if (x != x) throw InvalidArgument("positions must not be NaN");
        if (x < 0.0) {
            // the reference treats position < 0 as "SNP not used" (tools.h:16-24): never read, in nobody's window
            if (!allow_unused) throw InvalidArgument("positions must be non-negative");
            continue;
        }
        // unlike the reference, which walks whatever order it is given (undefined results), the band plan
        // needs the used SNPs in map order
        if (x < last) throw InvalidArgument("positions of the used SNPs must be non-decreasing (sort the .bim by position; the reference does not check this)");
        last = x;
    }
}

// SNPs with a negative position are "unused" (reference tools.h:16-24, stream.h:189-194, ldscalc.h:24,57): skipped in
// the .bed, members of no window, outputs left at NaN / -1 (MAF too: never decoded).  Here they stay rows of the
// band plan -- dropped rows, like SNPs that fail the MAF filter -- and get the position of the nearest used SNP before
// them so that the windows of the others can be found by bisection.  `eff` is empty when no SNP is unused.
struct UnusedView {
    std::vector<double> eff;
    std::vector<uint8_t> unused;
    bool any() const { return !unused.empty(); }
};
UnusedView mark_unused(const nldsc_ld_params* p) {
    UnusedView v;
    bool any = false;
    for (int32_t i = 0; i < p->n_snp && !any; ++i) any = p->positions[i] < 0.0;
    if (!any) return v;
    v.eff.assign(p->positions, p->positions + p->n_snp);
    v.unused.assign(static_cast<size_t>(p->n_snp), 0);
    double fill = 0.0;
    for (int32_t i = 0; i < p->n_snp; ++i) if (v.eff[i] >= 0.0) { fill = v.eff[i]; break; }
    for (int32_t i = 0; i < p->n_snp; ++i) {
        if (v.eff[i] < 0.0) { v.eff[i] = fill; v.unused[i] = 1; }
        else fill = v.eff[i];
    }
    return v;
}

template <class T>
T* dmalloc(size_t n) {          // parity / measurement hooks only; sessions allocate from their Arena
    void* p = nullptr;
    CUDA_CHECK(cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T)));
    return static_cast<T*>(p);
}

template <class T>
T* amalloc(Arena* a, size_t n) {
    return static_cast<T*>(a->alloc(std::max<size_t>(n, 1) * sizeof(T)));
}

int32_t floor64(int32_t x) { return x & ~63; }

cudaEvent_t ev_get(int device, bool timing);     // per-device pools, defined below
void preload_kernels(int device);

// NVTX range over a host phase of a call: nsys / ncu timelines show open / upload / plan / run /
// download per shard, and one range per chunk inside run.
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};

struct SpanScope {
    nldsc_ld_session* s;
    cudaStream_t st;
    size_t idx;
    SpanScope(nldsc_ld_session* s_, cudaStream_t st_, int family) : s(s_), st(st_) {
        if (s->spans_used == s->spans.size()) {
            nldsc_ld_session::Span sp;
            sp.a = ev_get(s->device, true);
            sp.b = ev_get(s->device, true);
            sp.family = family;
            s->spans.push_back(sp);
        }
        idx = s->spans_used++;
        s->spans[idx].family = family;
        CUDA_CHECK(cudaEventRecord(s->spans[idx].a, st));
    }
    ~SpanScope() { cudaEventRecord(s->spans[idx].b, st); }
};

// NLDSC_TRACE=1: host-side phase times on stderr
struct Tracer {
    bool on = getenv("NLDSC_TRACE") != nullptr;
    std::chrono::steady_clock::time_point t = std::chrono::steady_clock::now();
    void lap(const char* what) {
        if (!on) return;
        const auto n = std::chrono::steady_clock::now();
        fprintf(stderr, "[nldsc_ld]   %-18s %.3f ms\n", what, std::chrono::duration<double, std::milli>(n - t).count());
        t = n;
    }
};

size_t env_size(const char* name, size_t dflt) {
    const char* v = getenv(name);
    if (!v || !*v) return dflt;
    char* end = nullptr;
    const double d = strtod(v, &end);
    if (end == v || !(d > 0)) return dflt;
    return static_cast<size_t>(d);
}

}  // namespace

// ---- per-device pools of streams and events --------------------------------------------------------
// A calculate call creates ~100 events and three streams; creating and destroying them costs ~0.5 ms per
// call, which is 7 % of an 7 ms chromosome.  Sessions borrow them here and hand them back idle.
namespace {
struct DevPool {
    std::vector<cudaEvent_t> ev[2];          // [0] cudaEventDisableTiming, [1] timing
    std::vector<cudaStream_t> st[3];         // [0] high priority, [1] default, [2] low priority
};
DevPool g_pool[kMaxDevices];

cudaEvent_t ev_get(int device, bool timing) {
    {
        std::lock_guard<std::mutex> g(g_arena_mu);
        auto& v = g_pool[device & (kMaxDevices - 1)].ev[timing ? 1 : 0];
        if (!v.empty()) { cudaEvent_t e = v.back(); v.pop_back(); return e; }
    }
    cudaEvent_t e = nullptr;
    if (timing) CUDA_CHECK(cudaEventCreate(&e));
    else CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    return e;
}
void ev_put(int device, bool timing, cudaEvent_t e) {
    if (!e) return;
    std::lock_guard<std::mutex> g(g_arena_mu);
    g_pool[device & (kMaxDevices - 1)].ev[timing ? 1 : 0].push_back(e);
}
cudaStream_t stream_get(int device, int kind) {
    {
        std::lock_guard<std::mutex> g(g_arena_mu);
        auto& v = g_pool[device & (kMaxDevices - 1)].st[kind];
        if (!v.empty()) { cudaStream_t t = v.back(); v.pop_back(); return t; }
    }
    int prio_lo = 0, prio_hi = 0;
    CUDA_CHECK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    cudaStream_t t = nullptr;
    if (kind == 1) CUDA_CHECK(cudaStreamCreateWithFlags(&t, cudaStreamNonBlocking));
    else CUDA_CHECK(cudaStreamCreateWithPriority(&t, cudaStreamNonBlocking, kind == 0 ? prio_hi : prio_lo));
    return t;
}
void stream_put(int device, int kind, cudaStream_t t) {     // the stream must be idle
    if (!t) return;
    std::lock_guard<std::mutex> g(g_arena_mu);
    g_pool[device & (kMaxDevices - 1)].st[kind].push_back(t);
}
}  // namespace

// ---- session ------------------------------------------------------------------------------

static void session_free(nldsc_ld_session* s) {
    if (!s) return;
    cudaSetDevice(s->device);
    for (auto& t : s->readers) if (t.joinable()) t.join();
    if (s->fd_direct >= 0) close(s->fd_direct);
    // every buffer lives in the arena; make sure nothing queued still touches it before it is reused
    if (s->own_stream) cudaStreamSynchronize(s->own_stream);
    if (s->copy_stream) cudaStreamSynchronize(s->copy_stream);
    if (s->unpack_stream) cudaStreamSynchronize(s->unpack_stream);
    if (s->band2_stream) cudaStreamSynchronize(s->band2_stream);
    arena_release(s->arena);
    host_stage_release(s->stage);
    plan_stage_release(s->pstage);
    ev_put(s->device, false, s->ev_planned);
    ev_put(s->device, true, s->ev_up0);
    ev_put(s->device, true, s->ev_up1);
    if (s->maps) band_maps_destroy(s->maps);
    for (auto& sp : s->spans) { ev_put(s->device, true, sp.a); ev_put(s->device, true, sp.b); }
    for (auto e : s->ev_piece) ev_put(s->device, false, e);
    for (auto e : s->ev_unpacked) ev_put(s->device, false, e);
    for (auto e : s->ev_band) ev_put(s->device, false, e);
    ev_put(s->device, false, s->ev_fork);
    stream_put(s->device, 2, s->unpack_stream);
    stream_put(s->device, 0, s->band2_stream);
    ev_put(s->device, false, s->ev_join);
    ev_put(s->device, true, s->ev_run0);
    ev_put(s->device, true, s->ev_run1);
    stream_put(s->device, 0, s->own_stream);
    stream_put(s->device, 1, s->copy_stream);
    delete s;
}

// Stage 1 of a session: device, geometry, the packed-row buffer and the copy machinery -- everything
// the H2D upload needs, so the upload can start before the (host-side) band plan is built.
static nldsc_ld_session* session_open(int device, const nldsc_ld_params* p, int32_t load_begin, int32_t load_end) {
    if (load_begin < 0 || load_end > p->n_snp || load_begin >= load_end || (load_begin % kBlk) != 0)
        throw InvalidArgument("shard load range is empty, unaligned or outside [0, n_snp)");
    const int ndev = usable_devices();
    if (device < 0 || device >= ndev) throw CudaError("CUDA device index out of range");
    CUDA_CHECK(cudaSetDevice(device));
    require_sm100(device);
    preload_kernels(device);

    std::unique_ptr<nldsc_ld_session, void (*)(nldsc_ld_session*)> sp(new nldsc_ld_session, session_free);
    nldsc_ld_session* s = sp.get();
    s->device = device;
    s->arena = arena_acquire(device);
    s->pstage = plan_stage_acquire(device);
    s->sm_count = device_sm_count(device);
    s->n_org = p->n_org;
    s->n_snp = p->n_snp;
    s->maf_thr = p->maf;
    s->std_thr = p->std_thr;
    s->rsq_thr = p->rsq_thr;
    s->bps = (static_cast<int64_t>(p->n_org) + 3) / 4;
    s->npad = static_cast<int32_t>(((static_cast<int64_t>(p->n_org) + kKAlign - 1) / kKAlign) * kKAlign);    // int8 planes; fp4: below
    const char* band = getenv("NLDSC_BAND");
    s->use_simt = band && strcmp(band, "simt") == 0;
    // two j-blocks per supertile pay once the MMA time of a supertile outweighs the (then exposed) epilogue
    const char* wd = getenv("NLDSC_WIDE");
    // Measured: the wide kernel wins at every N tried (N = 2,504: 1.39 vs 1.60 ms per 100,000 SNPs; 5,000: 2.13 vs 2.41;
    // 10,000: 3.65 vs 4.33; N >= 100,000: 10-35 %), also where the epilogue dominates -- it has twice the epilogue warps.
    // NLDSC_WIDE=0 selects the narrow, double-buffered kernel (kept as the second implementation the tests compare).
    s->wide = !s->use_simt && (wd ? wd[0] == '1' : true);
    // missing-free chunks take the 2-plane kernel (a third of the MMAs); decided on the device, chunk by chunk
    const char* mf = getenv("NLDSC_MFREE");
    s->mfree = s->wide && !(mf && mf[0] == '0');
    // fp4 planes (e2m1 nibbles, two organisms per byte) for the wide / missing-free bodies: tcgen05.mma.kind::mxf4 multiplies
    // 64 organisms per instruction at twice the int8 rate from half the operand bytes and sums these 0 / 1 / 2 values exactly
    // (counts stay below 2 N < 2^22).  The narrow kernel and the SIMT yardstick keep int8 planes (NLDSC_FP4=0: so does the wide one).
    const char* f4 = getenv("NLDSC_FP4");
    s->fp4 = s->wide && !(f4 && f4[0] == '0') && p->n_org <= 2000000;
    // 160-column supertiles for the 6-plane work on fp4 planes: an mxf4 MMA costs the same whatever its N, and five 32-SNP column
    // groups take two of them (N = 256 + 224) where two 64-SNP j-blocks take two for 128 SNPs.  Measured against the 2 x 2-block
    // supertiles, same box, bitwise equal (profiles/r02za_x160_*.json): N=2,504 +3 % (its accumulator is one piece, the epilogue
    // fully exposed), N=5,000 -2.6 %, N=7,000 -5.7 %, N=10,000 -7.2 %, N=50,000 -7.5 %, N=315,599 -7 %, N=100,000 with a 10 cM
    // window -13 %.  Default from N = 4,096 (where the chunks shrink to 8,192 rows); NLDSC_X160=0 / 1 forces.
    const char* x16 = getenv("NLDSC_X160");
    s->x160 = s->fp4 && (x16 ? x16[0] == '1' : p->n_org >= 4096);
    if (s->fp4)      // plane row BYTES: a whole number of 128-byte pipeline stages = 256 organisms each
        s->npad = static_cast<int32_t>(((static_cast<int64_t>(p->n_org) + 2 * kKAlign - 1) / (2 * kKAlign)) * kKAlign);
    s->load_begin = load_begin;
    s->load_end = load_end;
    s->n_load = load_end - load_begin;

    s->d_packed = amalloc<uint8_t>(s->arena, static_cast<size_t>(s->n_load) * s->bps + 16);
    // the band kernel's CTAs go first; unpack blocks fill the thread slots that are left on every SM
    s->own_stream = stream_get(device, 0);
    s->copy_stream = stream_get(device, 1);
    s->unpack_stream = stream_get(device, 2);
    s->band2_stream = stream_get(device, 0);
    s->ev_join = ev_get(device, false);
    s->ev_fork = ev_get(device, false);
    // the 16 bytes of slack the word-wise kernels may read past the last row
    CUDA_CHECK(cudaMemsetAsync(s->d_packed + static_cast<size_t>(s->n_load) * s->bps, 0, 16, s->copy_stream));
    // upload granularity: ~8 MB pieces, a whole number of 64-row blocks each
    const int64_t rows8 = static_cast<int64_t>(env_size("NLDSC_PIECE_BYTES", size_t(8) << 20)) / s->bps;
    s->piece_rows = static_cast<int32_t>(std::max<int64_t>(kBlk, rows8 / kBlk * kBlk));
    s->ev_piece.resize((s->n_load + s->piece_rows - 1) / s->piece_rows);
    for (auto& e : s->ev_piece) e = ev_get(device, false);
    s->ev_run0 = ev_get(device, true);
    s->ev_run1 = ev_get(device, true);
    s->ev_up0 = ev_get(device, true);
    s->ev_up1 = ev_get(device, true);
    return sp.release();
}

static void note_pieces_issued(nldsc_ld_session* s, int32_t n, const char* error = nullptr) {
    {
        std::lock_guard<std::mutex> g(s->up_mu);
        s->pieces_issued = n;
        if (error) s->up_error = error;
    }
    s->up_cv.notify_all();
}

// H2D of the packed rows from host memory, one copy + event per piece so run() can start on the
// first chunk while the rest of the shard is still crossing PCIe.  host_rows points at row load_begin.
static void session_upload_mem(nldsc_ld_session* s, const uint8_t* host_rows, cudaStream_t st) {
    CUDA_CHECK(cudaSetDevice(s->device));
    cudaStream_t cs = st ? st : s->copy_stream;
    const int32_t n_pieces = static_cast<int32_t>(s->ev_piece.size());
    CUDA_CHECK(cudaEventRecord(s->ev_up0, cs));
    for (int32_t k = 0; k < n_pieces; ++k) {
        const int32_t r0 = k * s->piece_rows, r1 = std::min(s->n_load, r0 + s->piece_rows);
        const size_t off = static_cast<size_t>(r0) * s->bps, len = static_cast<size_t>(r1 - r0) * s->bps;
        CUDA_CHECK(cudaMemcpyAsync(s->d_packed + off, host_rows + off, len, cudaMemcpyHostToDevice, cs));
        CUDA_CHECK(cudaEventRecord(s->ev_piece[k], cs));
    }
    CUDA_CHECK(cudaEventRecord(s->ev_up1, cs));
    note_pieces_issued(s, n_pieces);
    s->upload_started = true;
}

// H2D straight from the .bed: a reader thread preads piece after piece into a small ring of pinned
// slots and queues the copy of each; the caller's thread meanwhile plans and enqueues the kernels,
// which wait (on the device) for the piece events.  Replaces the byte-at-a-time ifstream of
// BedStreamReader (reference stream.h:43-69) -- SURVEY section 8(f)-1.
static void session_upload_file(nldsc_ld_session* s, int fd) {
    CUDA_CHECK(cudaSetDevice(s->device));
    s->stage = host_stage_acquire(s->device);
    s->upload_started = true;
    const int32_t n_pieces = static_cast<int32_t>(s->ev_piece.size());
    s->piece_done.assign(n_pieces, 0);
    // NLDSC_READERS: reader threads of this shard (default min(8, cores / 2), at most 8).  A driver that keeps several GPUs busy
    // from one host sets it to cores / GPUs, so that the readers of all shards together do not oversubscribe the host.
    const size_t hw = std::max<size_t>(1, std::thread::hardware_concurrency());
    const size_t dflt_readers = std::min<size_t>(HostStage::kReaders, std::max<size_t>(2, hw / 2));
    const int want_readers = static_cast<int>(std::min<size_t>(HostStage::kReaders, std::max<size_t>(1, env_size("NLDSC_READERS", dflt_readers))));
    const int n_readers = std::max(1, std::min<int>(want_readers, n_pieces));
    // NLDSC_ODIRECT=1: read the .bed past the page cache (O_DIRECT), for files that are read once and are larger than
    // the host's memory can usefully cache (a 94.7 GB UK-Biobank-scale genome).  The byte range of a slot is widened
    // to 4 KB boundaries (the pinned slots are page-aligned and carry 8 KB of slack); when the file system refuses
    // O_DIRECT the buffered descriptor is used.
    int fd_direct = -1;
    if (env_size("NLDSC_ODIRECT", 0) != 0 && s->bed_path) fd_direct = open(s->bed_path, O_RDONLY | O_DIRECT);
    s->fd_direct = fd_direct;
    for (int t = 0; t < n_readers; ++t) {
        s->readers.emplace_back([s, fd, fd_direct, t, n_readers, n_pieces] {
            auto publish = [&](int32_t k, const char* error) {        // piece k is queued (or the upload has failed)
                {
                    std::lock_guard<std::mutex> g(s->up_mu);
                    if (error) {
                        if (s->up_error.empty()) s->up_error = error;
                        s->pieces_issued = n_pieces;                    // unblock run(); the error surfaces there / at download
                    } else {
                        s->piece_done[k] = 1;
                        while (s->pieces_issued < n_pieces && s->piece_done[s->pieces_issued]) ++s->pieces_issued;
                    }
                }
                s->up_cv.notify_all();
            };
            try {
                CUDA_CHECK(cudaSetDevice(s->device));
                const size_t slot_cap = s->stage->bytes / HostStage::kSlots;
                const size_t slot_bytes = slot_cap - 8192;       // payload; the rest is the O_DIRECT alignment slack
                int flip = 0;
                // test hook: hold reader 0 back so that later pieces are queued before earlier ones
                const useconds_t lag = t == 0 ? static_cast<useconds_t>(env_size("NLDSC_TEST_READER0_LAG_US", 0)) : 0;
                for (int32_t k = t; k < n_pieces; k += n_readers) {
                    const int32_t r0 = k * s->piece_rows, r1 = std::min(s->n_load, r0 + s->piece_rows);
                    size_t off = static_cast<size_t>(r0) * s->bps;
                    const size_t end = static_cast<size_t>(r1) * s->bps;
                    while (off < end) {                      // a piece may span several slots (very long rows)
                        const int slot = 2 * t + flip;
                        flip ^= 1;
                        const size_t len = std::min(slot_bytes, end - off);
                        char* buf = static_cast<char*>(s->stage->base) + slot * slot_cap;
                        CUDA_CHECK(cudaEventSynchronize(s->stage->ev[slot]));       // slot's previous copy has landed
                        if (lag) usleep(lag);
                        const off_t file_off = 3 + static_cast<off_t>(s->load_begin) * s->bps + static_cast<off_t>(off);
                        if (fd_direct >= 0) {
                            // aligned superset [a0, a1) of [file_off, file_off + len); the last block may end at EOF
                            const off_t a0 = file_off & ~static_cast<off_t>(4095);
                            const off_t a1 = (file_off + static_cast<off_t>(len) + 4095) & ~static_cast<off_t>(4095);
                            size_t got = 0;
                            const size_t need = static_cast<size_t>(file_off - a0) + len;
                            while (got < need) {
                                const ssize_t r = pread(fd_direct, buf + got, static_cast<size_t>(a1 - a0) - got, a0 + static_cast<off_t>(got));
                                if (r <= 0) throw CudaError("short O_DIRECT read on .bed file");
                                got += static_cast<size_t>(r);
                                if ((got & 4095) != 0 && got < need) throw CudaError("unaligned O_DIRECT read on .bed file");
                            }
                            buf += file_off - a0;
                        } else {
                            size_t got = 0;
                            while (got < len) {
                                const ssize_t r = pread(fd, buf + got, len - got, file_off + static_cast<off_t>(got));
                                if (r <= 0) throw CudaError("short read on .bed file");
                                got += static_cast<size_t>(r);
                            }
                        }
                        CUDA_CHECK(cudaMemcpyAsync(s->d_packed + off, buf, len, cudaMemcpyHostToDevice, s->copy_stream));
                        CUDA_CHECK(cudaEventRecord(s->stage->ev[slot], s->copy_stream));
                        off += len;
                    }
                    CUDA_CHECK(cudaEventRecord(s->ev_piece[k], s->copy_stream));
                    publish(k, nullptr);
                }
            } catch (const std::exception& e) {
                publish(-1, e.what());
            }
        });
    }
}

static void join_reader(nldsc_ld_session* s) {
    for (auto& t : s->readers) if (t.joinable()) t.join();
    if (s->fd_direct >= 0) { close(s->fd_direct); s->fd_direct = -1; }
}

// The 160-column supertiles (XTile) of one chunk = i-blocks [b, e) of the shard's block list: every i-pair's j-range
// [jb_lo[k], jb_hi[k+1]] (64-SNP blocks) is cut into runs of five 32-SNP groups from its own first column -- nothing is
// aligned to more than 32 SNPs, so only the last run of an i-pair is partly outside its range (n_grp < 5).  Same partial-sum
// slots as the 64-block tiles (slot0[p] = first slot of the chunk's p-th i-pair); executed in raster order: run m of `raster`
// consecutive i-pairs side by side, so that concurrent CTA pairs share operand rows in L2.  Host-only (CPU-tested through
// nldsc_ld_debug_plan_xtiles).
static void plan_xtiles_chunk(const int32_t* jb_lo, const int32_t* jb_hi, int32_t b, int32_t e, int32_t ib0, int64_t cap,
                              int32_t raster, const std::vector<int32_t>& slot0, std::vector<XTile>& out) {
    for (int32_t g0 = b; g0 < e; g0 += 2 * raster) {
        const int32_t g1 = std::min(e, g0 + 2 * raster);
        bool any = true;
        for (int32_t m = 0; any; ++m) {
            any = false;
            for (int32_t k = g0; k < g1; k += 2) {
                const bool two = k + 1 < e;
                const int32_t lo_k = jb_lo[k], hi_k = two ? jb_hi[k + 1] : jb_hi[k];
                const int32_t n32 = 2 * (hi_k - lo_k + 1);
                if (5 * m >= n32) continue;
                any = true;
                XTile t;
                t.i0[0] = (ib0 + k) * kBlk;
                t.i0[1] = two ? (ib0 + k + 1) * kBlk : kDeadBlock;
                t.a_row[0] = static_cast<int32_t>(t.i0[0] % cap);
                t.a_row[1] = two ? static_cast<int32_t>(t.i0[1] % cap) : t.a_row[0];
                t.j0 = lo_k * kBlk + 160 * m;
                t.slot = slot0[(k - b) / 2] + ((5 * m) >> 1);
                t.n_grp = std::min(5, n32 - 5 * m);
                t.odd = (5 * m) & 1;
                out.push_back(t);
            }
        }
    }
}

// Stage 2: windows -> owned i-blocks -> plane ring -> chunks and supertiles, and their device copies.
// `ramp`: start with small chunks and double up to the regular size -- for the calculate* entry points,
// where the upload runs concurrently and the first kernels should start after a few MB have landed.
// `whole_pos` (one shard = the whole file, load range [0, n_snp)): the per-row windows are computed ON the device
// straight into d_win_lo / d_win_hi and never visit the host; the host only needs the j-block range of every
// i-block, which it gets from two bisections per block with the same predicate.  `lo` / `hi` are not read then.
static void session_plan(nldsc_ld_session* s, int32_t row_begin, int32_t row_end, const std::vector<int32_t>& lo,
                         const std::vector<int32_t>& hi, bool ramp = false, const double* whole_pos = nullptr,
                         double wind = 0.0) {
    if (row_begin < 0 || row_end > s->n_snp || row_begin >= row_end)
        throw InvalidArgument("shard row range is empty or outside [0, n_snp)");
    const bool fast = whole_pos != nullptr;
    if (fast && (s->load_begin != 0 || s->load_end != s->n_snp || row_begin != 0 || row_end != s->n_snp))
        throw InvalidArgument("whole-file planning needs the shard to be the whole file");
    // rows this shard needs: owned range widened by the window halo, 64-aligned on the left
    if (!fast && (s->load_begin > floor64(lo[row_begin]) || s->load_end < hi[row_end - 1] + 1))
        throw InvalidArgument("shard load range does not cover the window halo of its rows");
    CUDA_CHECK(cudaSetDevice(s->device));
    Tracer tr;
    s->row_begin = row_begin;
    s->row_end = row_end;
    s->own_begin = row_begin - s->load_begin;
    s->own_end = row_end - s->load_begin;
    const int32_t n_load = s->n_load;
    const int32_t n_load_pad = (n_load + kBlk - 1) / kBlk * kBlk;
    cudaStream_t st = s->own_stream;

    // window bounds, shard-local; pad rows get an empty window.  Pinned-free staging: the vectors
    // live in the session until the async copies have been consumed.
    std::vector<int32_t>& wlo = s->h_win_lo;
    std::vector<int32_t>& whi = s->h_win_hi;
    if (!fast) {
        wlo.assign(n_load_pad, 1);
        whi.assign(n_load_pad, 0);
        for (int32_t i = 0; i < n_load; ++i) {
            wlo[i] = std::max(lo[i + s->load_begin] - s->load_begin, 0);
            whi[i] = std::min(hi[i + s->load_begin] - s->load_begin, n_load - 1);
        }
        for (int32_t i = s->own_begin; i < s->own_end; ++i) s->total_pairs += whi[i] - wlo[i];
    }
    // window edges of single rows on the host: the same bisections, the same IEEE expression as windows_kernel
    auto row_lo = [&](int32_t i) {
        int32_t x = 0, y = i;
        while (x < y) {
            const int32_t m = x + ((y - x) >> 1);
            if (std::abs(whole_pos[m] - whole_pos[i]) <= wind) y = m; else x = m + 1;
        }
        return x;
    };
    auto row_hi = [&](int32_t i) {
        int32_t x = i, y = n_load - 1;
        while (x < y) {
            const int32_t m = x + ((y - x + 1) >> 1);
            if (std::abs(whole_pos[m] - whole_pos[i]) <= wind) x = m; else y = m - 1;
        }
        return x;
    };

    // ---- owned i-blocks and the j-block range each one needs
    const int32_t ib0 = s->own_begin / kBlk, ib1 = (s->own_end - 1) / kBlk + 1;
    const int32_t n_ib = ib1 - ib0;
    std::vector<int32_t> jb_lo(n_ib), jb_hi(n_ib);
    int32_t max_span = 0;
    for (int32_t b = ib0; b < ib1; ++b) {
        const int32_t r0 = std::max(b * kBlk, s->own_begin), r1 = std::min(b * kBlk + kBlk, s->own_end) - 1;
        jb_lo[b - ib0] = (fast ? row_lo(r0) : wlo[r0]) / kBlk;
        jb_hi[b - ib0] = (fast ? row_hi(r1) : whi[r1]) / kBlk;
        max_span = std::max(max_span, (jb_hi[b - ib0] - jb_lo[b - ib0] + 1) * kBlk);
        // the exact count comes back from the device with the results; this estimate only steers the chunk schedule
        if (fast) s->total_pairs += static_cast<int64_t>(r1 - r0 + 1) * std::max(0, (jb_hi[b - ib0] - jb_lo[b - ib0]) * kBlk);
    }
    // a CTA pair scores two adjacent i-blocks against the union of their column ranges
    for (int32_t b = 0; b + 1 < n_ib; ++b)
        max_span = std::max(max_span, (jb_hi[b + 1] - jb_lo[b] + 1) * kBlk);
    tr.lap("windows");

    // ---- fixed-size device buffers
    s->d_counts = amalloc<int32_t>(s->arena, static_cast<size_t>(n_load_pad) * 3);
    s->d_coef = amalloc<SnpCoef>(s->arena, n_load_pad);
    s->d_flags = amalloc<uint8_t>(s->arena, n_load_pad);
    s->d_maf = amalloc<double>(s->arena, n_load_pad);
    s->d_rstd = amalloc<double>(s->arena, n_load_pad);
    s->d_win_lo = amalloc<int32_t>(s->arena, n_load_pad);
    s->d_win_hi = amalloc<int32_t>(s->arena, n_load_pad);
    if (!s->h_unused.empty()) {
        s->d_unused = amalloc<uint8_t>(s->arena, n_load_pad);
        CUDA_CHECK(cudaMemsetAsync(s->d_unused, 0, n_load_pad, st));
        CUDA_CHECK(cudaMemcpyAsync(s->d_unused, s->h_unused.data(), s->h_unused.size(), cudaMemcpyHostToDevice, st));
    }
    CUDA_CHECK(cudaMemsetAsync(s->d_flags, 0, n_load_pad, st));
    CUDA_CHECK(cudaMemsetAsync(s->d_coef, 0, sizeof(SnpCoef) * n_load_pad, st));
    const int32_t n_own = s->own_end - s->own_begin;
    s->d_l2 = amalloc<double>(s->arena, n_own);
    s->d_l2d = amalloc<double>(s->arena, n_own);
    s->d_wsa = amalloc<int32_t>(s->arena, n_own);
    s->d_wsd = amalloc<int32_t>(s->arena, n_own);
    s->d_wsde = amalloc<int32_t>(s->arena, n_own);
    int64_t tiles_bound = 0;                      // >= supertiles under any pairing: one per (i-block, j-block)
    for (int32_t b = 0; b < n_ib; ++b) tiles_bound += jb_hi[b] - jb_lo[b] + 2;
    auto pad16 = [](size_t b) { return (b + 15) & ~size_t(15); };
    const size_t b_pos = fast ? pad16(sizeof(double) * n_load) : 0;
    if (fast) {
        // positions -> pinned -> device (SM copy), windows straight into d_win_lo / d_win_hi; the pinned buffer is
        // sized once for everything this plan stages, because growing it would free memory the copy kernel reads
        s->pstage->reserve(b_pos + pad16(sizeof(Tile) * static_cast<size_t>(tiles_bound)) + pad16(sizeof(int32_t) * (2 * static_cast<size_t>(n_ib) + 4)) +
                           (s->mfree ? pad16(sizeof(QTile) * static_cast<size_t>(tiles_bound)) : 0) +
                           (s->x160 ? pad16(sizeof(XTile) * static_cast<size_t>(tiles_bound)) : 0));
        double* d_pos = amalloc<double>(s->arena, static_cast<size_t>(n_load) + 2);
        s->d_total_pairs = amalloc<unsigned long long>(s->arena, 1);
        CUDA_CHECK(cudaMemsetAsync(s->d_total_pairs, 0, sizeof(unsigned long long), st));
        memcpy(s->pstage->base, whole_pos, sizeof(double) * n_load);
        launch_copy16(s->pstage->base, d_pos, b_pos, st);
        launch_windows(d_pos, n_load, wind, s->d_win_lo, s->d_win_hi, st, n_load_pad, s->d_total_pairs);
    }
    tr.lap("fixed buffers");

    // ---- plane ring: as many SNP rows as needed, within the memory budget (4 planes of npad bytes each: fp4 or int8)
    const size_t row_bytes = static_cast<size_t>(kPlanes) * s->npad;
    // A block cached from an earlier call is reusable -- but the ring is ONE allocation, so only the largest free
    // block counts; when that does not hold the shard, budget from what the device has free (blocks this session
    // does not use are released by Arena::alloc before it gives up).
    const size_t reusable = s->arena->largest_free();
    size_t budget = reusable;
    if (static_cast<size_t>(n_load_pad) * row_bytes > reusable || getenv("NLDSC_PLANE_BYTES")) {
        size_t free_b = 0, total_b = 0, idle = 0;
        CUDA_CHECK(cudaMemGetInfo(&free_b, &total_b));
        for (const Arena::Block& b : s->arena->blocks) if (b.used == 0) idle += b.size;
        budget = std::max(reusable, static_cast<size_t>(0.85 * static_cast<double>(free_b + idle)));
        budget = std::min(budget, env_size("NLDSC_PLANE_BYTES", budget));
    }
    // Measured: past ~30 GB of planes the band kernel slows down (N = 315,599: 37 -> 50 ms per 40,000 SNPs; the
    // 120 GB of full-size configs[4]: 2.87 -> 2.70 s), so the ring is kept to 24 GB unless one block pair's
    // columns need more (3 x that, so that unpack can run a chunk ahead) or NLDSC_PLANE_BYTES says otherwise.
    if (!getenv("NLDSC_PLANE_BYTES")) {
        const size_t soft = std::max<size_t>(size_t(24) << 30, 3 * static_cast<size_t>(max_span) * row_bytes);
        budget = std::min(budget, soft);
    }
    int64_t cap = static_cast<int64_t>(budget / row_bytes) / kBlk * kBlk;
    cap = std::min<int64_t>(cap, n_load_pad);
    if (cap < max_span) {
        char buf[200];
        snprintf(buf, sizeof(buf),
                 "window too wide for device memory: one SNP block pair needs %d resident SNP rows of %zu bytes, budget %zu bytes",
                 max_span, row_bytes, budget);
        throw CudaError(buf);
    }
    s->plane_rows = cap;
    s->d_planes = amalloc<int8_t>(s->arena, static_cast<size_t>(cap) * row_bytes);
    tr.lap("plane ring");
    // Fused launch (ONE band launch per shard, fed by a device-side "rows ready" counter instead of one launch per
    // chunk): needs the ring to hold the whole shard (nothing is ever overwritten), the per-tile partial sums of ALL
    // chunks at once (7 KB per 64 x 64 SNP-pair block: bounded to 1.5 GB), and the wide kernel with short rows
    // (long rows: unpack and band take turns, and the per-chunk ramps are negligible beside 30 ms band launches).
    {
        // Measured (profiles/r02f_ab_step*.json, same box): the fused launch is SLOWER -- configs[1] 4.31 vs 3.82 ms per
        // step, N = 2,504: 1.91 vs 1.48 -- because the CTA pairs drift apart over 70 supertiles each and stop sharing
        // operand slices in L2, which the per-chunk launches re-align every 8,192 rows.  Opt-in (NLDSC_FUSED=1) only.
        const char* fu = getenv("NLDSC_FUSED");
        const size_t part_bytes = static_cast<size_t>(tiles_bound) * 4 * kBlk * 28;
        s->fused_ok = s->wide && (fu && fu[0] == '1') && cap == n_load_pad && part_bytes <= (size_t(3) << 29) &&
                      s->n_org < 65536;
    }

    // ---- chunks: greedy runs of i-blocks whose column span fits the ring
    // Chunk size: 8,192 rows measured best on configs[1] once consecutive band launches overlap (two streams): 3.65 ms
    // per step against 3.74 (8,192 / 16,384 / 32,768 ...) and 3.96 (16,384).  Earlier reasoning, still true for the ramp:
    // big chunks amortise the band kernel's launch and its partial last wave; with a concurrent
    // upload (`ramp`) a chunk cannot start before its last row has landed, so chunks start small, double,
    // and stop at half the regular size.
    const bool fixed_rows = getenv("NLDSC_CHUNK_ROWS") != nullptr;
    // With a concurrent upload a chunk also waits for all of its bytes, and the last chunk is scored after the
    // last byte has landed: cap it in BYTES (NLDSC_CHUNK_BYTES, default 256 MB = ~5 ms of PCIe) so that wait
    // stays short when rows are long (N = 315,599: 16,384 rows would be 1.3 GB = 24 ms).
    const size_t byte_rows = std::max<size_t>(2048, env_size("NLDSC_CHUNK_BYTES", size_t(256) << 20) / static_cast<size_t>(s->bps) / kBlk * kBlk);
    // Short rows (N < 4,096: reference panels such as 1000 Genomes) make a supertile so cheap that 8,192 rows are only a few
    // waves of them, and the per-launch ramp and tail dominate -- above all for the missing-free supertiles, which cover four
    // times the pairs each: 32,768 rows there (N = 2,504 missing-free: 0.72 vs 1.04 ms per 100,000 SNPs; with missing data 1.31
    // vs 1.33; at N = 10,000 the 6-plane path loses 1 % to 16,384-row chunks: profiles/r02k_ab_chunks_*.json).
    const size_t base_rows = s->n_org < 4096 ? 32768 : 8192;
    const int32_t chunk_rows = static_cast<int32_t>(
        std::min<size_t>(env_size("NLDSC_CHUNK_ROWS", ramp ? std::min<size_t>(base_rows, byte_rows) : base_rows),
                         static_cast<size_t>(std::numeric_limits<int32_t>::max() / 2)));
    // without a concurrent upload a short first chunk still pays: its unpack is the only one that runs alone
    const char* rp = getenv("NLDSC_RAMP");
    const bool ramp0 = rp ? rp[0] != '0' : true;
    // first chunk with a concurrent upload: what has landed by the time the plan is built and shipped
    // (~0.7 ms of PCIe, ~40 MB), so its kernels neither wait for data nor start needlessly small
    const int64_t landed = static_cast<int64_t>(env_size("NLDSC_FIRST_BYTES", size_t(40) << 20)) / std::max<int64_t>(1, s->bps);
    const int32_t first_rows = static_cast<int32_t>(std::min<int64_t>(chunk_rows, std::max<int64_t>(512, landed / kBlk * kBlk)));
    int32_t ramp_rows = (ramp0 && !fixed_rows) ? (ramp ? first_rows : static_cast<int32_t>(base_rows)) : chunk_rows;
    const char* tp = getenv("NLDSC_TAPER");
    // taper the chunk size towards the end of the shard only when the upload is clearly the longer pole
    // (narrow windows / few organisms): then the last chunk starts when the last byte lands and its length
    // is the tail.  Rough rates: 2.6e15 executed int8 op/s in the band kernel, 54 GB/s over PCIe gen5.
    const double n_own_rows = std::max(1, s->own_end - s->own_begin);
    const double compute_ns_row = (static_cast<double>(s->total_pairs) / n_own_rows + 3.0 * kBlk) * s->npad * 12.0 / 2.6e6;
    const double upload_ns_row = static_cast<double>(s->bps) / 54.0;
    const int32_t taper_min = static_cast<int32_t>(std::max<size_t>(kBlk, env_size("NLDSC_TAPER_MIN", 4096) / kBlk * kBlk));
    const bool taper = ramp && !fixed_rows && (tp ? tp[0] != '0' : compute_ns_row < 0.7 * upload_ns_row);
    const int32_t wave = std::max(1, s->sm_count / 2);       // supertiles scored concurrently (one per CTA pair)
    std::vector<Tile>& tiles = s->h_tiles;
    std::vector<int32_t>& tile_begin = s->h_tile_begin;
    tiles.clear();
    tile_begin.clear();
    s->h_qtiles.clear();
    s->h_xtiles.clear();
    int32_t unpacked_to = jb_lo[0] * kBlk;   // rows below the first needed column never enter the ring
    int32_t slot_base = 0;
    // supertiles of the i-block pair starting at block k when the chunk ends at e
    auto pair_tiles = [&](int32_t k, int32_t e) { return (k + 1 < e ? jb_hi[k + 1] : jb_hi[k]) - jb_lo[k] + 1; };
    // ... in the units the band kernel deals to its CTA pairs: runs of five 32-SNP groups with the 160-column body (the
    // 64-block bodies keep counting j-blocks, which is what their chunk sizes were measured with)
    auto pair_units = [&](int32_t k, int32_t e) {
        const int32_t nb = pair_tiles(k, e);
        return s->x160 ? (2 * nb + 4) / 5 : nb;
    };
    for (int32_t b = 0; b < n_ib;) {
        Chunk c{};
        c.iblock0 = ib0 + b;
        c.tile0 = static_cast<int64_t>(tiles.size());
        c.tb0 = static_cast<int32_t>(tile_begin.size());
        const int32_t col0 = jb_lo[b] * kBlk;
        int32_t target = std::min(ramp_rows, chunk_rows);
        // with a concurrent upload the last chunk cannot start before the last byte has landed: taper the
        // chunk size towards the end of the shard so that tail is short
        if (taper) target = std::min(target, std::max(taper_min, (n_ib - b) * kBlk / 2 / taper_min * taper_min));
        int32_t e = b;
        while (e < n_ib) {
            const int32_t col1 = (jb_hi[e] + 1) * kBlk;
            if (col1 - col0 > cap) break;
            if (e > b && (e - b) * kBlk >= target) break;
            ++e;
        }
        // The kernel deals supertiles round-robin to `wave` CTA pairs and every supertile costs the same,
        // so a launch takes ceil(n_tiles / wave) tile times: pull the chunk end back by up to 1/8 of the
        // chunk (whole i-block pairs) to where the last wave is fullest.
        if (!fixed_rows && e < n_ib && e - b >= 32 && (e - b) % 2 == 0) {
            int64_t nt = 0;
            for (int32_t k = b; k < e; k += 2) nt += pair_units(k, e);
            int32_t best_e = e;
            double best_fill = static_cast<double>(nt) / (static_cast<double>((nt + wave - 1) / wave) * wave);
            int64_t nt_k = nt;
            for (int32_t e2 = e - 2; e2 >= e - (e - b) / 8 && e2 - b >= 2; e2 -= 2) {
                nt_k -= pair_units(e2, e);                  // drop the pair (e2, e2 + 1)
                const double fill = static_cast<double>(nt_k) / (static_cast<double>((nt_k + wave - 1) / wave) * wave);
                if (fill > best_fill + 1e-9) { best_fill = fill; best_e = e2; }
            }
            e = best_e;
        }
        ramp_rows = std::min(chunk_rows, ramp_rows * 2);
        // e > b is guaranteed by cap >= max_span
        int32_t col1 = (jb_hi[e - 1] + 1) * kBlk;
        // fused launch: a band CTA reads per-SNP flags (1 byte each) of rows that are final while later rows are still
        // being written; with chunk ends on 128-row boundaries no 128-byte line ever holds both kinds
        if (s->fused_ok) col1 = std::min(n_load_pad, (col1 + 127) / 128 * 128);
        c.n_iblocks = e - b;
        c.new_row0 = std::max(unpacked_to, col0);
        c.new_row1 = std::max(col1, c.new_row0);
        unpacked_to = c.new_row1;
        // supertile rows: i-blocks (k, k+1) share their j-blocks; windows are monotone, so the union of
        // the two j ranges is [jb_lo[k], jb_hi[k+1]] (pairs outside a block's own window score nothing).
        // Slots (where partials go, what finalize walks) are i-pair-major; the EXECUTION order interleaves
        // the tiles of kRaster consecutive i-pairs by j-block: the ~74 supertiles in flight then cover
        // ~8 i-pairs x ~9 j-blocks, so every A slice is shared by ~9 CTA pairs and every B slice by ~8.
        const int32_t kRaster = static_cast<int32_t>(std::max<size_t>(1, std::min<size_t>(64, env_size("NLDSC_RASTER", 8))));
        std::vector<int32_t> slot0;                       // first slot of every i-pair of the chunk
        int32_t nslots_chunk = 0;
        for (int32_t k = b; k < e; k += 2) {
            tile_begin.push_back(nslots_chunk);
            slot0.push_back(nslots_chunk);
            nslots_chunk += pair_tiles(k, e);
        }
        for (int32_t g0 = b; g0 < e; g0 += 2 * kRaster) {
            const int32_t g1 = std::min(e, g0 + 2 * kRaster);
            const int32_t last = g1 - 1 - ((g1 - 1 - g0) % 2);                    // first block of the group's last i-pair
            const int32_t jb_first = jb_lo[g0], jb_last = (last + 1 < e ? jb_hi[last + 1] : jb_hi[last]);
            // wide supertiles (large N) carry two adjacent j-blocks, paired on even block indices; a block left alone
            // at either end of an i-pair's range is a supertile of its own, so nothing outside the band is scored
            const int32_t step = s->wide ? 2 : 1;
            for (int32_t jb = s->wide ? (jb_first & ~1) : jb_first; jb <= jb_last; jb += step) {
                for (int32_t k = g0; k < g1; k += 2) {
                    const bool two = k + 1 < e;
                    const int32_t lo_k = jb_lo[k], hi_k = two ? jb_hi[k + 1] : jb_hi[k];
                    const int32_t ja = std::max(jb, lo_k), jz = std::min(jb + step - 1, hi_k);
                    if (ja > jz) continue;
                    Tile t;
                    t.i0[0] = (ib0 + k) * kBlk;
                    t.i0[1] = two ? (ib0 + k + 1) * kBlk : kDeadBlock;
                    t.a_row[0] = static_cast<int32_t>(t.i0[0] % cap);
                    t.a_row[1] = two ? static_cast<int32_t>(t.i0[1] % cap) : t.a_row[0];
                    t.j0 = ja * kBlk;
                    t.b_row = static_cast<int32_t>(t.j0 % cap);
                    t.slot = (s->fused_ok ? slot_base : 0) + slot0[(k - b) / 2] + (ja - lo_k);
                    t.b_row2 = jz > ja ? static_cast<int32_t>((static_cast<int64_t>(ja + 1) * kBlk) % cap) : -1;
                    tiles.push_back(t);
                }
            }
        }
        // missing-free supertiles of the same chunk: four i-blocks (two i-pairs) x an even-aligned pair of j-blocks,
        // writing the same slots; rastered like the 6-plane tiles.  A j-block of the pair that lies outside an i-pair's
        // range has no slot (-1): all its pairs are out of window, the kernel scores them to nothing.
        c.blk0 = jb_lo[b];
        c.blk1 = jb_hi[e - 1] + 1;
        c.qtile0 = static_cast<int64_t>(s->h_qtiles.size());
        if (s->mfree) {
            const int32_t base = s->fused_ok ? slot_base : 0;
            const int32_t n_blocks_ring = n_load_pad / kBlk;
            for (int32_t g0 = b; g0 < e; g0 += 4 * kRaster) {
                const int32_t g1 = std::min(e, g0 + 4 * kRaster);
                for (int32_t jp = jb_lo[g0] >> 1; jp <= (jb_hi[g1 - 1] >> 1); ++jp) {
                    for (int32_t k = g0; k < g1; k += 4) {
                        const int32_t kend = std::min(e, k + 4);
                        if (2 * jp + 1 < jb_lo[k] || 2 * jp > jb_hi[kend - 1]) continue;
                        QTile t{};
                        for (int32_t m = 0; m < 4; ++m) {
                            const bool alive = k + m < kend;
                            t.i0[m] = alive ? (ib0 + k + m) * kBlk : kDeadBlock;
                            t.a_row[m] = alive ? static_cast<int32_t>(t.i0[m] % cap) : t.a_row[0];
                        }
                        t.j0 = 2 * jp * kBlk;
                        for (int32_t jj = 0; jj < 2; ++jj) {
                            const int32_t jb = 2 * jp + jj;
                            t.b_row[jj] = jb < n_blocks_ring ? static_cast<int32_t>((static_cast<int64_t>(jb) * kBlk) % cap) : t.b_row[0];
                        }
                        for (int32_t p = 0; p < 2; ++p) {
                            const int32_t first = k + 2 * p;
                            for (int32_t jj = 0; jj < 2; ++jj) {
                                t.slot[p][jj] = -1;
                                if (first >= kend) continue;
                                const int32_t lo_p = jb_lo[first], hi_p = first + 1 < kend ? jb_hi[first + 1] : jb_hi[first];
                                const int32_t jb = 2 * jp + jj;
                                if (jb >= lo_p && jb <= hi_p) t.slot[p][jj] = base + slot0[(first - b) / 2] + (jb - lo_p);
                            }
                        }
                        s->h_qtiles.push_back(t);
                    }
                }
            }
        }
        c.n_qtiles = static_cast<int32_t>(static_cast<int64_t>(s->h_qtiles.size()) - c.qtile0);
        // 160-column supertiles of the same chunk: every i-pair's j-range [jb_lo, jb_hi] cut into runs of five 32-SNP
        // groups from its own first column (nothing is aligned to more than 32 SNPs, so only the last run of an i-pair
        // is partly outside its range); same slots, same raster (run m of kRaster i-pairs side by side)
        c.xtile0 = static_cast<int64_t>(s->h_xtiles.size());
        if (s->x160 && !s->fused_ok) plan_xtiles_chunk(jb_lo.data(), jb_hi.data(), b, e, ib0, cap, kRaster, slot0, s->h_xtiles);
        c.n_xtiles = static_cast<int32_t>(static_cast<int64_t>(s->h_xtiles.size()) - c.xtile0);
        s->max_chunk_slots = std::max(s->max_chunk_slots, nslots_chunk);
        c.slot_base = slot_base;                         // fused plans: where the chunk's partials start
        slot_base += nslots_chunk;
        tile_begin.push_back(nslots_chunk);
        c.n_tiles = static_cast<int32_t>(tiles.size() - c.tile0);
        s->max_chunk_tiles = std::max(s->max_chunk_tiles, c.n_tiles);
        s->chunks.push_back(c);
        b = e;
    }
    tr.lap("tile plan");
    s->total_tiles = 0;                            // in 64-wide supertiles (= partial slots), whichever kernel scores them
    for (size_t k = 0; k + 1 < tile_begin.size(); ++k) s->total_tiles += std::max(0, tile_begin[k + 1] - tile_begin[k]);
    s->d_tiles = amalloc<Tile>(s->arena, tiles.size());
    if (s->mfree) s->d_qtiles = amalloc<QTile>(s->arena, s->h_qtiles.size());
    if (s->x160) s->d_xtiles = amalloc<XTile>(s->arena, s->h_xtiles.size() + 1);
    s->d_tile_begin = amalloc<int32_t>(s->arena, tile_begin.size());
    {
        // plan arrays -> pinned staging -> device by an SM copy: the H2D copy engine is busy with (or about to
        // be queued full of) the shard's packed rows, and these few MB must not wait behind them
        const size_t b_win = fast ? 0 : pad16(sizeof(int32_t) * n_load_pad), b_tiles = pad16(sizeof(Tile) * tiles.size()),
                     b_tb = pad16(sizeof(int32_t) * tile_begin.size()), b_qt = s->mfree ? pad16(sizeof(QTile) * s->h_qtiles.size()) : 0,
                     b_xt = s->x160 ? pad16(sizeof(XTile) * s->h_xtiles.size()) : 0;
        if (!fast) s->pstage->reserve(2 * b_win + b_tiles + b_tb + b_qt + b_xt);
        else if (b_pos + b_tiles + b_tb + b_qt + b_xt > s->pstage->bytes) throw CudaError("plan staging bound exceeded");
        char* h = s->pstage->base + b_pos;
        if (!fast) {
            memcpy(h, wlo.data(), sizeof(int32_t) * n_load_pad);
            memcpy(h + b_win, whi.data(), sizeof(int32_t) * n_load_pad);
            launch_copy16(h, s->d_win_lo, b_win, st);
            launch_copy16(h + b_win, s->d_win_hi, b_win, st);
        }
        memcpy(h + 2 * b_win, tiles.data(), sizeof(Tile) * tiles.size());
        memcpy(h + 2 * b_win + b_tiles, tile_begin.data(), sizeof(int32_t) * tile_begin.size());
        launch_copy16(h + 2 * b_win, s->d_tiles, b_tiles, st);
        launch_copy16(h + 2 * b_win + b_tiles, s->d_tile_begin, b_tb, st);
        if (s->mfree && !s->h_qtiles.empty()) {
            memcpy(h + 2 * b_win + b_tiles + b_tb, s->h_qtiles.data(), sizeof(QTile) * s->h_qtiles.size());
            launch_copy16(h + 2 * b_win + b_tiles + b_tb, s->d_qtiles, b_qt, st);
        }
        if (s->x160 && !s->h_xtiles.empty()) {
            memcpy(h + 2 * b_win + b_tiles + b_tb + b_qt, s->h_xtiles.data(), sizeof(XTile) * s->h_xtiles.size());
            launch_copy16(h + 2 * b_win + b_tiles + b_tb + b_qt, s->d_xtiles, b_xt, st);
        }
    }
    if (s->mfree) {
        s->d_blk_miss = amalloc<uint8_t>(s->arena, static_cast<size_t>(n_load_pad / kBlk) + 16);
        s->d_mfree_flags = amalloc<uint32_t>(s->arena, s->chunks.size() + 4);
        CUDA_CHECK(cudaMemsetAsync(s->d_mfree_flags, 0, sizeof(uint32_t) * (s->chunks.size() + 4), st));
        s->total_qtiles = static_cast<int64_t>(s->h_qtiles.size());
    }
    // [slot][rank][half][row]; fused plans keep every chunk's partials (chunk c at slot_base[c]), the others two chunks' worth
    const size_t np = static_cast<size_t>(s->fused_ok ? slot_base : s->max_chunk_slots) * 4 * kBlk;
    if (s->fused_ok) {
        s->d_ready = amalloc<uint32_t>(s->arena, 4);
        CUDA_CHECK(cudaMemsetAsync(s->d_ready, 0, 16, st));
    }
    s->part.add = amalloc<double>(s->arena, np);
    s->part.dom = amalloc<double>(s->arena, np);
    s->part.wsa = amalloc<int32_t>(s->arena, np);
    s->part.wsd = amalloc<int32_t>(s->arena, np);
    s->part.wsde = amalloc<int32_t>(s->arena, np);
    if (s->fused_ok) {
        s->part2 = s->part;                           // every chunk has its own region: no double buffer
    } else {
        s->part2.add = amalloc<double>(s->arena, np);
        s->part2.dom = amalloc<double>(s->arena, np);
        s->part2.wsa = amalloc<int32_t>(s->arena, np);
        s->part2.wsd = amalloc<int32_t>(s->arena, np);
        s->part2.wsde = amalloc<int32_t>(s->arena, np);
    }
    if (!s->use_simt) s->maps = band_maps_create(s->d_planes, s->plane_rows, s->npad);
    // the plan is consumed by kernels on the session's own stream or on a caller's stream
    s->ev_planned = ev_get(s->device, false);
    CUDA_CHECK(cudaEventRecord(s->ev_planned, st));
    s->planned = true;
    tr.lap("tile upload + maps");
}

static void unpack_rows(nldsc_ld_session* s, int32_t r0, int32_t r1, cudaStream_t st) {
    // rows [r0, r1) -> ring slots (row % plane_rows); rows at or past n_load are zero padding
    const int64_t plane_stride = s->plane_rows * static_cast<int64_t>(s->npad);
    while (r0 < r1) {
        const int32_t slot = static_cast<int32_t>(r0 % s->plane_rows);
        int32_t n = static_cast<int32_t>(std::min<int64_t>(r1 - r0, s->plane_rows - slot));
        if (r0 < s->n_load) {
            n = std::min(n, s->n_load - r0);
            (s->fp4 ? launch_unpack_fp4 : launch_unpack)(s->d_packed, s->bps, r0, n, s->n_org, s->npad, s->d_planes, plane_stride, slot,
                                                         s->d_counts, st);
        } else {
            launch_zero_rows(s->d_planes, plane_stride, s->npad, slot, n, st);
        }
        ++s->launches;
        r0 += n;
    }
}

static void session_run(nldsc_ld_session* s, cudaStream_t st_in) {
    if (!s->upload_started) throw CudaError("session_run called before session_upload");
    if (!s->planned) throw CudaError("session_run called before the band plan was built");
    CUDA_CHECK(cudaSetDevice(s->device));
    cudaStream_t st = st_in ? st_in : s->own_stream;
    s->spans_used = 0;
    s->launches = 0;
    if (st != s->own_stream) CUDA_CHECK(cudaStreamWaitEvent(st, s->ev_planned, 0));
    CUDA_CHECK(cudaEventRecord(s->ev_run0, st));
    const double n = static_cast<double>(s->n_org);
    // Two streams: unpack + stats of the next chunk (HBM-bound, no shared memory) run beside the band
    // kernel of the current one (tensor / L2-bound, one persistent CTA per SM) -- NLDSC_OVERLAP=0 puts
    // everything back on the compute stream.
    const char* ov = getenv("NLDSC_OVERLAP");
    const bool overlap = !(ov && ov[0] == '0');
    cudaStream_t us = overlap ? s->unpack_stream : st;
    // band launches alternate between two streams: every launch is a full persistent grid, so the CTA pairs of
    // chunk c+1 start on the SMs that chunk c's partial last wave leaves idle instead of after its slowest pair
    // With long rows (N >= 65,536) the wide band kernel loses more to a concurrent unpack than the overlap saves
    // (N = 315,599: 46.7 ms side by side, 45.6 ms in turn; 10 cM band at N = 100,000: 65.9 vs 60.2 ms; at
    // N = 10,000 it is the other way round: 4.09 vs 4.62 ms), so there the two kernels take turns.
    const char* wo = getenv("NLDSC_WIDE_OVERLAP");
    // On the fp4 pipe the band kernel moves half the operand bytes and the balance tips: side by side wins 13 % at N = 315,599
    // with a 1 cM window (11.6 vs 13.4 ms per 20,000 SNPs) and loses 2 % on the 10 cM band at N = 100,000
    // (profiles/r02s_ab_knobs_*.json), so the kernels take turns only with int8 planes.
    const bool turns = overlap && s->wide && (wo ? wo[0] == '0' : (!s->fp4 && s->n_org >= 65536));
    const int32_t commit2 = env_size("NLDSC_COMMIT2", 0) != 0 ? 1 : 0;
    const char* bsn = getenv("NLDSC_BAND_STREAMS");
    // ONE band launch for the whole shard (plan permitting): the persistent kernel starts now and every CTA pair waits,
    // tile by tile, for the rows it needs to be published by the unpack stream -- no per-chunk launch ramp, one
    // partial last wave per shard instead of one per chunk, and with a concurrent upload the tensor work follows
    // the bytes as they land.  Needs the second stream (the kernel would wait for work queued behind itself).
    const bool fused = s->fused_ok && overlap && !turns && !s->use_simt && s->wide;
    const bool two_band = !fused && overlap && !(bsn && bsn[0] == '1') && s->chunks.size() > 1;
    auto band_args = [&](const Tile* tiles, int32_t n_tiles, const Partials& part) {
        BandArgs a{};
        a.tiles = tiles;
        a.n_tiles = n_tiles;
        a.k_steps = s->npad / kKAlign;
        a.coef = s->d_coef;
        a.flags = s->d_flags;
        a.win_lo = s->d_win_lo;
        a.win_hi = s->d_win_hi;
        a.n_load = s->n_load;
        a.n = n;
        a.inv_n = 1.0 / n;
        a.adj_c = (n - 1.0) / (n - 2.0);
        a.rsq_thr = s->rsq_thr;
        a.part = part;
        a.dump = nullptr;
        a.commit2 = commit2;
        a.fp4 = s->fp4 ? 1 : 0;
        a.ring_rows = static_cast<int32_t>(s->plane_rows);
        { const char* ra = getenv("NLDSC_RUNAHEAD"); a.runahead = (ra && ra[0] == '0') ? 0 : 1; }
        a.no_skip = env_size("NLDSC_NOSKIP", 0) != 0 ? 1 : 0;
        return a;
    };
    auto finalize_chunk = [&](const Chunk& c, const Partials& part, cudaStream_t fs) {
        FinalizeArgs f{};
        f.tile_begin = s->d_tile_begin + c.tb0;
        f.iblock0 = c.iblock0;
        f.n_iblocks = c.n_iblocks;
        f.own_begin = s->own_begin;
        f.own_end = s->own_end;
        f.part = part;
        f.flags = s->d_flags;
        f.l2 = s->d_l2;
        f.l2d = s->d_l2d;
        f.wsa = s->d_wsa;
        f.wsd = s->d_wsd;
        f.wsde = s->d_wsde;
        SpanScope sc(s, fs, kFamFinalize);
        launch_finalize(f, fs);
        ++s->launches;
    };
    if (fused) CUDA_CHECK(cudaMemsetAsync(s->d_ready, 0, sizeof(uint32_t), st));    // before the fork: publishes come after it
    // missing-free path: per chunk, decided on the device after the chunk's stats (not with the fused launch, whose
    // single kernel cannot hand chunks to another kernel)
    const bool mfree = s->mfree && !fused && !s->use_simt && s->wide && s->d_blk_miss != nullptr;
    if (mfree) CUDA_CHECK(cudaMemsetAsync(s->d_blk_miss, 0, static_cast<size_t>((s->n_load + kBlk - 1) / kBlk), st));
    if (overlap) {
        CUDA_CHECK(cudaEventRecord(s->ev_fork, st));
        CUDA_CHECK(cudaStreamWaitEvent(us, s->ev_fork, 0));
        if (two_band) CUDA_CHECK(cudaStreamWaitEvent(s->band2_stream, s->ev_fork, 0));
        while (s->ev_unpacked.size() < s->chunks.size()) {
            s->ev_unpacked.push_back(ev_get(s->device, false));
            s->ev_band.push_back(ev_get(s->device, false));
        }
    }
    bool band_launched = false;
    auto launch_fused_band = [&] {
        BandArgs a = band_args(s->d_tiles, static_cast<int32_t>(s->h_tiles.size()), s->part);
        a.ready_rows = s->d_ready;
        SpanScope sc(s, st, kFamBand);
        launch_band_tc_wide(s->maps, a, s->sm_count, st);
        ++s->launches;
        band_launched = true;
    };
    int32_t next_piece = 0;                 // pieces below this one have been waited for on the unpack stream
    try {
    for (size_t ci = 0; ci < s->chunks.size(); ++ci) {
        const Chunk& c = s->chunks[ci];
        char chunk_tag[48];
        snprintf(chunk_tag, sizeof(chunk_tag), "nldsc_ld chunk %zu", ci);
        NvtxRange chunk_range(chunk_tag);
        const bool odd = two_band && (ci & 1);
        cudaStream_t bs = odd ? s->band2_stream : st;
        const int32_t r0 = std::min(c.new_row0, s->n_load), r1 = std::min(c.new_row1, s->n_load);
        if (r1 > r0) {
            // the chunk's kernels wait on the device for the last piece they read; the event must have
            // been recorded by then, so wait on the host for the reader to have issued that piece
            const int32_t piece = (r1 - 1) / s->piece_rows;
            {
                std::unique_lock<std::mutex> g(s->up_mu);
                s->up_cv.wait(g, [&] { return s->pieces_issued > piece; });
                if (!s->up_error.empty()) throw CudaError(s->up_error);
            }
            // every piece up to that one: the .bed readers queue their pieces on the copy stream out of order, so
            // the event of piece k says nothing about piece k - 1
            for (; next_piece <= piece; ++next_piece) CUDA_CHECK(cudaStreamWaitEvent(us, s->ev_piece[next_piece], 0));
        }
        // long rows: the unpack of chunk c waits for the band kernel of chunk c-1, i.e. the two kernels take turns
        if (turns && ci >= 1) CUDA_CHECK(cudaStreamWaitEvent(us, s->ev_band[ci - 1], 0));
        if (overlap && c.new_row1 > c.new_row0) {
            // ring hazard: rows [new_row0, new_row1) land in the slots of rows `plane_rows` earlier; the
            // latest earlier chunk that still reads any of those must have finished its band kernel
            const int64_t first_safe = static_cast<int64_t>(c.new_row1) - s->plane_rows;   // rows below are overwritten
            int64_t wait_for = -1;
            for (size_t k = 0; k < ci; ++k) {
                const Chunk& e = s->chunks[k];
                const int64_t col0 = static_cast<int64_t>(s->h_tiles[e.tile0].j0);          // first column the chunk reads
                if (col0 < first_safe) wait_for = static_cast<int64_t>(k);
            }
            if (wait_for >= 0) CUDA_CHECK(cudaStreamWaitEvent(us, s->ev_band[wait_for], 0));
            // band launches alternate between two streams: the one before it ran on the other stream
            if (wait_for >= 1 && two_band) CUDA_CHECK(cudaStreamWaitEvent(us, s->ev_band[wait_for - 1], 0));
        }
        {
            // unpack the chunk's new rows into the ring; the same pass counts their genotypes
            SpanScope sc(s, us, kFamUnpack);
            if (r1 > r0)
                CUDA_CHECK(cudaMemsetAsync(s->d_counts + static_cast<size_t>(r0) * 3, 0,
                                           sizeof(int32_t) * 3 * static_cast<size_t>(r1 - r0), us));
            unpack_rows(s, c.new_row0, c.new_row1, us);
        }
        if (r1 > r0) {
            SpanScope sc(s, us, kFamStats);
            launch_stats(s->d_counts, r0, r1 - r0, s->n_org, s->maf_thr, s->std_thr, s->d_coef, s->d_flags,
                         s->d_maf, s->d_rstd, us, s->d_unused, mfree ? s->d_blk_miss : nullptr);
            ++s->launches;
        }
        if (mfree) {
            // 1 = no kept SNP among the columns this chunk reads has a missing genotype
            launch_chunk_mfree_flag(s->d_blk_miss, c.blk0, std::min(c.blk1, (s->n_load + kBlk - 1) / kBlk), s->d_mfree_flags + ci, us);
            ++s->launches;
        }
        if (fused) {
            // the chunk's rows (planes, coefficients, flags) are final: let the waiting CTA pairs at them
            launch_publish_rows(s->d_ready, static_cast<uint32_t>(c.new_row1), us);
            ++s->launches;
            // the band kernel goes in once the first chunk's work is queued (it then never waits for anything that
            // is not already on its way to the device)
            if (!band_launched) launch_fused_band();
            continue;
        }
        if (overlap) {
            CUDA_CHECK(cudaEventRecord(s->ev_unpacked[ci], us));
            CUDA_CHECK(cudaStreamWaitEvent(bs, s->ev_unpacked[ci], 0));
        }
        const Partials& part = (odd && !s->fused_ok) ? s->part2 : s->part;
        BandArgs a = band_args(s->d_tiles + c.tile0, c.n_tiles, part);
        if (mfree) {
            a.qtiles = s->d_qtiles + c.qtile0;
            a.n_qtiles = c.n_qtiles;
            a.mfree_flag = s->d_mfree_flags + ci;
        }
        if (s->x160 && s->wide && !s->use_simt && c.n_xtiles > 0) {
            a.xtiles = s->d_xtiles + c.xtile0;
            a.n_xtiles = c.n_xtiles;
        }
        {
            SpanScope sc(s, bs, kFamBand);
            if (s->use_simt)
                launch_band_simt(s->d_planes, s->plane_rows * static_cast<int64_t>(s->npad), s->npad, a, bs);
            else
                (s->wide ? launch_band_tc_wide : launch_band_tc)(s->maps, a, s->sm_count, bs);
            ++s->launches;
        }
        if (overlap) CUDA_CHECK(cudaEventRecord(s->ev_band[ci], bs));
        // fused-capable plans number their slots shard-wide: the chunk's partials start at its slot_base
        Partials fp = part;
        if (s->fused_ok) {
            const size_t o = static_cast<size_t>(c.slot_base) * 4 * kBlk;
            fp.add += o; fp.dom += o; fp.wsa += o; fp.wsd += o; fp.wsde += o;
        }
        finalize_chunk(c, fp, bs);
    }
    } catch (...) {
        // a failed upload: the fused band kernel is already waiting for rows that will never be published
        if (fused && band_launched) launch_publish_rows(s->d_ready, 0xFFFFFFFFu, s->band2_stream);
        throw;
    }
    if (fused && !band_launched) launch_fused_band();
    if (fused) {
        for (const Chunk& c : s->chunks) {
            Partials fp = s->part;
            const size_t o = static_cast<size_t>(c.slot_base) * 4 * kBlk;
            fp.add += o; fp.dom += o; fp.wsa += o; fp.wsd += o; fp.wsde += o;
            finalize_chunk(c, fp, st);
        }
    }
    if (two_band) {
        CUDA_CHECK(cudaEventRecord(s->ev_join, s->band2_stream));
        CUDA_CHECK(cudaStreamWaitEvent(st, s->ev_join, 0));
    }
    CUDA_CHECK(cudaEventRecord(s->ev_run1, st));
    s->ran = true;
}

static void session_download(nldsc_ld_session* s, nldsc_ld_result* out, cudaStream_t st_in) {
    if (!s->ran) throw CudaError("session_download called before session_run");
    join_reader(s);
    if (!s->up_error.empty()) throw CudaError(s->up_error);
    if (!out || !out->l2 || !out->l2d || !out->maf || !out->residuals_std || !out->l2_ws || !out->l2d_ws ||
        !out->l2d_wse)
        throw InvalidArgument("result arrays must all be non-null");
    CUDA_CHECK(cudaSetDevice(s->device));
    cudaStream_t st = st_in ? st_in : s->own_stream;
    const size_t n = static_cast<size_t>(s->own_end - s->own_begin);
    CUDA_CHECK(cudaMemcpyAsync(out->l2, s->d_l2, n * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaMemcpyAsync(out->l2d, s->d_l2d, n * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaMemcpyAsync(out->maf, s->d_maf + s->own_begin, n * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaMemcpyAsync(out->residuals_std, s->d_rstd + s->own_begin, n * sizeof(double),
                               cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaMemcpyAsync(out->l2_ws, s->d_wsa, n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaMemcpyAsync(out->l2d_ws, s->d_wsd, n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaMemcpyAsync(out->l2d_wse, s->d_wsde, n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
}

}  // namespace nldsc

// ---- multi-GPU driver ----------------------------------------------------------------------

namespace {

// Contiguous row ranges with (roughly) equal band area = sum of window sizes.
std::vector<int32_t> split_rows(const std::vector<int32_t>& lo, const std::vector<int32_t>& hi, int parts) {
    const int32_t n = static_cast<int32_t>(lo.size());
    std::vector<int32_t> cut(parts + 1, n);
    cut[0] = 0;
    double total = 0;
    for (int32_t i = 0; i < n; ++i) total += hi[i] - lo[i] + 1;
    double acc = 0;
    int k = 1;
    for (int32_t i = 0; i < n && k < parts; ++i) {
        acc += hi[i] - lo[i] + 1;
        while (k < parts && acc >= total * k / parts) {
            // cut on a 64-row boundary so block-aligned shards waste no partial i-blocks
            int32_t c = std::min(n, (i + 1 + kBlk / 2) / kBlk * kBlk);
            cut[k] = std::max(c, cut[k - 1]);
            ++k;
        }
    }
    return cut;
}

struct ShardJob {
    int device;
    int32_t row_begin, row_end;
    int rc = NLDSC_OK;
    std::string msg;
    nldsc_ld_shard_report rep{};
};

std::atomic<int> g_report_device_times{0};       // nldsc_ld_report_device_times
// what the last calculate* call made on this thread did, shard by shard (nldsc_ld_last_call_report)
thread_local std::vector<nldsc_ld_shard_report> tl_report;

// placement chosen by the caller of nldsc_ld_calculate_on for the calls made on this thread
thread_local int tl_first_device = -1;
thread_local int tl_n_devices = 0;

int gpus_from_env(int available) {
    if (tl_n_devices > 0) return std::min(tl_n_devices, available);
    const char* v = getenv("NLDSC_GPUS");
    int want = 1;
    if (v && *v) {
        if (strcmp(v, "all") == 0) want = available;
        else want = atoi(v);
    }
    if (want < 1) want = 1;
    return std::min(want, available);
}

int first_device_from_env() {
    if (tl_first_device >= 0) return tl_first_device;
    const char* v = getenv("NLDSC_DEVICE");
    return (v && *v) ? atoi(v) : 0;
}

// compute_windows on the device (plan.cu).  Positions go up and the two int32 arrays come back through
// pinned staging and SM copies on `st`, never the copy engines: a shard's bulk upload may already be queued
// there.  Device scratch is cached per device.
void compute_windows_device(int device, PlanStage* ps, cudaStream_t st, const double* pos, int32_t n, double wind,
                            std::vector<int32_t>& lo, std::vector<int32_t>& hi) {
    static std::mutex mu;
    static void* scratch[64] = {nullptr};
    static size_t scratch_bytes[64] = {0};
    std::lock_guard<std::mutex> g(mu);
    CUDA_CHECK(cudaSetDevice(device));
    const size_t n2 = (static_cast<size_t>(n) + 1) & ~size_t(1), n4 = (static_cast<size_t>(n) + 3) & ~size_t(3);
    const size_t b_pos = n2 * sizeof(double), b_idx = n4 * sizeof(int32_t), need = b_pos + 2 * b_idx;
    void*& buf = scratch[device & 63];
    if (scratch_bytes[device & 63] < need) {
        if (buf) cudaFree(buf);
        buf = nullptr;
        scratch_bytes[device & 63] = 0;
        CUDA_CHECK(cudaMalloc(&buf, need + need / 4));
        scratch_bytes[device & 63] = need + need / 4;
    }
    char* d = static_cast<char*>(buf);
    ps->reserve(need);
    memcpy(ps->base, pos, sizeof(double) * n);
    launch_copy16(ps->base, d, b_pos, st);
    launch_windows(reinterpret_cast<const double*>(d), n, wind, reinterpret_cast<int32_t*>(d + b_pos),
                   reinterpret_cast<int32_t*>(d + b_pos + b_idx), st);
    launch_copy16(d + b_pos, ps->base + b_pos, 2 * b_idx, st);
    CUDA_CHECK(cudaStreamSynchronize(st));
    lo.resize(n);
    hi.resize(n);
    memcpy(lo.data(), ps->base + b_pos, sizeof(int32_t) * n);
    memcpy(hi.data(), ps->base + b_pos + b_idx, sizeof(int32_t) * n);
}

// `start_upload(session)` begins the asynchronous H2D of the session's packed rows [load_begin, load_end)
// (from host memory or from the .bed); it must not block on the transfer.
template <class StartUpload>
void run_sharded(const nldsc_ld_params* p_in, nldsc_ld_result* out, StartUpload&& start_upload) {
    Tracer tr;
    if (!p_in || p_in->n_snp <= 0 || p_in->n_org <= 0 || !p_in->positions || !(p_in->ld_wind >= 0.0))
        validate_params(p_in, /*allow_unused=*/true);                 // throws with the right message
    if (!out || !out->l2 || !out->l2d || !out->maf || !out->residuals_std || !out->l2_ws || !out->l2d_ws ||
        !out->l2d_wse)
        throw InvalidArgument("result arrays must all be non-null");
    UnusedView uv;
    nldsc_ld_params p_eff = *p_in;
    const nldsc_ld_params* p = &p_eff;
    // the O(n_snp) check of the positions (and the search for unused SNPs): before the shards are cut when there are several;
    // with ONE shard only after its upload has been started -- the bytes need none of it, and it is 0.1 ms of a 5.6 ms call
    auto check_positions = [&] {
        validate_params(p_in, /*allow_unused=*/true);
        uv = mark_unused(p_in);
        if (uv.any()) p_eff.positions = uv.eff.data();
    };
    const int ndev = usable_devices();
    const int dev0 = first_device_from_env();
    const int ndev_use = gpus_from_env(std::max(1, ndev - dev0));
    int parts = ndev_use;
    // Test hook (NLDSC_TEST_SHARDS=k): cut k shards whatever the number of GPUs and deal them round-robin to the
    // devices in use, so that a ONE-GPU box still runs the thread-per-shard / halo / gather code of the multi-GPU
    // path (several sessions then share a device: private arenas and staging buffers, pooled streams).
    const int test_shards = static_cast<int>(env_size("NLDSC_TEST_SHARDS", 0));
    if (test_shards > 0) parts = test_shards;
    parts = std::max(1, std::min<int>(parts, (p->n_snp + kBlk - 1) / kBlk));
    if (parts > 1) check_positions();
    tr.lap("validate+devices");

    const bool trace = tr.on;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
        return std::chrono::duration<double, std::milli>(b - a).count();
    };
    std::vector<int32_t> lo, hi;
    const bool whole = parts == 1;

    auto work = [&](ShardJob& j) {      // one shard on one GPU
        nldsc_ld_session* s = nullptr;
        const auto t0 = now();
        auto t1 = t0, t2 = t0, t3 = t0, t4 = t0, t5 = t0;
        try {
            // one shard = the whole file: its load range needs no windows, so the packed rows start crossing
            // PCIe before anything is planned (windows and plan travel by SM copies, not the copy engine)
            char tag[64];
            snprintf(tag, sizeof(tag), "nldsc_ld shard gpu%d rows %d-%d", j.device, j.row_begin, j.row_end);
            NvtxRange shard_range(tag);
            {
                NvtxRange r("nldsc_ld open");
                if (whole) s = session_open(j.device, p, 0, p->n_snp);
                else s = session_open(j.device, p, floor64(lo[j.row_begin]), hi[j.row_end - 1] + 1);
            }
            t1 = now();
            {
                NvtxRange r("nldsc_ld upload start");
                start_upload(s);
            }
            if (whole) check_positions();          // a bad position throws here: session_free waits for the copies in flight
            if (uv.any()) s->h_unused.assign(uv.unused.begin() + s->load_begin, uv.unused.begin() + s->load_end);
            t2 = now();
            {
                NvtxRange r("nldsc_ld plan");
                if (whole) session_plan(s, j.row_begin, j.row_end, lo, hi, /*ramp=*/true, p->positions, p->ld_wind);
                else session_plan(s, j.row_begin, j.row_end, lo, hi, /*ramp=*/true);
            }
            t3 = now();
            {
                NvtxRange r("nldsc_ld run (issue)");
                session_run(s, nullptr);
            }
            t4 = now();
            NvtxRange dl_range("nldsc_ld download + sync");
            nldsc_ld_result sub = *out;
            sub.l2 += j.row_begin; sub.l2d += j.row_begin; sub.maf += j.row_begin;
            sub.residuals_std += j.row_begin; sub.l2_ws += j.row_begin; sub.l2d_ws += j.row_begin;
            sub.l2d_wse += j.row_begin;
            session_download(s, &sub, nullptr);
            t5 = now();
            j.rep.load_begin = s->load_begin;
            j.rep.load_end = s->load_end;
            j.rep.chunks = static_cast<int32_t>(s->chunks.size());
            j.rep.pairs = s->total_pairs;
            j.rep.bytes_h2d = static_cast<int64_t>(s->n_load) * s->bps;
            // the device times of the call (a hundred event queries and two small device reads: 0.15-0.2 ms of a 6 ms call)
            // only when somebody asked for them: NLDSC_TRACE, or nldsc_ld_report_device_times(1)
            if (trace || g_report_device_times.load(std::memory_order_relaxed)) {
                nldsc_ld_stats st;
                nldsc_ld_session_stats(s, &st);
                float up = 0, gap = 0, lag = 0;
                if (s->readers.empty()) {                              // memory uploads only (the .bed readers record none)
                    cudaEventElapsedTime(&up, s->ev_up0, s->ev_up1);
                    cudaEventElapsedTime(&gap, s->ev_up0, s->ev_run0);
                    cudaEventElapsedTime(&lag, s->ev_up1, s->ev_run1);
                }
                j.rep.pairs = st.pairs;
                j.rep.ms_dev_upload = up;
                j.rep.ms_dev_run = st.ms_total;
                j.rep.ms_dev_band = st.ms_band;
                j.rep.ms_dev_unpack = st.ms_unpack;
                if (trace) fprintf(stderr, "[nldsc_ld]   device: upload %.2f ms, run %.2f ms (band %.2f unpack %.2f count %.2f stats %.2f "
                                "finalize %.2f), run starts %.2f after upload start, ends %.2f after upload end, %d chunks\n",
                        up, st.ms_total, st.ms_band, st.ms_unpack, st.ms_count, st.ms_stats, st.ms_finalize, gap, lag, st.chunks);
            }
        } catch (const std::invalid_argument& e) {
            j.rc = NLDSC_ERR_INVALID; j.msg = e.what();
        } catch (const std::exception& e) {
            j.rc = NLDSC_ERR_RUNTIME; j.msg = e.what();
        }
        session_free(s);
        j.rep.device = j.device;
        j.rep.row_begin = j.row_begin;
        j.rep.row_end = j.row_end;
        j.rep.ms_open = static_cast<float>(ms(t0, t1));
        j.rep.ms_upload_start = static_cast<float>(ms(t1, t2));
        j.rep.ms_plan = static_cast<float>(ms(t2, t3));
        j.rep.ms_run_issue = static_cast<float>(ms(t3, t4));
        j.rep.ms_download = static_cast<float>(ms(t4, t5));
        j.rep.ms_free = static_cast<float>(ms(t5, now()));
        j.rep.ms_wall = static_cast<float>(ms(t0, now()));
        if (trace)
            fprintf(stderr, "[nldsc_ld] gpu %d rows [%d,%d): open %.2f ms, upload-start %.2f, windows+plan %.2f, run-issue %.2f, "
                            "download+sync %.2f, free %.2f\n",
                    j.device, j.row_begin, j.row_end, ms(t0, t1), ms(t1, t2), ms(t2, t3), ms(t3, t4), ms(t4, t5),
                    ms(t5, now()));
    };

    require_sm100(dev0);
    std::vector<ShardJob> jobs;
    if (whole) {
        jobs.push_back({dev0, 0, p->n_snp});
    } else {
        PlanStage* ps = plan_stage_acquire(dev0);
        try {
            compute_windows_device(dev0, ps, nullptr, p->positions, p->n_snp, p->ld_wind, lo, hi);
        } catch (...) {
            plan_stage_release(ps);
            throw;
        }
        plan_stage_release(ps);
        tr.lap("windows (device)");
        const std::vector<int32_t> cut = split_rows(lo, hi, parts);
        for (int k = 0; k < parts; ++k)
            if (cut[k + 1] > cut[k]) jobs.push_back({dev0 + k % ndev_use, cut[k], cut[k + 1]});
    }
    if (jobs.size() == 1) {
        work(jobs[0]);
    } else {
        std::vector<std::thread> th;
        for (auto& j : jobs) th.emplace_back(work, std::ref(j));
        for (auto& t : th) t.join();
    }
    tl_report.clear();
    for (auto& j : jobs) tl_report.push_back(j.rep);
    for (auto& j : jobs) {
        if (j.rc == NLDSC_ERR_INVALID) throw InvalidArgument(j.msg);
        if (j.rc != NLDSC_OK) throw CudaError(j.msg);
    }
}

}  // namespace

// ---- extern "C" ----------------------------------------------------------------------------

extern "C" {

int nldsc_ld_device_count(int32_t* count, char* err, size_t errlen) {
    if (count) *count = 0;
    return guarded(err, errlen, [&] {
        const int n = usable_devices();
        if (count) *count = n;
    });
}

int nldsc_ld_plan_windows(const nldsc_ld_params* params, int32_t* lo, int32_t* hi, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        validate_params(params);
        if (!lo || !hi) throw InvalidArgument("lo / hi is null");
        std::vector<int32_t> l, h;
        compute_windows(params->positions, params->n_snp, params->ld_wind, l, h);
        std::copy(l.begin(), l.end(), lo);
        std::copy(h.begin(), h.end(), hi);
    });
}

int nldsc_ld_plan_shards(const nldsc_ld_params* params, int32_t parts, int32_t* cuts, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        validate_params(params);
        if (!cuts || parts < 1) throw InvalidArgument("cuts is null or parts < 1");
        std::vector<int32_t> l, h;
        compute_windows(params->positions, params->n_snp, params->ld_wind, l, h);
        const std::vector<int32_t> c = split_rows(l, h, parts);
        std::copy(c.begin(), c.end(), cuts);
    });
}

int nldsc_ld_calculate_packed(const uint8_t* packed, const nldsc_ld_params* params, nldsc_ld_result* out, char* err,
                              size_t errlen) {
    return guarded(err, errlen, [&] {
        if (!packed) throw InvalidArgument("packed is null");
        if (!params || params->n_snp <= 0 || params->n_org <= 0) validate_params(params, true);   // throws; the positions are checked once, in run_sharded
        const int64_t bps = (static_cast<int64_t>(params->n_org) + 3) / 4;
        run_sharded(params, out, [&](nldsc_ld_session* s) {
            session_upload_mem(s, packed + static_cast<size_t>(s->load_begin) * bps, nullptr);
        });
    });
}

int nldsc_ld_calculate(const nldsc_ld_params* params, nldsc_ld_result* out, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        if (!params || params->n_snp <= 0 || params->n_org <= 0 || !params->positions) validate_params(params, true);   // throws; full check in run_sharded
        if (!params->bedfile) throw InvalidArgument("bedfile is null");
        const int fd = open(params->bedfile, O_RDONLY);
        if (fd < 0) throw CudaError(std::string("cannot open ") + params->bedfile + ": " + strerror(errno));
        struct Closer { int fd; ~Closer() { close(fd); } } closer{fd};
        // PLINK magic, SNP-major mode (reference stream.h:88-102 -> std::invalid_argument)
        unsigned char magic[3] = {0, 0, 0};
        if (pread(fd, magic, 3, 0) != 3 || magic[0] != 0x6C || magic[1] != 0x1B || magic[2] != 0x01)
            throw InvalidArgument("Wrong magic number in .bed file (expected 6c 1b 01, SNP-major)");
        const int64_t bps = (static_cast<int64_t>(params->n_org) + 3) / 4;
        struct stat stt;
        if (fstat(fd, &stt) != 0) throw CudaError("fstat failed on .bed file");
        const int64_t want = 3 + bps * params->n_snp;
        if (stt.st_size != want) {
            char buf[200];
            snprintf(buf, sizeof(buf), ".bed file size %lld does not match 3 + n_snp * ceil(n_org / 4) = %lld",
                     static_cast<long long>(stt.st_size), static_cast<long long>(want));
            throw InvalidArgument(buf);
        }
        // every shard streams its own byte range: reader thread -> pinned slot ring -> H2D (session_upload_file)
        run_sharded(params, out, [&](nldsc_ld_session* s) {
            s->bed_path = params->bedfile;
            session_upload_file(s, fd);
        });
    });
}

int nldsc_ld_calculate_on(const nldsc_ld_params* params, nldsc_ld_result* out, int32_t first_device, int32_t n_devices,
                          char* err, size_t errlen) {
    if (first_device < 0 || n_devices < 1) {
        set_err(err, errlen, "first_device must be >= 0 and n_devices >= 1");
        return NLDSC_ERR_INVALID;
    }
    struct Placement {
        Placement(int d, int n) { tl_first_device = d; tl_n_devices = n; }
        ~Placement() { tl_first_device = -1; tl_n_devices = 0; }
    } placement(first_device, n_devices);
    int32_t have = 0;
    const int rc = nldsc_ld_device_count(&have, err, errlen);
    if (rc != NLDSC_OK) return rc;
    if (first_device >= have) {
        set_err(err, errlen, "first_device is not a usable CUDA device");
        return NLDSC_ERR_RUNTIME;
    }
    return nldsc_ld_calculate(params, out, err, errlen);
}

int nldsc_ld_calculate_packed_on(const uint8_t* packed, const nldsc_ld_params* params, nldsc_ld_result* out,
                                 int32_t first_device, int32_t n_devices, char* err, size_t errlen) {
    if (first_device < 0 || n_devices < 1) {
        set_err(err, errlen, "first_device must be >= 0 and n_devices >= 1");
        return NLDSC_ERR_INVALID;
    }
    struct Placement {
        Placement(int d, int n) { tl_first_device = d; tl_n_devices = n; }
        ~Placement() { tl_first_device = -1; tl_n_devices = 0; }
    } placement(first_device, n_devices);
    int32_t have = 0;
    const int rc = nldsc_ld_device_count(&have, err, errlen);
    if (rc != NLDSC_OK) return rc;
    if (first_device >= have) {
        set_err(err, errlen, "first_device is not a usable CUDA device");
        return NLDSC_ERR_RUNTIME;
    }
    return nldsc_ld_calculate_packed(packed, params, out, err, errlen);
}

int nldsc_ld_warmup(int32_t first_device, int32_t n_devices, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        const int ndev = usable_devices();
        if (first_device < 0 || n_devices < 1 || first_device >= ndev) throw InvalidArgument("device range outside the usable devices");
        const int last = std::min(ndev, first_device + n_devices);
        // CUDA context creation, the pinned .bed staging ring and the stream pool cost 0.3-1 s per device when the
        // first chromosome pays for them one device after the other; here every device is prepared by its own thread
        std::vector<std::thread> th;
        std::vector<std::string> errs(static_cast<size_t>(last - first_device));
        for (int d = first_device; d < last; ++d) {
            th.emplace_back([d, first_device, &errs] {
                try {
                    CUDA_CHECK(cudaSetDevice(d));
                    CUDA_CHECK(cudaFree(nullptr));
                    require_sm100(d);
                    host_stage_release(host_stage_acquire(d));
                    PlanStage* ps = plan_stage_acquire(d);
                    ps->reserve(size_t(8) << 20);
                    plan_stage_release(ps);
                    for (int k = 0; k < 3; ++k) stream_put(d, k, stream_get(d, k));
                    install_band_kernels();
                } catch (const std::exception& e) {
                    errs[static_cast<size_t>(d - first_device)] = e.what();
                }
            });
        }
        for (auto& t : th) t.join();
        for (const auto& e : errs) if (!e.empty()) throw CudaError(e);
    });
}

int nldsc_ld_session_create(nldsc_ld_session** out, int device, const nldsc_ld_params* params, int32_t row_begin,
                            int32_t row_end, char* err, size_t errlen) {
    if (out) *out = nullptr;
    return guarded(err, errlen, [&] {
        if (!out) throw InvalidArgument("session out pointer is null");
        validate_params(params);
        std::vector<int32_t> lo, hi;
        compute_windows(params->positions, params->n_snp, params->ld_wind, lo, hi);
        if (row_begin < 0 || row_end > params->n_snp || row_begin >= row_end)
            throw InvalidArgument("shard row range is empty or outside [0, n_snp)");
        nldsc_ld_session* s = session_open(device, params, floor64(lo[row_begin]), hi[row_end - 1] + 1);
        try {
            session_plan(s, row_begin, row_end, lo, hi);
        } catch (...) {
            session_free(s);
            throw;
        }
        *out = s;
    });
}

int nldsc_ld_session_upload(nldsc_ld_session* s, const uint8_t* packed_rows, void* stream, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        if (!s || !packed_rows) throw InvalidArgument("null session or buffer");
        session_upload_mem(s, packed_rows, static_cast<cudaStream_t>(stream));
    });
}

int nldsc_ld_session_run(nldsc_ld_session* s, void* stream, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        if (!s) throw InvalidArgument("null session");
        session_run(s, static_cast<cudaStream_t>(stream));
    });
}

int nldsc_ld_session_download(nldsc_ld_session* s, nldsc_ld_result* out, void* stream, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        if (!s) throw InvalidArgument("null session");
        session_download(s, out, static_cast<cudaStream_t>(stream));
    });
}

int nldsc_ld_session_sync(nldsc_ld_session* s, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        if (!s) throw InvalidArgument("null session");
        CUDA_CHECK(cudaSetDevice(s->device));
        CUDA_CHECK(cudaStreamSynchronize(s->copy_stream));
        CUDA_CHECK(cudaStreamSynchronize(s->own_stream));
    });
}

int nldsc_ld_session_stats(nldsc_ld_session* s, nldsc_ld_stats* o) {
    if (!s || !o) return NLDSC_ERR_INVALID;
    memset(o, 0, sizeof(*o));
    o->pairs = s->total_pairs;
    if (s->d_total_pairs && s->ran) {             // whole-file plan: the exact count was summed on the device
        unsigned long long v = 0;
        cudaSetDevice(s->device);
        if (cudaEventSynchronize(s->ev_run1) == cudaSuccess &&
            cudaMemcpy(&v, s->d_total_pairs, sizeof(v), cudaMemcpyDeviceToHost) == cudaSuccess)
            o->pairs = static_cast<int64_t>(v);
    }
    o->tiles = s->total_tiles;
    o->chunks = static_cast<int32_t>(s->chunks.size());
    o->kernel_launches = s->launches;
    o->load_begin = s->load_begin;
    o->load_end = s->load_end;
    o->npad = s->npad;
    o->band_impl = s->use_simt ? 0 : (s->wide ? 2 : 1);
    o->qtiles = s->total_qtiles;
    o->fp4 = s->fp4 ? 1 : 0;
    o->xtiles = static_cast<int64_t>(s->h_xtiles.size());
    o->xtiles_run = o->xtiles;
    if (s->ran && s->mfree && s->d_mfree_flags) {
        std::vector<uint32_t> fl(s->chunks.size());
        cudaSetDevice(s->device);
        if (cudaEventSynchronize(s->ev_run1) == cudaSuccess &&
            cudaMemcpy(fl.data(), s->d_mfree_flags, sizeof(uint32_t) * fl.size(), cudaMemcpyDeviceToHost) == cudaSuccess) {
            for (size_t k = 0; k < fl.size(); ++k) {
                if (!fl[k]) continue;
                ++o->mfree_chunks;
                o->mfree_qtiles += s->chunks[k].n_qtiles;
                o->mfree_tiles_skipped += s->chunks[k].n_tiles;
                o->xtiles_run -= s->chunks[k].n_xtiles;
            }
        }
    }
    if (s->ran) {
        cudaSetDevice(s->device);
        if (cudaEventSynchronize(s->ev_run1) != cudaSuccess) return NLDSC_ERR_RUNTIME;
        float fam[kFamN] = {0, 0, 0, 0, 0};
        // band launches of consecutive chunks overlap (two streams): their time is the union of the spans, not the sum
        std::vector<std::pair<float, float>> band;
        for (size_t i = 0; i < s->spans_used; ++i) {
            float ms = 0;
            if (s->spans[i].family == kFamBand) {
                float t0 = 0, t1 = 0;
                if (cudaEventElapsedTime(&t0, s->ev_run0, s->spans[i].a) == cudaSuccess &&
                    cudaEventElapsedTime(&t1, s->ev_run0, s->spans[i].b) == cudaSuccess)
                    band.emplace_back(t0, t1);
            } else if (cudaEventElapsedTime(&ms, s->spans[i].a, s->spans[i].b) == cudaSuccess) {
                fam[s->spans[i].family] += ms;
            }
        }
        std::sort(band.begin(), band.end());
        float cur0 = 0, cur1 = -1;
        for (const auto& iv : band) {
            if (cur1 < cur0 || iv.first > cur1) {
                if (cur1 >= cur0) fam[kFamBand] += cur1 - cur0;
                cur0 = iv.first; cur1 = iv.second;
            } else {
                cur1 = std::max(cur1, iv.second);
            }
        }
        if (cur1 >= cur0) fam[kFamBand] += cur1 - cur0;
        o->ms_count = fam[kFamCount];
        o->ms_stats = fam[kFamStats];
        o->ms_unpack = fam[kFamUnpack];
        o->ms_band = fam[kFamBand];
        o->ms_finalize = fam[kFamFinalize];
        cudaEventElapsedTime(&o->ms_total, s->ev_run0, s->ev_run1);
    }
    return NLDSC_OK;
}

void nldsc_ld_session_destroy(nldsc_ld_session* s) { session_free(s); }

int nldsc_ld_last_call_report(nldsc_ld_shard_report* out, int32_t max, int32_t* n) {
    if (!n) return NLDSC_ERR_INVALID;
    *n = static_cast<int32_t>(tl_report.size());
    for (int32_t k = 0; out && k < max && k < *n; ++k) out[k] = tl_report[k];
    return NLDSC_OK;
}

int nldsc_ld_report_device_times(int32_t on) {
    g_report_device_times.store(on ? 1 : 0, std::memory_order_relaxed);
    return NLDSC_OK;
}

int nldsc_ld_arena_trim(int32_t device) {
    arena_trim(device);
    return NLDSC_OK;
}

// ---- parity hooks ----------------------------------------------------------------------------

int nldsc_ld_debug_unpack(const uint8_t* packed, int32_t n_snp, int32_t n_org, int8_t* planes_out, int32_t* counts_out,
                          int32_t* npad_out, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        if (!packed || n_snp <= 0 || n_org <= 0) throw InvalidArgument("bad arguments");
        usable_devices();
        CUDA_CHECK(cudaSetDevice(first_device_from_env()));
        require_sm100(first_device_from_env());
        const int64_t bps = (static_cast<int64_t>(n_org) + 3) / 4;
        const int32_t npad = static_cast<int32_t>(((static_cast<int64_t>(n_org) + kKAlign - 1) / kKAlign) * kKAlign);
        if (npad_out) *npad_out = npad;
        if (!planes_out && !counts_out) return;
        uint8_t* d_packed = dmalloc<uint8_t>(static_cast<size_t>(n_snp) * bps + 16);
        int32_t* d_counts = dmalloc<int32_t>(static_cast<size_t>(n_snp) * 3);
        const size_t plane_bytes = static_cast<size_t>(n_snp) * npad;
        int8_t* d_planes = dmalloc<int8_t>(plane_bytes * kPlanes);
        struct Free { void* a; void* b; void* c; ~Free() { cudaFree(a); cudaFree(b); cudaFree(c); } } fr{d_packed, d_counts, d_planes};
        CUDA_CHECK(cudaMemset(d_packed + static_cast<size_t>(n_snp) * bps, 0, 16));
        CUDA_CHECK(cudaMemcpy(d_packed, packed, static_cast<size_t>(n_snp) * bps, cudaMemcpyHostToDevice));
        CUDA_CHECK(cudaMemset(d_counts, 0, sizeof(int32_t) * 3 * static_cast<size_t>(n_snp)));
        launch_unpack(d_packed, bps, 0, n_snp, n_org, npad, d_planes, static_cast<int64_t>(plane_bytes), 0, d_counts, nullptr);
        CUDA_CHECK(cudaDeviceSynchronize());
        if (counts_out)
            CUDA_CHECK(cudaMemcpy(counts_out, d_counts, sizeof(int32_t) * 3 * n_snp, cudaMemcpyDeviceToHost));
        if (planes_out) CUDA_CHECK(cudaMemcpy(planes_out, d_planes, plane_bytes * kPlanes, cudaMemcpyDeviceToHost));
    });
}

int nldsc_ld_debug_unpack_fp4(const uint8_t* packed, int32_t n_snp, int32_t n_org, int8_t* planes_out, int32_t* counts_out,
                          int32_t* npad_out, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        if (!packed || n_snp <= 0 || n_org <= 0) throw InvalidArgument("bad arguments");
        usable_devices();
        CUDA_CHECK(cudaSetDevice(first_device_from_env()));
        require_sm100(first_device_from_env());
        const int64_t bps = (static_cast<int64_t>(n_org) + 3) / 4;
        const int32_t npad = static_cast<int32_t>(((static_cast<int64_t>(n_org) + 2 * kKAlign - 1) / (2 * kKAlign)) * kKAlign);   // bytes
        if (npad_out) *npad_out = npad;
        if (!planes_out && !counts_out) return;
        uint8_t* d_packed = dmalloc<uint8_t>(static_cast<size_t>(n_snp) * bps + 16);
        int32_t* d_counts = dmalloc<int32_t>(static_cast<size_t>(n_snp) * 3);
        const size_t plane_bytes = static_cast<size_t>(n_snp) * npad;
        int8_t* d_planes = dmalloc<int8_t>(plane_bytes * kPlanes);
        struct Free { void* a; void* b; void* c; ~Free() { cudaFree(a); cudaFree(b); cudaFree(c); } } fr{d_packed, d_counts, d_planes};
        CUDA_CHECK(cudaMemset(d_packed + static_cast<size_t>(n_snp) * bps, 0, 16));
        CUDA_CHECK(cudaMemcpy(d_packed, packed, static_cast<size_t>(n_snp) * bps, cudaMemcpyHostToDevice));
        CUDA_CHECK(cudaMemset(d_counts, 0, sizeof(int32_t) * 3 * static_cast<size_t>(n_snp)));
        launch_unpack_fp4(d_packed, bps, 0, n_snp, n_org, npad, d_planes, static_cast<int64_t>(plane_bytes), 0, d_counts, nullptr);
        CUDA_CHECK(cudaDeviceSynchronize());
        if (counts_out)
            CUDA_CHECK(cudaMemcpy(counts_out, d_counts, sizeof(int32_t) * 3 * n_snp, cudaMemcpyDeviceToHost));
        if (planes_out) CUDA_CHECK(cudaMemcpy(planes_out, d_planes, plane_bytes * kPlanes, cudaMemcpyDeviceToHost));
    });
}

int nldsc_ld_debug_cross_counts(const uint8_t* packed, int32_t n_snp, int32_t n_org, int32_t i_block, int32_t j_block,
                                int32_t* acc_out, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        if (!packed || !acc_out || n_snp <= 0 || n_org <= 0) throw InvalidArgument("bad arguments");
        const int32_t n_blocks = (n_snp + kBlk - 1) / kBlk;
        if (i_block < 0 || j_block < 0 || i_block >= n_blocks || j_block >= n_blocks)
            throw InvalidArgument("block index out of range");
        usable_devices();
        const int dev = first_device_from_env();
        CUDA_CHECK(cudaSetDevice(dev));
        require_sm100(dev);
        int sm_count = 0;
        CUDA_CHECK(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
        const int64_t bps = (static_cast<int64_t>(n_org) + 3) / 4;
        const int32_t npad = static_cast<int32_t>(((static_cast<int64_t>(n_org) + kKAlign - 1) / kKAlign) * kKAlign);
        const int32_t rows = n_blocks * kBlk;
        uint8_t* d_packed = dmalloc<uint8_t>(static_cast<size_t>(n_snp) * bps + 16);
        const size_t plane_bytes = static_cast<size_t>(rows) * npad;
        int8_t* d_planes = dmalloc<int8_t>(plane_bytes * kPlanes);
        int32_t* d_dump = dmalloc<int32_t>(2 * 128 * 192);
        Tile* d_tile = dmalloc<Tile>(1);
        SnpCoef* d_coef = dmalloc<SnpCoef>(rows);
        uint8_t* d_flags = dmalloc<uint8_t>(rows);
        int32_t* d_w = dmalloc<int32_t>(static_cast<size_t>(rows) * 2);
        double* d_pd = dmalloc<double>(8 * kBlk);             // one slot: [rank][half][row] per field
        int32_t* d_pi = dmalloc<int32_t>(12 * kBlk);
        struct Free {
            std::vector<void*> v;
            ~Free() { for (void* q : v) cudaFree(q); }
        } fr{{d_packed, d_planes, d_dump, d_tile, d_coef, d_flags, d_w, d_pd, d_pi}};
        CUDA_CHECK(cudaMemset(d_packed + static_cast<size_t>(n_snp) * bps, 0, 16));
        CUDA_CHECK(cudaMemcpy(d_packed, packed, static_cast<size_t>(n_snp) * bps, cudaMemcpyHostToDevice));
        CUDA_CHECK(cudaMemset(d_planes, 0, plane_bytes * kPlanes));
        CUDA_CHECK(cudaMemset(d_coef, 0, sizeof(SnpCoef) * rows));
        CUDA_CHECK(cudaMemset(d_flags, 0, rows));
        CUDA_CHECK(cudaMemset(d_w, 0, sizeof(int32_t) * 2 * rows));
        launch_unpack(d_packed, bps, 0, n_snp, n_org, npad, d_planes, static_cast<int64_t>(plane_bytes), 0, nullptr, nullptr);
        Tile t{{i_block * kBlk, i_block * kBlk}, {i_block * kBlk, kDeadBlock}, j_block * kBlk, j_block * kBlk};
        CUDA_CHECK(cudaMemcpy(d_tile, &t, sizeof(t), cudaMemcpyHostToDevice));
        BandArgs a{};
        a.tiles = d_tile;
        a.n_tiles = 1;
        a.k_steps = npad / kKAlign;
        a.coef = d_coef;
        a.flags = d_flags;
        a.win_lo = d_w;
        a.win_hi = d_w + rows;
        a.n_load = rows;
        a.n = n_org;
        a.inv_n = 1.0 / n_org;
        a.adj_c = 1.0;
        a.rsq_thr = 0.0;
        a.part.add = d_pd;
        a.part.dom = d_pd + 4 * kBlk;
        a.part.wsa = d_pi;
        a.part.wsd = d_pi + 4 * kBlk;
        a.part.wsde = d_pi + 8 * kBlk;
        a.dump = d_dump;
        const char* band = getenv("NLDSC_BAND");
        if (band && strcmp(band, "simt") == 0) {
            launch_band_simt(d_planes, static_cast<int64_t>(plane_bytes), npad, a, nullptr);
        } else {
            BandMaps* maps = band_maps_create(d_planes, rows, npad);
            struct MapFree { BandMaps* m; ~MapFree() { band_maps_destroy(m); } } mf{maps};
            launch_band_tc(maps, a, sm_count, nullptr);
        }
        CUDA_CHECK(cudaDeviceSynchronize());
        CUDA_CHECK(cudaMemcpy(acc_out, d_dump, sizeof(int32_t) * 128 * 192, cudaMemcpyDeviceToHost));
    });
}

int nldsc_ld_debug_plan_xtiles(const int32_t* jb_lo, const int32_t* jb_hi, int32_t n_iblocks, int32_t raster, int32_t* tiles_out,
                               int32_t max_tiles, int32_t* n_tiles, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        if (!jb_lo || !jb_hi || !n_tiles || n_iblocks <= 0 || raster <= 0 || (max_tiles > 0 && !tiles_out))
            throw InvalidArgument("bad arguments");
        for (int32_t k = 0; k < n_iblocks; ++k)
            if (jb_lo[k] < 0 || jb_hi[k] < jb_lo[k] || (k && (jb_lo[k] < jb_lo[k - 1] || jb_hi[k] < jb_hi[k - 1])))
                throw InvalidArgument("block ranges must be non-empty and non-decreasing");
        // one chunk holding every i-block; slots as session_plan numbers them: i-pair-major, one per 64-SNP j-block of the pair's range
        std::vector<int32_t> slot0;
        int32_t n_slots = 0;
        for (int32_t k = 0; k < n_iblocks; k += 2) {
            slot0.push_back(n_slots);
            n_slots += (k + 1 < n_iblocks ? jb_hi[k + 1] : jb_hi[k]) - jb_lo[k] + 1;
        }
        std::vector<XTile> t;
        plan_xtiles_chunk(jb_lo, jb_hi, 0, n_iblocks, 0, int64_t(1) << 40, raster, slot0, t);
        *n_tiles = static_cast<int32_t>(t.size());
        static_assert(sizeof(XTile) == 8 * sizeof(int32_t), "tiles_out rows are the eight int32 fields of an XTile");
        if (tiles_out) memcpy(tiles_out, t.data(), sizeof(XTile) * std::min<size_t>(t.size(), static_cast<size_t>(std::max(0, max_tiles))));
    });
}

int nldsc_ld_debug_plan_windows_device(const nldsc_ld_params* params, int32_t* lo, int32_t* hi, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        validate_params(params);
        if (!lo || !hi) throw InvalidArgument("lo / hi is null");
        usable_devices();
        const int dev = first_device_from_env();
        require_sm100(dev);
        std::vector<int32_t> l, h;
        PlanStage* ps = plan_stage_acquire(dev);
        try {
            compute_windows_device(dev, ps, nullptr, params->positions, params->n_snp, params->ld_wind, l, h);
        } catch (...) {
            plan_stage_release(ps);
            throw;
        }
        plan_stage_release(ps);
        std::copy(l.begin(), l.end(), lo);
        std::copy(h.begin(), h.end(), hi);
    });
}

int nldsc_ld_debug_band_probe(int32_t mode, int32_t n_org, int32_t n_tiles, int32_t reps, float* ms_out, char* err,
                              size_t errlen) {
    return guarded(err, errlen, [&] {
        if (mode < 0 || mode > 6 || n_org <= 0 || n_tiles <= 0 || reps <= 0 || !ms_out) throw InvalidArgument("bad arguments");
        usable_devices();
        const int dev = first_device_from_env();
        CUDA_CHECK(cudaSetDevice(dev));
        require_sm100(dev);
        int sm_count = 0;
        CUDA_CHECK(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
        const int64_t bps = (static_cast<int64_t>(n_org) + 3) / 4;
        const int32_t npad = static_cast<int32_t>(((static_cast<int64_t>(n_org) + kKAlign - 1) / kKAlign) * kKAlign);
        const int32_t rows = 4096, nb = rows / kBlk;
        uint8_t* d_packed = dmalloc<uint8_t>(static_cast<size_t>(rows) * bps + 16);
        const size_t plane_bytes = static_cast<size_t>(rows) * npad;
        int8_t* d_planes = dmalloc<int8_t>(plane_bytes * kPlanes);
        Tile* d_tiles = dmalloc<Tile>(n_tiles);
        SnpCoef* d_coef = dmalloc<SnpCoef>(rows);
        uint8_t* d_flags = dmalloc<uint8_t>(rows);
        int32_t* d_w = dmalloc<int32_t>(static_cast<size_t>(rows) * 2);
        const size_t n_slots = static_cast<size_t>(n_tiles) * (mode >= 5 ? 2 : 1);
        double* d_pd = dmalloc<double>(n_slots * 8 * kBlk);
        int32_t* d_pi = dmalloc<int32_t>(n_slots * 12 * kBlk);
        struct Free {
            std::vector<void*> v;
            ~Free() { for (void* q : v) cudaFree(q); }
        } fr{{d_packed, d_planes, d_tiles, d_coef, d_flags, d_w, d_pd, d_pi}};
        CUDA_CHECK(cudaMemset(d_packed, 0xB9, static_cast<size_t>(rows) * bps + 16));   // codes 01 10 11 10 per byte
        CUDA_CHECK(cudaMemset(d_coef, 0, sizeof(SnpCoef) * rows));
        // modes 0-2: every pair rejected (cheapest epilogue); 3 = epilogue only and 4 = whole kernel, every pair scored
        const bool wide = mode >= 5;                 // 6 = wide kernel, whole, cheap epilogue (5, its MMA-only variant, deadlocks: rejected above)
        const bool full_epi = mode == 3 || mode == 4;
        CUDA_CHECK(cudaMemset(d_flags, full_epi ? (kFlagKeep | kFlagDom) : 0, rows));
        CUDA_CHECK(cudaMemset(d_w, 0, sizeof(int32_t) * 2 * rows));
        if (full_epi) {
            std::vector<int32_t> whi(rows, rows - 1);
            CUDA_CHECK(cudaMemcpy(d_w + rows, whi.data(), sizeof(int32_t) * rows, cudaMemcpyHostToDevice));
        }
        launch_unpack(d_packed, bps, 0, rows, n_org, npad, d_planes, static_cast<int64_t>(plane_bytes), 0, nullptr, nullptr);
        std::vector<Tile> tiles(n_tiles);
        for (int32_t t = 0; t < n_tiles; ++t) {       // band-like walk: 22 j-blocks per i-block pair
            const int32_t p = (t / 22) % (nb / 2), jb = (2 * p + t % 22) % nb;
            tiles[t] = Tile{{2 * p * kBlk, (2 * p + 1) * kBlk}, {2 * p * kBlk, (2 * p + 1) * kBlk}, jb * kBlk, jb * kBlk,
                            wide ? 2 * t : t, wide ? ((jb + 1) % nb) * kBlk : -1};
        }
        CUDA_CHECK(cudaMemcpy(d_tiles, tiles.data(), sizeof(Tile) * n_tiles, cudaMemcpyHostToDevice));
        BandArgs a{};
        a.tiles = d_tiles;
        a.n_tiles = n_tiles;
        a.k_steps = npad / kKAlign;
        a.coef = d_coef;
        a.flags = d_flags;
        a.win_lo = d_w;
        a.win_hi = d_w + rows;
        a.n_load = rows;
        a.n = n_org;
        a.inv_n = 1.0 / n_org;
        a.adj_c = 1.0;
        a.part.add = d_pd;
        a.part.dom = d_pd + n_slots * 4 * kBlk;
        a.part.wsa = d_pi;
        a.part.wsd = d_pi + n_slots * 4 * kBlk;
        a.part.wsde = d_pi + n_slots * 8 * kBlk;
        a.probe = mode == 4 ? 0 : (mode == 5 ? 1 : (mode == 6 ? 0 : mode));
        BandMaps* maps = band_maps_create(d_planes, rows, npad);
        struct MapFree { BandMaps* m; ~MapFree() { band_maps_destroy(m); } } mf{maps};
        cudaEvent_t e0, e1;
        CUDA_CHECK(cudaEventCreate(&e0));
        CUDA_CHECK(cudaEventCreate(&e1));
        auto launch = wide ? launch_band_tc_wide : launch_band_tc;
        launch(maps, a, sm_count, nullptr);                                              // warm-up
        CUDA_CHECK(cudaEventRecord(e0, nullptr));
        for (int r = 0; r < reps; ++r) launch(maps, a, sm_count, nullptr);
        CUDA_CHECK(cudaEventRecord(e1, nullptr));
        CUDA_CHECK(cudaEventSynchronize(e1));
        float ms = 0;
        CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        *ms_out = ms / reps;
    });
}

int nldsc_ld_debug_fp4_probe(int32_t n_mma, int32_t n_cols, int32_t per_commit, int32_t all_pairs, float* out, float* ms_out,
                             char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        if (n_mma < 1 || n_cols < 16 || n_cols > 256 || (n_cols % 16) || per_commit < 1 || !out || !ms_out) throw InvalidArgument("bad arguments");
        usable_devices();
        const int dev = first_device_from_env();
        CUDA_CHECK(cudaSetDevice(dev));
        require_sm100(dev);
        const int pairs = all_pairs ? device_sm_count(dev) / 2 : 1;
        float* d = dmalloc<float>(static_cast<size_t>(pairs) * 2 * 128 * 16);
        struct Free { void* p; ~Free() { cudaFree(p); } } fr{d};
        launch_fp4_probe(n_mma, n_cols, per_commit, pairs, d, nullptr);                   // warm-up
        *ms_out = launch_fp4_probe(n_mma, n_cols, per_commit, pairs, d, nullptr);
        CUDA_CHECK(cudaMemcpy(out, d, sizeof(float) * 2 * 128 * 16, cudaMemcpyDeviceToHost));   // the first pair's rows
    });
}

int nldsc_ld_debug_h2d_probe(int64_t bytes, int32_t write_combined, int32_t first_device, int32_t n_devices, int32_t reps,
                             double* gbs_out, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        const int ndev = usable_devices();
        if (bytes <= 0 || n_devices < 1 || first_device < 0 || first_device + n_devices > ndev || reps < 1 || !gbs_out)
            throw InvalidArgument("bad arguments");
        std::vector<std::thread> th;
        std::vector<std::string> errs(static_cast<size_t>(n_devices));
        std::mutex mu;
        std::condition_variable cv;
        int ready = 0;
        for (int k = 0; k < n_devices; ++k) {
            th.emplace_back([&, k] {
                void* h = nullptr;
                void* d = nullptr;
                cudaStream_t st = nullptr;
                cudaEvent_t e0 = nullptr, e1 = nullptr;
                try {
                    CUDA_CHECK(cudaSetDevice(first_device + k));
                    CUDA_CHECK(cudaHostAlloc(&h, static_cast<size_t>(bytes), write_combined ? cudaHostAllocWriteCombined : cudaHostAllocDefault));
                    memset(h, 0x5A, static_cast<size_t>(bytes));
                    CUDA_CHECK(cudaMalloc(&d, static_cast<size_t>(bytes)));
                    CUDA_CHECK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
                    CUDA_CHECK(cudaEventCreate(&e0));
                    CUDA_CHECK(cudaEventCreate(&e1));
                    CUDA_CHECK(cudaMemcpyAsync(d, h, static_cast<size_t>(bytes), cudaMemcpyHostToDevice, st));
                    CUDA_CHECK(cudaStreamSynchronize(st));
                    {   // all devices start their timed copies together
                        std::unique_lock<std::mutex> g(mu);
                        if (++ready == n_devices) cv.notify_all();
                        else cv.wait(g, [&] { return ready >= n_devices; });
                    }
                    CUDA_CHECK(cudaEventRecord(e0, st));
                    for (int r = 0; r < reps; ++r) CUDA_CHECK(cudaMemcpyAsync(d, h, static_cast<size_t>(bytes), cudaMemcpyHostToDevice, st));
                    CUDA_CHECK(cudaEventRecord(e1, st));
                    CUDA_CHECK(cudaEventSynchronize(e1));
                    float ms = 0;
                    CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
                    gbs_out[k] = static_cast<double>(bytes) * reps / (ms * 1e-3) / 1e9;
                } catch (const std::exception& e) {
                    errs[static_cast<size_t>(k)] = e.what();
                    std::lock_guard<std::mutex> g(mu);
                    ++ready;
                    cv.notify_all();
                }
                if (e0) cudaEventDestroy(e0);
                if (e1) cudaEventDestroy(e1);
                if (st) cudaStreamDestroy(st);
                if (d) cudaFree(d);
                if (h) cudaFreeHost(h);
            });
        }
        for (auto& t : th) t.join();
        for (const auto& e : errs) if (!e.empty()) throw CudaError(e);
    });
}

int nldsc_ld_debug_i8_peak(int32_t n_cols, int32_t per_commit, int32_t depth, int32_t iters, int32_t reps, float* ms_out,
                           double* tops_out, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        if (n_cols < 16 || n_cols > 256 || (n_cols % 16) != 0 || (per_commit != 4 && per_commit != 8 && per_commit != 16) || (depth & 0xFF) < 1 ||
            (depth & 0xFF) > 8 || iters < 1 || reps < 1 || !ms_out)
            throw InvalidArgument("bad arguments");
        usable_devices();
        const int dev = first_device_from_env();
        CUDA_CHECK(cudaSetDevice(dev));
        require_sm100(dev);
        const int sms = device_sm_count(dev);
        cudaEvent_t e0, e1;
        CUDA_CHECK(cudaEventCreate(&e0));
        CUDA_CHECK(cudaEventCreate(&e1));
        launch_i8_peak(iters, n_cols, per_commit, depth, sms, nullptr);                         // warm-up
        CUDA_CHECK(cudaEventRecord(e0, nullptr));
        for (int r = 0; r < reps; ++r) launch_i8_peak(iters, n_cols, per_commit, depth, sms, nullptr);
        CUDA_CHECK(cudaEventRecord(e1, nullptr));
        CUDA_CHECK(cudaEventSynchronize(e1));
        float ms = 0;
        CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        *ms_out = ms / reps;
        // int8 ops of one launch: CTA pairs x commits x MMAs x (2 x 256 x n_cols x 32)
        if (tops_out)
            *tops_out = static_cast<double>(sms / 2) * iters * per_commit * 2.0 * 256.0 * n_cols * 32.0 / (*ms_out * 1e-3) / 1e12;
    });
}

}  // extern "C"
