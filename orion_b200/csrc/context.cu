// context.cu -- parameter generation and HBM-resident tables.
#include "context.h"
#include <algorithm>
#include <map>
#include <unordered_map>
#include <cuda.h>

namespace ob {

unsigned long long g_launches = 0;
double g_alg_bytes = 0;
static cudaStream_t g_stream = nullptr;
static bool g_dev_ok = false;

void ensure_device() {
    if (g_dev_ok) return;
    setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);      // lanes + side streams: no channel sharing (no effect once a context exists)
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        OB_FATAL("no CUDA device visible: liborion_b200 has no CPU path (" + std::string(cudaGetErrorString(e)) + ")");
    int dev = 0;
    if (const char* s = std::getenv("ORION_B200_DEVICE")) dev = std::atoi(s);
    else if (const char* s2 = std::getenv("LOCAL_RANK")) dev = std::atoi(s2) % n;
    OB_CUDA(cudaSetDevice(dev));
    OB_CUDA(cudaStreamCreateWithFlags(&g_stream, cudaStreamNonBlocking));
    cudaMemPool_t pool;
    OB_CUDA(cudaDeviceGetDefaultMemPool(&pool, dev));
    uint64_t thr = UINT64_MAX;
    OB_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr));
    // lanes and side streams share this pool: never let an allocation on one stream wait for a free queued on another
    // (that would serialise independent lanes); reuse across streams only once the free is known to have completed
    int off = 0;
    if (const char* e = std::getenv("ORION_B200_POOL_DEPS")) off = std::atoi(e);
    OB_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolReuseAllowInternalDependencies, &off));
    g_dev_ok = true;
}
static cudaStream_t g_cur = nullptr;                      // current stream while forked (single host thread by contract)
cudaStream_t stream() { ensure_device(); return g_cur ? g_cur : g_stream; }

// ---- fork / join ----
static std::vector<cudaStream_t>& g_side_pool = *new std::vector<cudaStream_t>();   // free side streams (leaked, like every global owner)
static bool g_side_pool_init = false;
static std::vector<cudaEvent_t>& g_event_pool = *new std::vector<cudaEvent_t>();
static cudaEvent_t take_event() {
    if (g_event_pool.empty()) { cudaEvent_t e; OB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); return e; }
    cudaEvent_t e = g_event_pool.back(); g_event_pool.pop_back(); return e;
}
cudaStream_t main_stream() { ensure_device(); return g_stream; }
void set_current_stream(cudaStream_t s) { g_cur = (s == g_stream) ? nullptr : s; }
DevEvent::DevEvent() : e(take_event()) {}
DevEvent::~DevEvent() { if (e) g_event_pool.push_back(e); }      // waits already enqueued keep the state they captured
EvRef record_event(cudaStream_t s) {
    OB_HP("record_event");
    EvRef r = std::make_shared<DevEvent>();
    OB_CUDA(cudaEventRecord(r->e, s));
    return r;
}
// PDL policy: measured +3 % on the logN=16 key switch (tail of one kernel overlaps the head of the next), -30 % on a
// bare rescale chain, and on the ResNet-20 replay at logN=13: 0.373 -> 0.339 s/image with the dataflow lanes (before the
// lanes it cost 9 % there: early-resident dependents took SM slots from the side streams' kernels; with eight lanes the
// shorter kernel-to-kernel gap of every chain wins).  Default: on for every ring; ORION_B200_PDL=0/1 overrides.
static bool g_pdl_auto = true;
int g_pdl_suppress = 0;
bool pdl_enabled() {
    static int forced = -2;
    if (forced == -2) { const char* s = std::getenv("ORION_B200_PDL"); forced = s ? (std::atoi(s) != 0) : -1; }
    if (g_pdl_suppress) return false;
    return forced >= 0 ? forced == 1 : g_pdl_auto;
}
namespace {
int g_probe_mode = 0;
typedef std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ProbeList;
ProbeList* probe_events() { static auto* v = new ProbeList[PROBE_SLOTS]; return v; }
bool probe_active(int slot) { return g_probe_mode == 2 || (g_probe_mode == 1 && slot == PROBE_IP_F64); }
}
const char* probe_slot_name(int slot) {
    static const char* names[PROBE_SLOTS] = {"fused_ntt2_inner_product_f64", "fused_ntt2_inner_product_int", "intt_c1", "modup_base_conversion",
                                             "ntt_first_pass_digits", "intt_p", "moddown_base_conversion", "moddown_ntt_first_pass",
                                             "moddown_ntt2_submul"};
    return slot >= 0 && slot < PROBE_SLOTS ? names[slot] : "";
}
void probe_enable(int mode) { g_probe_mode = mode; }
void probe_begin(int slot) {
    if (!probe_active(slot)) return;
    cudaEvent_t a, b;
    OB_CUDA(cudaEventCreate(&a)); OB_CUDA(cudaEventCreate(&b));
    OB_CUDA(cudaEventRecord(a, stream()));
    probe_events()[slot].push_back({a, b});
}
void probe_end(int slot) {
    if (!probe_active(slot)) return;
    OB_CUDA(cudaEventRecord(probe_events()[slot].back().second, stream()));
}
double probe_read_ms(long* launches, int slot) {
    double tot = 0;
    if (slot < 0 || slot >= PROBE_SLOTS) { if (launches) *launches = 0; return 0; }
    auto& ev = probe_events()[slot];
    for (auto& e : ev) {
        OB_CUDA(cudaEventSynchronize(e.second));
        float ms = 0; OB_CUDA(cudaEventElapsedTime(&ms, e.first, e.second));
        tot += ms;
        cudaEventDestroy(e.first); cudaEventDestroy(e.second);
    }
    if (launches) *launches = (long)ev.size();
    ev.clear();
    return tot;
}
namespace {
struct HpRow { const char* name; unsigned long long ns = 0, n = 0; };
std::vector<HpRow>& hp_rows() { static auto* v = new std::vector<HpRow>(); return *v; }
inline unsigned long long now_ns() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return (unsigned long long)ts.tv_sec * 1000000000ull + ts.tv_nsec; }
}
bool hostprof_on() { static int on = -1; if (on < 0) { const char* s = std::getenv("ORION_B200_HOSTPROF"); on = (s && std::atoi(s) != 0) ? 1 : 0; } return on == 1; }
int hostprof_slot(const char* name) { hp_rows().push_back(HpRow{name}); return (int)hp_rows().size() - 1; }
HostProf::HostProf(int i) : id(i), t0(hostprof_on() ? now_ns() : 0) {}
HostProf::~HostProf() { if (t0) { auto& r = hp_rows()[id]; r.ns += now_ns() - t0; r.n++; } }
void hostprof_dump() {
    std::map<std::string, std::pair<unsigned long long, unsigned long long>> agg;       // template instantiations share a name
    for (auto& r : hp_rows()) { agg[r.name].first += r.ns; agg[r.name].second += r.n; r.ns = 0; r.n = 0; }
    for (auto& kv : agg)
        if (kv.second.second)
            std::fprintf(stderr, "hostprof %-28s n=%8llu  %9.2f ms  %7.2f us/call\n", kv.first.c_str(), kv.second.second,
                         kv.second.first / 1e6, kv.second.first / 1e3 / kv.second.second);
}
// ---- driver-API launch path (common.h) ----
namespace {
typedef CUresult (*CuLaunchKernelExFn)(const CUlaunchConfig*, CUfunction, void**, void**);
CuLaunchKernelExFn g_cu_launch = nullptr;
std::unordered_map<const void*, CUfunction>& fn_cache() { static auto* m = new std::unordered_map<const void*, CUfunction>(); return *m; }
}
bool driver_launch_enabled() {
    static int on = -1;
    if (on < 0) {
        const char* s = std::getenv("ORION_B200_DRIVER_LAUNCH");
        on = (s && std::atoi(s) != 0) ? 1 : 0;
        if (on) {
            void* fp = nullptr;
            cudaDriverEntryPointQueryResult qr;
            if (cudaGetDriverEntryPoint("cuLaunchKernelEx", &fp, cudaEnableDefault, &qr) != cudaSuccess || !fp ||
                qr != cudaDriverEntryPointSuccess) { cudaGetLastError(); on = 0; }
            else g_cu_launch = (CuLaunchKernelExFn)fp;
        }
    }
    return on == 1;
}
void driver_launch(const void* kernel, dim3 grid, dim3 block, size_t smem, cudaStream_t st, void** params) {
    auto& cache = fn_cache();
    auto it = cache.find(kernel);
    if (it == cache.end()) {
        cudaFunction_t f = nullptr;
        OB_CUDA(cudaGetFuncBySymbol(&f, kernel));
        it = cache.emplace(kernel, (CUfunction)f).first;
    }
    CUlaunchConfig cfg;
    std::memset(&cfg, 0, sizeof(cfg));
    cfg.gridDimX = grid.x; cfg.gridDimY = grid.y; cfg.gridDimZ = grid.z;
    cfg.blockDimX = block.x; cfg.blockDimY = block.y; cfg.blockDimZ = block.z;
    cfg.sharedMemBytes = (unsigned)smem; cfg.hStream = (CUstream)st;
    CUlaunchAttribute attr;
    std::memset(&attr, 0, sizeof(attr));
    attr.id = CU_LAUNCH_ATTRIBUTE_PROGRAMMATIC_STREAM_SERIALIZATION;
    attr.value.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = &attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
    const CUresult r = g_cu_launch(&cfg, it->second, params, nullptr);
    if (r != CUDA_SUCCESS) OB_FATAL("cuLaunchKernelEx failed with CUresult " + std::to_string((int)r));
}
bool dry_run() {
    static int on = -1;
    if (on < 0) { const char* s = std::getenv("ORION_B200_DRYRUN"); on = (s && std::atoi(s) != 0) ? 1 : 0; }
    return on == 1;
}
bool stream_fork_enabled() {
    static int on = -1;
    if (on < 0) { const char* s = std::getenv("ORION_B200_STREAMS"); on = (s && std::atoi(s) == 0) ? 0 : 1; }
    return on == 1;
}
StreamFork::StreamFork(int n) : parent_(stream()) {
    OB_HP("StreamFork");
    if (!stream_fork_enabled() || n <= 1) { joined_ = true; return; }
    cudaEvent_t e = take_event();
    OB_CUDA(cudaEventRecord(e, parent_));
    for (int i = 0; i < n; ++i) {
        cudaStream_t s;
        if (!g_side_pool_init) {                              // pre-created, handed out FIFO: forks of operations that run
            g_side_pool_init = true;                          // concurrently on different lanes get different side streams
            for (int k = 0; k < 16; ++k) { cudaStream_t t; OB_CUDA(cudaStreamCreateWithFlags(&t, cudaStreamNonBlocking)); g_side_pool.push_back(t); }
        }
        if (g_side_pool.empty()) OB_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
        else { s = g_side_pool.front(); g_side_pool.erase(g_side_pool.begin()); }
        OB_CUDA(cudaStreamWaitEvent(s, e, 0));
        side_.push_back(s);
    }
    g_event_pool.push_back(e);                            // recorded events may be reused once waited on (stream-ordered)
}
void StreamFork::branch(int i) {
    if (side_.empty()) return;
    g_cur = side_[i % side_.size()];
}
void StreamFork::join() {
    if (joined_) return;
    OB_HP("StreamFork::join");
    joined_ = true;
    for (cudaStream_t s : side_) {
        cudaEvent_t e = take_event();
        OB_CUDA(cudaEventRecord(e, s));
        OB_CUDA(cudaStreamWaitEvent(parent_, e, 0));
        g_event_pool.push_back(e);
        g_side_pool.push_back(s);
    }
    g_cur = (parent_ == g_stream) ? nullptr : parent_;
}

// Measured and rejected: a per-stream free list in front of the pool (cudaMallocAsync / cudaFreeAsync cost 2.4 + 0.5 us
// of host time per buffer, 52 k buffers per ResNet-20 inference).  It saved 10 ms of a 390 ms inference and held 30+ GB
// of cached blocks over the 25 lanes and side streams, so DevBuf stays directly on the stream-ordered pool.
namespace {
std::map<int, Reclaimer>& reclaimers() { static auto* m = new std::map<int, Reclaimer>(); return *m; }
int g_reclaimer_id = 0, g_pin_depth = 0;
unsigned long long g_use_clock = 1, g_pin_floor = ~0ull, g_reclaimed = 0;
bool g_reclaiming = false;
size_t g_dev_bytes = 0, g_dev_peak = 0;                    // bytes currently held through DevBuf, and their high-water mark
size_t g_dev_peak2 = 0, g_dev_mark = 0;                    // high-water mark since device_bytes_mark() (scratch need of one call)
// ORION_B200_MEM_LIMIT_MB: treat DevBuf allocations beyond this total as out of memory (exercises the reclaim path on a small ring)
size_t mem_limit() { static const size_t l = std::getenv("ORION_B200_MEM_LIMIT_MB") ? (size_t)std::atoll(std::getenv("ORION_B200_MEM_LIMIT_MB")) << 20 : 0; return l; }
cudaError_t try_alloc(void** p, size_t n) {
    if (mem_limit() && g_dev_bytes + n > mem_limit()) return cudaErrorMemoryAllocation;
    const cudaError_t e = cudaMallocAsync(p, n, stream());
    if (e != cudaSuccess) cudaGetLastError();
    return e;
}
}
int register_reclaimer(Reclaimer r) { reclaimers()[++g_reclaimer_id] = std::move(r); return g_reclaimer_id; }
void unregister_reclaimer(int id) { reclaimers().erase(id); }
unsigned long long use_tick() { return ++g_use_clock; }
unsigned long long pin_floor() { return g_pin_depth ? g_pin_floor : ~0ull; }
unsigned long long reclaimed_bytes() { return g_reclaimed; }
unsigned long long g_graph_epoch = 1;
int g_capturing = 0;
bool g_capture_failed = false;
bool graphs_enabled(int logN) {
    static const int forced = [] { const char* e = std::getenv("ORION_B200_GRAPHS"); return e ? (std::atoi(e) != 0 ? 1 : 0) : -1; }();
    if (forced >= 0) return forced == 1;
    return logN <= 14;
}
size_t device_bytes(bool peak) { return peak ? g_dev_peak : g_dev_bytes; }
void device_bytes_mark() { g_dev_mark = g_dev_bytes; g_dev_peak2 = g_dev_bytes; }
size_t device_bytes_since_mark() { return g_dev_peak2 - g_dev_mark; }
OpPin::OpPin() { if (g_pin_depth++ == 0) g_pin_floor = ++g_use_clock; }
OpPin::~OpPin() { --g_pin_depth; }

// the pool is out of memory: device-wide synchronisation (frees queued on any lane complete, nothing in flight reads what is
// about to be evicted), retry; then evict through the reclaimers, a few GB at a time, until the allocation fits
static void* alloc_under_pressure(size_t n) {
    capture_abort();
    void* p = nullptr;
    OB_CHECK(!g_reclaiming, "out of device memory while reclaiming");
    g_reclaiming = true;
    static const bool verbose = std::getenv("ORION_B200_DEBUG") != nullptr;
    for (int round = 0; round < 64; ++round) {
        OB_CUDA(cudaDeviceSynchronize());
        if (try_alloc(&p, n) == cudaSuccess) { g_reclaiming = false; return p; }
        const size_t want = std::max<size_t>(n + (n >> 2), mem_limit() ? mem_limit() / 8 : (size_t)6 << 30);
        size_t freed = 0;
        for (auto& kv : reclaimers()) { if (freed >= want) break; freed += kv.second(want - freed); }
        g_reclaimed += freed;
        if (freed) graphs_invalidate();
        if (verbose) std::fprintf(stderr, "[mem] allocation of %.2f GB failed: reclaimed %.2f GB (round %d)\n", n / 1073741824.0, freed / 1073741824.0, round);
        if (freed == 0) {
            // last resort: hand the pool's unused blocks back to the driver so that it can re-map them into one range
            cudaMemPool_t pool; int dev = 0;
            OB_CUDA(cudaGetDevice(&dev)); OB_CUDA(cudaDeviceGetDefaultMemPool(&pool, dev));
            OB_CUDA(cudaDeviceSynchronize());
            OB_CUDA(cudaMemPoolTrimTo(pool, 0));
            if (try_alloc(&p, n) == cudaSuccess) { g_reclaiming = false; return p; }
            break;
        }
    }
    g_reclaiming = false;
    size_t fr = 0, tot = 0; cudaMemGetInfo(&fr, &tot);
    OB_FATAL("CUDA error: out of memory allocating " + std::to_string(n >> 20) + " MiB (device free " + std::to_string(fr >> 20) + " of " +
             std::to_string(tot >> 20) + " MiB, nothing left to reclaim)");
    return nullptr;
}
CaptureArena* g_arena = nullptr;
static cudaStream_t g_capture_origin = nullptr;
bool capture_begin(cudaStream_t origin, CaptureArena* arena) {
    if (g_capturing) return false;
    if (cudaStreamBeginCapture(origin, cudaStreamCaptureModeRelaxed) != cudaSuccess) { cudaGetLastError(); return false; }
    g_capturing = 1; g_capture_failed = false; g_capture_origin = origin; g_arena = arena;
    return true;
}
cudaGraph_t capture_end() {
    if (!g_capturing) return nullptr;                          // aborted on the way
    g_capturing = 0; g_arena = nullptr;
    cudaGraph_t graph = nullptr;
    const cudaError_t e = cudaStreamEndCapture(g_capture_origin, &graph);
    if (e != cudaSuccess || g_capture_failed) { cudaGetLastError(); if (graph) cudaGraphDestroy(graph); return nullptr; }
    return graph;
}
void capture_abort() {
    if (!g_capturing) return;
    g_capturing = 0; g_capture_failed = true; g_arena = nullptr;
    cudaGraph_t graph = nullptr;
    cudaStreamEndCapture(g_capture_origin, &graph);
    if (graph) cudaGraphDestroy(graph);
    cudaGetLastError();
}
void* CaptureArena::alloc(size_t n, cudaStream_t s) {
    n = (n + 255) & ~(size_t)255;
    auto& fl = free_[s];
    auto it = fl.lower_bound(n);
    if (it != fl.end() && it->first <= n + (n >> 1)) {                // best fit, at most 50 % waste
        const size_t off = it->second;
        fl.erase(it);
        return base + off;
    }
    if (top + n > cap) { overflow = true; return nullptr; }          // the capture is discarded (DevBuf::alloc falls back to the pool)
    void* p = base + top;
    top += n;
    return p;
}
void CaptureArena::free(void* p, size_t n, cudaStream_t s) {
    if (overflow) return;
    n = (n + 255) & ~(size_t)255;
    free_[s].emplace(n, (size_t)((char*)p - base));
}
void DevBuf::alloc(size_t n) {
    release();
    if (n == 0) return;
    if (g_arena) {
        void* q = g_arena->alloc(n, stream());
        if (!g_arena->overflow) { p = q; bytes = n; arena = true; return; }
        capture_abort();                                       // scratch does not fit: the capture is over, the rest runs from the pool
    }
    OB_HP("cudaMallocAsync");
    const cudaError_t e = try_alloc(&p, n);
    if (e == cudaErrorMemoryAllocation) p = alloc_under_pressure(n);
    else OB_CUDA(e);
    bytes = n; g_dev_bytes += n; if (g_dev_bytes > g_dev_peak) g_dev_peak = g_dev_bytes;
    if (g_dev_bytes > g_dev_peak2) g_dev_peak2 = g_dev_bytes;
}
void DevBuf::release() {
    if (p && arena) { if (g_arena) g_arena->free(p, bytes, g_cur ? g_cur : g_stream); p = nullptr; bytes = 0; arena = false; return; }
    if (p) { OB_HP("cudaFreeAsync"); cudaFreeAsync(p, g_cur ? g_cur : g_stream); p = nullptr; g_dev_bytes -= bytes; bytes = 0; }
}

// ------------------------------------------------------------------------------------------------
// host number theory
// ------------------------------------------------------------------------------------------------
u64 mulmod_u64(u64 a, u64 b, u64 q) { return (u64)(((u128)a * b) % q); }
u64 powmod_u64(u64 a, u64 e, u64 q) {
    u64 r = 1 % q; a %= q;
    for (; e; e >>= 1) { if (e & 1) r = mulmod_u64(r, a, q); a = mulmod_u64(a, a, q); }
    return r;
}
u64 invmod_u64(u64 a, u64 q) { return powmod_u64(a % q, q - 2, q); }
u64 shoup_companion(u64 w, u64 q) { return (u64)(((u128)w << 64) / q); }

bool is_prime_u64(u64 n) {
    if (n < 2) return false;
    for (u64 p : {2ull, 3ull, 5ull, 7ull, 11ull, 13ull, 17ull, 19ull, 23ull, 29ull, 31ull, 37ull}) {
        if (n == p) return true;
        if (n % p == 0) return false;
    }
    u64 d = n - 1; int s = 0;
    while ((d & 1) == 0) { d >>= 1; ++s; }
    for (u64 a : {2ull, 3ull, 5ull, 7ull, 11ull, 13ull, 17ull, 19ull, 23ull, 29ull, 31ull, 37ull}) {
        u64 x = powmod_u64(a, d, n);
        if (x == 1 || x == n - 1) continue;
        bool composite = true;
        for (int r = 1; r < s && composite; ++r) { x = mulmod_u64(x, x, n); if (x == n - 1) composite = false; }
        if (composite) return false;
    }
    return true;
}

// NTT-friendly primes (= 1 mod nthroot) alternating above / below 2^bits, nearest first
// (the order Lattigo's NextAlternatingPrime produces; LG/scheme.go:57-65 reaches it via GenModuli).
std::vector<u64> gen_ntt_primes(int bits, u64 nthroot, int count, const std::vector<u64>& skip) {
    std::vector<u64> out;
    u64 up = (1ull << bits) + 1, down = (1ull << bits) + 1 - nthroot;
    bool up_ok = true, down_ok = true;
    auto take = [&](u64 c) {
        if (!is_prime_u64(c)) return;
        if (std::find(skip.begin(), skip.end(), c) != skip.end()) return;
        if (std::find(out.begin(), out.end(), c) != out.end()) return;
        out.push_back(c);
    };
    while ((int)out.size() < count) {
        OB_CHECK(up_ok || down_ok, "ran out of NTT-friendly primes of this size");
        if (up_ok) {
            if (up >> (bits + 1)) up_ok = false;
            else { u64 c = up; up += nthroot; take(c); }
        }
        if ((int)out.size() >= count) break;
        if (down_ok) {
            if (down < nthroot || (down >> (bits - 1)) == 0) down_ok = false;
            else { u64 c = down; down -= nthroot; take(c); }
        }
    }
    return out;
}

void gen_moduli(const std::vector<int>& logQ, const std::vector<int>& logP, int log_nthroot,
                std::vector<u64>& Q, std::vector<u64>& P) {
    std::map<int, int> need;
    for (int b : logQ) need[b]++;
    for (int b : logP) need[b]++;
    std::map<int, std::vector<u64>> pool;
    for (auto& kv : need) {
        OB_CHECK(kv.first >= 10 && kv.first <= 61, "prime size must be within 10..61 bits");
        pool[kv.first] = gen_ntt_primes(kv.first, 1ull << log_nthroot, kv.second, {});
    }
    std::map<int, int> used;
    Q.clear(); P.clear();
    for (int b : logQ) Q.push_back(pool[b][used[b]++]);
    for (int b : logP) P.push_back(pool[b][used[b]++]);
}

static void factor_distinct(u64 n, std::vector<u64>& f) {
    auto push = [&](u64 p) { if (std::find(f.begin(), f.end(), p) == f.end()) f.push_back(p); };
    for (u64 p = 2; p < 4096 && p * p <= n; ++p)
        if (n % p == 0) { push(p); while (n % p == 0) n /= p; }
    if (n == 1) return;
    if (is_prime_u64(n)) { push(n); return; }
    // Pollard rho (Brent-free, Floyd) for the remaining cofactor
    u64 d = n;
    for (u64 c = 1; d == n; ++c) {
        u64 x = 2, y = 2; d = 1;
        while (d == 1) {
            x = (mulmod_u64(x, x, n) + c) % n;
            y = (mulmod_u64(y, y, n) + c) % n; y = (mulmod_u64(y, y, n) + c) % n;
            u64 a = x > y ? x - y : y - x, b = n;
            while (b) { u64 t = a % b; a = b; b = t; }
            d = a;
        }
    }
    factor_distinct(d, f); factor_distinct(n / d, f);
}
u64 primitive_root(u64 q) {
    std::vector<u64> f; factor_distinct(q - 1, f);
    for (u64 g = 3;; ++g) {
        bool ok = true;
        for (u64 p : f) if (powmod_u64(g, (q - 1) / p, q) == 1) { ok = false; break; }
        if (ok) return g;
    }
}
u64 find_psi(u64 q, u64 nthroot) { return powmod_u64(primitive_root(q), (q - 1) / nthroot, q); }

ModConst make_modconst(u64 q) {
    ModConst m; std::memset(&m, 0, sizeof(m));
    OB_CHECK(q < (1ull << 62) && (q & 1), "modulus must be odd and below 2^62");
    m.q = q; m.bits = 64 - __builtin_clzll(q);
    m.mu = (u64)((((u128)1) << (62 + m.bits)) / q);
    u64 inv = 1;                                   // Newton iteration for q^-1 mod 2^64
    for (int i = 0; i < 6; ++i) inv *= 2 - q * inv;
    m.qinv_neg = (u64)(0 - inv);
    m.r1 = (u64)((((u128)1) << 64) % q); m.r1_sh = shoup_companion(m.r1, q);
    m.r2 = mulmod_u64(m.r1, m.r1, q);    m.r2_sh = shoup_companion(m.r2, q);
    m.qd = (double)q;
    m.qinv_d = 1.0 / (double)q;
    m.rinv_d = (double)invmod_u64(m.r1, q);
    m.one_sh = shoup_companion(1, q);
    return m;
}

// ------------------------------------------------------------------------------------------------
// Context
// ------------------------------------------------------------------------------------------------
static uint32_t bitrev_u32(uint32_t x, int bits) {
    uint32_t r = 0;
    for (int i = 0; i < bits; ++i) { r = (r << 1) | (x & 1); x >>= 1; }
    return r;
}

template <class T>
static void upload(DevBuf& d, const std::vector<T>& h) {
    capture_abort();                                           // a table built lazily: for real, from the pool
    d.alloc(h.size() * sizeof(T));
    if (!h.empty()) {
        OB_CUDA(cudaMemcpyAsync(d.p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, stream()));
        OB_CUDA(cudaStreamSynchronize(stream()));     // h may be a temporary
    }
}

Context::Context(int logN_, const std::vector<u64>& Q, const std::vector<u64>& P) {
    ensure_device();
    logN = logN_; N = 1 << logN; nQ = (int)Q.size(); nP = (int)P.size(); nmod = nQ + nP;
    g_pdl_auto = true;
    OB_CHECK(logN >= 10 && logN <= 17, "logN must be within 10..17");
    OB_CHECK(nmod <= 64, "at most 64 moduli");
    mod = Q; mod.insert(mod.end(), P.begin(), P.end());
    std::vector<ulonglong2> twf((size_t)nmod * N), twi((size_t)nmod * N), ninv(nmod);
    std::vector<double> twdf((size_t)nmod * N), twdi((size_t)nmod * N);
    for (int m = 0; m < nmod; ++m) {
        u64 q = mod[m];
        OB_CHECK((q - 1) % (2ull * N) == 0, "modulus is not 1 mod 2N");
        mc.push_back(make_modconst(q));
        u64 ps = find_psi(q, 2ull * N), psi_inv = invmod_u64(ps, q);
        psi.push_back(ps);
        u64 pf = 1, pi = 1;
        for (int i = 0; i < N; ++i) {
            uint32_t j = bitrev_u32((uint32_t)i, logN);
            twf[(size_t)m * N + j] = make_ulonglong2(pf, shoup_companion(pf, q));
            twi[(size_t)m * N + j] = make_ulonglong2(pi, shoup_companion(pi, q));
            twdf[(size_t)m * N + j] = (double)pf; twdi[(size_t)m * N + j] = (double)pi;   // exact below 2^53
            pf = mulmod_u64(pf, ps, q); pi = mulmod_u64(pi, psi_inv, q);
        }
        u64 ni = invmod_u64((u64)N % q, q);
        ninv[m] = make_ulonglong2(ni, shoup_companion(ni, q));
    }
    upload(d_mc, mc); upload(d_tw_fwd, twf); upload(d_tw_inv, twi); upload(d_ninv, ninv);
    upload(d_twd_fwd, twdf); upload(d_twd_inv, twdi);

    std::vector<ulonglong2> pinv(nQ), pmodq(nQ), qlinv((size_t)nQ * nQ);
    for (int i = 0; i < nQ; ++i) {
        u64 q = mod[i], pm = 1;
        for (int k = 0; k < nP; ++k) pm = mulmod_u64(pm, mod[nQ + k] % q, q);
        u64 pi = invmod_u64(pm, q);
        pinv[i] = make_ulonglong2(pi, shoup_companion(pi, q));
        pmodq[i] = make_ulonglong2(pm, shoup_companion(pm, q));
        for (int l = 0; l < nQ; ++l) {
            u64 v = (l == i) ? 0 : invmod_u64(mod[l] % q, q);
            qlinv[(size_t)l * nQ + i] = make_ulonglong2(v, shoup_companion(v, q));
        }
    }
    upload(d_pinv, pinv); upload(d_pmodq, pmodq); upload(d_qlinv, qlinv);
}

Context::~Context() {}

const LimbDesc* Context::limb_list(const std::vector<LimbDesc>& v) {
    OB_HP("limb_list");
    std::vector<uint32_t> key; key.reserve(v.size() * 4);
    for (auto& d : v) { key.push_back(d.in_off); key.push_back(d.out_off); key.push_back(d.mod); key.push_back(d.aux); }
    auto it = limb_cache_.find(key);
    if (it == limb_cache_.end()) {
        auto buf = std::make_unique<DevBuf>();
        upload(*buf, v);
        it = limb_cache_.emplace(std::move(key), std::move(buf)).first;
    }
    return it->second->as<LimbDesc>();
}

const BaseConv& Context::base_conv(const std::vector<int>& src, const std::vector<int>& dst) {
    std::vector<int> key = src; key.push_back(-1); key.insert(key.end(), dst.begin(), dst.end());
    auto it = bc_cache_.find(key);
    if (it != bc_cache_.end()) return *it->second;
    auto bc = std::make_unique<BaseConv>();
    int ns = (int)src.size(), nd = (int)dst.size();
    bc->ns = ns; bc->nd = nd; bc->h_src = src; bc->h_dst = dst;
    std::vector<u64> inv(2 * ns), mat((size_t)ns * nd), negvq((size_t)nd * (ns + 1)), hs(ns), hd(nd);
    std::vector<double> matd((size_t)ns * nd), negd((size_t)nd * (ns + 1));
    bc->f64_src = true;
    for (int i = 0; i < ns; ++i) if (mc[src[i]].bits > 46) bc->f64_src = false;
    for (int i = 0; i < ns; ++i) {
        u64 qi = mod[src[i]], prod = 1;
        for (int k = 0; k < ns; ++k) if (k != i) prod = mulmod_u64(prod, mod[src[k]] % qi, qi);
        inv[2 * i] = invmod_u64(prod, qi); inv[2 * i + 1] = shoup_companion(inv[2 * i], qi);
        hs[i] = mulmod_u64(qi - 1, (qi + 1) / 2, qi);                 // (Q-1)/2 mod q_i, Q = 0 mod q_i
    }
    for (int d = 0; d < nd; ++d) {
        u64 p = mod[dst[d]], all = 1;
        for (int i = 0; i < ns; ++i) {
            u64 prod = 1;
            for (int k = 0; k < ns; ++k) if (k != i) prod = mulmod_u64(prod, mod[src[k]] % p, p);
            mat[(size_t)i * nd + d] = mulmod_u64(prod, mc[dst[d]].r1, p);
            matd[(size_t)i * nd + d] = (double)prod;
            all = mulmod_u64(all, mod[src[i]] % p, p);
        }
        for (int v = 0; v <= ns; ++v) {
            negvq[(size_t)d * (ns + 1) + v] = (p - mulmod_u64((u64)v, all, p)) % p;
            negd[(size_t)d * (ns + 1) + v] = (double)negvq[(size_t)d * (ns + 1) + v];
        }
        hd[d] = mulmod_u64((all + p - 1) % p, (p + 1) / 2, p);
    }
    upload(bc->src_mod, src); upload(bc->dst_mod, dst);
    upload(bc->inv, inv); upload(bc->mat, mat); upload(bc->negvq, negvq);
    upload(bc->half_src, hs); upload(bc->half_dst, hd);
    upload(bc->mat_d, matd); upload(bc->negvq_d, negd);
    static const bool tc = !(std::getenv("ORION_B200_BC_TC") && std::atoi(std::getenv("ORION_B200_BC_TC")) == 0);   // on by default
    if (tc && ns <= 8) {
        // INT8 tensor-core contraction (ORION_B200_BC_TC=0 falls back to the scalar kernel).  y_i = sum_a y_{i,a} 2^(8a) (bytes), so
        //   sum_i y_i M[i][t] = sum_{(i,a)} y_{i,a} W[i][a][t] (mod p_t),  W[i][a][t] = M[i][t] 2^(8a) mod p_t  (Montgomery form kept),
        // and with W split into bytes b the contraction is 8 (or 4) exact u8 x u8 -> s32 GEMM columns per destination:
        //   S_b[x][t] = sum_{(i,a)} y_{i,a}[x] * byte_b(W[i][a][t]) <= 64 * 255^2 < 2^22,   out = REDC(sum_b S_b 2^(8b)).
        // Column c of n8-tile tau of destination block blk holds (t, b) = (4 blk + (c >> 1), 2 tau + (c & 1)), so that after the
        // MMAs lane j of a quad owns all eight byte sums of destination 4 blk + j.  K index: chunk ch, lane-quad position jB,
        // register half -> (i, a) = (jB + 4 ch, 0..3 | 4..7), matching the A fragments the kernel builds from the words of y.
        const int nblk = (nd + 3) / 4;
        std::vector<uint32_t> fr((size_t)nblk * 4 * 2 * 32 * 2, 0u);
        for (int blk = 0; blk < nblk; ++blk)
            for (int tau = 0; tau < 4; ++tau)
                for (int ch = 0; ch < 2; ++ch)
                    for (int lane = 0; lane < 32; ++lane) {
                        const int col = lane >> 2, jb = lane & 3, t = 4 * blk + (col >> 1), b = 2 * tau + (col & 1), i = jb + 4 * ch;
                        if (t >= nd || i >= ns) continue;
                        const u64 p = mod[dst[t]];
                        u64 w = mat[(size_t)i * nd + t];
                        uint32_t r[2] = {0u, 0u};
                        for (int a = 0; a < 8; ++a) {
                            r[a >> 2] |= (uint32_t)((w >> (8 * b)) & 0xFF) << (8 * (a & 3));
                            w = mulmod_u64(w, 256 % p, p);
                        }
                        const size_t o = ((((size_t)blk * 4 + tau) * 2 + ch) * 32 + lane) * 2;
                        fr[o] = r[0]; fr[o + 1] = r[1];
                    }
        upload(bc->tc_frag, fr);
    }
    return *bc_cache_.emplace(std::move(key), std::move(bc)).first->second;
}

u64 Context::galois_element(int rot) const {
    int ord = N / 2, k = ((rot % ord) + ord) % ord;
    u64 g = 1, mask = 2ull * N - 1;
    for (int i = 0; i < k; ++i) g = (g * 5) & mask;
    return g;
}

const uint32_t* Context::galois_index(u64 galEl) {
    auto it = gal_cache_.find(galEl);
    if (it == gal_cache_.end()) {
        std::vector<uint32_t> idx(N);
        u64 mask = 2ull * N - 1;
        for (int i = 0; i < N; ++i) {
            u64 e = (galEl * (2ull * bitrev_u32(i, logN) + 1)) & mask;
            idx[i] = bitrev_u32((uint32_t)((e - 1) >> 1), logN);
        }
        auto buf = std::make_unique<DevBuf>();
        upload(*buf, idx);
        it = gal_cache_.emplace(galEl, std::move(buf)).first;
    }
    return it->second->as<uint32_t>();
}

// the gather table of g^-1 is the scatter table of g (position of exponent e -> position of exponent g*e)
const uint32_t* Context::galois_scatter(u64 galEl) {
    const u64 mask = 2ull * N - 1;
    u64 inv = 1, b = galEl & mask;
    for (u64 e = (u64)N - 1; e; e >>= 1) { if (e & 1) inv = (inv * b) & mask; b = (b * b) & mask; }   // g^(phi(2N)-1), phi = N
    OB_CHECK(((inv * galEl) & mask) == 1, "Galois element is not a unit mod 2N");
    return galois_index(inv);
}

}  // namespace ob
