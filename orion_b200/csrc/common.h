// common.h -- shared host-side declarations for liborion_b200.so (B200 / sm_100a CKKS backend).
#pragma once
#include <functional>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>
#include <cuda_runtime.h>
#include "modarith.cuh"

namespace ob {

// The reference's Go shim panics (process abort) on any error (LG/evaluator.go:52-54); this backend
// does the same: print and abort.  There is no CPU fallback anywhere in this library.
[[noreturn]] inline void fatal(const char* file, int line, const std::string& msg) {
    std::fprintf(stderr, "orion_b200 fatal: %s  (%s:%d)\n", msg.c_str(), file, line);
    std::fflush(stderr);
    std::abort();
}
#define OB_FATAL(msg) ::ob::fatal(__FILE__, __LINE__, (msg))
#define OB_CHECK(cond, msg) do { if (!(cond)) OB_FATAL(std::string("check failed: " #cond " -- ") + (msg)); } while (0)
// While a launch sequence is being captured into a CUDA graph (g_capturing), a failing call does not abort: it marks the capture
// failed and the sequence carries on (nothing executes during capture); the capturer then discards the graph and runs eagerly.
extern int g_capturing;
extern bool g_capture_failed;
#define OB_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { \
    if (::ob::g_capturing) { ::ob::g_capture_failed = true; cudaGetLastError(); } \
    else OB_FATAL(std::string("CUDA error: ") + cudaGetErrorString(e__) + " in " #call); } } while (0)

// launch counter: every kernel of this library goes through OB_LAUNCH so bench.py can report gpu_launches
extern unsigned long long g_launches;
// algorithmic-byte counter (compulsory HBM traffic -- operands read once, results written once): every
// evaluator-level operation adds its figure, so a whole inference's roofline is (sum over the ops it ran) / time / peak
extern double g_alg_bytes;
inline void alg_bytes(double b) { g_alg_bytes += b; }
// Launches can be programmatic dependent launches (PDL): the next kernel of the stream may be scheduled while this one
// drains, and blocks in OB_PDL_SYNC() -- the first statement of EVERY kernel of this library -- until its predecessor has
// completed and flushed.  Overlaps the tail wave of large kernels with the head of the next; policy in context.cu
// (pdl_enabled: on for logN >= 15, ORION_B200_PDL=0/1 overrides).
bool pdl_enabled();
// Launch sequences whose descriptors are written by a KERNEL (csrc/noalloc.cu) must not be programmatic dependent launches: the
// consumers read their descriptors before griddepcontrol.wait (OB_PDL_LAUNCH / OB_PDL_WAIT below).  While a PdlSuppress is alive
// every launch is an ordinary, fully ordered one.
extern int g_pdl_suppress;
struct PdlSuppress { PdlSuppress() { ++g_pdl_suppress; } ~PdlSuppress() { --g_pdl_suppress; } };
// Kernel probe: while enabled, the launcher of the key switch's dominant kernel brackets it with CUDA events on its own
// stream (bench.py's roofline.achieved); probe_read sums the elapsed times and clears the list.
// Slots name the stages of a key switch (PROBE_* below); slot 0 is the dominant kernel.  Events between launches keep the
// next kernel from starting early (no PDL overlap across a probe), so bench.py enables only slot 0 inside the timed region
// and all slots in a separate pass for the per-stage split.
enum ProbeSlot { PROBE_IP_F64 = 0, PROBE_IP_INT, PROBE_INTT_C1, PROBE_MODUP, PROBE_NTT1_DIGITS, PROBE_INTT_P, PROBE_MODDOWN_BC,
                 PROBE_MODDOWN_NTT1, PROBE_MODDOWN_FIN, PROBE_SLOTS };
const char* probe_slot_name(int slot);
void probe_enable(int mask_or_on);            // 0 off, 1 slot 0 only, 2 every slot
void probe_begin(int slot = PROBE_IP_F64);
void probe_end(int slot = PROBE_IP_F64);
double probe_read_ms(long* launches, int slot = PROBE_IP_F64);
struct ProbeScope { int s; explicit ProbeScope(int slot) : s(slot) { probe_begin(s); } ~ProbeScope() { probe_end(s); } };
// Host-side enqueue profiler (ORION_B200_HOSTPROF=1): scoped wall-clock timers around the host work of the hot path
// (launch API, allocator, plan lookups, events); B200_HostProfDump prints the table.  Off: one predictable branch.
struct HostProf { int id; unsigned long long t0; explicit HostProf(int i); ~HostProf(); };
int hostprof_slot(const char* name);
bool hostprof_on();
void hostprof_dump();
#define OB_HP(name) static const int hp_slot__ = ::ob::hostprof_slot(name); ::ob::HostProf hp__(hp_slot__)
bool dry_run();        // ORION_B200_DRYRUN=1: skip the launches (host-side enqueue cost measurement only; results are garbage)
// Driver-API launch path (ORION_B200_DRIVER_LAUNCH=1; default off): the CUfunction of every kernel is resolved once
// (cudaGetFuncBySymbol) and launched with cuLaunchKernelEx obtained through cudaGetDriverEntryPoint -- no libcuda link
// dependency, no per-launch symbol lookup inside the runtime.  Measured on the ResNet-20 replay (58 k launches per image):
// 3.4 -> 2.7 us of host time per launch, and no change of the 0.331 s/image -- with the dataflow lanes the replay is bounded
// by the GPU-side dependency chains, not by the enqueue rate -- so the runtime path stays the default.
bool driver_launch_enabled();
void driver_launch(const void* kernel, dim3 grid, dim3 block, size_t smem, cudaStream_t st, void** params);
template <class... KArgs, class... Args>
inline void launch_kernel(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    if (dry_run()) return;
    OB_HP("cudaLaunchKernelEx");
    if (driver_launch_enabled()) {
        auto go = [&](KArgs... k) {
            void* params[] = {(void*)&k...};
            driver_launch((const void*)kernel, grid, block, smem, st, params);
        };
        go(KArgs(args)...);
        return;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
    OB_CUDA(cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...));
}
#define OB_LAUNCH(kernel, grid, block, smem, stream, ...) do { \
    ::ob::launch_kernel(kernel, dim3(grid), dim3(block), (size_t)(smem), (stream), __VA_ARGS__); ::ob::g_launches++; } while (0)
#ifdef __CUDACC__
// griddepcontrol.launch_dependents: the dependent grid may start launching once every CTA of this grid got here;
// griddepcontrol.wait: block until the grid this one depends on has completed (memory visible).
#define OB_PDL_SYNC() do { asm volatile("griddepcontrol.launch_dependents;"); asm volatile("griddepcontrol.wait;" ::: "memory"); } while (0)
// The two halves on their own: a kernel may read what NO KERNEL writes -- launch descriptors, per-modulus constants, conversion
// matrices (host-built tables, uploaded by memcpy before any launch that uses them) -- between OB_PDL_LAUNCH and OB_PDL_WAIT,
// so that the dependent loads of its prologue overlap the tail of its predecessor; everything a kernel may have produced
// (ciphertext limbs, scratch) is touched only after OB_PDL_WAIT.
#define OB_PDL_LAUNCH() asm volatile("griddepcontrol.launch_dependents;")
#define OB_PDL_WAIT() asm volatile("griddepcontrol.wait;" ::: "memory")
#endif

cudaStream_t stream();          // the stream work is currently issued on (the library's main stream unless forked)
cudaStream_t main_stream();
void set_current_stream(cudaStream_t s);     // nullptr = back to the main stream
// Pooled CUDA event (timing disabled) with shared ownership: ciphertext handles keep the event of their last writer and of
// readers on other streams (dataflow lanes, capi.cu); the event returns to the pool when the last holder drops it.
struct DevEvent {
    cudaEvent_t e = nullptr;
    DevEvent();
    ~DevEvent();
    DevEvent(const DevEvent&) = delete;
    DevEvent& operator=(const DevEvent&) = delete;
};
typedef std::shared_ptr<DevEvent> EvRef;
EvRef record_event(cudaStream_t s);          // a fresh pooled event recorded on s

// Fork / join over side streams: independent sub-computations (the two EvalMod halves of a bootstrap, the independent
// products of a polynomial evaluation) are issued on side streams so their launch-latency-bound kernel chains overlap
// on the GPU.  Single host thread; branch(i) switches the current stream, join() makes the parent wait for all.
class StreamFork {
  public:
    explicit StreamFork(int n);
    ~StreamFork() { join(); }
    void branch(int i);          // subsequent work goes to side stream i (0 <= i < n)
    void join();                 // parent stream waits for every side stream; current stream = parent again
    int size() const { return (int)side_.size(); }
  private:
    cudaStream_t parent_;
    std::vector<cudaStream_t> side_;
    bool joined_ = false;
};
bool stream_fork_enabled();      // ORION_B200_STREAMS=0 disables forking (everything on one stream)
void ensure_device();           // aborts loudly when no CUDA device is usable

// stream-ordered device buffer (cudaMallocAsync pool with an unlimited release threshold -> frees are cached)
// ---- memory pressure ----
// A logN = 16 network does not fit in 180 GB with everything resident (ResNet-20: 115 rotation keys x 595 MB, 8 381 encoded
// diagonals = 50 GB, three bootstrappers' keys and DFT factors).  When the stream-ordered pool cannot satisfy an allocation,
// DevBuf::alloc synchronises the device, retries, and then asks the registered reclaimers -- least-recently-used Galois keys
// that can be regenerated from the resident secret key, encoded bootstrap factor sets that can be re-encoded -- to free memory
// until the allocation succeeds; only when nothing is left to give back is the error fatal.  State an operation is still
// reading on the host side is protected by OpPin: objects touched since the outermost live OpPin was taken are not evictable.
typedef std::function<size_t(size_t want)> Reclaimer;       // frees (stream-ordered) up to ~want bytes, returns what it freed
int register_reclaimer(Reclaimer r);
void unregister_reclaimer(int id);
unsigned long long use_tick();                              // monotonic use clock (LRU stamps)
unsigned long long pin_floor();                             // stamps >= this belong to the operation in progress
struct OpPin { OpPin(); ~OpPin(); OpPin(const OpPin&) = delete; OpPin& operator=(const OpPin&) = delete; };
// Captured CUDA graphs (bootstrap.cu, polyeval.cu) bake device addresses of shared state: whoever frees or rebuilds such state
// (evicted Galois keys, re-encoded / dropped bootstrap factor sets, deleted polynomials or schemes) bumps the epoch, and a
// graph captured under an older epoch is discarded and captured again.
extern unsigned long long g_graph_epoch;
inline void graphs_invalidate() { ++g_graph_epoch; }
bool graphs_enabled(int logN);                             // ORION_B200_GRAPHS=0/1; default: on for logN <= 14
unsigned long long reclaimed_bytes();                       // diagnostics: total bytes given back under pressure
size_t device_bytes(bool peak);
void device_bytes_mark();                                   // start measuring the scratch need of a call ...
size_t device_bytes_since_mark();                           // ... its high-water mark above the level at the mark                             // bytes held through DevBuf now / at their high-water mark

// While a sequence is captured into a CUDA graph, its scratch buffers come from a CaptureArena instead of the pool: a persistent
// device range owned by the graph, carved by a host-side allocator with one free list per stream (a block freed on a stream is
// reused only by later allocations on that stream -- the ordering contract of cudaFreeAsync / cudaMallocAsync), so the graph holds
// kernels and copies only (memory nodes made a 1 700-node bootstrap graph cost 10 ms per launch on the host).
struct CaptureArena {
    char* base = nullptr;
    size_t cap = 0, top = 0;
    bool overflow = false;
    std::map<cudaStream_t, std::multimap<size_t, size_t>> free_;      // stream -> (size -> offset)
    void* alloc(size_t n, cudaStream_t s);
    void free(void* p, size_t n, cudaStream_t s);
};
extern CaptureArena* g_arena;                                 // non-null while capturing
// capture_begin / capture_end bracket a launch sequence on `origin`; capture_end returns the graph, or nullptr when anything in
// between failed.  capture_abort() is called wherever PERSISTENT device state is created lazily (tables, plans, keys, encoded
// factors): such state must be built for real, from the pool, so the capture is ended on the spot (the rest of the sequence then
// runs eagerly on scratch addresses -- garbage in, garbage out, confined to the arena) and the capturer repeats the whole call
// eagerly.  Steady-state sequences never get there: the capturer runs a key eagerly once before it captures it.
bool capture_begin(cudaStream_t origin, CaptureArena* arena);
cudaGraph_t capture_end();
void capture_abort();
struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    bool arena = false;                                        // carved from a CaptureArena: never goes back to the pool
    DevBuf() {}
    explicit DevBuf(size_t n) { alloc(n); }
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    DevBuf(DevBuf&& o) noexcept : p(o.p), bytes(o.bytes), arena(o.arena) { o.p = nullptr; o.bytes = 0; o.arena = false; }
    DevBuf& operator=(DevBuf&& o) noexcept { if (this != &o) { release(); p = o.p; bytes = o.bytes; arena = o.arena; o.p = nullptr; o.bytes = 0; o.arena = false; } return *this; }
    ~DevBuf() { release(); }
    void alloc(size_t n);
    void release();
    u64* u() const { return (u64*)p; }
    template <class T> T* as() const { return (T*)p; }
};

// one limb of a batched kernel launch: where it is read, where it is written, which modulus
struct LimbDesc { uint32_t in_off, out_off, mod, aux; };   // offsets in units of N words

}  // namespace ob
