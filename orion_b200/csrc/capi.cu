// capi.cu -- the Orion-facing C ABI: the export names, argument order and conventions of the Go shim
// orion/backend/lattigo/*.go (bound by orion/backend/lattigo/bindings.py:126-746), backed by the device-resident
// Scheme.  One process-global scheme, integer handles from smallest-free-int tables, arrays returned by value as
// {Data, Length} and released with FreeCArray, abort on error -- all as in the reference.
// Declarations with per-function reference citations: include/orion_b200_scheme.h.
#include "scheme.h"
#include "lintrans.h"
#include "polyeval.h"
#include "bootstrap.h"
#include "../../include/orion_b200_scheme.h"
#include <cmath>
#include <random>

using namespace ob;

namespace ob {
std::unique_ptr<Scheme>& g_scheme = *new std::unique_ptr<Scheme>();   // heap-allocated and never destroyed: no static-destruction-order hazards (and no CUDA calls) at process exit
static u64 g_seed = 0;
static bool g_seed_set = false;
Scheme& S() {
    if (!g_scheme) OB_FATAL("no scheme: call NewScheme first");
    return *g_scheme;
}
}  // namespace ob

namespace {
template <class T, class A> A make_array(const std::vector<T>& v) {
    A a;
    a.Length = (unsigned long)v.size();
    a.Data = (decltype(a.Data))std::malloc(std::max<size_t>(1, v.size()) * sizeof(*a.Data));
    for (size_t i = 0; i < v.size(); ++i) a.Data[i] = (std::remove_pointer_t<decltype(a.Data)>)v[i];
    return a;
}
int push_ct(std::unique_ptr<Ciphertext> c) { return S().cts.add(std::move(c)); }
int push_pt(std::unique_ptr<Plaintext> p) { return S().pts.add(std::move(p)); }
Ciphertext& CT(int id) { return S().cts.get(id, "ciphertext"); }
Plaintext& PT(int id) { return S().pts.get(id, "plaintext"); }
std::unique_ptr<Ciphertext> fresh() { return std::make_unique<Ciphertext>(); }

// ---- dataflow lanes -------------------------------------------------------------------------------------------
// The reference's Python layer issues one backend call per ciphertext (`for id in tensor.ids: EvaluatePolynomial(id)`,
// PY/poly_evaluator.py:28-31, PY/lt_evaluator.py:172-193); the calls on different ciphertexts of a tensor are
// independent, and at logN = 13 one call cannot fill a B200.  Every evaluator entry point therefore runs on the stream
// its operands live on ("follow the data"); a heavy call (bootstrap, polynomial, linear transform, key switch) whose
// operands are all settled on the main stream opens the next lane.  A handle remembers the event of its last writer
// (waited for by consumers on other lanes) and the events of readers on other lanes (waited for by the next writer and
// by the free).  The main stream never waits on a lane, so it stays a cheap fork point; host-visible results
// (Decrypt, Export*) synchronise the lane they ran on.  ORION_B200_ASYNC=0 keeps everything on the main stream.
constexpr int NLANES = 8;
cudaStream_t g_lanes[NLANES];
int g_lanes_ready = 0, g_next_lane = 0;
bool lanes_enabled() {
    static int on = -1;
    if (on < 0) { const char* e = std::getenv("ORION_B200_ASYNC"); on = (e && std::atoi(e) == 0) ? 0 : 1; if (!stream_fork_enabled()) on = 0; }
    return on == 1;
}
cudaStream_t lane(int i) {
    if (!g_lanes_ready) {
        ensure_device();
        for (int k = 0; k < NLANES; ++k) OB_CUDA(cudaStreamCreateWithFlags(&g_lanes[k], cudaStreamNonBlocking));
        g_lanes_ready = 1;
    }
    return g_lanes[i % NLANES];
}
EvRef* const g_tail = new EvRef[NLANES];                  // last event recorded on each lane (leaked: no destructor order at exit)
int lane_index(cudaStream_t s) { for (int k = 0; k < NLANES; ++k) if (g_lanes_ready && g_lanes[k] == s) return k; return -1; }
cudaStream_t pick_lane() {                                // an idle lane if there is one, else round-robin
    lane(0);
    for (int k = 0; k < NLANES; ++k) {
        const int c = (g_next_lane + k) % NLANES;
        if (!g_tail[c] || cudaEventQuery(g_tail[c]->e) == cudaSuccess) { g_next_lane = c + 1; return g_lanes[c]; }
    }
    (void)cudaGetLastError();
    return g_lanes[g_next_lane++ % NLANES];
}
void wait_on(cudaStream_t s, const EvRef& e) { if (e) OB_CUDA(cudaStreamWaitEvent(s, e->e, 0)); }
void prune(std::vector<EvRef>& v) {                       // drop reader events that have completed
    size_t k = 0;
    for (auto& e : v) if (cudaEventQuery(e->e) != cudaSuccess) v[k++] = e;
    v.resize(k);
    (void)cudaGetLastError();                             // cudaErrorNotReady is not an error
}

struct Op {
    cudaStream_t s = nullptr;                             // nullptr: the main stream
    std::vector<Track*> reads, writes;
    Op(bool heavy, std::initializer_list<Track*> w, std::initializer_list<Track*> r) : reads(r), writes(w) {
        if (lanes_enabled()) {
            // follow the data while the operand is the last thing queued on its lane (a true chain); once that lane has
            // moved on (a sibling consumer of the same operand was queued there) take another lane and wait for the operand
            Track* first = nullptr;
            for (Track* t : writes) if (!first && t->ready) first = t;
            for (Track* t : reads) if (!first && t->ready) first = t;
            if (first) {
                const int h = lane_index(first->home);
                s = (h >= 0 && g_tail[h] == first->ready) ? first->home : pick_lane();
            } else if (heavy) {
                s = pick_lane();
            }
        }
        cudaStream_t run = s ? s : main_stream();
        if (s) wait_on(s, record_event(main_stream()));   // everything already queued on the main stream comes first
        for (Track* t : writes) {
            if (t->ready && t->home != run) wait_on(run, t->ready);
            for (auto& b : t->busy) wait_on(run, b);      // readers on other lanes must be done before the overwrite
            t->busy.clear();
        }
        for (Track* t : reads) if (t->ready && t->home != run) wait_on(run, t->ready);
        set_current_stream(s);
    }
    void done(std::initializer_list<Track*> outs) {
        if (s) {
            EvRef e = record_event(s);
            g_tail[lane_index(s)] = e;
            for (Track* t : writes) { t->ready = e; t->home = s; }
            for (Track* t : outs) { t->ready = e; t->home = s; t->busy.clear(); }
            for (Track* t : reads) {
                bool written = false;
                for (Track* w : writes) if (w == t) written = true;
                if (written || t->home == s) continue;    // same lane: stream order is enough
                if (t->busy.size() >= 8) prune(t->busy);
                t->busy.push_back(e);
            }
        } else {
            for (Track* t : writes) { t->ready.reset(); t->home = nullptr; }
            for (Track* t : outs) { t->ready.reset(); t->home = nullptr; t->busy.clear(); }
        }
        set_current_stream(nullptr);
    }
};
// new ciphertext from f() reading the given handles
template <class F> int op_new(bool heavy, std::initializer_list<Track*> reads, F&& f) {
    Op op(heavy, {}, reads);
    std::unique_ptr<Ciphertext> o = f();
    op.done({&o->trk});
    return push_ct(std::move(o));
}
// f() overwrites the handle `w` (and reads the others)
template <class F> void op_inplace(bool heavy, Track* w, std::initializer_list<Track*> reads, F&& f) {
    Op op(heavy, {w}, reads);
    f();
    op.done({});
}
// run f() where the handle lives and leave the host synchronised with it (Decrypt, Export*)
template <class F> void op_host(Track* t, F&& f) {
    Op op(false, {}, {t});
    f();
    OB_CUDA(cudaStreamSynchronize(stream()));
    op.done({});
}
// free a tracked buffer without making the main stream wait: the free is queued where the buffer lives, behind its readers
template <class T> void release_tracked(HandleTable<T>& tab, int id) {
    auto it = tab.live.find(id);
    if (it == tab.live.end()) return;
    Track& t = it->second->trk;
    if (t.ready || !t.busy.empty()) {
        cudaStream_t s = t.home ? t.home : lane(NLANES - 1);
        for (auto& b : t.busy) wait_on(s, b);
        set_current_stream(s);
        tab.del(id);
        set_current_stream(nullptr);
    } else {
        tab.del(id);
    }
}
void settle_all() {                                       // before shared device state is torn down or rebuilt
    cudaDeviceSynchronize();
    for (int k = 0; k < NLANES; ++k) g_tail[k].reset();
}
}  // namespace

extern "C" {

// ---------------------------------------------------------------- scheme
void B200_SetSeed(uint64_t seed) { g_seed = seed; g_seed_set = true; }

void NewScheme(int logN, int* logQ, int lenQ, int* logP, int lenP, int logScale, int h, char* ringType, char* keysPath, char* ioMode) {
    (void)keysPath; (void)ioMode;
    SchemeParams p;
    p.logN = logN; p.log_scale = logScale; p.h = h;
    p.logQ.assign(logQ, logQ + lenQ); p.logP.assign(logP, logP + lenP);
    std::string rt = ringType ? ringType : "standard";
    for (auto& ch : rt) ch = (char)std::tolower(ch);
    p.conjugate_invariant = (rt != "standard");
    // default: ChaCha20 keyed with 2 x 256 bits of OS randomness (Lattigo: blake2b XOF keyed from crypto/rand);
    // B200_SetSeed switches to the oracle's invertible test generator -- parity tests only
    const SchemeRng rng = g_seed_set ? SchemeRng::test_mode(g_seed) : SchemeRng::from_os();
    settle_all();
    lt_reset(); poly_reset(); boot_reset();
    g_scheme.reset();
    g_scheme = std::make_unique<Scheme>(p, rng);
}
void DeleteScheme(void) {
    settle_all();
    lt_reset(); poly_reset(); boot_reset();
    g_scheme.reset();
}
void FreeCArray(void* p) { std::free(p); }

// ---------------------------------------------------------------- tensors
// After DeleteScheme the tables are gone and a late Delete* is a no-op, as in the Go shim (LG/scheme.go:89-100 resets the heaps,
// LG/tensors.go:35-42 deletes from a map): the reference's Python objects free their handles from __del__, in any order, also
// at interpreter exit (PY/tensors.py:16-22,112-118).
void DeletePlaintext(int id) { if (!g_scheme) return; release_tracked(S().pts, id); }
void DeleteCiphertext(int id) { if (!g_scheme) return; release_tracked(S().cts, id); }
// scales above 2^64 (a product before its rescale) saturate instead of overflowing the cast
static unsigned long scale_to_ulong(scale_t s) { return s >= 18446744073709551615.0L ? ~0ul : (s <= 0 ? 0ul : (unsigned long)s); }
unsigned long GetPlaintextScale(int id) { return scale_to_ulong(PT(id).scale); }
unsigned long GetCiphertextScale(int id) { return scale_to_ulong(CT(id).scale); }
void SetPlaintextScale(int id, unsigned long scale) { PT(id).scale = (scale_t)scale; }
void SetCiphertextScale(int id, unsigned long scale) { CT(id).scale = (scale_t)scale; }
int GetPlaintextLevel(int id) { return PT(id).level; }
int GetCiphertextLevel(long id) { return CT((int)id).level; }
int GetPlaintextSlots(long id) { return 1 << PT((int)id).log_slots; }
int GetCiphertextSlots(long id) { return 1 << CT((int)id).log_slots; }
int GetCiphertextDegree(long id) { return CT((int)id).degree; }
ArrayResultUInt64 GetModuliChain(void) {
    std::vector<u64> q(S().ctx->mod.begin(), S().ctx->mod.begin() + S().nQ);
    return make_array<u64, ArrayResultUInt64>(q);
}
ArrayResultInt GetLivePlaintexts(void) { return make_array<int, ArrayResultInt>(S().pts.ids()); }
ArrayResultInt GetLiveCiphertexts(void) { return make_array<int, ArrayResultInt>(S().cts.ids()); }

// ---------------------------------------------------------------- key generator
void NewKeyGenerator(void) { S(); }
void GenerateSecretKey(void) { S().gen_secret_key(); }
void GeneratePublicKey(void) { S().gen_public_key(); }
void GenerateRelinearizationKey(void) { S().gen_relin_key(); }
void GenerateEvaluationKeys(void) { S(); }     // the reference only wraps the relin key in an (empty) key set here
ArrayResultByte SerializeSecretKey(void) {
    // own container, not Lattigo's MarshalBinary: "OB2SK1\0\0" | u64 N | N x int8 ternary coefficients
    Scheme& s = S();
    OB_CHECK(s.has_sk(), "no secret key to serialize");
    std::vector<unsigned char> b(16 + s.N);
    std::memcpy(b.data(), "OB2SK1\0\0", 8);
    u64 n = s.N; std::memcpy(b.data() + 8, &n, 8);
    for (int i = 0; i < s.N; ++i) b[16 + i] = (unsigned char)(signed char)s.sk_coeffs[i];
    return make_array<unsigned char, ArrayResultByte>(b);
}
void LoadSecretKey(unsigned char* data, unsigned long len) {
    Scheme& s = S();
    OB_CHECK(len == (unsigned long)(16 + s.N) && std::memcmp(data, "OB2SK1\0\0", 8) == 0, "LoadSecretKey: not an orion_b200 secret key blob");
    std::vector<int64_t> c(s.N);
    for (int i = 0; i < s.N; ++i) c[i] = (int64_t)(signed char)data[16 + i];
    s.set_secret_key(c);
}

// ---------------------------------------------------------------- encoder / encryptor
void NewEncoder(void) { S(); }
int Encode(float* values, int lenValues, int level, unsigned long scale) {
    Scheme& s = S();
    std::vector<double> v(values, values + lenValues);
    return push_pt(s.encode(v.data(), lenValues, level, (scale_t)scale, s.log_max_slots));
}
ArrayResultFloat Decode(int ptID) {
    std::vector<double2> v = S().decode(PT(ptID));          // plaintexts are always settled: created synchronously or on the main stream
    std::vector<float> out(1 << S().log_max_slots, 0.0f);
    for (size_t i = 0; i < v.size() && i < out.size(); ++i) out[i] = (float)v[i].x;
    return make_array<float, ArrayResultFloat>(out);
}
void NewEncryptor(void) { S(); }
void NewDecryptor(void) { S(); }
int Encrypt(int ptID) { return op_new(false, {&PT(ptID).trk}, [&] { return S().encrypt(PT(ptID)); }); }
int Decrypt(int ctID) {
    std::unique_ptr<Plaintext> p;
    op_host(&CT(ctID).trk, [&] { p = S().decrypt(CT(ctID)); });
    return push_pt(std::move(p));
}

// ---------------------------------------------------------------- evaluator
void B200_NoteEvaluatorKey(uint64_t g);
void AddRotationKey(int rot) {
    u64 g = S().ctx->galois_element(rot); B200_NoteEvaluatorKey(g);
    if (!S().has_galois_key(g)) { S().galois_key(g); OB_CUDA(cudaStreamSynchronize(stream())); }   // shared by every lane from now on
}
void NewEvaluator(void) {
    Scheme& s = S();
    if (!s.has_sk()) return;
    for (int i = 1; i < (1 << s.log_max_slots); i *= 2) AddRotationKey(i);                        // LG/evaluator.go:25-31
}
#define TRK(id) (&CT(id).trk)
#define PTRK(id) (&PT(id).trk)
int Negate(int ct) { return op_new(false, {TRK(ct)}, [&] { auto o = fresh(); S().mul_int(CT(ct), -1, *o); return o; }); }
int Rotate(int ct, int rot) { AddRotationKey(rot); op_inplace(true, TRK(ct), {}, [&] { S().rotate(CT(ct), rot, CT(ct)); }); return ct; }
int RotateNew(int ct, int rot) { AddRotationKey(rot); return op_new(true, {TRK(ct)}, [&] { auto o = fresh(); S().rotate(CT(ct), rot, *o); return o; }); }
int Rescale(int ct) { op_inplace(false, TRK(ct), {}, [&] { S().rescale(CT(ct), CT(ct)); }); return ct; }
int RescaleNew(int ct) {                                                                    // LG/evaluator.go:91-99: rescales the input too
    op_inplace(false, TRK(ct), {}, [&] { S().rescale(CT(ct), CT(ct)); });
    return op_new(false, {TRK(ct)}, [&] { return S().clone(CT(ct)); });
}
int AddScalar(int ct, float c) { op_inplace(false, TRK(ct), {}, [&] { S().add_const(CT(ct), (double)c, CT(ct)); }); return ct; }
int AddScalarNew(int ct, float c) { return op_new(false, {TRK(ct)}, [&] { auto o = fresh(); S().add_const(CT(ct), (double)c, *o); return o; }); }
int SubScalar(int ct, float c) { op_inplace(false, TRK(ct), {}, [&] { S().add_const(CT(ct), -(double)c, CT(ct)); }); return ct; }
int SubScalarNew(int ct, float c) { return op_new(false, {TRK(ct)}, [&] { auto o = fresh(); S().add_const(CT(ct), -(double)c, *o); return o; }); }
int MulScalarInt(int ct, int k) { op_inplace(false, TRK(ct), {}, [&] { S().mul_int(CT(ct), k, CT(ct)); }); return ct; }
int MulScalarIntNew(int ct, int k) { return op_new(false, {TRK(ct)}, [&] { auto o = fresh(); S().mul_int(CT(ct), k, *o); return o; }); }
int MulScalarFloat(int ct, float c) { op_inplace(false, TRK(ct), {}, [&] { S().mul_float(CT(ct), (double)c, CT(ct)); }); return ct; }
int MulScalarFloatNew(int ct, float c) { return op_new(false, {TRK(ct)}, [&] { auto o = fresh(); S().mul_float(CT(ct), (double)c, *o); return o; }); }
int AddPlaintext(int ct, int pt) { op_inplace(false, TRK(ct), {PTRK(pt)}, [&] { S().add_pt(CT(ct), PT(pt), CT(ct), false); }); return ct; }
int AddPlaintextNew(int ct, int pt) { return op_new(false, {TRK(ct), PTRK(pt)}, [&] { auto o = fresh(); S().add_pt(CT(ct), PT(pt), *o, false); return o; }); }
int SubPlaintext(int ct, int pt) { op_inplace(false, TRK(ct), {PTRK(pt)}, [&] { S().add_pt(CT(ct), PT(pt), CT(ct), true); }); return ct; }
int SubPlaintextNew(int ct, int pt) { return op_new(false, {TRK(ct), PTRK(pt)}, [&] { auto o = fresh(); S().add_pt(CT(ct), PT(pt), *o, true); return o; }); }
int MulPlaintext(int ct, int pt) { op_inplace(false, TRK(ct), {PTRK(pt)}, [&] { S().mul_pt(CT(ct), PT(pt), CT(ct)); }); return ct; }
int MulPlaintextNew(int ct, int pt) { return op_new(false, {TRK(ct), PTRK(pt)}, [&] { auto o = fresh(); S().mul_pt(CT(ct), PT(pt), *o); return o; }); }
int AddCiphertext(int a, int b) { op_inplace(false, TRK(a), {TRK(b)}, [&] { S().add_ct(CT(a), CT(b), CT(a), false); }); return a; }
int AddCiphertextNew(int a, int b) { return op_new(false, {TRK(a), TRK(b)}, [&] { auto o = fresh(); S().add_ct(CT(a), CT(b), *o, false); return o; }); }
int SubCiphertext(int a, int b) { op_inplace(false, TRK(a), {TRK(b)}, [&] { S().add_ct(CT(a), CT(b), CT(a), true); }); return a; }
int SubCiphertextNew(int a, int b) { return op_new(false, {TRK(a), TRK(b)}, [&] { auto o = fresh(); S().add_ct(CT(a), CT(b), *o, true); return o; }); }
int MulRelinCiphertext(int a, int b) { op_inplace(true, TRK(a), {TRK(b)}, [&] { S().mul_relin(CT(a), CT(b), CT(a)); }); return a; }
int MulRelinCiphertextNew(int a, int b) { return op_new(true, {TRK(a), TRK(b)}, [&] { auto o = fresh(); S().mul_relin(CT(a), CT(b), *o); return o; }); }

// ---------------------------------------------------------------- extras required by the (drifted) Python layer
int CloneCiphertext(int ct) { return op_new(false, {TRK(ct)}, [&] { return S().clone(CT(ct)); }); }                    // PY/tensors.py:229
int ModDropCiphertext(int ct) { Ciphertext& c = CT(ct); OB_CHECK(c.level >= 1, "mod drop at level 0"); op_inplace(false, TRK(ct), {}, [&] { S().drop_level(c, c.level - 1); }); return ct; }  // PY/evaluator.py:36-40

// ---------------------------------------------------------------- interop: raw limbs in / out (parity tests, e2e bench, multi-GPU gather)
int B200_ImportCiphertext(const uint64_t* data, int level, int degree, double scale, int logSlots) {
    Scheme& s = S();
    OB_CHECK(logSlots >= 0 && logSlots <= s.log_max_slots, "B200_ImportCiphertext: logSlots out of range");
    auto c = s.new_ct(level, degree, (scale_t)scale, logSlots > 0 ? logSlots : s.log_max_slots);   // 0 = full slots
    OB_CUDA(cudaMemcpyAsync(c->data.p, data, c->data.bytes, cudaMemcpyHostToDevice, stream()));
    OB_CUDA(cudaStreamSynchronize(stream()));
    return push_ct(std::move(c));
}
void B200_ExportCiphertext(int ct, uint64_t* out) {
    Ciphertext& c = CT(ct);
    op_host(&c.trk, [&] {
        OB_CUDA(cudaMemcpyAsync(out, c.data.p, (size_t)(c.degree + 1) * (c.level + 1) * S().N * 8, cudaMemcpyDeviceToHost, stream()));
    });
}
void* B200_CiphertextDevicePtr(int ct) {                  // the caller touches the buffer from its own streams: settle it first
    Ciphertext& c = CT(ct);
    if (c.trk.ready || !c.trk.busy.empty()) { OB_CUDA(cudaDeviceSynchronize()); c.trk = Track(); }
    return c.data.p;
}
// ---- multi-GPU exchange on device memory (orion_b200/parallel.py): NCCL reads / writes a ciphertext's limbs where they live.
// B200_NewCiphertext allocates an empty ciphertext with the given metadata (the receive buffer of a remote ciphertext);
// the scale travels as two doubles (hi + lo) so the 64-bit mantissa of the library's long double survives exactly.
// B200_CiphertextAcquire orders the library's MAIN stream after the handle's last writer (and, for_write, after its readers
// on other lanes) with device-side event waits only -- no host synchronisation -- and marks the handle as settled on the
// main stream: external work enqueued on that stream next (torch.distributed with B200_Stream() as the current stream)
// is ordered after the producers, and every later library call on the handle is ordered after that work.
int B200_NewCiphertext(int level, int degree, double scaleHi, double scaleLo, int logSlots) {
    Scheme& s = S();
    OB_CHECK(level >= 0 && level <= s.max_level && degree >= 1 && degree <= 2, "B200_NewCiphertext: bad level or degree");
    auto c = s.new_ct(level, degree, (scale_t)scaleHi + (scale_t)scaleLo, logSlots > 0 ? logSlots : s.log_max_slots);
    return push_ct(std::move(c));
}
void* B200_CiphertextAcquire(int ct, int for_write) {
    Ciphertext& c = CT(ct);
    Track& t = c.trk;
    cudaStream_t m = main_stream();
    if (t.ready && t.home != m) wait_on(m, t.ready);
    if (for_write) { for (auto& b : t.busy) wait_on(m, b); t.busy.clear(); }
    t.ready.reset(); t.home = nullptr;
    return c.data.p;
}
void B200_GetCiphertextScaleParts(int ct, double* hi, double* lo) {
    const scale_t sc = CT(ct).scale;
    *hi = (double)sc; *lo = (double)(sc - (scale_t)*hi);
}
int B200_GetCiphertextLogSlots(int ct) { return CT(ct).log_slots; }
double B200_GetCiphertextScaleF64(int ct) { return (double)CT(ct).scale; }
void B200_SetCiphertextScaleF64(int ct, double s) { CT(ct).scale = (scale_t)s; }
void B200_ExportPlaintext(int pt, uint64_t* out) {
    Plaintext& p = PT(pt);
    OB_CUDA(cudaMemcpyAsync(out, p.data.p, (size_t)(p.level + 1) * S().N * 8, cudaMemcpyDeviceToHost, stream()));
    OB_CUDA(cudaStreamSynchronize(stream()));
}
int B200_ImportPlaintext(const uint64_t* data, int level, double scale) {
    Scheme& s = S();
    auto p = std::make_unique<Plaintext>();
    p->level = level; p->scale = (scale_t)scale; p->log_slots = s.log_max_slots;
    p->data.alloc((size_t)(level + 1) * s.N * 8);
    OB_CUDA(cudaMemcpyAsync(p->data.p, data, p->data.bytes, cudaMemcpyHostToDevice, stream()));
    OB_CUDA(cudaStreamSynchronize(stream()));
    return push_pt(std::move(p));
}
void B200_ExportSecretKey(int64_t* coeffs) { std::memcpy(coeffs, S().sk_coeffs.data(), (size_t)S().N * 8); }
void B200_ImportSecretKey(const int64_t* coeffs) { S().set_secret_key(std::vector<int64_t>(coeffs, coeffs + S().N)); }
// evaluation key in standard (non-Montgomery) residues: [beta][2][nQ+nP][N].  which: 0 = relin, else Galois element
static void export_evk(const EvalKey& k, uint64_t* out) {
    Scheme& s = S();
    const int beta = s.ctx->num_digits(s.max_level), nmod = s.ctx->nmod;
    DevBuf tmp(k.data.bytes);
    from_montgomery(*s.ctx, k.data.u(), tmp.u(), beta * 2, nmod, s.nQ);
    OB_CUDA(cudaMemcpyAsync(out, tmp.p, tmp.bytes, cudaMemcpyDeviceToHost, stream()));
    OB_CUDA(cudaStreamSynchronize(stream()));
}
void B200_ExportRelinKey(uint64_t* out) { OB_CHECK(S().rlk != nullptr, "no relin key"); export_evk(*S().rlk, out); }
void B200_ExportGaloisKey(uint64_t galEl, uint64_t* out) { export_evk(S().galois_key(galEl), out); }
void B200_ExportPublicKey(uint64_t* out) {
    Scheme& s = S();
    OB_CUDA(cudaMemcpyAsync(out, s.pk.p, s.pk.bytes, cudaMemcpyDeviceToHost, stream()));
    OB_CUDA(cudaStreamSynchronize(stream()));
}
int B200_SchemeInfo(int* logN, int* nQ, int* nP, uint64_t* moduli /* nQ+nP, may be NULL */) {
    Scheme& s = S();
    *logN = s.ctx->logN; *nQ = s.nQ; *nP = s.nP;
    if (moduli) for (int i = 0; i < s.ctx->nmod; ++i) moduli[i] = s.ctx->mod[i];
    return 0;
}
int B200_EncodeF64(const double* values, int n, int level, double scale) {
    Scheme& s = S();
    return push_pt(s.encode(values, n, level, (scale_t)scale, s.log_max_slots));
}
void B200_DecodeF64(int pt, double* out_re, double* out_im) {
    std::vector<double2> v = S().decode(PT(pt));
    for (size_t i = 0; i < v.size(); ++i) { out_re[i] = v[i].x; if (out_im) out_im[i] = v[i].y; }
}
int B200_Conjugate(int ct) { return op_new(true, {TRK(ct)}, [&] { auto o = fresh(); S().conjugate(CT(ct), *o); return o; }); }

}  // extern "C"

// ---------------------------------------------------------------- linear transforms (LG/lineartransform.go:26-211)
namespace {
std::set<u64> g_eval_keys;      // Galois elements owned by the evaluator (power-of-two rotations, AddRotationKey)
std::set<u64> g_lt_keys;        // Galois elements generated for linear transforms (dropped by RemoveRotationKeys)
}
extern "C" {

void B200_NoteEvaluatorKey(uint64_t g) { g_eval_keys.insert(g); }
void NewLinearTransformEvaluator(void) { S(); }

int GenerateLinearTransform(int* diagIdx, int nIdx, float* diagData, int nData, int level, float bsgsRatio, char* ioMode) {
    Scheme& s = S();
    const int slots = 1 << s.log_max_slots;
    // io_mode "load" (LG/lineartransform.go:76-84): the diagonals were serialized earlier and come back through
    // LoadPlaintextDiagonal; only the plan is built here and diagData may be empty
    const bool load = ioMode != nullptr && std::strcmp(ioMode, "load") == 0;
    OB_CHECK(load || (long long)nData == (long long)nIdx * slots, "GenerateLinearTransform: diagData must hold nIdx * slots values");
    std::map<int, std::vector<double2>> diags;
    for (int k = 0; k < nIdx; ++k) {
        std::vector<double2>& v = diags[diagIdx[k]];
        if (load) continue;
        v.resize(slots);
        for (int x = 0; x < slots; ++x) v[x] = make_double2((double)diagData[(size_t)k * slots + x], 0.0);
    }
    const int log_ratio = (int)std::log((double)bsgsRatio);      // natural log, as LG/lineartransform.go:67
    return lt_push(lt_build(s, diags, level, s.q_at(level), s.log_max_slots, log_ratio, !load));
}
int EvaluateLinearTransform(int transform, int ct) { return op_new(true, {&CT(ct).trk}, [&] { return lt_apply(S(), lt_get(transform), CT(ct)); }); }
void DeleteLinearTransform(int id) { if (!g_scheme) return; settle_all(); lt_delete(id); }
ArrayResultInt GetLinearTransformRotationKeys(int transform) {
    return make_array<u64, ArrayResultInt>(lt_galois_elements(S(), lt_get(transform)));
}
void GenerateLinearTransformRotationKey(int galEl) {
    if (!S().has_galois_key((u64)galEl)) g_lt_keys.insert((u64)galEl);
    S().galois_key((u64)galEl);
}
// own containers (not Lattigo MarshalBinary): "OB2GK1\0\0" | u64 galEl | u64 words | data (Montgomery form)
ArrayResultByte GenerateAndSerializeRotationKey(int galEl) {
    Scheme& s = S();
    const bool had = s.has_galois_key((u64)galEl);
    const EvalKey& k = s.galois_key((u64)galEl);
    ArrayResultByte a;
    a.Length = 24 + k.data.bytes;
    a.Data = (unsigned char*)std::malloc(a.Length);
    std::memcpy(a.Data, "OB2GK1\0\0", 8);
    u64 g = (u64)galEl, w = k.data.bytes / 8;
    std::memcpy(a.Data + 8, &g, 8); std::memcpy(a.Data + 16, &w, 8);
    OB_CUDA(cudaMemcpyAsync(a.Data + 24, k.data.p, k.data.bytes, cudaMemcpyDeviceToHost, stream()));
    OB_CUDA(cudaStreamSynchronize(stream()));
    if (!had) s.galois_keys.erase((u64)galEl);                  // the reference does not keep the key it serialises
    return a;
}
void LoadRotationKey(unsigned char* data, unsigned long len, unsigned long galEl) {
    Scheme& s = S();
    OB_CHECK(len > 24 && std::memcmp(data, "OB2GK1\0\0", 8) == 0, "LoadRotationKey: not an orion_b200 rotation key blob");
    u64 w; std::memcpy(&w, data + 16, 8);
    OB_CHECK(len == 24 + w * 8, "LoadRotationKey: truncated blob");
    auto k = std::make_unique<EvalKey>();
    k->data.alloc(w * 8);
    OB_CUDA(cudaMemcpyAsync(k->data.p, data + 24, w * 8, cudaMemcpyHostToDevice, stream()));
    OB_CUDA(cudaStreamSynchronize(stream()));
    if (!s.has_galois_key(galEl)) g_lt_keys.insert(galEl);
    s.galois_keys[galEl] = std::move(k);
}
ArrayResultByte SerializeDiagonal(int transform, int diag) {
    Scheme& s = S();
    LinearTransform& lt = lt_get(transform);
    auto it = lt.slot_of.find(diag & ((1 << lt.log_slots) - 1));
    OB_CHECK(it != lt.slot_of.end(), "SerializeDiagonal: no such diagonal");
    const size_t bytes = (size_t)lt.W * s.N * 8;
    ArrayResultByte a;
    a.Length = 16 + bytes;
    a.Data = (unsigned char*)std::malloc(a.Length);
    std::memcpy(a.Data, "OB2DG1\0\0", 8);
    u64 w = bytes / 8; std::memcpy(a.Data + 8, &w, 8);
    OB_CUDA(cudaMemcpyAsync(a.Data + 16, lt.diags.u() + (size_t)it->second * lt.W * s.N, bytes, cudaMemcpyDeviceToHost, stream()));
    OB_CUDA(cudaStreamSynchronize(stream()));
    return a;
}
void LoadPlaintextDiagonal(unsigned char* data, unsigned long len, int transform, unsigned long diag) {
    Scheme& s = S();
    LinearTransform& lt = lt_get(transform);
    auto it = lt.slot_of.find((int)diag & ((1 << lt.log_slots) - 1));
    OB_CHECK(it != lt.slot_of.end(), "LoadPlaintextDiagonal: no such diagonal");
    const size_t bytes = (size_t)lt.W * s.N * 8;
    OB_CHECK(len == 16 + bytes && std::memcmp(data, "OB2DG1\0\0", 8) == 0, "LoadPlaintextDiagonal: bad blob");
    if (!lt.has_data()) lt.diags.alloc((size_t)lt.diag_idx.size() * bytes);
    OB_CUDA(cudaMemcpyAsync(lt.diags.u() + (size_t)it->second * lt.W * s.N, data + 16, bytes, cudaMemcpyHostToDevice, stream()));
    OB_CUDA(cudaStreamSynchronize(stream()));
}
void RemovePlaintextDiagonals(int transform) { settle_all(); lt_get(transform).diags.release(); }
void RemoveRotationKeys(void) {
    Scheme& s = S();
    settle_all();
    for (u64 g : g_lt_keys) if (!g_eval_keys.count(g)) s.galois_keys.erase(g);
    g_lt_keys.clear();
}
// the drifted Python layer passes what GetLinearTransformRotationKeys returned (PY/lt_evaluator.py:77 <- HG/bindings.py:1879)
void GenerateConsolidatedRotationKeys(int* galEls, int n) { for (int i = 0; i < n; ++i) GenerateLinearTransformRotationKey(galEls[i]); }

}  // extern "C"

// ---------------------------------------------------------------- polynomial evaluation (LG/polyeval.go:33-84)
extern "C" {
void NewPolynomialEvaluator(void) { S(); }
static int make_poly(float* coeffs, int n, bool cheb) {
    OB_CHECK(n >= 1, "polynomial needs at least one coefficient");
    auto p = std::make_unique<Polynomial>();
    p->chebyshev = cheb;
    p->coeffs.assign(coeffs, coeffs + n);
    return poly_push(std::move(p));
}
int GenerateMonomial(float* coeffs, int n) { return make_poly(coeffs, n, false); }
int GenerateChebyshev(float* coeffs, int n) { return make_poly(coeffs, n, true); }
int B200_GeneratePolynomialF64(const double* coeffs, int n, int chebyshev) {
    auto p = std::make_unique<Polynomial>();
    p->chebyshev = chebyshev != 0;
    p->coeffs.assign(coeffs, coeffs + n);
    return poly_push(std::move(p));
}
int EvaluatePolynomial(int ct, int poly, unsigned long outScale) {
    return op_new(true, {&CT(ct).trk}, [&] { return poly_evaluate(S(), CT(ct), poly_get(poly), (scale_t)outScale); });
}
int GetPolyDepth(int poly) { return poly_depth(poly_get(poly)); }

// ---------------------------------------------------------------- bootstrapping (LG/bootstrapper.go:19-93)
void NewBootstrapper(int* logPs, int n, int numSlots) { boot_new(S(), std::vector<int>(logPs, logPs + n), numSlots); }
int Bootstrap(int ct, int numSlots) { return op_new(true, {&CT(ct).trk}, [&] { return boot_apply_graphed(S(), CT(ct), numSlots); }); }
void DeleteBootstrappers(void) { settle_all(); boot_reset(); }
}  // extern "C"
