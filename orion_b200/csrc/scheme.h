// scheme.h -- the CKKS scheme object behind the Orion-facing C ABI: parameters, keys, handle tables and the
// evaluator operations with the semantics of the Go shim (LG/*.go) it replaces.
#pragma once
#include <set>
#include "evaluator.h"
#include "scheme_kernels.h"

namespace ob {

typedef long double scale_t;    // 64-bit mantissa: holds every uint64 scale of the C ABI exactly

// Dataflow tracking of a handle's device buffer (capi.cu lanes): the stream and event of its last writer, and the events
// of readers that ran on other streams (a later writer / the free must wait for them).  Empty = settled on the main stream.
struct Track {
    cudaStream_t home = nullptr;
    EvRef ready;
    std::vector<EvRef> busy;
};
struct Plaintext {
    DevBuf data;                // [level+1][N], NTT domain
    int level = 0, log_slots = 0;
    scale_t scale = 1;
    Track trk;
};
struct Ciphertext {
    DevBuf data;                // [degree+1][level+1][N], NTT domain
    int level = 0, degree = 1, log_slots = 0;
    scale_t scale = 1;
    Track trk;
    u64* poly(int p, int N) const { return data.u() + (size_t)p * (level + 1) * N; }
};
struct EvalKey { DevBuf data; unsigned long long last_use = 0; };   // [beta(max level)][2][nQ+nP][N], NTT domain, Montgomery form; LRU stamp (common.h: memory pressure)

// smallest-free-integer handle table: the semantics of LG/minheap.go:27-98
template <class T>
struct HandleTable {
    std::map<int, std::unique_ptr<T>> live;
    std::set<int> freed;
    int next = 0;
    int add(std::unique_ptr<T> obj) {
        int id;
        if (!freed.empty()) { id = *freed.begin(); freed.erase(freed.begin()); } else id = next++;
        live[id] = std::move(obj);
        return id;
    }
    T& get(int id, const char* what) {
        auto it = live.find(id);
        if (it == live.end()) OB_FATAL(std::string("Heap object not found for integer: ") + std::to_string(id) + " (" + what + ")");
        return *it->second;
    }
    void del(int id) { auto it = live.find(id); if (it != live.end()) { live.erase(it); freed.insert(id); } }
    void reset() { live.clear(); freed.clear(); next = 0; }
    std::vector<int> ids() const { std::vector<int> v; for (auto& kv : live) v.push_back(kv.first); return v; }
};

struct SchemeParams {
    int logN = 0, log_scale = 0, h = 0;
    std::vector<int> logQ, logP;
    bool conjugate_invariant = false;
};

class Scheme {
  public:
    Scheme(const SchemeParams& p, const SchemeRng& rng);
    // a scheme over an explicit chain sharing nothing with the parameter generator (bootstrapping contexts)
    Scheme(int logN, const std::vector<u64>& Q, const std::vector<u64>& P, int log_scale, int h, const SchemeRng& rng);
    ~Scheme();

    SchemeParams params;
    std::unique_ptr<Context> ctx;
    std::unique_ptr<PolyEvaluator> pev;
    int N, nQ, nP, max_level, log_max_slots;
    scale_t default_scale;

    // ---- randomness: every sampling event takes a fresh block of streams ----
    SchemeRng rng;                                       // see scheme_kernels.h: ChaCha20 from OS keys, or the oracle's test generator
    u64 next_event = 0;
    u64 new_event() { return (next_event++) * 16384ull; }

    // ---- keys ----
    std::vector<int64_t> sk_coeffs;                 // ternary secret, host copy
    DevBuf sk;                                      // [nQ+nP][N] NTT, standard form
    DevBuf pk;                                      // [2][nQ+nP][N] NTT
    std::unique_ptr<EvalKey> rlk;
    std::map<u64, std::unique_ptr<EvalKey>> galois_keys;
    bool has_sk() const { return sk.p != nullptr; }
    void gen_secret_key();
    void set_secret_key(const std::vector<int64_t>& coeffs);
    void gen_public_key();
    void gen_relin_key();
    const EvalKey& galois_key(u64 galEl);           // generated on demand (LG/evaluator.go:34-46)
    bool has_galois_key(u64 galEl) const { return galois_keys.count(galEl) != 0; }
    void drop_galois_keys() { galois_keys.clear(); }
    size_t reclaim_galois_keys(size_t want);        // under memory pressure: drop LRU keys (regenerated on next use; needs the secret key)
    int reclaimer_id = 0;
    // gadget key switching skIn -> skOut; sk_in / sk_out are [nQ+nP][N] NTT polynomials
    std::unique_ptr<EvalKey> gen_eval_key(const u64* sk_in, const u64* sk_out);

    // ---- handle tables ----
    HandleTable<Plaintext> pts;
    HandleTable<Ciphertext> cts;

    // ---- encoder / encryptor ----
    std::unique_ptr<Plaintext> encode(const double* values, int n, int level, scale_t scale, int log_slots);
    std::unique_ptr<Plaintext> encode_complex(const double2* values, int n, int level, scale_t scale, int log_slots);
    // batch of full-slot real vectors -> [batch][limbs][N] NTT residues over the chain indices `mods`
    // batch of slot vectors already on the device ([batch][2^log_slots] complex, destroyed) -> out [batch][W][N] NTT residues
    // over Q_0..Q_level (+ P when with_p), optionally in Montgomery form.  The bulk path behind lintrans.Encode.
    void encode_device_batch(double2* dvals, int log_slots, int batch, scale_t scale, int level, bool with_p,
                             u64* out, bool montgomery);
    std::vector<double2> decode(const Plaintext& pt);
    std::unique_ptr<Ciphertext> encrypt(const Plaintext& pt);
    std::unique_ptr<Plaintext> decrypt(const Ciphertext& ct);

    // ---- evaluator (LG/evaluator.go semantics) ----
    std::unique_ptr<Ciphertext> new_ct(int level, int degree, scale_t scale, int log_slots);
    std::unique_ptr<Ciphertext> clone(const Ciphertext& a);
    void drop_level(Ciphertext& a, int new_level);                          // keep limbs 0..new_level
    void add_ct(const Ciphertext& a, const Ciphertext& b, Ciphertext& out, bool sub);
    void add_pt(const Ciphertext& a, const Plaintext& b, Ciphertext& out, bool sub);
    void mul_pt(const Ciphertext& a, const Plaintext& b, Ciphertext& out);
    void add_const(const Ciphertext& a, double c, Ciphertext& out);         // c0 += round(c*scale)
    void mul_int(const Ciphertext& a, long long k, Ciphertext& out);
    void mul_float(const Ciphertext& a, double c, Ciphertext& out);         // * round(c*q_level), scale *= q_level
    void mul_const_scaled(const Ciphertext& a, double c, scale_t cscale, Ciphertext& out);  // * round(c*cscale)
    // out = sum_i round(c_i * S / scale(a_i)) * a_i  (+ round(c0 * S)) at `level` (operands may sit higher), scale exactly S:
    // the same residues as mul_const_scaled + add_ct + add_const term by term, in one launch per 16 terms
    // out = sum_i k_i * a_i (+ c0) with explicit signed integer multipliers (integer-valued long doubles), scale `S` as given
    void lin_comb_int(const std::vector<const Ciphertext*>& a, const std::vector<long double>& k, long double c0, int level,
                      scale_t S, Ciphertext& out);
    void lin_comb_scaled(const std::vector<const Ciphertext*>& a, const std::vector<double>& c, double c0, int level, scale_t S,
                         Ciphertext& out);
    void mul_relin(const Ciphertext& a, const Ciphertext& b, Ciphertext& out, bool relin = true);
    void relinearize(Ciphertext& a);
    void rescale(const Ciphertext& a, Ciphertext& out);
    void rotate(const Ciphertext& a, int rot, Ciphertext& out);
    void conjugate(const Ciphertext& a, Ciphertext& out);
    void automorphism(const Ciphertext& a, u64 galEl, Ciphertext& out);
    scale_t q_at(int level) const { return (scale_t)ctx->mod[level]; }

  private:
    void init_common(const SchemeRng& rng);
    void match_scale(const Ciphertext& a, scale_t sa, const u64* b_data, int b_polys, int b_level, scale_t sb, int level,
                     Ciphertext& out, bool sub);
    void assign(Ciphertext& out, std::unique_ptr<Ciphertext> tmp);
};

std::vector<u64> const_to_rns(const Context& c, long double value_scaled_abs, bool neg, int limbs, int L);

}  // namespace ob
