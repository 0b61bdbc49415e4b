// scheme.cu -- parameters, key generation, encode/encrypt/decrypt.  Evaluator operations: scheme_eval.cu.
// Replaces the state and calls behind LG/scheme.go:35-100, LG/keygenerator.go:13-59, LG/encoder.go:11-42,
// LG/encryptor.go:10-38 (all of which forward to Lattigo); everything is generated and kept on the device.
#include "scheme.h"
#include <cmath>

namespace ob {

namespace {
constexpr int TPB = 256;

// evaluation-key assembly for one digit: b = e - a*s_out (+ (P mod q)*s_in on the digit's own limbs);
// both halves are written in Montgomery form.  t = chain index.
__global__ void __launch_bounds__(TPB)
evk_assemble_kernel(const u64* __restrict__ a, const u64* __restrict__ e, const u64* __restrict__ s_out,
                    const u64* __restrict__ s_in, const ulonglong2* __restrict__ pmodq, int own_lo, int own_hi,
                    u64* __restrict__ out_b, u64* __restrict__ out_a, int N, const ModConst* __restrict__ mc) {
    OB_PDL_SYNC();
    const int t = blockIdx.y, x = blockIdx.x * TPB + threadIdx.x;
    const ModConst m = mc[t];
    const size_t o = (size_t)t * N + x;
    const u64 av = a[o];
    u64 b = sub_mod(e[o], barrett_mul(av, s_out[o], m), m.q);
    if (t >= own_lo && t < own_hi) {
        const ulonglong2 w = pmodq[t];
        b = add_mod(b, shoup_mul(s_in[o], w.x, w.y, m.q), m.q);
    }
    out_b[o] = shoup_mul(b, m.r1, m.r1_sh, m.q);
    out_a[o] = shoup_mul(av, m.r1, m.r1_sh, m.q);
}
// b = e - a*s   (public key, standard form)
__global__ void __launch_bounds__(TPB)
pk_assemble_kernel(const u64* __restrict__ a, const u64* __restrict__ e, const u64* __restrict__ s,
                   u64* __restrict__ out_b, int N, const ModConst* __restrict__ mc) {
    OB_PDL_SYNC();
    const int t = blockIdx.y, x = blockIdx.x * TPB + threadIdx.x;
    const ModConst m = mc[t];
    const size_t o = (size_t)t * N + x;
    out_b[o] = sub_mod(e[o], barrett_mul(a[o], s[o], m), m.q);
}
}  // namespace

Scheme::Scheme(const SchemeParams& p, const SchemeRng& rng_) : params(p) {
    // RingType ConjugateInvariant (configs/lola.yml; LG/scheme.go:46-55): Lattigo folds Z[X+X^-1]/(X^2N+1) into N
    // coefficients, N real slots, primes = 1 mod 4N.  This backend serves it with the standard ring of degree 2N: the
    // same prime chain (NthRoot 4N), the same N slots (orion/backend/python/parameters.py:33-35), every operation
    // unchanged; real inputs stay real.  Cost: ciphertexts are twice the size of Lattigo's folded representation.
    const int ring_logN = p.logN + (p.conjugate_invariant ? 1 : 0);
    std::vector<u64> Q, P;
    gen_moduli(p.logQ, p.logP, ring_logN + 1, Q, P);
    ctx = std::make_unique<Context>(ring_logN, Q, P);
    init_common(rng_);
}
Scheme::Scheme(int logN, const std::vector<u64>& Q, const std::vector<u64>& P, int log_scale, int h, const SchemeRng& rng_) {
    params.logN = logN; params.log_scale = log_scale; params.h = h;
    for (u64 q : Q) params.logQ.push_back(64 - __builtin_clzll(q));
    for (u64 q : P) params.logP.push_back(64 - __builtin_clzll(q));
    ctx = std::make_unique<Context>(logN, Q, P);
    init_common(rng_);
}
void Scheme::init_common(const SchemeRng& rng_) {
    pev = std::make_unique<PolyEvaluator>(*ctx);
    N = ctx->N; nQ = ctx->nQ; nP = ctx->nP; max_level = nQ - 1; log_max_slots = ctx->logN - 1;
    default_scale = std::ldexp((scale_t)1, params.log_scale);
    rng = rng_;
    reclaimer_id = register_reclaimer([this](size_t want) { return reclaim_galois_keys(want); });
}
size_t Scheme::reclaim_galois_keys(size_t want) {
    if (!has_sk()) return 0;                                  // loaded keys without the secret key cannot be rebuilt
    size_t freed = 0;
    const unsigned long long floor = pin_floor();
    while (freed < want) {
        auto old = galois_keys.end();
        for (auto i = galois_keys.begin(); i != galois_keys.end(); ++i)
            if (i->second->last_use < floor && (old == galois_keys.end() || i->second->last_use < old->second->last_use)) old = i;
        if (old == galois_keys.end()) break;
        freed += old->second->data.bytes;
        galois_keys.erase(old);
    }
    if (std::getenv("ORION_B200_DEBUG"))
        std::fprintf(stderr, "[mem] scheme %p (nQ %d): dropped %.2f GB of Galois keys, %zu keys left (pin floor %llu)\n", (void*)this, nQ,
                     freed / 1073741824.0, galois_keys.size(), floor);
    return freed;
}
Scheme::~Scheme() {
    unregister_reclaimer(reclaimer_id);
    graphs_invalidate();
    cudaStreamSynchronize(stream());
    drop_scheme_tables(*ctx);
}

// ---- keys ----
void Scheme::set_secret_key(const std::vector<int64_t>& coeffs) {
    OB_CHECK((int)coeffs.size() == N, "secret key length");
    sk_coeffs = coeffs;
    DevBuf small((size_t)N * 8);
    OB_CUDA(cudaMemcpyAsync(small.p, coeffs.data(), (size_t)N * 8, cudaMemcpyHostToDevice, stream()));
    std::vector<int> mods(ctx->nmod);
    std::vector<LimbDesc> l(ctx->nmod);
    for (int t = 0; t < ctx->nmod; ++t) { mods[t] = t; l[t] = LimbDesc{(uint32_t)t, (uint32_t)t, (uint32_t)t, 0}; }
    sk.alloc((size_t)ctx->nmod * N * 8);
    small_to_rns(*ctx, small.as<int64_t>(), sk.u(), mods);
    ntt_forward(*ctx, sk.u(), sk.u(), make_ntt_plan(*ctx, l));
    OB_CUDA(cudaStreamSynchronize(stream()));
}
void Scheme::gen_secret_key() {
    std::vector<int64_t> s;
    sample_ternary_hw_host(N, params.h, rng.sec, new_event(), s);
    set_secret_key(s);
}

static NttPlan full_chain_plan(Context& c) {
    std::vector<LimbDesc> l(c.nmod);
    for (int t = 0; t < c.nmod; ++t) l[t] = LimbDesc{(uint32_t)t, (uint32_t)t, (uint32_t)t, 0};
    return make_ntt_plan(c, l);
}
static std::vector<int> full_chain(const Context& c) {
    std::vector<int> m(c.nmod);
    for (int t = 0; t < c.nmod; ++t) m[t] = t;
    return m;
}

void Scheme::gen_public_key() {
    OB_CHECK(has_sk(), "GeneratePublicKey needs a secret key");
    const u64 ev = new_event();
    const size_t W = (size_t)ctx->nmod * N;
    pk.alloc(2 * W * 8);
    DevBuf small((size_t)N * 8), e(W * 8);
    u64* a = pk.u() + W;
    sample_uniform(*ctx, a, full_chain(*ctx), rng.pub, ev);
    sample_gaussian(*ctx, small.as<int64_t>(), rng.sec, ev + 512);
    small_to_rns(*ctx, small.as<int64_t>(), e.u(), full_chain(*ctx));
    ntt_forward(*ctx, e.u(), e.u(), full_chain_plan(*ctx));
    dim3 grid(N / TPB, ctx->nmod);
    OB_LAUNCH(pk_assemble_kernel, grid, TPB, 0, stream(), a, e.u(), sk.u(), pk.u(), N, ctx->dmc());
}

std::unique_ptr<EvalKey> Scheme::gen_eval_key(const u64* sk_in, const u64* sk_out) {
    OB_CHECK(nP > 0, "evaluation keys need special primes (LogP)");
    const u64 ev = new_event();
    const int beta = ctx->num_digits(max_level), nmod = ctx->nmod;
    const size_t W = (size_t)nmod * N;
    auto key = std::make_unique<EvalKey>();
    key->data.alloc((size_t)beta * 2 * W * 8);
    DevBuf a(W * 8), e(W * 8), small((size_t)N * 8);
    const NttPlan plan = full_chain_plan(*ctx);
    const std::vector<int> mods = full_chain(*ctx);
    dim3 grid(N / TPB, nmod);
    for (int j = 0; j < beta; ++j) {
        sample_uniform(*ctx, a.u(), mods, rng.pub, ev + (u64)j * 128);
        sample_gaussian(*ctx, small.as<int64_t>(), rng.sec, ev + (u64)j * 128 + 64);
        small_to_rns(*ctx, small.as<int64_t>(), e.u(), mods);
        ntt_forward(*ctx, e.u(), e.u(), plan);
        const int lo = j * nP, hi = std::min((j + 1) * nP, nQ);
        OB_LAUNCH(evk_assemble_kernel, grid, TPB, 0, stream(), a.u(), e.u(), sk_out, sk_in, ctx->d_pmodq.as<ulonglong2>(),
                  lo, hi, key->data.u() + ((size_t)j * 2 + 0) * W, key->data.u() + ((size_t)j * 2 + 1) * W, N, ctx->dmc());
    }
    return key;
}

void Scheme::gen_relin_key() {
    OB_CHECK(has_sk(), "GenerateRelinearizationKey needs a secret key");
    DevBuf s2((size_t)ctx->nmod * N * 8);
    ew_binary(*ctx, EW_MUL, sk.u(), ctx->nmod, sk.u(), ctx->nmod, 1, s2.u(), ctx->nmod, 1, ctx->nmod, nQ);
    rlk = gen_eval_key(s2.u(), sk.u());
}

const EvalKey& Scheme::galois_key(u64 galEl) {
    auto it = galois_keys.find(galEl);
    if (it != galois_keys.end()) { it->second->last_use = use_tick(); return *it->second; }
    OB_CHECK(has_sk(), "Galois key generation needs a secret key");
    capture_abort();                                          // a key generated lazily: for real, from the pool
    // key switches s -> sigma_g^-1(s); the automorphism sigma_g applied afterwards maps it back to s
    u64 inv = 1, twoN = 2ull * N;
    for (u64 e = galEl, k = twoN / 2 - 1; k; k >>= 1, e = (e * e) & (twoN - 1)) if (k & 1) inv = (inv * e) & (twoN - 1);
    OB_CHECK(((inv * galEl) & (twoN - 1)) == 1, "Galois element is not invertible");
    DevBuf s_out((size_t)ctx->nmod * N * 8);
    automorphism_ntt(*ctx, sk.u(), s_out.u(), ctx->nmod, ctx->galois_index(inv));
    auto key = gen_eval_key(sk.u(), s_out.u());
    // the key is shared by every dataflow lane: it is published only once it is complete (a lazily generated key is built on
    // whichever lane asked for it first, and another lane could otherwise read it half written)
    OB_CUDA(cudaStreamSynchronize(stream()));
    key->last_use = use_tick();
    return *galois_keys.emplace(galEl, std::move(key)).first->second;
}

// ---- encoder (LG/encoder.go:16-42) ----
static std::vector<int> qp_mods(const Context& c, int level, bool with_p) {
    std::vector<int> m;
    for (int i = 0; i <= level; ++i) m.push_back(i);
    if (with_p) for (int k = 0; k < c.nP; ++k) m.push_back(c.nQ + k);
    return m;
}

void Scheme::encode_device_batch(double2* dvals, int log_slots, int batch, scale_t scale, int level, bool with_p,
                                 u64* out, bool montgomery) {
    OB_CHECK(log_slots >= 1 && log_slots <= log_max_slots, "slot count out of range");
    const std::vector<int> mods = qp_mods(*ctx, level, with_p);
    const int W = (int)mods.size();
    special_ifft(*ctx, dvals, log_slots, batch);
    encode_round_to_rns(*ctx, dvals, log_slots, (double)scale, out, mods, batch);
    std::vector<LimbDesc> l((size_t)batch * W);
    for (int b = 0; b < batch; ++b)
        for (int t = 0; t < W; ++t)
            l[(size_t)b * W + t] = LimbDesc{(uint32_t)(b * W + t), (uint32_t)(b * W + t), (uint32_t)mods[t], 0};
    ntt_forward(*ctx, out, out, make_ntt_plan(*ctx, l));
    if (montgomery) to_montgomery(*ctx, out, out, batch, W, level + 1);
}

std::unique_ptr<Plaintext> Scheme::encode_complex(const double2* values, int n, int level, scale_t scale, int log_slots) {
    OB_CHECK(level >= 0 && level <= max_level, "Encode: level out of range");
    const int slots = 1 << log_slots;
    OB_CHECK(n <= slots, "Encode: more values than slots");
    std::vector<double2> h(slots, make_double2(0, 0));
    for (int i = 0; i < n; ++i) h[i] = values[i];
    DevBuf d((size_t)slots * sizeof(double2));
    OB_CUDA(cudaMemcpyAsync(d.p, h.data(), d.bytes, cudaMemcpyHostToDevice, stream()));
    auto pt = std::make_unique<Plaintext>();
    pt->level = level; pt->scale = scale; pt->log_slots = log_slots;
    pt->data.alloc((size_t)(level + 1) * N * 8);
    encode_device_batch(d.as<double2>(), log_slots, 1, scale, level, false, pt->data.u(), false);
    return pt;
}
std::unique_ptr<Plaintext> Scheme::encode(const double* values, int n, int level, scale_t scale, int log_slots) {
    std::vector<double2> h(n);
    for (int i = 0; i < n; ++i) h[i] = make_double2(values[i], 0.0);
    return encode_complex(h.data(), n, level, scale, log_slots);
}

std::vector<double2> Scheme::decode(const Plaintext& pt) {
    const int L = pt.level + 1, slots = 1 << pt.log_slots;
    DevBuf coef((size_t)L * N * 8), vals((size_t)slots * sizeof(double2));
    std::vector<LimbDesc> l(L);
    for (int i = 0; i < L; ++i) l[i] = LimbDesc{(uint32_t)i, (uint32_t)i, (uint32_t)i, 0};
    ntt_inverse(*ctx, pt.data.u(), coef.u(), make_ntt_plan(*ctx, l));
    decode_crt_to_complex(*ctx, coef.u(), pt.level, (double)pt.scale, vals.as<double2>(), pt.log_slots);
    special_fft(*ctx, vals.as<double2>(), pt.log_slots, 1);
    std::vector<double2> out(slots);
    OB_CUDA(cudaMemcpyAsync(out.data(), vals.p, vals.bytes, cudaMemcpyDeviceToHost, stream()));
    OB_CUDA(cudaStreamSynchronize(stream()));
    return out;
}

// ---- encryptor (LG/encryptor.go:20-38): public-key encryption in QP, ModDown by P, add the message ----
std::unique_ptr<Ciphertext> Scheme::new_ct(int level, int degree, scale_t scale, int log_slots) {
    auto ct = std::make_unique<Ciphertext>();
    ct->level = level; ct->degree = degree; ct->scale = scale; ct->log_slots = log_slots;
    ct->data.alloc((size_t)(degree + 1) * (level + 1) * N * 8);
    return ct;
}

std::unique_ptr<Ciphertext> Scheme::encrypt(const Plaintext& pt) {
    OB_CHECK(pk.p != nullptr, "Encrypt needs a public key");
    const int level = pt.level, L = level + 1, K = nP, W = L + K, nmod = ctx->nmod;
    const u64 ev = new_event();
    auto ct = new_ct(level, 1, pt.scale, pt.log_slots);
    const std::vector<int> mods = qp_mods(*ctx, level, true);
    std::vector<LimbDesc> l(W);
    for (int t = 0; t < W; ++t) l[t] = LimbDesc{(uint32_t)t, (uint32_t)t, (uint32_t)mods[t], 0};
    const NttPlan plan = make_ntt_plan(*ctx, l);
    // u: ternary with the secret's Hamming weight
    std::vector<int64_t> uh;
    sample_ternary_hw_host(N, params.h, rng.sec, ev, uh);
    DevBuf small((size_t)N * 8), u((size_t)W * N * 8), e((size_t)W * N * 8), acc((size_t)2 * W * N * 8);
    OB_CUDA(cudaMemcpyAsync(small.p, uh.data(), (size_t)N * 8, cudaMemcpyHostToDevice, stream()));
    small_to_rns(*ctx, small.as<int64_t>(), u.u(), mods);
    ntt_forward(*ctx, u.u(), u.u(), plan);
    for (int c = 0; c < 2; ++c) {
        sample_gaussian(*ctx, small.as<int64_t>(), rng.sec, ev + 1 + c);
        small_to_rns(*ctx, small.as<int64_t>(), e.u(), mods);
        ntt_forward(*ctx, e.u(), e.u(), plan);
        u64* a = acc.u() + (size_t)c * W * N;
        const u64* pkc = pk.u() + (size_t)c * nmod * N;
        ew_binary(*ctx, EW_MUL, u.u(), W, pkc, nmod, 1, a, W, 1, L, L);                                      // Q part
        if (K) ew_binary(*ctx, EW_MUL, u.u() + (size_t)L * N, W, pkc + (size_t)nQ * N, nmod, 1, a + (size_t)L * N, W, 1, K, 0);
        ew_binary(*ctx, EW_ADD, a, W, e.u(), W, 1, a, W, 1, W, L);
    }
    if (K) {
        pev->moddown(level, acc.u(), 2, pt.data.u(), L, 1, ct->data.u(), L);
    } else {
        ew_binary(*ctx, EW_ADD, acc.u(), W, pt.data.u(), L, 1, ct->data.u(), L, 1, L, L);
        ew_binary(*ctx, EW_COPY, acc.u() + (size_t)W * N, W, nullptr, 0, 1, ct->data.u() + (size_t)L * N, L, 1, L, L);
    }
    OB_CUDA(cudaStreamSynchronize(stream()));      // uh (host) was copied with a pageable async memcpy
    return ct;
}

std::unique_ptr<Plaintext> Scheme::decrypt(const Ciphertext& ct) {
    OB_CHECK(has_sk(), "Decrypt needs a secret key");
    OB_CHECK(ct.degree == 1, "Decrypt expects a degree-1 ciphertext");
    const int L = ct.level + 1;
    auto pt = std::make_unique<Plaintext>();
    pt->level = ct.level; pt->scale = ct.scale; pt->log_slots = ct.log_slots;
    pt->data.alloc((size_t)L * N * 8);
    ew_binary(*ctx, EW_MUL, ct.poly(1, N), L, sk.u(), ctx->nmod, 1, pt->data.u(), L, 1, L, L);
    ew_binary(*ctx, EW_ADD, pt->data.u(), L, ct.poly(0, N), L, 1, pt->data.u(), L, 1, L, L);
    return pt;
}

}  // namespace ob
