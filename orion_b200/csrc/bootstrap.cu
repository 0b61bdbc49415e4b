// bootstrap.cu -- CKKS bootstrapping on the device: ScaleUp, ModRaise, (SubSum,) CoeffsToSlots, EvalMod, SlotsToCoeffs.
// Replaces LG/bootstrapper.go:19-93 (bootstrapping.NewParametersFromLiteral / Evaluator.Bootstrap in Lattigo);
// SURVEY.md 8a rows a15, a16.  HEonGPU analogue for structure: HE/src/heongpu/lib/host/ckks/operator.cu:3894-4026.
//
// Same contract as the reference: input at any level >= 0 with the default scale, output at the top residual level
// (len(LogQ)-1) with the default scale, sparse slot counts supported; the circuit consumes its levels from moduli
// appended to the residual chain (never from LogQ), with `boot_params.LogP` as the key-switching primes.
// The circuit itself is this backend's own (Lattigo passes only LogN/LogP/Xs/LogSlots and defaults the rest):
//   chain   = Q_residual | 3 x StC primes | (6 + r) x 60-bit EvalMod primes | 4 x 56-bit CtS primes
//   ScaleUp : at level 0 multiply by k = round(q0 / (2^10 * scale)) so that the message sits 2^-10 below q0
//   ModRaise: centred lift of both polynomials from q0 to the whole chain: phase = k*scale*m + e + q0*I, |I| <= K
//   SubSum  : sparse slots only -- sum of the Galois conjugates that fix the slot subring (factor N/(2n))
//   CtS     : 4 hoisted-BSGS transforms = the inverse special FFT without its bit-reversal, scaled so that the slots
//             hold u/(2K'), u = phase_j/q0; real / imaginary parts split with one conjugation
//   EvalMod : sin(2 pi u) = cos(2 pi (u - 1/4)): degree-63 Chebyshev interpolant of cos(2 pi (K' x - 1/4) / 2^r)
//             followed by r double-angle steps c <- 2c^2 - 1
//   StC     : 3 transforms = the forward special FFT without bit-reversal, scaled by g q0/(2 pi k scale), g = MaxSlots/slots
//             (the shim's post-factor, LG/bootstrapper.go:73-74: a zero-padded input x comes back as replicate(x)); the last
//             plaintext scale is chosen so that the output scale is exactly the default scale.
#include "bootstrap.h"
#include "lintrans.h"
#include "polyeval.h"
#include "graphs.h"
#include <algorithm>
#include <cmath>
#include <ctime>
#include <complex>

namespace ob {

namespace {
typedef std::complex<double> cd;
typedef std::map<int, std::vector<cd>> DiagMat;       // rotation -> diagonal (length n)

constexpr int TPB = 256;
constexpr int LOG_MSG_RATIO = 10;
// levels (= merged DFT factor matrices) spent on CoeffsToSlots / SlotsToCoeffs; ORION_B200_CTS_DEPTH / _STC_DEPTH override
static int env_depth(const char* name, int dflt) { const char* s = std::getenv(name); const int v = s ? std::atoi(s) : dflt; return v >= 1 && v <= 8 ? v : dflt; }
static const int CTS_DEPTH = env_depth("ORION_B200_CTS_DEPTH", 4), STC_DEPTH = env_depth("ORION_B200_STC_DEPTH", 3);
// EvalMod: Chebyshev interpolant of degree 2^CHEB_DEPTH - 1 on an argument range of at most CHEB_RANGE radians, then r
// double-angle steps.  Depth 6 (degree 63, 27 rad) or depth 5 (degree 31, 13.5 rad, one more double-angle step): the same
// number of levels; ORION_B200_CHEB_DEPTH overrides.
static const int CHEB_DEPTH = std::min(6, std::max(5, env_depth("ORION_B200_CHEB_DEPTH", 6)));
static const double CHEB_RANGE = CHEB_DEPTH == 6 ? 27.0 : 13.5;

// out[p][i][x] = centred(in[p][x] mod q0) mod q_i      (coefficient domain)
__global__ void __launch_bounds__(TPB)
mod_raise_kernel(const u64* __restrict__ in, u64* __restrict__ out, int limbs, int N, const ModConst* __restrict__ mc) {
    OB_PDL_SYNC();
    const int i = blockIdx.y, p = blockIdx.z, x = blockIdx.x * TPB + threadIdx.x;
    const u64 q0 = mc[0].q, v = in[(size_t)p * N + x];
    const ModConst m = mc[i];
    u64 r;
    if (v > (q0 >> 1)) { const u64 a = barrett_red(0, q0 - v, m); r = a ? m.q - a : 0; }
    else r = barrett_red(0, v, m);
    out[((size_t)p * limbs + i) * N + x] = r;
}

// (M2 * M1) in diagonal form: diag_{k1+k2}[s] += d2_{k2}[s] * d1_{k1}[s + k2]
DiagMat diag_mul(const DiagMat& m2, const DiagMat& m1, int n) {
    DiagMat out;
    for (auto& a : m2)
        for (auto& b : m1) {
            const int k = (a.first + b.first) & (n - 1);
            std::vector<cd>& d = out[k];
            if (d.empty()) d.assign(n, cd(0, 0));
            for (int s = 0; s < n; ++s) d[s] += a.second[s] * b.second[(s + a.first) & (n - 1)];
        }
    return out;
}

// one stage of the special FFT (forward, DIT) / IFFT (inverse, DIF) on n slots, as a 3-diagonal matrix
DiagMat fft_stage(int n, int len, int M, const std::vector<double2>& roots, bool inverse) {
    const int lenh = len >> 1, lenq = len << 2, gap = M / lenq;
    DiagMat m;
    std::vector<cd>& d0 = m[0]; d0.assign(n, cd(0, 0));
    std::vector<cd>& dp = m[lenh & (n - 1)]; if (dp.empty()) dp.assign(n, cd(0, 0));
    std::vector<cd>& dm = m[(n - lenh) & (n - 1)]; if (dm.empty()) dm.assign(n, cd(0, 0));
    std::vector<int> rot(lenh);
    long long p = 1;
    for (int j = 0; j < lenh; ++j) { rot[j] = (int)p; p = (p * 5) % M; }
    for (int s = 0; s < n; ++s) {
        const int r = s & (len - 1);
        if (r < lenh) {
            const int idx = (rot[r] % lenq) * gap;
            const cd w(roots[idx].x, roots[idx].y);
            m[0][s] += cd(1, 0);
            m[lenh & (n - 1)][s] += inverse ? cd(1, 0) : w;
        } else {
            const int j = r - lenh;
            const int idx = inverse ? (lenq - (rot[j] % lenq)) * gap : (rot[j] % lenq) * gap;
            const cd w(roots[idx].x, roots[idx].y);
            m[0][s] += -w;
            m[(n - lenh) & (n - 1)][s] += inverse ? w : cd(1, 0);
        }
    }
    return m;
}

// the (log n) stages grouped into `depth` merged matrices, in the order they are applied
std::vector<DiagMat> dft_factors(int log_n, int M, int depth, bool inverse, const std::vector<double2>& roots) {
    const int n = 1 << log_n;
    std::vector<int> lens;                     // application order
    if (inverse) for (int len = n; len >= 2; len >>= 1) lens.push_back(len);
    else for (int len = 2; len <= n; len <<= 1) lens.push_back(len);
    const int groups = std::min(depth, log_n);
    std::vector<DiagMat> out;
    int pos = 0;
    for (int g = 0; g < groups; ++g) {
        const int cnt = log_n / groups + (g < log_n % groups ? 1 : 0);
        DiagMat acc = fft_stage(n, lens[pos], M, roots, inverse);
        for (int k = 1; k < cnt; ++k) acc = diag_mul(fft_stage(n, lens[pos + k], M, roots, inverse), acc, n);
        pos += cnt;
        out.push_back(std::move(acc));
    }
    return out;
}

std::map<int, std::vector<double2>> to_lt_diags(const DiagMat& m, double factor) {
    std::map<int, std::vector<double2>> out;
    for (auto& kv : m) {
        bool nz = false;
        for (auto& v : kv.second) if (std::abs(v) > 1e-300) { nz = true; break; }
        if (!nz) continue;
        std::vector<double2>& d = out[kv.first];
        d.resize(kv.second.size());
        for (size_t s = 0; s < d.size(); ++s) d[s] = make_double2(kv.second[s].real() * factor, kv.second[s].imag() * factor);
    }
    return out;
}

struct Bootstrapper {
    int log_slots = 0, r = 3, K1 = 0, top_level = 0, res_levels = 0;
    std::shared_ptr<Scheme> bs;                          // scheme over the extended chain, same secret (one per slot count; at
                                                         // logN >= 15 ONE for all slot counts: the chain does not depend on it and a key is 0.4 GB)
    std::vector<int> log_p;
    int busy = 0;                                        // boot_apply in progress: its factor set is not reclaimable
    std::vector<std::unique_ptr<LinearTransform>> cts, stc;
    std::vector<DiagMat> cts_m, stc_m;                   // host factors (encoded lazily: their scales depend on the input)
    Polynomial cheb;
    std::unique_ptr<Plaintext> times_i;                  // X^(N/2) at the top level, scale 1
    scale_t built_for_scale = 0;
    u64 built_for_k = 0;
    // encoded factors of previously seen input scales (a network bootstraps at a handful of distinct scales)
    struct Encoded { std::vector<std::unique_ptr<LinearTransform>> cts, stc; u64 last_use = 0; };
    std::map<scale_t, Encoded> encoded;
    u64 use_clock = 0;
};
size_t lt_set_bytes(const std::vector<std::unique_ptr<LinearTransform>>& v) { size_t b = 0; for (auto& l : v) b += l->diags.bytes; return b; }
std::map<int, std::unique_ptr<Bootstrapper>>& g_boot = *new std::map<int, std::unique_ptr<Bootstrapper>>();     // keyed by slot count (LG/bootstrapper.go:16)

// ORION_B200_BOOT_DUMP=<dir>: every stage of a bootstrap writes its ciphertext words (and the host-side tables) there, for
// the stage-by-stage comparison with oracle/bootstrap.py (tools/boot_lockstep.py)
const char* dump_dir() { static const char* d = std::getenv("ORION_B200_BOOT_DUMP"); return d; }
void dump_raw(const std::string& tag, const void* p, size_t bytes) {
    const std::string path = std::string(dump_dir()) + "/" + tag + ".bin";
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return;
    std::fwrite(p, 1, bytes, f); std::fclose(f);
}
void dump_ct(const std::string& tag, const Ciphertext& c, int N) {
    if (!dump_dir()) return;
    const size_t bytes = (size_t)(c.degree + 1) * (c.level + 1) * N * 8;
    std::vector<unsigned char> h(bytes);
    OB_CUDA(cudaMemcpyAsync(h.data(), c.data.p, bytes, cudaMemcpyDeviceToHost, stream()));
    OB_CUDA(cudaStreamSynchronize(stream()));
    dump_raw(tag, h.data(), bytes);
    const long double sc = c.scale; const double meta[4] = {(double)c.level, (double)c.degree, (double)sc, (double)(sc - (long double)(double)sc)};
    dump_raw(tag + ".meta", meta, sizeof(meta));
}
void dump_diag(const std::string& tag, const DiagMat& m) {
    if (!dump_dir()) return;
    std::vector<double> flat;
    for (auto& kv : m) { flat.push_back((double)kv.first); for (auto& v : kv.second) { flat.push_back(v.real()); flat.push_back(v.imag()); } }
    dump_raw(tag, flat.data(), flat.size() * 8);
}

std::vector<double> cheb_interp(int deg, double K1, int r) {
    const int n = deg + 1;
    std::vector<double> f(n), c(n);
    const double PI = 3.141592653589793238462643383279502884;
    for (int k = 0; k < n; ++k) {
        const double x = std::cos(PI * (k + 0.5) / n);
        f[k] = std::cos(2.0 * PI * (K1 * x - 0.25) / (double)(1 << r));
    }
    for (int j = 0; j < n; ++j) {
        long double s = 0;
        for (int k = 0; k < n; ++k) s += (long double)f[k] * cosl((long double)PI * j * (k + 0.5L) / n);
        c[j] = (double)(s * (j == 0 ? 1.0L : 2.0L) / n);
    }
    return c;
}
}  // namespace

void boot_graphs_reset();
void boot_reset() { if (!g_boot.empty()) cudaStreamSynchronize(stream()); boot_graphs_reset(); g_boot.clear(); graphs_invalidate(); }

// memory pressure (common.h): encoded DFT factor sets are re-encodable from the host factors -- parked sets first (LRU), then
// the current set of every bootstrapper that is not inside boot_apply
static size_t boot_reclaim(size_t want) {
    size_t freed = 0;
    while (freed < want) {
        Bootstrapper* ob = nullptr; std::map<scale_t, Bootstrapper::Encoded>::iterator old;
        for (auto& kv : g_boot)
            for (auto i = kv.second->encoded.begin(); i != kv.second->encoded.end(); ++i)
                if (!ob || i->second.last_use < old->second.last_use) { ob = kv.second.get(); old = i; }
        if (!ob) break;
        freed += lt_set_bytes(old->second.cts) + lt_set_bytes(old->second.stc);
        ob->encoded.erase(old);
    }
    for (auto& kv : g_boot) {
        Bootstrapper& b = *kv.second;
        if (freed >= want) break;
        if (b.busy || b.cts.empty()) continue;
        freed += lt_set_bytes(b.cts) + lt_set_bytes(b.stc);
        b.cts.clear(); b.stc.clear(); b.built_for_scale = 0;
    }
    return freed;
}

void boot_new(Scheme& s, const std::vector<int>& logP, int num_slots) {
    if (g_boot.count(num_slots)) return;                                   // cached per slot count, as the reference
    OB_CHECK(s.has_sk(), "NewBootstrapper needs the secret key (bootstrapping keys are generated here)");
    int log_slots = 0;
    while ((1 << log_slots) < num_slots) ++log_slots;
    OB_CHECK((1 << log_slots) == num_slots && log_slots >= 1 && log_slots <= s.log_max_slots, "NewBootstrapper: bad slot count");
    auto b = std::make_unique<Bootstrapper>();
    b->log_slots = log_slots; b->res_levels = s.nQ;
    // |I| bound for a ternary secret of Hamming weight h: 8 sigma of a sum of h uniform terms
    const int h = std::min(s.params.h, s.N);
    const int K = (int)std::ceil(8.0 * std::sqrt((double)h / 12.0));
    b->K1 = K + 1;
    b->r = 3;
    while (2.0 * 3.14159265358979 * b->K1 / (double)(1 << b->r) > CHEB_RANGE) ++b->r;
    const int logd = (int)std::lround(std::log2((double)s.default_scale));
    const int stc_bits = std::max(39, std::min(56, logd + 12));
    std::vector<u64> Q(s.ctx->mod.begin(), s.ctx->mod.begin() + s.nQ), used = s.ctx->mod, P;
    auto take = [&](int bits, int count, std::vector<u64>& dst) {
        std::vector<u64> p = gen_ntt_primes(bits, 2ull * s.N, count, used);
        for (u64 q : p) { dst.push_back(q); used.push_back(q); }
    };
    take(stc_bits, STC_DEPTH, Q);
    take(60, CHEB_DEPTH + b->r, Q);
    take(56, CTS_DEPTH, Q);
    std::map<int, int> np;
    for (int bits : logP) np[bits]++;
    for (auto& kv : np) take(kv.first, kv.second, P);
    OB_CHECK(!P.empty(), "NewBootstrapper: boot_params.LogP is empty");
    static const int reclaimer = register_reclaimer(boot_reclaim); (void)reclaimer;
    b->log_p = logP;
    static const char* share_env = std::getenv("ORION_B200_BOOT_SHARE");
    const bool share = share_env ? std::atoi(share_env) != 0 : s.ctx->logN >= 15;
    if (share)
        for (auto& kv : g_boot) if (kv.second->log_p == logP) { b->bs = kv.second->bs; break; }
    const bool fresh = !b->bs;
    if (fresh) b->bs = std::make_shared<Scheme>(s.ctx->logN, Q, P, s.params.log_scale, s.params.h, s.rng.derive(0xB007B007ull ^ (u64)num_slots));
    Scheme& bs = *b->bs;
    if (fresh) {
        bs.default_scale = s.default_scale;
        bs.set_secret_key(s.sk_coeffs);
        bs.gen_relin_key();
    }
    b->top_level = bs.max_level;
    std::vector<double2> roots; host_roots(2 * s.N, roots);
    b->cts_m = dft_factors(log_slots, 2 * s.N, CTS_DEPTH, true, roots);
    b->stc_m = dft_factors(log_slots, 2 * s.N, STC_DEPTH, false, roots);
    b->cheb.chebyshev = true;
    b->cheb.coeffs = cheb_interp((1 << CHEB_DEPTH) - 1, (double)b->K1, b->r);
    if (dump_dir()) {
        std::vector<u64> chain = Q; chain.insert(chain.end(), P.begin(), P.end());
        dump_raw("chain", chain.data(), chain.size() * 8);
        dump_raw("cheb", b->cheb.coeffs.data(), b->cheb.coeffs.size() * 8);
        for (size_t k = 0; k < b->cts_m.size(); ++k) dump_diag("cts_m" + std::to_string(k), b->cts_m[k]);
        for (size_t k = 0; k < b->stc_m.size(); ++k) dump_diag("stc_m" + std::to_string(k), b->stc_m[k]);
    }
    // X^(N/2): i in every slot
    std::vector<double2> iv(1 << bs.log_max_slots, make_double2(0.0, 1.0));
    b->times_i = bs.encode_complex(iv.data(), (int)iv.size(), b->top_level, 1, bs.log_max_slots);
    // keys: conjugation, SubSum rotations, and every rotation of the DFT factors (index sets do not depend on scales)
    bs.galois_key(bs.ctx->galois_conj());
    for (int i = log_slots; i < bs.log_max_slots; ++i) bs.galois_key(bs.ctx->galois_element(1 << i));
    for (const std::vector<DiagMat>* ms : {&b->cts_m, &b->stc_m})
        for (const DiagMat& m : *ms) {
            std::vector<int> idx;
            for (auto& kv : m) idx.push_back(kv.first);
            const int n1 = find_best_bsgs_ratio(idx, 1 << log_slots, 0);
            std::map<int, std::vector<int>> index; std::vector<int> gs, bb;
            bsgs_index(idx, 1 << log_slots, n1, index, gs, bb);
            for (int g : gs) if (g) bs.galois_key(bs.ctx->galois_element(g));
            for (int i : bb) if (i) bs.galois_key(bs.ctx->galois_element(i));
        }
    g_boot[num_slots] = std::move(b);
}

std::unique_ptr<Ciphertext> boot_apply(Scheme& s, const Ciphertext& in, int num_slots) {
    OB_HP("boot_apply (total)");
    auto it = g_boot.find(num_slots);
    OB_CHECK(it != g_boot.end(), "Bootstrap: no bootstrapper for this slot count (call NewBootstrapper)");
    OB_CHECK(in.degree == 1, "Bootstrap expects a degree-1 ciphertext");
    Bootstrapper& b = *it->second;
    struct Busy { int& n; explicit Busy(int& x) : n(x) { ++n; } ~Busy() { --n; } } busy(b.busy);
    Scheme& bs = *b.bs;
    Context& c = *bs.ctx;
    const int N = bs.N, n_log = b.log_slots, top = b.top_level;
    const scale_t q0 = bs.q_at(0);
    const double PI = 3.141592653589793238462643383279502884;

    // ---- ScaleUp at level 0 ----
    auto low = s.clone(in);
    s.drop_level(*low, 0);
    long double kf = floorl((long double)q0 / (ldexpl(1.0L, LOG_MSG_RATIO) * low->scale) + 0.5L);
    if (kf < 1) kf = 1;
    if (kf > 1) s.mul_const_scaled(*low, 1.0, kf, *low);                    // scale becomes k * scale
    const scale_t eff = low->scale;

    // ---- ModRaise ----
    auto ct = bs.new_ct(top, 1, q0, n_log);                                 // declared scale S_r = q0
    {
        DevBuf coef((size_t)2 * N * 8);
        std::vector<LimbDesc> l0 = {LimbDesc{0, 0, 0, 0}, LimbDesc{1, 1, 0, 0}}, all;
        ntt_inverse(*s.ctx, low->data.u(), coef.u(), make_ntt_plan(*s.ctx, l0));
        OB_LAUNCH(mod_raise_kernel, dim3(N / TPB, top + 1, 2), TPB, 0, stream(), coef.u(), ct->data.u(), top + 1, N, c.dmc());
        for (int p = 0; p < 2; ++p)
            for (int i = 0; i <= top; ++i)
                all.push_back(LimbDesc{(uint32_t)(p * (top + 1) + i), (uint32_t)(p * (top + 1) + i), (uint32_t)i, 0});
        ntt_forward(c, ct->data.u(), ct->data.u(), make_ntt_plan(c, all));
    }
    dump_ct("00_low", *low, N); dump_ct("01_raised", *ct, N);
    // ---- SubSum (sparse packing): project onto the slot subring; multiplies the kept coefficients by g ----
    double g = 1.0;
    for (int i = n_log; i < bs.log_max_slots; ++i) {
        Ciphertext rot;
        bs.rotate(*ct, 1 << i, rot);
        bs.add_ct(*ct, rot, *ct, false);
        g *= 2.0;
    }

    // ---- (re)encode the DFT factors when the scale situation changed ----
    const int cts_groups = (int)b.cts_m.size(), stc_groups = (int)b.stc_m.size();
    const int lvl_after_cts = top - CTS_DEPTH, lvl_after_mod = lvl_after_cts - (CHEB_DEPTH + b.r);
    const scale_t s_mod = bs.q_at(lvl_after_cts);                            // EvalMod input scale
    const scale_t s_cheb = bs.q_at(lvl_after_cts - CHEB_DEPTH);              // Chebyshev output scale
    if (b.cts.empty() || b.built_for_scale != eff) {
        if (!b.cts.empty()) {                                 // park the current set, keep at most 6
            auto& e = b.encoded[b.built_for_scale];
            e.cts = std::move(b.cts); e.stc = std::move(b.stc); e.last_use = ++b.use_clock;
            while (b.encoded.size() > 6) {
                capture_abort(); graphs_invalidate();             // a captured bootstrap may point at the set that goes
                auto old = b.encoded.begin();
                for (auto i = b.encoded.begin(); i != b.encoded.end(); ++i) if (i->second.last_use < old->second.last_use) old = i;
                OB_CUDA(cudaDeviceSynchronize());                 // bootstraps on other lanes may still read these factors
                b.encoded.erase(old);
            }
        }
        b.cts.clear(); b.stc.clear();
        auto hit = b.encoded.find(eff);
        if (hit != b.encoded.end()) {
            b.cts = std::move(hit->second.cts); b.stc = std::move(hit->second.stc);
            b.encoded.erase(hit);
            b.built_for_scale = eff;
        }
    }
    if (b.cts.empty()) {
        if (std::getenv("ORION_B200_DEBUG")) std::fprintf(stderr, "[boot] (re)encoding DFT factors for scale %.6Lg\n", (long double)eff);
        // decoded slots after CtS must be u/(2 K'), u = coeff/(g q0); the raised ciphertext decodes to SF(coeffs)/S_r
        const double lambda = (double)((long double)q0 / (2.0L * b.K1 * g * (long double)q0));     // S_r = q0
        const double n = (double)(1 << n_log);
        const double f_cts = std::pow(lambda / n, 1.0 / cts_groups);
        scale_t cur = q0;
        for (int k = 0; k < cts_groups; ++k) {
            const int lvl = top - k;
            scale_t ps = bs.q_at(lvl);
            if (k == cts_groups - 1) ps = bs.q_at(lvl) * (s_mod / cur);      // land exactly on s_mod
            b.cts.push_back(lt_build(bs, to_lt_diags(b.cts_m[k], f_cts), lvl, ps, n_log, 0));
            cur = cur * ps / bs.q_at(lvl);
        }
        // EvalMod output scale (tracked exactly below) -> StC plaintext scales so that the result has the default scale
        scale_t s_out = s_cheb;
        for (int k = 0; k < b.r; ++k) s_out = s_out * s_out / bs.q_at(lvl_after_cts - CHEB_DEPTH - k);
        // The Go shim multiplies Lattigo's output by 2^(LogMaxSlots - LogSlots) (LG/bootstrapper.go:73-74): Orion hands over a
        // ZERO-PADDED full-slot vector (orion/nn/operations.py:59-60,85-87), whose projection onto the slot subring is
        // replicate(x)/g, and expects replicate(x) back.  The factor g is folded into SlotsToCoeffs (no extra level).
        const double mu = (double)((long double)q0 / (2.0L * PI * (long double)eff)) * g;
        const double f_stc = std::pow(mu, 1.0 / stc_groups);
        const scale_t target = s.default_scale;
        const long double per = powl((long double)(target / s_out), 1.0L / stc_groups);
        cur = s_out;
        for (int k = 0; k < stc_groups; ++k) {
            const int lvl = lvl_after_mod - k;
            scale_t ps = bs.q_at(lvl) * per;
            if (k == stc_groups - 1) ps = bs.q_at(lvl) * (target / cur);
            b.stc.push_back(lt_build(bs, to_lt_diags(b.stc_m[k], f_stc), lvl, ps, n_log, 0));
            cur = cur * ps / bs.q_at(lvl);
        }
        b.built_for_scale = eff;
        OB_CUDA(cudaStreamSynchronize(stream()));             // encoded on this lane, read by every later bootstrap on any lane
    }

    dump_ct("02_subsum", *ct, N);
    // ---- CoeffsToSlots ----
    for (int k = 0; k < cts_groups; ++k) {
        auto t = lt_apply(bs, *b.cts[k], *ct);
        bs.rescale(*t, *t);
        ct = std::move(t);
        dump_ct("03_cts" + std::to_string(k), *ct, N);
    }
    if (ct->level > lvl_after_cts) bs.drop_level(*ct, lvl_after_cts);
    ct->scale = s_mod;
    Ciphertext cj, re, im;
    bs.conjugate(*ct, cj);
    bs.add_ct(*ct, cj, re, false);                                           // 2 Re(y) = u_re / K'
    bs.add_ct(*ct, cj, im, true);                                            // 2i Im(y)
    bs.mul_pt(im, *b.times_i, im);                                           // * i  ->  -2 Im(y)
    bs.mul_int(im, -1, im);                                                  //      ->  u_im / K'
    dump_ct("04_re", re, N); dump_ct("05_im", im, N);

    // ---- EvalMod on both halves ----
    // ORION_B200_BOOT_DEBUG=1: decrypt the EvalMod intermediates (the bootstrapping scheme holds the secret key) and print
    // their distance from the exact function of the decrypted input -- where the precision of a circuit variant goes
    static const bool boot_debug = std::getenv("ORION_B200_BOOT_DEBUG") != nullptr;
    auto dbg_vals = [&](const Ciphertext& c) { auto p = bs.decrypt(c); return bs.decode(*p); };
    auto eval_mod = [&](const Ciphertext& x) {
        std::vector<double2> xin;
        if (boot_debug) xin = dbg_vals(x);
        auto report = [&](const char* what, int k, const Ciphertext& c) {
            if (!boot_debug) return;
            std::vector<double2> v = dbg_vals(c);
            double worst = 0, worst_im = 0, xmax = 0;
            for (size_t i = 0; i < v.size(); ++i) {
                const double want = std::cos(2.0 * PI * ((double)b.K1 * xin[i].x - 0.25) / (double)(1 << (b.r - k)));
                worst = std::max(worst, std::fabs(v[i].x - want)); worst_im = std::max(worst_im, std::fabs(v[i].y));
                xmax = std::max(xmax, std::fabs(xin[i].x));
            }
            std::fprintf(stderr, "[boot] %s %d: level %d, log2 scale %.3f, max |err| %.3e, max |imag| %.3e, max |x| %.4f\n", what, k,
                         c.level, (double)std::log2((double)c.scale), worst, worst_im, xmax);
        };
        auto y = poly_evaluate(bs, x, b.cheb, s_cheb);
        report("chebyshev", 0, *y);
        if (dump_dir()) dump_ct(std::string("06_cheb_") + (&x == &re ? "re" : "im"), *y, N);
        for (int k = 0; k < b.r; ++k) {
            Ciphertext sq;
            bs.mul_relin(*y, *y, sq);
            bs.lin_comb_int({&sq}, {2.0L}, -roundl(sq.scale), sq.level, sq.scale, sq);      // 2 c^2 - 1 in one launch
            bs.rescale(sq, *y);
            report("double-angle", k + 1, *y);
        }
        return y;
    };
    std::unique_ptr<Ciphertext> vre, vim;
    {
        StreamFork f(2);                                                     // the two halves are independent chains
        f.branch(0); vre = eval_mod(re);
        f.branch(1); vim = eval_mod(im);
        f.join();
    }
    bs.mul_pt(*vim, *b.times_i, *vim);
    bs.add_ct(*vre, *vim, *ct, false);                                       // sin(2 pi u_re) + i sin(2 pi u_im)
    dump_ct("07_vre", *vre, N); dump_ct("08_evalmod", *ct, N);

    // ---- SlotsToCoeffs ----
    for (int k = 0; k < stc_groups; ++k) {
        auto t = lt_apply(bs, *b.stc[k], *ct);
        bs.rescale(*t, *t);
        ct = std::move(t);
        dump_ct("09_stc" + std::to_string(k), *ct, N);
    }
    const int out_level = s.max_level;
    if (ct->level > out_level) bs.drop_level(*ct, out_level);
    OB_CHECK(ct->level == out_level, "Bootstrap: level accounting error");
    auto out = s.new_ct(out_level, 1, s.default_scale, in.log_slots);
    OB_CUDA(cudaMemcpyAsync(out->data.p, ct->data.p, out->data.bytes, cudaMemcpyDeviceToDevice, stream()));
    return out;
}

// ---- Bootstrap as a captured CUDA graph (graphs.h) ----------------------------------------------------------------------------
namespace {
typedef std::tuple<int, int, long double, int> BootGraphKey;       // slot count, input level, input scale, input slots
std::map<BootGraphKey, UnaryGraphSet>& boot_graphs() { static auto* m = new std::map<BootGraphKey, UnaryGraphSet>(); return *m; }
}  // namespace
void boot_graphs_reset() { cudaDeviceSynchronize(); boot_graphs().clear(); }

std::unique_ptr<Ciphertext> boot_apply_graphed(Scheme& s, const Ciphertext& in, int num_slots) {
    if (!graphs_enabled(s.ctx->logN) || dump_dir() || std::getenv("ORION_B200_BOOT_DEBUG")) return boot_apply(s, in, num_slots);
    UnaryGraphSet& set = boot_graphs()[BootGraphKey(num_slots, in.level, (long double)in.scale, in.log_slots)];
    // scratch of one bootstrap of the ResNet-20 chain at N = 8192: just under 2 GB with the per-stream reuse of the arena
    return run_unary_graphed(s, set, in, std::max<size_t>((size_t)1 << 30, (size_t)s.N << 18), "bootstrap",
                             [&](const Ciphertext& x) { return boot_apply(s, x, num_slots); });
}

}  // namespace ob
