// scheme_eval.cu -- ciphertext-level evaluator operations with the semantics of LG/evaluator.go:48-317
// (which forwards to Lattigo's ckks.Evaluator): level alignment by dropping limbs, scale bookkeeping,
// integer-ratio scale matching for additions, relinearisation and rotations through hybrid key switching.
#include "scheme.h"
#include <cmath>

namespace ob {

// residues of (+/-) an integer-valued long double, one per limb of the QP-at-level layout
std::vector<u64> const_to_rns(const Context& c, long double mag, bool neg, int limbs, int L) {
    OB_HP("const_to_rns");
    std::vector<u64> r(limbs);
    for (int t = 0; t < limbs; ++t) {
        const u64 q = c.mod[t < L ? t : c.nQ + (t - L)];
        u64 v = (u64)fmodl(mag, (long double)q);
        if (v >= q) v %= q;
        r[t] = (neg && v) ? q - v : v;
    }
    return r;
}

void Scheme::assign(Ciphertext& out, std::unique_ptr<Ciphertext> tmp) {
    out.data = std::move(tmp->data);
    out.level = tmp->level; out.degree = tmp->degree; out.scale = tmp->scale; out.log_slots = tmp->log_slots;
}

std::unique_ptr<Ciphertext> Scheme::clone(const Ciphertext& a) {
    auto c = new_ct(a.level, a.degree, a.scale, a.log_slots);
    OB_CUDA(cudaMemcpyAsync(c->data.p, a.data.p, (size_t)(a.degree + 1) * (a.level + 1) * N * 8, cudaMemcpyDeviceToDevice, stream()));
    return c;
}

void Scheme::drop_level(Ciphertext& a, int new_level) {
    OB_CHECK(new_level >= 0 && new_level <= a.level, "drop_level: bad target level");
    if (new_level == a.level) return;
    auto t = new_ct(new_level, a.degree, a.scale, a.log_slots);
    ew_binary(*ctx, EW_COPY, a.data.u(), a.level + 1, nullptr, 0, 1, t->data.u(), new_level + 1, a.degree + 1, new_level + 1, new_level + 1);
    assign(a, std::move(t));
}

static bool scales_close(scale_t a, scale_t b) {
    scale_t r = a / b;
    return r > 1 - 1e-9L && r < 1 + 1e-9L;
}

// out = a (+/-) b, b given as raw polynomials.  Lattigo aligns levels by dropping limbs and, when the scales differ by
// an integer factor > 1, multiplies the smaller-scale operand by that factor first (ckks.Evaluator.evaluateInPlace).
void Scheme::match_scale(const Ciphertext& a, scale_t sa, const u64* b_data, int b_polys, int b_level, scale_t sb, int level,
                         Ciphertext& out, bool sub) {
    const int L = level + 1, pa = a.degree + 1, pb = b_polys, pmax = std::max(pa, pb), pcommon = std::min(pa, pb);
    auto tmp = new_ct(level, pmax - 1, std::max(sa, sb), a.log_slots);
    const u64* A = a.data.u(); int a_stride = a.level + 1;
    const u64* B = b_data;     int b_stride = b_level + 1;
    DevBuf scaled;
    if (!scales_close(sa, sb)) {
        const bool a_small = sa < sb;
        const scale_t ratio = floorl(a_small ? sb / sa : sa / sb);
        if (ratio > 1) {
            const int polys = a_small ? pa : pb;
            scaled.alloc((size_t)polys * L * N * 8);
            std::vector<u64> sc = const_to_rns(*ctx, ratio, false, L, L);
            ew_mul_scalar(*ctx, a_small ? A : B, a_small ? a_stride : b_stride, scaled.u(), L, polys, L, L, sc);
            if (a_small) { A = scaled.u(); a_stride = L; } else { B = scaled.u(); b_stride = L; }
        }
    }
    ew_binary(*ctx, sub ? EW_SUB : EW_ADD, A, a_stride, B, b_stride, pcommon, tmp->data.u(), L, pcommon, L, L);
    if (pa > pb)
        ew_binary(*ctx, EW_COPY, A + (size_t)pb * a_stride * N, a_stride, nullptr, 0, 1, tmp->data.u() + (size_t)pb * L * N, L, pa - pb, L, L);
    if (pb > pa)
        ew_binary(*ctx, sub ? EW_NEG : EW_COPY, B + (size_t)pa * b_stride * N, b_stride, nullptr, 0, 1,
                  tmp->data.u() + (size_t)pa * L * N, L, pb - pa, L, L);
    assign(out, std::move(tmp));
}

void Scheme::add_ct(const Ciphertext& a, const Ciphertext& b, Ciphertext& out, bool sub) {
    match_scale(a, a.scale, b.data.u(), b.degree + 1, b.level, b.scale, std::min(a.level, b.level), out, sub);
}
void Scheme::add_pt(const Ciphertext& a, const Plaintext& b, Ciphertext& out, bool sub) {
    match_scale(a, a.scale, b.data.u(), 1, b.level, b.scale, std::min(a.level, b.level), out, sub);
}

void Scheme::mul_pt(const Ciphertext& a, const Plaintext& b, Ciphertext& out) {
    const int level = std::min(a.level, b.level), L = level + 1;
    auto tmp = new_ct(level, a.degree, a.scale * b.scale, a.log_slots);
    ew_binary(*ctx, EW_MUL, a.data.u(), a.level + 1, b.data.u(), b.level + 1, 1, tmp->data.u(), L, a.degree + 1, L, L);
    assign(out, std::move(tmp));
}

void Scheme::add_const(const Ciphertext& a, double c, Ciphertext& out) {
    const int L = a.level + 1;
    const long double v = roundl(fabsl((long double)c * a.scale));
    std::vector<u64> sc = const_to_rns(*ctx, v, c < 0, L, L);
    auto tmp = new_ct(a.level, a.degree, a.scale, a.log_slots);
    ew_add_scalar(*ctx, a.data.u(), L, tmp->data.u(), L, a.degree + 1, L, L, sc, 0);
    assign(out, std::move(tmp));
}

void Scheme::mul_const_scaled(const Ciphertext& a, double c, scale_t cscale, Ciphertext& out) {
    const int L = a.level + 1;
    const long double v = roundl(fabsl((long double)c * cscale));
    std::vector<u64> sc = const_to_rns(*ctx, v, c < 0, L, L);
    auto tmp = new_ct(a.level, a.degree, a.scale * cscale, a.log_slots);
    ew_mul_scalar(*ctx, a.data.u(), L, tmp->data.u(), L, a.degree + 1, L, L, sc);
    assign(out, std::move(tmp));
}
void Scheme::lin_comb_scaled(const std::vector<const Ciphertext*>& a, const std::vector<double>& c, double c0, int level,
                             scale_t S, Ciphertext& out) {
    OB_CHECK(!a.empty() && a.size() == c.size(), "lin_comb_scaled: bad operand list");
    std::vector<long double> k(a.size());
    for (size_t i = 0; i < a.size(); ++i) {
        const long double v = roundl(fabsl((long double)c[i] * (S / a[i]->scale)));       // as mul_const_scaled(a_i, c_i, S/s_i)
        k[i] = c[i] < 0 ? -v : v;
    }
    const long double v0 = roundl(fabsl((long double)c0 * S));                            // as add_const at scale S
    lin_comb_int(a, k, c0 < 0 ? -v0 : v0, level, S, out);
}
void Scheme::lin_comb_int(const std::vector<const Ciphertext*>& a, const std::vector<long double>& kk, long double c0, int level,
                          scale_t S, Ciphertext& out) {
    OB_CHECK(!a.empty() && a.size() == kk.size(), "lin_comb_int: bad operand list");
    const int L = level + 1, polys = a[0]->degree + 1;
    auto tmp = new_ct(level, a[0]->degree, S, a[0]->log_slots);
    std::unique_ptr<Ciphertext> part;                                    // running sum when there are more than 15 terms
    for (size_t i0 = 0; i0 < a.size(); i0 += LINCOMB_MAX - 1) {
        const size_t m = std::min<size_t>(LINCOMB_MAX - 1, a.size() - i0);
        std::vector<const u64*> src; std::vector<int> stride; std::vector<std::vector<u64>> k;
        if (part) { src.push_back(part->data.u()); stride.push_back(L); k.push_back(const_to_rns(*ctx, 1.0L, false, L, L)); }
        for (size_t i = i0; i < i0 + m; ++i) {
            OB_CHECK(a[i]->level >= level && a[i]->degree + 1 == polys, "lin_comb: operand below the target level");
            src.push_back(a[i]->data.u()); stride.push_back(a[i]->level + 1);
            k.push_back(const_to_rns(*ctx, fabsl(kk[i]), kk[i] < 0, L, L));
        }
        const bool last = i0 + m >= a.size();
        std::vector<u64> k0;
        if (last && c0 != 0.0L) k0 = const_to_rns(*ctx, fabsl(c0), c0 < 0, L, L);
        auto dst = last ? nullptr : new_ct(level, a[0]->degree, S, a[0]->log_slots);
        lin_comb(*ctx, (int)src.size(), src.data(), stride.data(), k, (last && c0 != 0.0L) ? &k0 : nullptr,
                 last ? tmp->data.u() : dst->data.u(), polys, L);
        if (!last) part = std::move(dst);
    }
    assign(out, std::move(tmp));
}
void Scheme::mul_int(const Ciphertext& a, long long k, Ciphertext& out) {
    mul_const_scaled(a, (double)k, 1, out);
}
void Scheme::mul_float(const Ciphertext& a, double c, Ciphertext& out) {
    if (c == std::floor(c) && std::fabs(c) < 9e15) { mul_int(a, (long long)c, out); return; }   // Lattigo: integers keep the scale
    mul_const_scaled(a, c, q_at(a.level), out);
}

void Scheme::mul_relin(const Ciphertext& a, const Ciphertext& b, Ciphertext& out, bool relin) {
    OB_CHECK(a.degree == 1 && b.degree == 1, "MulRelin expects degree-1 operands");
    const int level = std::min(a.level, b.level), L = level + 1;
    auto t3 = new_ct(level, 2, a.scale * b.scale, a.log_slots);
    tensor_product(*ctx, a.data.u(), a.level + 1, b.data.u(), b.level + 1, t3->data.u(), L);
    if (relin) relinearize(*t3);
    assign(out, std::move(t3));
}

void Scheme::relinearize(Ciphertext& a) {
    if (a.degree < 2) return;
    OB_CHECK(rlk != nullptr, "relinearisation key missing (GenerateRelinearizationKey)");
    const int L = a.level + 1;
    auto tmp = new_ct(a.level, 1, a.scale, a.log_slots);
    pev->keyswitch(a.level, a.poly(2, N), rlk->data.u(), a.data.u(), L, 2, tmp->data.u(), L);
    assign(a, std::move(tmp));
}

void Scheme::rescale(const Ciphertext& a, Ciphertext& out) {
    OB_CHECK(a.level >= 1, "cannot rescale: input Ciphertext already at level 0");
    OB_CHECK(a.degree == 1, "rescale expects a degree-1 ciphertext");
    auto tmp = new_ct(a.level - 1, a.degree, a.scale / q_at(a.level), a.log_slots);
    pev->rescale(a.level, a.data.u(), a.level + 1, tmp->data.u(), a.level, a.degree + 1);
    assign(out, std::move(tmp));
}

void Scheme::automorphism(const Ciphertext& a, u64 galEl, Ciphertext& out) {
    OB_CHECK(a.degree == 1, "automorphism expects a degree-1 ciphertext");
    const int L = a.level + 1;
    OpPin pin;
    const EvalKey& key = galois_key(galEl);
    DevBuf ks((size_t)2 * L * N * 8);
    pev->keyswitch(a.level, a.poly(1, N), key.data.u(), a.data.u(), L, 1, ks.u(), L);
    auto tmp = new_ct(a.level, 1, a.scale, a.log_slots);
    automorphism_ntt(*ctx, ks.u(), tmp->data.u(), 2 * L, ctx->galois_index(galEl));
    assign(out, std::move(tmp));
}
void Scheme::rotate(const Ciphertext& a, int rot, Ciphertext& out) {
    const u64 g = ctx->galois_element(rot);
    if (g == 1) { if (&a != &out) assign(out, clone(a)); return; }
    automorphism(a, g, out);
}
void Scheme::conjugate(const Ciphertext& a, Ciphertext& out) { automorphism(a, ctx->galois_conj(), out); }

}  // namespace ob
